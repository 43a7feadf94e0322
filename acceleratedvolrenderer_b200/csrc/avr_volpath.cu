// avr_volpath.cu — kernel 1b (see avr_volpath.cuh) and its C entry points.
//
// Persistent kernel, one CTA per SM.  The CTA owns a pool of paths in shared memory (slots of 16-byte quads,
// SlotIO) and four MPMC queues of slot ids; a warp serves ONE queue at a time, so every op has one call site and
// runs convergently:
//   R  regenerate : flush the finished sample to the film, pull new (pixel, sample) work with one
//                   warp-aggregated atomic, generate the camera ray                         (queue Q_DEAD)
//   E1 scatter    : NEE preparation -> shadow ray pending                                   (queue Q_SCATTER)
//   E2 shadow end : NEE accumulate + phase sampling -> next camera segment pending          (queue Q_SHADOW_END)
//                   (Traits::FUSED: E1 and E2 are one op on Q_SCATTER, both rays are set up together and the
//                   tracking loop switches from the shadow ray to the waiting camera ray by itself)
//   S  setup      : boundary hit, RNG seeding, ray -> medium space, slab test, DDA init; runs at the end of
//                   every event batch (camera segment or NEE shadow ray: same code, a mode bit apart)
//   tracking loop (queue Q_TRACK, track_mode): every lane holds a path in registers and alternates
//   A  walk       : DDA steps to the next majorant segment with sigma_maj != 0, each followed by
//   A' sample     : the draw of the next tentative-collision distance, and
//   B  collide    : one tentative collision = one 32-byte density fetch, then the delta-tracking (camera) or
//                   ratio-tracking (shadow) update;
//   a lane whose path reaches an event parks it in its slot and is refilled from Q_TRACK.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <atomic>
#include <thread>

#include "avr_render_host.h"

#ifndef AVR_VOLPATH_PART
#define AVR_VOLPATH_PART 0
#endif

using namespace avr;

namespace avr {

constexpr int kBlock = 128;

// Lane states.  Values >= ST_WALK are the tracking states (the path sits in Q_TRACK or in a lane of the
// tracking loop); the others name the event op the path waits for.
enum {
    ST_DEAD = 0,           // R: needs a new path (a finished sample may be pending)
    ST_SETUP = 1,          // S: pending ray in fe.o / fe.d
    ST_EV_SCATTER = 2,     // E1: real scattering event
    ST_EV_SHADOW_END = 3,  // E2: shadow ray finished
    ST_EV_PATH_END = 4,    // transient: becomes ST_DEAD + pending at the end of the trip
    ST_RETIRED = 5,        // no work left
    ST_WALK = 8,           // A: between majorant segments, DDA step next
    ST_SAMPLE = 9,         // A': inside a non-empty majorant segment, next tentative-collision distance to draw
    ST_COLLIDE = 10        // B: tentative collision at cur_tMin, medium lookup + event next
};

// film / sample write of a finished path: cpu/integrators.cpp:272-297, film.h:95-100,239-255
template <typename TR>
AVR_DEV void write_sample(const RenderParams &rp, const PathLane<TR> &ln, double *film, float *samples_out) {
    float L[4] = {ln.L[0], ln.L[1], ln.L[2], ln.L[3]};
    if (TR::LEAN) {
#pragma unroll
        for (int i = 0; i < 4; ++i) L[i] = ln.Lsum * sample_dense1(rp.light_spec, ln.lambda[i]);
    }
    bool bad = false;
#pragma unroll
    for (int i = 0; i < 4; ++i) bad = bad || isnan(L[i]) || isinf(L[i]);
    if (bad) L[0] = L[1] = L[2] = L[3] = 0.f;  // NaN / infinite radiance -> black
    float lam[4], pdf[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) lam[i] = ln.lambda[i];
    visible_wavelengths_pdf<TR::EXACT>(lam, pdf);
    if (rp.dump_samples) {
        float *o = samples_out + sample_slot(rp, ln.px, ln.py, ln.samp) * 12;
#pragma unroll
        for (int i = 0; i < 4; ++i) { o[i] = L[i]; o[4 + i] = lam[i]; o[8 + i] = pdf[i]; }
        return;
    }
    // PixelSensor::ToSensorRGB film.h:95-100 ; RGBFilm::AddSample film.h:239-255 (weight 1: box filter)
    float Ld[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) Ld[i] = pdf[i] != 0 ? L[i] / pdf[i] : 0.f;
    float rgb[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float s = sample_dense1(rp.sensor + c * kNSpec, lam[0]) * Ld[0];
#pragma unroll
        for (int i = 1; i < 4; ++i) s += sample_dense1(rp.sensor + c * kNSpec, lam[i]) * Ld[i];
        rgb[c] = rp.imaging_ratio * (s / 4);
    }
    float m = fmaxf(rgb[0], fmaxf(rgb[1], rgb[2]));
    if (m > rp.max_component) {
        float k = rp.max_component / m;
        rgb[0] *= k; rgb[1] *= k; rgb[2] *= k;
    }
    double *px = film + 4 * ((size_t)ln.py * rp.width + ln.px);
    // RGBFilm::AddSample: rgbSum[c] += weight * rgb[c] (a float product), weightSum += weight.  Box filter: weight 1
    // (adding +-0 leaves a sum unchanged); the tabulated-filter case is a branch of its own so that the common one
    // carries no weight at all
    if (__builtin_expect(rp.fweighted != 0, 0)) {
        const float w = rp.fsamples[sample_slot(rp, ln.px, ln.py, ln.samp)].z;
        atomicAdd(px + 0, (double)(w * rgb[0]));
        atomicAdd(px + 1, (double)(w * rgb[1]));
        atomicAdd(px + 2, (double)(w * rgb[2]));
        atomicAdd(px + 3, (double)w);
        return;
    }
    if (rgb[0] != 0) atomicAdd(px + 0, (double)(1.f * rgb[0]));
    if (rgb[1] != 0) atomicAdd(px + 1, (double)(1.f * rgb[1]));
    if (rgb[2] != 0) atomicAdd(px + 2, (double)(1.f * rgb[2]));
    atomicAdd(px + 3, 1.0);
}

// RayIntegrator::EvaluatePixelSample prologue (cpu/integrators.cpp:235-262): sampler start,
// wavelengths, box-filter camera sample, PerspectiveCamera::GenerateRayDifferential, entry into the
// boundary.  Leaves the pending camera ray in ln.fe.o / ln.fe.d.  Returns false for a padding item.
template <typename TR>
AVR_DEV bool start_path(const MediumView &mv, const RenderParams &rp, PathLane<TR> &ln, long long item,
                        LambdaCtx<TR::NCH> &lc) {
    long long npix_pad = rp.total_items / (rp.s1 - rp.s0);
    int s = rp.s0 + int(item / npix_pad);
    long long pi = item % npix_pad;
    int tile = int(pi >> 5), within = int(pi & 31);
    int px = rp.x0 + (tile % rp.tiles_x) * 8 + (within & 7);
    int py = rp.y0 + (tile / rp.tiles_x) * 4 + (within >> 3);
    if (px >= rp.x1 || py >= rp.y1) return false;  // padding lane of an edge tile
    ln.px = px; ln.py = py; ln.samp = s;
    PathSamplerRef<TR::ZSOBOL> smp{rp.smp, ln.srng, px, py, s};
    smp.start_pixel_sample();
    float lu, fx, fy;
    const bool pre = rp.fsamples != nullptr;
    if (TR::ZSOBOL && pre) {
        // zsobol: the pre-pass drew the wavelength sample and the film offset of every pixel sample of the launch
        // (three sample-index computations per path, at full lanes there instead of here)
        lu = rp.fsamples[sample_slot(rp, px, py, s)].w;
        smp.skip(3);
    } else
        lu = smp.get1d();
    float pdf[4];
    sample_visible_wavelengths<TR::EXACT>(lu, ln.lambda, pdf);
    lc = make_lambda_ctx<TR::NCH>(mv, ln.lambda);
    if (!(TR::ZSOBOL && pre)) {
        float f0, f1;
        smp.get2d(&f0, &f1);  // GetCameraSample: GetPixel2D, samplers.h:797-814
        fx = lerp_exact(f0, -rp.filter_rx, rp.filter_rx);  // BoxFilter::Sample, filters.h:67-70
        fy = lerp_exact(f1, -rp.filter_ry, rp.filter_ry);
    }
    if (__builtin_expect(pre, TR::ZSOBOL)) {
        // tabulated / triangle filter (FilterSampler::Sample), or zsobol's box offsets: sampled by the pre-pass from
        // these same two draws (a second read of the entry: cheaper than two registers held across the wavelength code)
        const float4 fs = rp.fsamples[sample_slot(rp, px, py, s)];
        fx = fs.x;
        fy = fs.y;
    }
    float pfx = (float(px) + fx) + 0.5f, pfy = (float(py) + fy) + 0.5f;
    smp.skip(3);          // time (Get1D), lens (Get2D): a pinhole camera reads neither
    // cameraFromRaster(pFilm): Transform::operator()(Point3f), util/transform.h:314-323
    const float *c = rp.cfr;
    float xp = c[0] * pfx + c[1] * pfy + c[2] * 0.f + c[3];
    float yp = c[4] * pfx + c[5] * pfy + c[6] * 0.f + c[7];
    float zp = c[8] * pfx + c[9] * pfy + c[10] * 0.f + c[11];
    float wp = c[12] * pfx + c[13] * pfy + c[14] * 0.f + c[15];
    V3 pc = v3(xp, yp, zp);
    if (wp != 1) pc = pc / wp;
    float len = sqrtf(sqr(pc.x) + sqr(pc.y) + sqr(pc.z));
    V3 dir = pc / len;
    V3 ro, rd;
    xform_ray_exact(rp.rfc, true, v3(0, 0, 0), dir, nullptr, &ro, &rd);
#pragma unroll
    for (int i = 0; i < 4; ++i) ln.L[i] = 0.f;
    ln.beta = sconst<TR::NCH>(1.f);
    ln.r_u = sconst<TR::NCH>(1.f);
    ln.r_l = sconst<TR::NCH>(1.f);
    ln.depth = 0;
    ln.shadow = false;
    ln.cam_wait = false;
    ln.Lsum = 0.f;
    ln.rr_k = 0;
    if (!TR::EXACT && !TR::ZSOBOL) {
        // FAST: one tracking stream per path (raw draws as PCG32 sequence / seed give correlated streams:
        // they go through the hash)
        uint64_t h0 = hash_u32(ln.srng.u32()), h1 = hash_u32(ln.srng.u32());
        ln.trng.set_sequence(h0, h1);
    }
    // entry into the boundary (Intersect + SkipIntersection of the "interface" surface)
    float t0, t1;
    if (!(boundary_hit(rp, ro, rd, &t0, &t1) && t1 > 0)) {
        ln.state = ST_EV_PATH_END;  // L = 0; retired by the event phase of this trip
        return true;
    }
    if (t0 > 0) ro = v3(ro.x + rd.x * t0, ro.y + rd.y * t0, ro.z + rd.z * t0);
    ln.fe.o = ro;
    ln.fe.d = rd;
    ln.state = ST_SETUP;
    return true;
}

// ---- CTA path pool ------------------------------------------------------------------------------------
// A CTA owns `pool` paths whose state lives in shared-memory slots; a lane copies a slot into
// registers, runs one op on it and copies it back.  Between ops a path waits in one of four queues
// (slot ids), so ANY lane of the CTA can pick it up: event ops run on full 32-path batches and the
// tracking loop refills a lane the moment its path hits an event.

enum { Q_DEAD = 0, Q_SCATTER = 1, Q_SHADOW_END = 2, Q_TRACK = 3, kNumQueues = 4 };

// Bounded MPMC ring of slot ids.  A cell word is (round << 16) | slot while it holds an entry and
// (round << 16) | 0xffff while it is free for the producer of that round (Vyukov's per-cell sequence,
// folded into one word so that publishing an entry is a single store).  tail - head counts entries
// that are at least reserved; a consumer that claims a reserved cell spins until its producer's store.
struct PoolQueue {
    unsigned head, tail;  // monotonic positions
};

struct PoolView {
    uint4 *slots;              // pool * kSlotQuads
    volatile unsigned *cells;  // kNumQueues * cap
    PoolQueue *q;              // kNumQueues
    volatile int *live;        // slots not yet retired
    unsigned cap_mask;         // cap - 1 (cap = power of two >= pool)
    unsigned cap_log2;
};

AVR_DEV int pool_fill(const PoolView &pv, int qi) {
    unsigned h = *(volatile unsigned *)&pv.q[qi].head, t = *(volatile unsigned *)&pv.q[qi].tail;
    return (int)(t - h);
}

// Pop up to popc(takers) slot ids from queue qi.  Returns n; the k-th taker lane (k < n) receives an
// id in *slot.  All 32 lanes must call.  A __threadfence_block() must follow before the slots are read.
AVR_DEV int pool_pop(const PoolView &pv, int qi, unsigned takers, unsigned lane, int *slot) {
    int want = __popc(takers);
    int n = 0;
    unsigned pos0 = 0;
    if (lane == 0 && want > 0) {
        unsigned h = *(volatile unsigned *)&pv.q[qi].head;
        while (true) {
            int avail = (int)(*(volatile unsigned *)&pv.q[qi].tail - h);
            n = avail < want ? avail : want;
            if (n <= 0) { n = 0; break; }
            unsigned old = atomicCAS(&pv.q[qi].head, h, h + (unsigned)n);
            if (old == h) break;
            h = old;
        }
        pos0 = h;
    }
    n = __shfl_sync(0xffffffffu, n, 0);
    pos0 = __shfl_sync(0xffffffffu, pos0, 0);
    if (takers >> lane & 1u) {
        int k = __popc(takers & ((1u << lane) - 1));
        if (k < n) {
            unsigned pos = pos0 + k;
            unsigned cell = ((unsigned)qi << pv.cap_log2) + (pos & pv.cap_mask);
            unsigned round = (pos >> pv.cap_log2) << 16;
            unsigned w;
            while (((w = pv.cells[cell]) ^ round) >= 0xffffu) {}  // other round, or still free
            *slot = int(w & 0xffffu);
            pv.cells[cell] = ((round + 0x10000u) | 0xffffu);
        }
    }
    return n;
}

// Slot <-> register copy of a path, in three groups of 16-byte quads so that an op moves only what
// it touches:
//   A  event-only state   : pixel, sample index, sampler stream, wo, f_hat, ray o/d, wavelengths, L
//   B  shared small state : flags, throughputs, shadow-ray weights, T_maj, current segment, sampled spectra
//   C  tracking input     : sample-space ray, DDA, tracking RNG
// Events load A+B and store A+B+C; the tracking loop loads B+C (plus A where the variant needs the
// render-space ray or accumulates emission) and parks a finished path by storing B alone.
enum { GRP_A = 1, GRP_B = 2, GRP_C = 4, GRP_ALL = 7 };

template <typename TR>
struct SlotIO {
    static constexpr int N = TR::NCH;
    static constexpr int kQuadsA = TR::LEAN ? 4 : 6;
    // FAST kernels with the independent sampler seed the tracking RNG once per path (start_path) instead of
    // once per ray: it then belongs to the state every op touches (group B), not to the per-ray group C
    static constexpr bool kTrngInB = !TR::EXACT && !TR::ZSOBOL;
    static constexpr int kWordsB = (TR::LEAN ? 8 : 4 + 10 * N) + (kTrngInB ? 4 : 0);
    static constexpr int kQuadsB = (kWordsB + 3) / 4;
    static constexpr int kQuadsC = kTrngInB ? 4 : 5;
    // Traits::FUSED: a second ray block (the camera ray that waits while the shadow ray is tracked)
    static constexpr int kQuadsC2 = TR::FUSED ? kQuadsC : 0;
    static constexpr int kOffC = kQuadsA + kQuadsB;
    static constexpr int kSlotQuads = (kQuadsA + kQuadsB + kQuadsC + kQuadsC2) | 1;  // odd: consecutive slots start in different bank groups
    // groups the tracking loop needs besides B and C, and what it must write back when parking
    static constexpr int kTrackLoad = GRP_B | GRP_C | ((TR::EXACT || TR::GENERAL) ? GRP_A : 0);
    static constexpr int kTrackPark = GRP_B | (TR::GENERAL ? GRP_A : 0);

    template <int Q>
    struct Packer {
        unsigned w[Q * 4];
        int k = 0;
        AVR_DEV void operator()(unsigned &x) { w[k++] = x; }
        AVR_DEV void operator()(float &x) { w[k++] = __float_as_uint(x); }
        AVR_DEV void flush(uint4 *sp) {
#pragma unroll
            for (int i = k; i < Q * 4; ++i) w[i] = 0;
#pragma unroll
            for (int q = 0; q < Q; ++q) sp[q] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
        }
    };
    template <int Q>
    struct Unpacker {
        unsigned w[Q * 4];
        int k = 0;
        AVR_DEV void fill(const uint4 *sp) {
#pragma unroll
            for (int q = 0; q < Q; ++q) {
                uint4 v = sp[q];
                w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
            }
        }
        AVR_DEV void operator()(unsigned &x) { x = w[k++]; }
        AVR_DEV void operator()(float &x) { x = __uint_as_float(w[k++]); }
    };
    AVR_DEV static void split64(uint64_t v, unsigned &lo, unsigned &hi) { lo = (unsigned)v; hi = (unsigned)(v >> 32); }
    AVR_DEV static uint64_t join64(unsigned lo, unsigned hi) { return ((uint64_t)hi << 32) | lo; }

    template <typename F>
    AVR_DEV static void visit_a(PathLane<TR> &ln, F &f) {
        unsigned pxy = (unsigned)ln.px | ((unsigned)ln.py << 16), sm = (unsigned)ln.samp, lo, hi, lo2, hi2;
        split64(ln.srng.state, lo, hi);
        split64(ln.srng.inc, lo2, hi2);
        f(pxy); f(sm); f(lo); f(hi); f(lo2); f(hi2);
        ln.px = int(pxy & 0xffffu); ln.py = int(pxy >> 16); ln.samp = (int)sm;
        ln.srng.state = join64(lo, hi); ln.srng.inc = join64(lo2, hi2);
        if (TR::LEAN) {
            // the render-space camera ray (wo = -d; f_hat and the radiance sum live in group B)
            f(ln.fe.o.x); f(ln.fe.o.y); f(ln.fe.o.z); f(ln.fe.d.x); f(ln.fe.d.y); f(ln.fe.d.z);
#pragma unroll
            for (int i = 0; i < 4; ++i) f(ln.lambda[i]);
            return;
        }
        f(ln.wo.x); f(ln.wo.y); f(ln.wo.z); f(ln.f_hat);
        f(ln.fe.o.x); f(ln.fe.o.y); f(ln.fe.o.z); f(ln.fe.d.x); f(ln.fe.d.y); f(ln.fe.d.z);
#pragma unroll
        for (int i = 0; i < 4; ++i) f(ln.lambda[i]);
#pragma unroll
        for (int i = 0; i < 4; ++i) f(ln.L[i]);
    }
    template <typename F>
    AVR_DEV static void visit_b(PathLane<TR> &ln, LambdaCtx<N> &lc, bool &pending, F &f) {
        unsigned fl = (unsigned)ln.state | ((unsigned)ln.shadow << 8) | ((unsigned)pending << 9) | ((unsigned)ln.depth << 16);
        if (TR::FUSED) fl |= (unsigned)ln.cam_wait << 10;
        if (TR::LEAN) fl |= (unsigned)min(ln.rr_k, 31) << 11;
        f(fl);
        ln.state = int(fl & 0xffu); ln.shadow = (fl >> 8) & 1u; pending = (fl >> 9) & 1u; ln.depth = int(fl >> 16);
        if (TR::FUSED) ln.cam_wait = (fl >> 10) & 1u;
        if (TR::LEAN) ln.rr_k = int((fl >> 11) & 31u);
        f(ln.cur_tMin); f(ln.seg_tMax); f(ln.u);
        if (kTrngInB) visit_rng(ln.trng, f);
        if (TR::LEAN) {
            // the medium's sigma_a / sigma_s are constants of the MediumView, the path weights are not carried
            f(ln.T_ray.v[0]); f(ln.sigma_maj.v[0]); f(ln.f_hat); f(ln.Lsum);
            return;
        }
#pragma unroll
        for (int i = 0; i < N; ++i) {
            f(ln.beta.v[i]); f(ln.r_u.v[i]); f(ln.r_l.v[i]);
            f(ln.T_ray.v[i]); f(ln.s_rl.v[i]); f(ln.s_ru.v[i]);
            f(ln.T_maj.v[i]); f(ln.sigma_maj.v[i]);
            f(lc.sig_a.v[i]); f(lc.sig_s.v[i]);
        }
    }
    template <typename F>
    AVR_DEV static void visit_c(PathLane<TR> &ln, F &f) {
        f(ln.fe.so.x); f(ln.fe.so.y); f(ln.fe.so.z); f(ln.fe.sd.x); f(ln.fe.sd.y); f(ln.fe.sd.z);
        f(ln.dda.tMin); f(ln.dda.tMax);
#pragma unroll
        for (int i = 0; i < 3; ++i) { f(ln.dda.nct[i]); f(ln.dda.dTs[i]); }
        unsigned dl = (unsigned)ln.dda.lin;
        f(dl); f(ln.dda.rem);
        ln.dda.lin = (int)dl;
        if (!kTrngInB) visit_rng(ln.trng, f);
    }
    template <typename F>
    AVR_DEV static void visit_rng(RNG &r, F &f) {
        unsigned lo, hi, lo2, hi2;
        split64(r.state, lo, hi);
        split64(r.inc, lo2, hi2);
        f(lo); f(hi); f(lo2); f(hi2);
        r.state = join64(lo, hi); r.inc = join64(lo2, hi2);
    }

    template <int GROUPS>
    AVR_DEV static void store(uint4 *sp, PathLane<TR> &ln, LambdaCtx<N> &lc, bool pending) {
        if (GROUPS & GRP_A) { Packer<kQuadsA> p; visit_a(ln, p); p.flush(sp); }
        if (GROUPS & GRP_B) { Packer<kQuadsB> p; visit_b(ln, lc, pending, p); p.flush(sp + kQuadsA); }
        if (GROUPS & GRP_C) { Packer<kQuadsC> p; visit_c(ln, p); p.flush(sp + kOffC); }
    }
    // the ray block alone: `second` selects the waiting-ray block of Traits::FUSED
    AVR_DEV static void store_ray(uint4 *sp, PathLane<TR> &ln, bool second) {
        Packer<kQuadsC> p; visit_c(ln, p); p.flush(sp + kOffC + (second ? kQuadsC : 0));
    }
    AVR_DEV static void load_waiting_ray(const uint4 *sp, PathLane<TR> &ln) {
        Unpacker<kQuadsC> u; u.fill(sp + kOffC + kQuadsC); visit_c(ln, u);
    }
    template <int GROUPS>
    AVR_DEV static void load(const uint4 *sp, const MediumView &mv, PathLane<TR> &ln, LambdaCtx<N> &lc, bool &pending) {
        if (GROUPS & GRP_A) { Unpacker<kQuadsA> u; u.fill(sp); visit_a(ln, u); }
        if (GROUPS & GRP_B) {
            Unpacker<kQuadsB> u;
            u.fill(sp + kQuadsA);
            visit_b(ln, lc, pending, u);
            // derived: sigma_t and the wavelengths of the batch (make_lambda_ctx, avr_tmaj.cuh)
#pragma unroll
            for (int i = 0; i < 4; ++i) lc.lambda[i] = ln.lambda[i];
            if (TR::LEAN) {
                lc.sig_a = sconst<N>(mv.ach_sig_a);
                lc.sig_s = sconst<N>(mv.ach_sig_s);
            }
            if (mv.kind == KIND_RGB) lc.sigma_t = sconst<N>(1.f);
            else lc.sigma_t = lc.sig_a + lc.sig_s;
        }
        if (GROUPS & GRP_C) { Unpacker<kQuadsC> u; u.fill(sp + kOffC); visit_c(ln, u); }
    }
};

// shared-memory layout, computed identically on host and device
struct PoolLayout {
    unsigned off_slots, off_cells, off_q, off_live, total, cap, cap_log2;
};
__host__ __device__ inline PoolLayout pool_layout(unsigned maj_bytes, unsigned pool, unsigned slot_quads) {
    PoolLayout l;
    l.cap = 32;
    l.cap_log2 = 5;
    while (l.cap < pool) { l.cap <<= 1; l.cap_log2++; }
    l.off_slots = (maj_bytes + 15u) & ~15u;
    l.off_cells = l.off_slots + pool * slot_quads * 16u;
    l.off_q = l.off_cells + kNumQueues * l.cap * 4u;
    l.off_live = l.off_q + kNumQueues * (unsigned)sizeof(PoolQueue);
    l.total = l.off_live + 16u;
    return l;
}

#ifndef AVR_LEAP_MIN
#define AVR_LEAP_MIN 4  // FAST: smallest empty-cube radius (in cells) worth a DDA::leap instead of single steps
#endif
#ifndef AVR_CTA_THREADS
#define AVR_CTA_THREADS 512  // one 512-thread CTA per SM with one pool: 1.6 % faster than two of 256 (fuller batches)
#endif
#ifndef AVR_CTA_THREADS_FAST
// FAST scalar kernels (either sampler) fit 96 registers without a spill: 5 warps per scheduler instead
// of 4 hide more of the fetch latency (8 spp: 512 threads 19.1 ms, 576 17.9, 640 17.05, 704 17.9 (spills), 768 17.9);
// the EXACT kernels spill below 123 registers and lose (512: 52.1 ms, 576: 55.7, 640: 53.6, 704: 59.3)
#define AVR_CTA_THREADS_FAST 640
#endif
template <typename TR>
struct KernelCfg {
    // achromatic scalar grids: AVR_MIN_BLOCKS CTAs of 256 threads (<= 128 registers); the chromatic and general
    // kernels need 144..181 registers: one CTA per SM with as many warps as the register file holds
    static constexpr bool kLean = TR::NCH == 1 && !TR::GENERAL;
    static constexpr int kLeanThreads = !TR::EXACT ? AVR_CTA_THREADS_FAST : AVR_CTA_THREADS;
    static constexpr int kThreads = kLean ? kLeanThreads : 384;  // (not lean) 3 warps per scheduler: 168-register cap
#ifdef AVR_LEAN_MIN_BLOCKS
    static constexpr int kMinBlocks = kLean ? AVR_LEAN_MIN_BLOCKS : 1;
#else
    static constexpr int kMinBlocks =
        (kLean && AVR_MIN_BLOCKS * 256 >= kLeanThreads) ? AVR_MIN_BLOCKS * 256 / kLeanThreads : 1;
#endif
};

struct LaneCounters {
    unsigned rays, cam, sh, seg, sc;
};

// After an op: hand every path of the batch to the queue of the op it waits for next.
template <typename TR, int GROUPS>
AVR_DEV void requeue(const PoolView &pv, unsigned lane, bool have, int slot, PathLane<TR> &ln,
                     LambdaCtx<TR::NCH> &lc, bool pending) {
    if (have && ln.state == ST_EV_PATH_END) {
        // absorbed / escaped / max depth / zero throughput: the film write happens in the next R op
        ln.state = ST_DEAD;
        pending = true;
    }
    if (have && ln.state != ST_RETIRED)
        SlotIO<TR>::template store<GROUPS>(pv.slots + slot * SlotIO<TR>::kSlotQuads, ln, lc, pending);
    __threadfence_block();  // slot contents before the queue cells that publish them
    // queue the path of this lane goes to (4: none).  All queues are served by ONE shared-memory atomic
    // instruction: the first lane of every group of lanes with the same target reserves the group's cells
    // (four pool_push calls in a row cost four dependent atomic round trips per requeue)
    int tq = 4;
    if (have) {
        const int s = ln.state;
        tq = s >= ST_WALK ? Q_TRACK : (s == ST_EV_SCATTER ? Q_SCATTER : (s == ST_EV_SHADOW_END ? Q_SHADOW_END : (s == ST_DEAD ? Q_DEAD : 5)));
    }
    const unsigned peers = __match_any_sync(0xffffffffu, tq);
    const int leader = __ffs(peers) - 1;
    unsigned pos0 = 0;
    if ((int)lane == leader) {
        if (tq < 4) pos0 = atomicAdd(&pv.q[tq].tail, (unsigned)__popc(peers));
        else if (tq == 5) atomicSub((int *)pv.live, __popc(peers));  // retired slots
    }
    pos0 = __shfl_sync(0xffffffffu, pos0, leader);
    if (tq < 4) {
        unsigned pos = pos0 + __popc(peers & ((1u << lane) - 1));
        unsigned cell = ((unsigned)tq << pv.cap_log2) + (pos & pv.cap_mask);
        unsigned round = (pos >> pv.cap_log2) << 16;
        while (pv.cells[cell] != (round | 0xffffu)) {}
        pv.cells[cell] = round | (unsigned)slot;
    }
}

// One batch of an event op (R, E1 or E2, chosen warp-uniformly by the queue served), followed by the
// ray setup S for every path of the batch that now has a pending ray.
template <typename TR>
AVR_DEV void event_batch(const MediumView &mv, const RenderParams &rp, const PoolView &pv, int qi, unsigned lane,
                         PathLane<TR> &ln, LambdaCtx<TR::NCH> &lc, LaneCounters &cn, unsigned long long *counters,
                         double *film, float *samples_out) {
    constexpr int N = TR::NCH;
    int slot = 0;
    int n = pool_pop(pv, qi, 0xffffffffu, lane, &slot);
    if (n == 0) return;
    __threadfence_block();
    bool have = (int)lane < n;
    bool pending = false;
    bool sh_ray = false;        // Traits::FUSED: a shadow ray to set up (and track) before the pending camera ray
    V3 sh_d = v3(0.f, 0.f, 0.f);
    if (have) SlotIO<TR>::template load<GRP_A | GRP_B>(pv.slots + slot * SlotIO<TR>::kSlotQuads, mv, ln, lc, pending);

    if (qi == Q_DEAD) {
        // ---- R. regenerate: flush the finished sample, fetch work with one aggregated atomic ---------
        if (have && pending) {
            write_sample(rp, ln, film, samples_out);
            pending = false;
        }
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(&counters[CNT_WORK], (unsigned long long)n);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (have) {
            long long idx = (long long)base + lane;
            if (idx < rp.local_items) {
                long long chunk = idx >> 5;
                long long item = ((chunk * rp.world + rp.rank) << 5) + (idx & 31);
                if (start_path(mv, rp, ln, item, lc)) cn.rays++;
                else ln.state = ST_DEAD;  // padding item of an edge tile: nothing pending, fetch again
            } else
                ln.state = ST_RETIRED;
        }
    } else if (TR::FUSED && qi == Q_SCATTER) {
        // ---- E1 + E2 in one op (Traits::FUSED): SampleLd prologue (cpu/integrators.cpp:1300-1338) and the phase
        //      sampling that follows SampleLd (:1041-1060).  Leaves the shadow ray in sh_d and the next camera
        //      segment pending in fe.o / fe.d; the sampler draws keep the reference's order (light, then phase,
        //      then the camera ray's three) --------------------------------------------------------------------
        if (have) {
            cn.sc++;
            const V3 p = ln.fe.point(ln.cur_tMin);
            const V3 wo = -ln.fe.d;
            PathSamplerRef<TR::ZSOBOL> smp{rp.smp, ln.srng, ln.px, ln.py, ln.samp};
            float ua, ub;
            smp.skip(3);  // light choice (Get1D: one light, p = 1) and uLight (Get2D: unused by a delta light)
            const V3 wi = v3(rp.light_dir[0], rp.light_dir[1], rp.light_dir[2]);
            const float two_r = 2 * rp.scene_radius;
            bool lit = false;
#pragma unroll
            for (int i = 0; i < 4; ++i) lit = lit || sample_dense1(rp.light_spec, ln.lambda[i]) != 0;
            const float ph = hg_phase<TR::EXACT>(dot(wo, wi), mv.g);
            sh_d = v3(wi.x * two_r, wi.y * two_r, wi.z * two_r);  // pLight - p
            sh_ray = lit && ph != 0 && two_r != 0;
            if (sh_ray) {
                ln.f_hat = ph;
                ln.T_ray = sconst<N>(1.f);
                ln.rr_k = 0;
            }
            smp.get2d(&ua, &ub);
            float pdf;
            const V3 wi2 = sample_hg<TR::EXACT>(wo, mv.g, ua, ub, &pdf);
            ln.fe.o = p;  // scatter point: origin of both rays
            ln.cam_wait = false;
            if (pdf == 0) {
                ln.state = ST_EV_PATH_END;  // no camera continuation (the shadow ray may still contribute)
            } else {
                ln.fe.d = wi2;
                ln.state = ST_SETUP;
            }
        }
    } else if (!TR::FUSED && qi == Q_SCATTER) {
        // ---- E1. real scattering: SampleLd prologue (cpu/integrators.cpp:1300-1338) -> shadow ray ----
        if (have) {
            cn.sc++;
            V3 p = ln.fe.point(ln.cur_tMin);
            {
                PathSamplerRef<TR::ZSOBOL> smp{rp.smp, ln.srng, ln.px, ln.py, ln.samp};
                smp.skip(3);  // light choice (Get1D: one light, p = 1) and uLight (Get2D: unused by a delta light)
            }
            V3 wi = v3(rp.light_dir[0], rp.light_dir[1], rp.light_dir[2]);
            float two_r = 2 * rp.scene_radius;
            V3 pLight = v3(p.x + wi.x * two_r, p.y + wi.y * two_r, p.z + wi.z * two_r);
            bool lit = false;
#pragma unroll
            for (int i = 0; i < 4; ++i) lit = lit || sample_dense1(rp.light_spec, ln.lambda[i]) != 0;
            float ph = hg_phase<TR::EXACT>(dot(ln.wo, wi), mv.g);
            V3 rd = pLight - p;
            ln.fe.o = p;  // scatter point: origin of the shadow ray and of the next camera segment
            if (lit && ph != 0 && !(rd.x == 0 && rd.y == 0 && rd.z == 0)) {
                ln.f_hat = ph;
                ln.T_ray = sconst<N>(1.f);
                ln.s_rl = sconst<N>(1.f);
                ln.s_ru = sconst<N>(1.f);
                ln.fe.d = rd;
                ln.shadow = true;
                ln.state = ST_SETUP;
            } else {
                ln.T_ray = sconst<N>(0.f);  // Ld = 0: straight to phase sampling
                ln.T_maj = sconst<N>(1.f);
                ln.state = ST_EV_SHADOW_END;
            }
        }
    } else if (!TR::FUSED) {
        // ---- E2. end of the shadow ray: SampleLd epilogue (cpu/integrators.cpp:1382-1398), then phase
        //      sampling and the next camera segment (cpu/integrators.cpp:1041-1060) -------------------
        if (have) {
            Spec<N> T_ray = ln.T_ray * (ln.T_maj / ln.T_maj[0]);
            if (nonzero(T_ray)) {
                Spec<N> r_l = ln.s_rl * (ln.T_maj / ln.T_maj[0]);
                r_l = r_l * (ln.r_u * 1.f);  // r_l *= r_p * p_l ; p_l = 1 (single delta light)
                float avg = average(r_l);
                Spec<N> bf = ln.beta * ln.f_hat;
                bf = bf * T_ray;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float Ll = sample_dense1(rp.light_spec, ln.lambda[i]);
                    ln.L[i] += (bf[i % N] * Ll) / avg;
                }
            }
        }
        if (have) {
            float ua, ub;
            PathSamplerRef<TR::ZSOBOL>{rp.smp, ln.srng, ln.px, ln.py, ln.samp}.get2d(&ua, &ub);
            float pdf;
            V3 wi = sample_hg<TR::EXACT>(ln.wo, mv.g, ua, ub, &pdf);
            if (pdf == 0) {
                ln.state = ST_EV_PATH_END;
            } else {
                if (!TR::LEAN) {
                    ln.beta = ln.beta * (pdf / pdf);
                    ln.r_l = ln.r_u / pdf;
                }
                ln.fe.d = wi;  // fe.o still holds the scatter point
                ln.shadow = false;
                ln.state = ST_SETUP;
            }
        }
    }

    // ---- S. setup of the pending ray (camera segment or shadow ray).  Traits::FUSED: two passes over ONE copy of
    //      the code, the shadow ray of a scattering event first (it is the ray the tracking loop walks first), then
    //      the pending camera ray; a ray block goes to the slot as soon as it is complete -------------------------
    uint4 *const sp = pv.slots + slot * SlotIO<TR>::kSlotQuads;
#pragma unroll 1
    for (int pass = (TR::FUSED && qi == Q_SCATTER) ? 0 : 1; pass < 2; ++pass) {
        if (have && (pass == 0 ? sh_ray : ln.state == ST_SETUP)) {
            const V3 cam_o = ln.fe.o, cam_d = ln.fe.d;
            const int cam_state = ln.state;
            if (TR::FUSED) ln.shadow = pass == 0;
            V3 o = ln.fe.o, d = (TR::FUSED && pass == 0) ? sh_d : ln.fe.d;
            float t0, t1, tMax;
            bool hit;
            if (TR::EXACT) {
                hit = boundary_hit(rp, o, d, &t0, &t1) && t1 > 0;
                if (ln.shadow) {
                    // cpu/integrators.cpp:1336-1350: RNG(Hash(o), Hash(d)); tMax = exit or 1 - ShadowEpsilon
                    ln.trng.set_sequence(hash_v3(o), hash_v3(d));
                    tMax = 1 - kShadowEps;
                    if (hit && t1 < tMax) tMax = t1;
                    ln.u = ln.trng.uniform();
                    hit = true;
                } else {
                    // cpu/integrators.cpp:976-988: tMax = boundary exit, tracking RNG seeded from the sampler
                    ln.wo = -d;
                    tMax = t1;
                    if (hit) {
                        PathSamplerRef<TR::ZSOBOL> smp{rp.smp, ln.srng, ln.px, ln.py, ln.samp};
                        uint64_t hash0 = hash_float(smp.get1d());
                        uint64_t hash1 = hash_float(smp.get1d());
                        ln.trng.set_sequence(hash0, hash1);
                        ln.u = smp.get1d();
                    }
                }
            } else {
                // FAST: same estimator; the tracking stream of either ray kind is seeded from two hashed sampler draws
                hit = boundary_hit_fast(rp, o, d, &t0, &t1) && t1 > 0;
                // (raw draws as PCG32 sequence / seed give correlated streams: they must go through the hash)
                uint64_t hash0, hash1;
                if (TR::ZSOBOL) {
                    // zsobol keeps the reference's dimension layout: camera rays take the three draws of
                    // cpu/integrators.cpp:985-988, shadow rays hash their origin and direction (:1338)
                    if (ln.shadow) {
                        hash0 = hash_v3(o);
                        hash1 = hash_v3(d);
                    } else {
                        // the reference seeds the tracking RNG with Hash() of two sampler draws and takes u from a
                        // third (:985-988): the hash throws the draws' stratification away, so FAST hashes the
                        // path's identity and dimension instead of paying three ZSobol sample-index computations
                        PathSamplerRef<TR::ZSOBOL> smp{rp.smp, ln.srng, ln.px, ln.py, ln.samp};
                        const uint64_t id = ((uint64_t)(uint32_t)ln.py << 44) ^ ((uint64_t)(uint32_t)ln.px << 24) ^
                                            (uint64_t)(uint32_t)ln.samp ^ ((uint64_t)(uint32_t)rp.smp.seed << 52);
                        hash0 = mix_bits(id ^ (ln.srng.state * 0x9e3779b97f4a7c15ull));
                        hash1 = mix_bits(hash0 + 0xda942042e4dd58b5ull);
                        smp.skip(3);
                    }
                    ln.trng.set_sequence(hash0, hash1);
                }
                // (independent sampler: the path's tracking stream, seeded in start_path, simply continues)
                if (!(TR::FUSED && pass == 1 && sh_ray)) ln.u = neg_log1m(ln.trng.uniform());  // remaining optical depth to the first tentative collision
                if (ln.shadow) {
                    tMax = 1 - kShadowEps;
                    if (hit && t1 < tMax) tMax = t1;
                    hit = true;
                } else {
                    ln.wo = -d;
                    tMax = t1;
                }
            }
            if (!hit) {
                ln.state = ST_EV_PATH_END;  // the ray leaves without touching the boundary
            } else if (TR::EXACT) {
                // SampleT_maj prologue (media.h:744-749) + SampleRay (media.h:322-337)
                float len = sqrtf(sqr(d.x) + sqr(d.y) + sqr(d.z));
                tMax *= len;
                d = d / len;
                MRay mr = ray_to_medium(mv, o, d, tMax);
                if (mr.hit) ln.dda.template init<true>(mv, mr.o, mr.d, mr.tMin, mr.tMax); else ln.dda.invalidate();
                ln.fe.bind(mv, o, d);
                ln.T_maj = sconst<N>(1.f);
                ln.state = ST_WALK;
            } else {
                float l2 = __fmaf_rn(d.x, d.x, __fmaf_rn(d.y, d.y, d.z * d.z));
                float rl = rsqrtf(l2);
                tMax *= l2 * rl;
                d = d * rl;
                const float *a = mv.minv;
                V3 mo = v3(__fmaf_rn(a[0], o.x, __fmaf_rn(a[1], o.y, __fmaf_rn(a[2], o.z, a[3]))),
                           __fmaf_rn(a[4], o.x, __fmaf_rn(a[5], o.y, __fmaf_rn(a[6], o.z, a[7]))),
                           __fmaf_rn(a[8], o.x, __fmaf_rn(a[9], o.y, __fmaf_rn(a[10], o.z, a[11]))));
                V3 md = v3(__fmaf_rn(a[0], d.x, __fmaf_rn(a[1], d.y, a[2] * d.z)),
                           __fmaf_rn(a[4], d.x, __fmaf_rn(a[5], d.y, a[6] * d.z)),
                           __fmaf_rn(a[8], d.x, __fmaf_rn(a[9], d.y, a[10] * d.z)));
                V3 inv = v3(ftz_rcp(md.x), ftz_rcp(md.y), ftz_rcp(md.z));
                float h0 = 0.f, h1 = tMax;
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    float tn = (mv.bmin[i] - comp(mo, i)) * comp(inv, i), tf = (mv.bmax[i] - comp(mo, i)) * comp(inv, i);
                    h0 = fmaxf(h0, fminf(tn, tf));
                    h1 = fminf(h1, fmaxf(tn, tf));
                }
                if (h0 <= h1) ln.dda.template init_fast<true>(mv, mo, inv, md, h0, h1); else ln.dda.invalidate();
                ln.fe.bind(mv, o, d);
                ln.T_maj = sconst<N>(1.f);
                ln.state = ST_WALK;
            }
            if (TR::FUSED) {
                if (pass == 0) {
                    SlotIO<TR>::store_ray(sp, ln, false);
                    ln.fe.o = cam_o;  // the camera ray is still pending
                    ln.fe.d = cam_d;
                    ln.state = cam_state;
                } else if (ln.state == ST_WALK) {
                    SlotIO<TR>::store_ray(sp, ln, sh_ray);
                    ln.cam_wait = sh_ray;
                }
            }
        }
    }
    if (TR::FUSED) {
        if (have && sh_ray) {
            // tracked first: the shadow ray (its u was drawn in pass 0; a camera ray that left without touching the
            // boundary, or pdf = 0, simply leaves nothing waiting)
            ln.shadow = true;
            ln.state = ST_WALK;
        }
        requeue<TR, GRP_A | GRP_B>(pv, lane, have, slot, ln, lc, pending);
    } else
        requeue<TR, GRP_ALL>(pv, lane, have, slot, ln, lc, pending);
}

// Inside the tracking loop a shadow-ray lane carries (T_ray, r_u, r_l of the shadow ray) in the fields
// (beta, r_u, r_l), so that camera and shadow lanes share the weight-update instructions of op B.
// Applied after a slot load and undone before a slot store.
template <typename TR>
AVR_DEV void swap_shadow_weights(PathLane<TR> &ln) {
    if (TR::LEAN || !ln.shadow) return;
    Spec<TR::NCH> t;
    t = ln.beta; ln.beta = ln.T_ray; ln.T_ray = t;
    t = ln.r_u; ln.r_u = ln.s_ru; ln.s_ru = t;
    t = ln.r_l; ln.r_l = ln.s_rl; ln.s_rl = t;
}

// Tracking loop of one warp: every lane holds a path in registers and walks SampleT_maj (ops A and B);
// a lane whose path reaches an event parks it in its slot, queues it for the event op and is refilled
// from Q_TRACK.  The warp leaves when it runs dry, or yields its few remaining paths back to Q_TRACK
// when event batches are waiting.
template <typename TR>
AVR_DEV void track_mode(const MediumView &mv, const RenderParams &rp, const PoolView &pv, const float *maj,
                        unsigned lane, PathLane<TR> &ln, LambdaCtx<TR::NCH> &lc, LaneCounters &cn) {
    constexpr int N = TR::NCH;
    constexpr bool EX = TR::EXACT;
    bool have = false, pending = false;  // have: the lane holds a path (tracking, or finished and not yet parked)
    int slot = 0;
    ln.state = ST_DEAD;
    while (true) {
        if (TR::FUSED && ln.state == ST_EV_SHADOW_END) {
            // end of the NEE shadow ray, SampleLd epilogue (cpu/integrators.cpp:1382-1398; one channel: every weight
            // but f_hat * T_ray is 1, the light's spectrum is applied in write_sample), then on to the camera ray
            // that was set up together with it -- without a trip through the queues
            if (ln.T_ray[0] != 0) ln.Lsum = __fmaf_rn(ln.f_hat, ln.T_ray[0], ln.Lsum);
            if (ln.cam_wait) {
                SlotIO<TR>::load_waiting_ray(pv.slots + slot * SlotIO<TR>::kSlotQuads, ln);
                ln.cam_wait = false;
                ln.shadow = false;
                ln.u = neg_log1m(ln.trng.uniform());
                ln.state = ST_WALK;
                ln.mj = ln.dda.tMin < ln.dda.tMax ? maj[ln.dda.lin] : 0.f;
            } else
                ln.state = ST_EV_PATH_END;
        }
        // ---- park finished paths and refill from Q_TRACK, both in batches of >= refill_lanes -------
        bool fin = have && ln.state < ST_WALK;
        unsigned idle = __ballot_sync(0xffffffffu, !have || fin);
        if (__popc(idle) >= rp.refill_lanes || idle == 0xffffffffu) {
            if (__any_sync(0xffffffffu, fin)) {
                if (fin) swap_shadow_weights(ln);
                requeue<TR, SlotIO<TR>::kTrackPark>(pv, lane, fin, slot, ln, lc, pending);
                if (fin) {
                    have = false;
                    ln.state = ST_DEAD;
                }
            }
            int got = -1;
            int n = pool_pop(pv, Q_TRACK, idle, lane, &got);
            __threadfence_block();
            if (!have && __popc(idle & ((1u << lane) - 1)) < n) {
                slot = got;
                SlotIO<TR>::template load<SlotIO<TR>::kTrackLoad>(pv.slots + slot * SlotIO<TR>::kSlotQuads, mv, ln, lc, pending);
                swap_shadow_weights(ln);
                have = true;
                ln.mj = ln.dda.tMin < ln.dda.tMax ? maj[ln.dda.lin] : 0.f;
                // a path that was given back with its tentative collision pending: request the taps again
                if (Fetcher<TR>::kSplitFetch && ln.state == ST_COLLIDE) ln.fe.request(mv, ln.cur_tMin, ln.tapA, ln.tapB);
            }
            int active = 32 - __popc(idle) + n;
            if (active == 0) return;
            if (n == 0) {
                // nothing to refill with: give the remaining paths back if event work is waiting
                int ev = 0;
                if (lane == 0) {
                    int a = pool_fill(pv, Q_SCATTER), b = pool_fill(pv, Q_SHADOW_END), c = pool_fill(pv, Q_DEAD);
                    ev = a > b ? a : b;
                    ev = c > ev ? c : ev;
                }
                ev = __shfl_sync(0xffffffffu, ev, 0);
                if ((ev >= 32 && active <= rp.yield_lanes_full) || (ev > 0 && active <= rp.yield_lanes)) {
                    if (have) swap_shadow_weights(ln);
                    requeue<TR, SlotIO<TR>::kTrackLoad>(pv, lane, have, slot, ln, lc, pending);
                    return;
                }
            }
        }

        // ---- A. DDA steps to the next majorant segment with sigma_maj[0] != 0, each followed by A': the
        //      draw of the next tentative-collision distance inside the current segment.  Leaving a
        //      segment without a collision costs no B op -----------------------------------------------
        // (a lane without a path is ST_DEAD, so the state alone tells whether it walks)
        for (int k = 0; k < rp.walk_steps && __any_sync(0xffffffffu, (unsigned)(ln.state - ST_WALK) < 2u); ++k) {
            if (!EX && ln.state == ST_WALK && ln.mj < -(AVR_LEAP_MIN - 0.5f)) {
                // FAST: the walk stands inside an empty cube of the majorant grid (MediumView::maj_fast):
                // one leap to the face of the cube instead of a step per cell
                ln.dda.leap(mv, (int)(-ln.mj));
                cn.seg++;
#ifdef AVR_SEGSTAT
                cn.sc++;  // leaps
#endif
                if (ln.dda.tMin < ln.dda.tMax) ln.mj = maj[ln.dda.lin];
                else ln.state = ln.shadow ? ST_EV_SHADOW_END : ST_EV_PATH_END;
            } else if (ln.state == ST_WALK) {
                float a, b;
                int vox;
                if (!ln.dda.template next<!EX>(mv, &a, &b, &vox)) {
                    // iterator exhausted: SampleT_maj returns T_maj.  An unscattered camera ray
                    // leaves medium and boundary; no infinite lights in the scene class
                    // (cpu/integrators.cpp:1081-1106) -> path done
                    ln.state = ln.shadow ? ST_EV_SHADOW_END : ST_EV_PATH_END;
                } else if (!EX) {
                    // the majorant of this segment was requested when the DDA entered the voxel; request the
                    // next one now, a whole loop trip before it is needed
                    const float mj = ln.mj;
                    if (ln.dda.tMin < ln.dda.tMax) ln.mj = maj[ln.dda.lin];
                    // FAST: the distance to the next tentative collision is carried across segments as a
                    // remaining optical depth tau = -ln(1 - u) (ln.u holds it): by the memorylessness of
                    // the exponential this samples the same free-flight distribution as redrawing u in
                    // every segment (media.h:779-803) with one logarithm and one PCG32 draw per
                    // collision instead of one per segment.  With one channel T_maj / T_maj[0] == 1
                    // everywhere it is used, so T_maj is not tracked at all.
                    cn.seg++;
#ifdef AVR_SEGSTAT
                    if (!(mj > 0)) cn.rays++;  // empty cell next to a non-empty one (d = 1)
                    else if (lc.sigma_t[0] * mj * (b - a) < 0.25f) cn.cam++;  // thin segment
#endif
                    if (mj > 0) {
                        Spec<N> sm = lc.sigma_t * mj;
                        float dtau = sm[0] * (b - a);
                        bool hit = ln.u <= dtau && sm[0] != 0;
                        float tt = b;
                        if (hit) tt = __fmaf_rn(ln.u, ftz_rcp(sm[0]), a);
                        if (N > 1) ln.T_maj = ln.T_maj * exp_neg<false>(tt - a, sm);
                        if (hit) {
                            ln.sigma_maj = sm;
                            ln.cur_tMin = tt;
                            ln.seg_tMax = b;
                            ln.state = ST_COLLIDE;
                            if (Fetcher<TR>::kSplitFetch) ln.fe.request(mv, tt, ln.tapA, ln.tapB);
                            if (Fetcher<TR>::kPrefetch) ln.fe.prefetch(mv, tt);
                        } else
                            ln.u -= dtau;
                    }
                } else {
                    cn.seg++;
                    const float mj = ln.mj;
                    if (ln.dda.tMin < ln.dda.tMax) ln.mj = maj[ln.dda.lin];
                    Spec<N> sm = lc.sigma_t * mj;
                    if (mj == 0) {
                        // sigma_maj = 0 in every channel: T_maj *= FastExp(-0) = 1 exactly, nothing to do
                    } else if (sm[0] == 0) {
                        float dt = b - a;
                        if (isinf(dt)) dt = kFltMax;
                        ln.T_maj = ln.T_maj * exp_neg<EX>(dt, sm);
                    } else {
                        ln.sigma_maj = sm;
                        ln.cur_tMin = a;
                        ln.seg_tMax = b;
                        ln.state = ST_SAMPLE;
                    }
                }
            }
            if (EX && ln.state == ST_SAMPLE) {
                // media.h:779-803: t = tMin + SampleExponential(u, sigma_maj[0]); u = rng.Uniform();
                // T_maj *= FastExp(-(t - tMin) sigma_maj) at a collision, FastExp(-(tMax - tMin) sigma_maj) at the exit
                float t = ln.cur_tMin + sample_exponential<EX>(ln.u, ln.sigma_maj[0]);
                ln.u = ln.trng.uniform();
                bool hit = t < ln.seg_tMax;
                float dt = (hit ? t : ln.seg_tMax) - ln.cur_tMin;
                if (!hit && isinf(dt)) dt = kFltMax;
                ln.T_maj = ln.T_maj * exp_neg<EX>(dt, ln.sigma_maj);
                if (hit) ln.cur_tMin = t;
                ln.state = hit ? ST_COLLIDE : ST_WALK;
            }
        }

        // ---- B. the tentative collision at t = cur_tMin (medium lookup, event, weight update): runs when
        //      enough lanes wait for it, or when no lane is left walking -------------------------------
        {
            unsigned mc = __ballot_sync(0xffffffffu, ln.state == ST_COLLIDE);
            if (mc == 0) continue;
            if (__popc(mc) < rp.collide_min && __any_sync(0xffffffffu, ln.state == ST_WALK)) continue;
        }
        if (TR::LEAN && ln.state == ST_COLLIDE) {
            // one channel, fast arithmetic: the event probabilities are all that is left of
            // cpu/integrators.cpp:990-1074 (camera) and :1352-1380 (shadow) — see Traits::LEAN
            const float t = ln.cur_tMin;
            const float dens = ln.fe.density(mv, t);
            const float smj = ln.sigma_maj[0], inv = ftz_rcp(smj);
            const float sa = mv.ach_sig_a * dens, ss = mv.ach_sig_s * dens;
            if (!ln.shadow) {
                cn.cam++;
                const float pA = sa * inv, pS = ss * inv;
                const float sum = pA + pS + fmaxf(0.f, 1 - pA - pS);
                const float up = ln.trng.uniform() * sum;
                if (up < pA) {
                    ln.state = ST_EV_PATH_END;  // absorbed
                } else if (up < pA + pS) {
                    ln.state = ln.depth++ >= rp.max_depth ? ST_EV_PATH_END : ST_EV_SCATTER;  // cur_tMin = t already
                } else
                    ln.state = ST_SAMPLE;  // null collision: tMin = t already
            } else {
                cn.sh++;
                // ratio tracking: T_ray and r_u take sigma_n / sigma_maj, r_l stays 1; Russian roulette on
                // Tr = T_ray / Average(r_l + r_u)
                const float ratio = fmaxf(0.f, smj - sa - ss) * inv;
                float T = ln.T_ray[0] * ratio;
                const float ru = T * __int_as_float((127 - 2 * ln.rr_k) << 23);  // T / 4^rr_k
                if (T * ftz_rcp(1 + ru) < 0.05f) {
                    if (ln.trng.uniform() < 0.75f) T = 0.f;
                    else {
                        T *= 4.f;
                        ln.rr_k = min(ln.rr_k + 1, 31);
                    }
                }
                ln.T_ray.v[0] = T;
                ln.state = T != 0 ? ST_SAMPLE : ST_EV_SHADOW_END;
            }
        }
        if (!TR::LEAN && ln.state == ST_COLLIDE) {
            {
                const float t = ln.cur_tMin;
                Spec<N> T_maj = ln.T_maj;
                PropsN<N> mp;
                if (Fetcher<TR>::kSplitFetch) {
                    float dens = ln.fe.resolve(t, ln.tapA, ln.tapB);
                    mp.sigma_a = lc.sig_a * dens;
                    mp.sigma_s = lc.sig_s * dens;
                    mp.Le = sconst<4>(0.f);
                } else
                    mp = ln.fe.template props<N>(mv, lc, t);
                Spec<N> sigma_maj = ln.sigma_maj;
                Spec<N> sigma_n = clamp_zero(sigma_maj - mp.sigma_a - mp.sigma_s);
                // One instruction stream for both ray kinds.  Inside this loop a shadow lane keeps its
                // ratio-tracking weights in the camera fields (swap_shadow_weights): beta = T_ray,
                // r_u = its r_u, r_l = its r_l, so the weight update below is the same code.
                bool cont = false, alive = true;
                int mode = 2;  // 0 absorb, 1 scatter, 2 null collision (the only event of a shadow ray)
                if (!ln.shadow) {
                    // VolPathIntegrator::Li callback: event selection, cpu/integrators.cpp:990-1022
                    if (!nonzero(ln.beta)) {
                        alive = false;
                    } else {
                        cn.cam++;
                        if (TR::GENERAL && ln.depth < rp.max_depth && nonzero(mp.Le)) {
                            float pdf = sigma_maj[0] * T_maj[0];
                            Spec<N> betap = ln.beta * T_maj / pdf;
                            Spec<N> r_e = ln.r_u * sigma_maj * T_maj / pdf;
                            if (nonzero(r_e)) {
                                float avg = average(r_e);
#pragma unroll
                                for (int i = 0; i < 4; ++i)
                                    ln.L[i] += ((betap[i % N] * mp.sigma_a[i % N]) * mp.Le[i]) / avg;
                            }
                        }
                        float pAbsorb = fdiv<EX>(mp.sigma_a[0], sigma_maj[0]);
                        float pScatter = fdiv<EX>(mp.sigma_s[0], sigma_maj[0]);
                        float pNull = fmaxf(0.f, 1 - pAbsorb - pScatter);
                        float um = ln.trng.uniform();
                        mode = sample_discrete3(pAbsorb, pScatter, pNull, um);
                        if (mode == 0) alive = false;
                        else if (mode == 1 && ln.depth++ >= rp.max_depth) alive = false;
                    }
                } else {
                    cn.sh++;
                }
                if (!alive) {
                    ln.state = ST_EV_PATH_END;
                } else {
                    // weight update: camera scatter (:1031-1035) and null (:1066-1074), shadow ratio
                    // tracking (:1356-1367); pdf is the probability of the event that was taken
                    Spec<N> sx = mode == 1 ? mp.sigma_s : sigma_n;
                    float pdf = T_maj[0] * (ln.shadow ? sigma_maj[0] : sx[0]);
                    Spec<N> fx = sdiv<EX>(T_maj * sx, pdf);
                    ln.beta = ln.beta * fx;
                    if (!ln.shadow && mode == 2 && pdf == 0) ln.beta = sconst<N>(0.f);
                    ln.r_u = ln.r_u * fx;
                    if (mode == 2) ln.r_l = ln.r_l * sdiv<EX>(T_maj * sigma_maj, pdf);
                    if (ln.shadow) {
                        // Russian roulette on the ratio-tracking estimate, cpu/integrators.cpp:1369-1378
                        Spec<N> Tr = sdiv<EX>(ln.beta, average(ln.r_l + ln.r_u));
                        if (max_component(Tr) < 0.05f) {
                            float q = 0.75f;
                            if (ln.trng.uniform() < q) ln.beta = sconst<N>(0.f);
                            else ln.beta = ln.beta / (1 - q);
                        }
                        cont = nonzero(ln.beta);
                        if (!cont) {
                            ln.T_maj = sconst<N>(1.f);  // SampleT_maj returns 1 when the callback stops
                            ln.state = ST_EV_SHADOW_END;
                        }
                    } else if (mode == 1) {
                        if (nonzero(ln.beta) && nonzero(ln.r_u)) {
                            ln.cur_tMin = t;  // scatter position parameter for E1
                            ln.state = ST_EV_SCATTER;
                        } else
                            ln.state = ST_EV_PATH_END;
                    } else {
                        cont = nonzero(ln.beta) && nonzero(ln.r_u);
                        if (!cont) ln.state = ST_EV_PATH_END;
                    }
                }
                if (cont) {
                    ln.T_maj = sconst<N>(1.f);
                    ln.state = ST_SAMPLE;  // tMin = t already
                }
            }
        }
        if (!EX && ln.state == ST_SAMPLE) {
            // FAST, right after a null collision (same lanes, same trip): fresh tau, tested against the rest of the current
            // segment.  A lane that goes back to walking has all the walk rounds of the next trip
            float tau = neg_log1m(ln.trng.uniform());
            float dtau = ln.sigma_maj[0] * (ln.seg_tMax - ln.cur_tMin);
            bool hit = tau <= dtau;
            float tt = ln.seg_tMax;
            if (hit) tt = __fmaf_rn(tau, ftz_rcp(ln.sigma_maj[0]), ln.cur_tMin);
            if (N > 1) ln.T_maj = ln.T_maj * exp_neg<false>(tt - ln.cur_tMin, ln.sigma_maj);
            if (hit) {
                ln.cur_tMin = tt;
                if (Fetcher<TR>::kSplitFetch) ln.fe.request(mv, tt, ln.tapA, ln.tapB);
                if (Fetcher<TR>::kPrefetch) ln.fe.prefetch(mv, tt);
            } else
                ln.u = tau - dtau;
            ln.state = hit ? ST_COLLIDE : ST_WALK;
        }
    }
}

template <typename TR>
__global__ void __launch_bounds__(KernelCfg<TR>::kThreads, KernelCfg<TR>::kMinBlocks)
k_volpath(MediumView mv, RenderParams rp, unsigned long long *counters, double *film, float *samples_out) {
    constexpr int N = TR::NCH;
    using IO = SlotIO<TR>;
    extern __shared__ uint4 smem_q[];
    unsigned char *smem = reinterpret_cast<unsigned char *>(smem_q);
    const float *maj = TR::EXACT ? mv.maj : mv.maj_fast;
    const unsigned maj_bytes = !rp.maj_in_smem ? 0u
                               : (TR::EXACT ? unsigned(mv.mres[0] * mv.mres[1] * mv.mres[2]) : unsigned(mv.fmres[0] * mv.fmres[1] * mv.fmres[2])) * 4u;
    const PoolLayout lay = pool_layout(maj_bytes, (unsigned)rp.pool, IO::kSlotQuads);
    PoolView pv;
    pv.slots = reinterpret_cast<uint4 *>(smem + lay.off_slots);
    pv.cells = reinterpret_cast<volatile unsigned *>(smem + lay.off_cells);
    pv.q = reinterpret_cast<PoolQueue *>(smem + lay.off_q);
    pv.live = reinterpret_cast<volatile int *>(smem + lay.off_live);
    pv.cap_log2 = lay.cap_log2;
    pv.cap_mask = lay.cap - 1;
    {
        // The majorant grid, when it is small enough to live in shared memory (<= 16 KB: the 16^3 grid of GridMedium /
        // RGBGridMedium), is staged by ONE bulk asynchronous copy (TMA, cp.async.bulk) that completes on an mbarrier
        // while all threads initialise the pool; a size that is not a multiple of 16 bytes takes the plain loop.
        __shared__ __align__(8) unsigned long long s_maj_bar;
        const bool bulk = rp.maj_in_smem && (maj_bytes & 15u) == 0 && ((size_t)maj & 15u) == 0;
        if (rp.maj_in_smem) {
            float *s_maj = reinterpret_cast<float *>(smem);
            if (bulk) {
                const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_maj_bar);
                if (threadIdx.x == 0) {
                    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(1));
                    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(maj_bytes) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                     (unsigned)__cvta_generic_to_shared(s_maj)),
                                 "l"(maj), "r"(maj_bytes), "r"(bar)
                                 : "memory");
                }
            } else
                for (unsigned i = threadIdx.x; i < maj_bytes / 4; i += blockDim.x) s_maj[i] = maj[i];
            maj = s_maj;
        }
        // all slots start dead (state word 0 = ST_DEAD, nothing pending) and queued in Q_DEAD
        for (unsigned i = threadIdx.x; i < (unsigned)rp.pool * IO::kSlotQuads; i += blockDim.x)
            pv.slots[i] = make_uint4(0, 0, 0, 0);
        for (unsigned i = threadIdx.x; i < kNumQueues * lay.cap; i += blockDim.x) {
            unsigned qi = i >> lay.cap_log2, c = i & (lay.cap - 1);
            bool filled = qi == Q_DEAD && c < (unsigned)rp.pool;
            pv.cells[i] = filled ? c : 0xffffu;  // round 0: entry c, or free
        }
        if (threadIdx.x < kNumQueues) {
            pv.q[threadIdx.x].head = 0;
            pv.q[threadIdx.x].tail = threadIdx.x == Q_DEAD ? (unsigned)rp.pool : 0u;
        }
        if (threadIdx.x == 0) *pv.live = rp.pool;
        __syncthreads();  // (also orders thread 0's mbarrier.init before everybody's wait)
        if (bulk) {
            const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_maj_bar);
            asm volatile(
                "{\n"
                ".reg .pred p;\n"
                "AVR_MAJ_WAIT:\n"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                "@!p bra AVR_MAJ_WAIT;\n"
                "}\n" ::"r"(bar),
                "r"(0)
                : "memory");
        }
    }
    const unsigned lane = threadIdx.x & 31;
    PathLane<TR> ln;
    LambdaCtx<N> lc;
    LaneCounters cn = {0, 0, 0, 0, 0};

    // warps of one sub-partition share its L0 instruction cache: with a mask other than 15 the event code
    // stays on the chosen sub-partitions as long as both kinds of work are queued
    const bool ev_first = (rp.event_smsp_mask >> ((threadIdx.x >> 5) & 3)) & 1;
    while (true) {
        // ---- warp-level dispatch on the queue fill levels ---------------------------------------------
        int c_dead = 0, c_sc = 0, c_se = 0, c_tr = 0, live = 0;
        if (lane == 0) {
            c_dead = pool_fill(pv, Q_DEAD);
            c_sc = pool_fill(pv, Q_SCATTER);
            c_se = pool_fill(pv, Q_SHADOW_END);
            c_tr = pool_fill(pv, Q_TRACK);
            live = *pv.live;
        }
        int choice;  // queue to serve, -1: sleep, -2: done
        if (lane == 0) {
            if (live <= 0) choice = -2;
            else if (!ev_first && c_tr >= rp.track_min) choice = Q_TRACK;
            else if (c_sc >= 32) choice = Q_SCATTER;
            else if (c_se >= 32) choice = Q_SHADOW_END;
            else if (c_dead >= 32) choice = Q_DEAD;
            else if (c_tr >= rp.track_min) choice = Q_TRACK;
            else if (c_sc > 0 || c_se > 0 || c_dead > 0)
                choice = (c_sc >= c_se && c_sc >= c_dead) ? Q_SCATTER : (c_se >= c_dead ? Q_SHADOW_END : Q_DEAD);
            else if (c_tr > 0) choice = Q_TRACK;
            else choice = -1;
        }
        choice = __shfl_sync(0xffffffffu, choice, 0);
        if (choice == -2) break;
        if (choice == -1) {
            __nanosleep(200);
            continue;
        }
        if (choice == Q_TRACK) track_mode<TR>(mv, rp, pv, maj, lane, ln, lc, cn);
        else event_batch<TR>(mv, rp, pv, choice, lane, ln, lc, cn, counters, film, samples_out);
    }

    // ---- stats: warp reduce, one atomic per counter per warp ------------------------------------------
    unsigned vals[5] = {cn.rays, cn.cam, cn.sh, cn.seg, cn.sc};
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        unsigned long long v = vals[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0 && v) atomicAdd(&counters[CNT_RAYS + k], v);
    }
}

}  // namespace avr

#include "avr_volpath_host.inc"
