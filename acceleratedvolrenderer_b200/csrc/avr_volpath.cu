// avr_volpath.cu — kernel 1b (see avr_volpath.cuh) and its C entry points.
//
// Per-lane state machine under a warp-level scheduler: every lane holds one path and waits for
// exactly one of six ops; each trip the warp counts the waiting lanes per op (one REDUX) and runs the
// op with the best lanes x weight score for the lanes that want it.  Every op has ONE call site, so
// lanes execute it convergently and the SASS stays small:
//   R  regenerate : flush the finished sample to the film, pull new (pixel, sample) work with one
//                   warp-aggregated atomic, generate the camera ray
//   S  setup      : boundary hit, RNG seeding, ray -> medium space, slab test, DDA init
//                   (camera segment or NEE shadow ray: same code, a mode bit apart)
//   A  walk       : DDA steps to the next majorant segment with sigma_maj != 0
//   B  collide    : one tentative collision = one log, one exp, one 32-byte density fetch, then the
//                   delta-tracking (camera) or ratio-tracking (shadow) update
//   E1 scatter    : NEE preparation -> shadow ray pending
//   E2 shadow end : NEE accumulate + phase sampling -> next camera segment pending
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "avr_render_host.h"

using namespace avr;

namespace avr {

constexpr int kBlock = 128;

// Lane states.  The first six double as the op a lane waits for (see the scheduler in k_volpath).
enum {
    ST_DEAD = 0,           // R: needs a new path (a finished sample may be pending)
    ST_SETUP = 1,          // S: pending ray in fe.o / fe.d
    ST_WALK = 2,           // A: between majorant segments, DDA step next
    ST_COLLIDE = 3,        // B: inside a non-empty majorant segment, tentative collision next
    ST_EV_SCATTER = 4,     // E1: real scattering event
    ST_EV_SHADOW_END = 5,  // E2: shadow ray finished
    kNumOps = 6,
    ST_EV_PATH_END = 6,    // transient: becomes ST_DEAD + pending at the end of the trip
    ST_RETIRED = 7         // no work left
};

// film / sample write of a finished path: cpu/integrators.cpp:272-297, film.h:95-100,239-255
template <typename TR>
AVR_DEV void write_sample(const RenderParams &rp, const PathLane<TR> &ln, double *film, float *samples_out) {
    float L[4] = {ln.L[0], ln.L[1], ln.L[2], ln.L[3]};
    bool bad = false;
#pragma unroll
    for (int i = 0; i < 4; ++i) bad = bad || isnan(L[i]) || isinf(L[i]);
    if (bad) L[0] = L[1] = L[2] = L[3] = 0.f;  // NaN / infinite radiance -> black
    float lam[4], pdf[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        lam[i] = ln.lambda[i];
        float l = lam[i];
        pdf[i] = (l < 360 || l > 830) ? 0.f : 0.0039398042f / sqr(coshf(0.0072f * (l - 538)));
    }
    if (rp.dump_samples) {
        long long nsamp = rp.s1 - rp.s0;
        long long idx = ((long long)(ln.py - rp.y0) * (rp.x1 - rp.x0) + (ln.px - rp.x0)) * nsamp + (ln.samp - rp.s0);
        float *o = samples_out + idx * 12;
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
    if (rgb[0] != 0) atomicAdd(px + 0, (double)(1.f * rgb[0]));  // adding +-0 leaves the sum unchanged
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
    // IndependentSampler::StartPixelSample, samplers.h:457-460
    ln.srng.set_sequence(hash_3u32((uint32_t)px, (uint32_t)py, (uint32_t)rp.seed));
    ln.srng.advance((long long)s * 65536ll);
    float lu = ln.srng.uniform();
    float pdf[4];
    sample_visible_wavelengths(lu, ln.lambda, pdf);
    lc = make_lambda_ctx<TR::NCH>(mv, ln.lambda);
    float f0 = ln.srng.uniform(), f1 = ln.srng.uniform();
    float fx = lerp_exact(f0, -rp.filter_rx, rp.filter_rx), fy = lerp_exact(f1, -rp.filter_ry, rp.filter_ry);
    float pfx = (float(px) + fx) + 0.5f, pfy = (float(py) + fy) + 0.5f;
    ln.srng.uniform();  // time
    ln.srng.uniform();  // lens
    ln.srng.uniform();
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

template <typename TR>
__global__ void __launch_bounds__(kBlock)
k_volpath(MediumView mv, RenderParams rp, unsigned long long *counters, double *film, float *samples_out) {
    constexpr int N = TR::NCH;
    constexpr bool EX = TR::EXACT;
    extern __shared__ float s_maj[];
    const float *maj = mv.maj;
    {
        int nm = mv.mres[0] * mv.mres[1] * mv.mres[2];
        if (nm * 4 <= 48 * 1024) {
            for (int i = threadIdx.x; i < nm; i += blockDim.x) s_maj[i] = mv.maj[i];
            __syncthreads();
            maj = s_maj;
        }
    }
    const unsigned lane = threadIdx.x & 31;
    PathLane<TR> ln;
    ln.state = ST_DEAD;
    ln.n_cam = ln.n_sh = ln.n_seg = ln.n_sc = 0;
    ln.shadow = false;
    unsigned n_rays = 0;
    LambdaCtx<N> lc;
    bool pending = false;  // pending: a finished sample waits for its film write

    while (true) {
        // ---- scheduler: count the lanes waiting for each op with one warp reduction, run the op
        //      with the best score = lanes * weight (weights ~ 1 / op cost: a cheap op may run for few
        //      lanes, an expensive one waits until enough lanes want it) ------------------------
        int best;
        {
            // 5-bit fields hold counts up to 31; a full warp in one op overflows into the next field,
            // which the all-equal test below catches first
            int s0 = __shfl_sync(0xffffffffu, ln.state, 0);
            if (__all_sync(0xffffffffu, ln.state == s0)) {
                if (ln.state == ST_RETIRED) break;
                best = ln.state;
            } else {
                unsigned packed = ln.state < kNumOps ? (1u << (5 * ln.state)) : 0u;
                packed = __reduce_add_sync(0xffffffffu, packed);
                int bs = -1;
                best = 0;
#pragma unroll
                for (int op = 0; op < kNumOps; ++op) {
                    int sc = int((packed >> (5 * op)) & 31u) * rp.op_weight[op];
                    if (sc > bs) { bs = sc; best = op; }
                }
            }
        }

        if (best == ST_DEAD) {
            // ---- R. regenerate: flush finished samples, fetch work with one aggregated atomic ------
            bool dead = ln.state == ST_DEAD;
            unsigned md = __ballot_sync(0xffffffffu, dead);
            if (dead && pending) {
                write_sample(rp, ln, film, samples_out);
                pending = false;
            }
            unsigned long long base = 0;
            int leader = __ffs(md) - 1;
            if ((int)lane == leader) base = atomicAdd(&counters[CNT_WORK], (unsigned long long)__popc(md));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (dead) {
                long long idx = (long long)base + __popc(md & ((1u << lane) - 1));
                if (idx < rp.local_items) {
                    long long chunk = idx >> 5;
                    long long item = ((chunk * rp.world + rp.rank) << 5) + (idx & 31);
                    if (start_path(mv, rp, ln, item, lc)) n_rays++;
                } else
                    ln.state = ST_RETIRED;
            }
        } else if (best == ST_SETUP) {
            // ---- S. setup of the next tracking ray (camera segment or shadow ray) ---------------
            if (ln.state == ST_SETUP) {
                V3 o = ln.fe.o, d = ln.fe.d;
                float t0, t1;
                bool hit = boundary_hit(rp, o, d, &t0, &t1) && t1 > 0;
                float tMax;
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
                        uint64_t hash0 = hash_float(ln.srng.uniform());
                        uint64_t hash1 = hash_float(ln.srng.uniform());
                        ln.trng.set_sequence(hash0, hash1);
                        ln.u = ln.srng.uniform();
                    }
                }
                if (!hit) {
                    ln.state = ST_EV_PATH_END;  // the ray leaves without touching the boundary
                } else {
                    // SampleT_maj prologue (media.h:744-749) + SampleRay (media.h:322-337)
                    float len = sqrtf(sqr(d.x) + sqr(d.y) + sqr(d.z));
                    tMax *= len;
                    d = d / len;
                    MRay mr = ray_to_medium(mv, o, d, tMax);
                    if (mr.hit) ln.dda.init(mv, mr.o, mr.d, mr.tMin, mr.tMax); else ln.dda.invalidate();
                    ln.fe.bind(mv, o, d);
                    ln.T_maj = sconst<N>(1.f);
                    ln.state = ST_WALK;
                }
            }
        } else if (best == ST_WALK) {
            // ---- A. DDA steps towards the next majorant segment with sigma_maj[0] != 0 (at most
            //      walk_steps per trip, so colliding lanes are not held up by long empty stretches) ---
            for (int k = 0; k < rp.walk_steps; ++k) {
                if (ln.state == ST_WALK) {
                    float a, b;
                    int vox;
                    if (!ln.dda.next(mv, &a, &b, &vox)) {
                        // iterator exhausted: SampleT_maj returns T_maj.  An unscattered camera ray
                        // leaves medium and boundary; no infinite lights in the scene class
                        // (cpu/integrators.cpp:1081-1106) -> path done
                        ln.state = ln.shadow ? ST_EV_SHADOW_END : ST_EV_PATH_END;
                    } else {
                        ln.n_seg++;
                        Spec<N> sm = lc.sigma_t * maj[vox];
                        if (sm[0] == 0) {
                            float dt = b - a;
                            if (isinf(dt)) dt = kFltMax;
                            ln.T_maj = ln.T_maj * exp_neg<EX>(dt, sm);
                        } else {
                            ln.sigma_maj = sm;
                            ln.cur_tMin = a;
                            ln.seg_tMax = b;
                            ln.state = ST_COLLIDE;
                        }
                    }
                }
                if (!__any_sync(0xffffffffu, ln.state == ST_WALK)) break;
            }
        } else if (best == ST_COLLIDE) {
            // ---- B. tentative collisions along the current segment; repeats while enough lanes
            //      stay in the segment loop ---------------------------------------------------------
            do {
                if (ln.state == ST_COLLIDE) {
                    float t = ln.cur_tMin + sample_exponential<EX>(ln.u, ln.sigma_maj[0]);
                    ln.u = ln.trng.uniform();
                    if (t < ln.seg_tMax) {
                        Spec<N> T_maj = ln.T_maj * exp_neg<EX>(t - ln.cur_tMin, ln.sigma_maj);
                        PropsN<N> mp = ln.fe.template props<N>(mv, lc, t);
                        Spec<N> sigma_maj = ln.sigma_maj;
                        bool cont;
                        if (ln.shadow) {
                            // ratio tracking + Russian roulette, cpu/integrators.cpp:1356-1378
                            ln.n_sh++;
                            Spec<N> sigma_n = clamp_zero(sigma_maj - mp.sigma_a - mp.sigma_s);
                            float pdf = T_maj[0] * sigma_maj[0];
                            ln.T_ray = ln.T_ray * (T_maj * sigma_n / pdf);
                            ln.s_rl = ln.s_rl * (T_maj * sigma_maj / pdf);
                            ln.s_ru = ln.s_ru * (T_maj * sigma_n / pdf);
                            Spec<N> Tr = ln.T_ray / average(ln.s_rl + ln.s_ru);
                            if (max_component(Tr) < 0.05f) {
                                float q = 0.75f;
                                if (ln.trng.uniform() < q) ln.T_ray = sconst<N>(0.f);
                                else ln.T_ray = ln.T_ray / (1 - q);
                            }
                            cont = nonzero(ln.T_ray);
                            if (!cont) {
                                ln.T_maj = sconst<N>(1.f);  // SampleT_maj returns 1 when the callback stops
                                ln.state = ST_EV_SHADOW_END;
                            }
                        } else {
                            // VolPathIntegrator::Li callback, cpu/integrators.cpp:990-1078
                            cont = false;
                            if (!nonzero(ln.beta)) {
                                ln.state = ST_EV_PATH_END;
                            } else {
                                ln.n_cam++;
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
                                float pAbsorb = mp.sigma_a[0] / sigma_maj[0];
                                float pScatter = mp.sigma_s[0] / sigma_maj[0];
                                float pNull = fmaxf(0.f, 1 - pAbsorb - pScatter);
                                float um = ln.trng.uniform();
                                int mode = sample_discrete3(pAbsorb, pScatter, pNull, um);
                                if (mode == 0) {
                                    ln.state = ST_EV_PATH_END;
                                } else if (mode == 1) {
                                    if (ln.depth++ >= rp.max_depth) {
                                        ln.state = ST_EV_PATH_END;
                                    } else {
                                        float pdf = T_maj[0] * mp.sigma_s[0];
                                        ln.beta = ln.beta * (T_maj * mp.sigma_s / pdf);
                                        ln.r_u = ln.r_u * (T_maj * mp.sigma_s / pdf);
                                        if (nonzero(ln.beta) && nonzero(ln.r_u)) {
                                            ln.cur_tMin = t;  // scatter position parameter for E1
                                            ln.state = ST_EV_SCATTER;
                                        } else
                                            ln.state = ST_EV_PATH_END;
                                    }
                                } else {
                                    Spec<N> sigma_n = clamp_zero(sigma_maj - mp.sigma_a - mp.sigma_s);
                                    float pdf = T_maj[0] * sigma_n[0];
                                    ln.beta = ln.beta * (T_maj * sigma_n / pdf);
                                    if (pdf == 0) ln.beta = sconst<N>(0.f);
                                    ln.r_u = ln.r_u * (T_maj * sigma_n / pdf);
                                    ln.r_l = ln.r_l * (T_maj * sigma_maj / pdf);
                                    cont = nonzero(ln.beta) && nonzero(ln.r_u);
                                    if (!cont) ln.state = ST_EV_PATH_END;
                                }
                            }
                        }
                        if (cont) {
                            ln.T_maj = sconst<N>(1.f);
                            ln.cur_tMin = t;
                        }
                    } else {
                        float dt = ln.seg_tMax - ln.cur_tMin;
                        if (isinf(dt)) dt = kFltMax;
                        ln.T_maj = ln.T_maj * exp_neg<EX>(dt, ln.sigma_maj);
                        ln.state = ST_WALK;
                    }
                }
            } while (__popc(__ballot_sync(0xffffffffu, ln.state == ST_COLLIDE)) >= rp.collide_repeat);
        } else if (best == ST_EV_SCATTER) {
            // ---- E1. real scattering: SampleLd prologue (cpu/integrators.cpp:1300-1338) -> shadow ray
            if (ln.state == ST_EV_SCATTER) {
                ln.n_sc++;
                V3 p = ln.fe.point(ln.cur_tMin);
                ln.srng.uniform();  // light choice (single light, p = 1)
                ln.srng.uniform();  // uLight
                ln.srng.uniform();
                V3 wi = v3(rp.light_dir[0], rp.light_dir[1], rp.light_dir[2]);
                float two_r = 2 * rp.scene_radius;
                V3 pLight = v3(p.x + wi.x * two_r, p.y + wi.y * two_r, p.z + wi.z * two_r);
                bool lit = false;
#pragma unroll
                for (int i = 0; i < 4; ++i) lit = lit || sample_dense1(rp.light_spec, ln.lambda[i]) != 0;
                float ph = henyey_greenstein(dot(ln.wo, wi), mv.g);
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
        } else {
            // ---- E2. end of the shadow ray: SampleLd epilogue (cpu/integrators.cpp:1382-1398), then
            //      phase sampling and the next camera segment (cpu/integrators.cpp:1041-1060) --------
            if (ln.state == ST_EV_SHADOW_END) {
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
                float ua = ln.srng.uniform(), ub = ln.srng.uniform();
                float pdf;
                V3 wi = sample_hg(ln.wo, mv.g, ua, ub, &pdf);
                if (pdf == 0) {
                    ln.state = ST_EV_PATH_END;
                } else {
                    ln.beta = ln.beta * (pdf / pdf);
                    ln.r_l = ln.r_u / pdf;
                    ln.fe.d = wi;  // fe.o still holds the scatter point
                    ln.shadow = false;
                    ln.state = ST_SETUP;
                }
            }
        }
        if (ln.state == ST_EV_PATH_END) {
            // absorbed / escaped / max depth / zero throughput: the film write waits for R
            ln.state = ST_DEAD;
            pending = true;
        }
    }

    // ---- stats: warp reduce, one atomic per counter per warp ------------------------------------------
    unsigned vals[5] = {n_rays, ln.n_cam, ln.n_sh, ln.n_seg, ln.n_sc};
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        unsigned long long v = vals[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0 && v) atomicAdd(&counters[CNT_RAYS + k], v);
    }
}

}  // namespace avr

#include "avr_volpath_host.inc"
