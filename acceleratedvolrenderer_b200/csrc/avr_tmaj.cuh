// avr_tmaj.cuh — SampleT_maj<ConcreteMedium,F> (reference src/pbrt/media.h:741-806) for the device,
// in the straight nested-loop form (one thread = one ray).  The wavefront render kernel
// (avr_volpath.cu) runs the same arithmetic as a per-lane state machine.
#pragma once
#include "avr_medium.cuh"

#ifndef AVR_PREFETCH
#define AVR_PREFETCH 0  // build knob: 1 / 2 = prefetch.global.L1 / .L2 of the density octet when the collision distance is drawn
#endif
#ifndef AVR_SPLIT_FETCH
#define AVR_SPLIT_FETCH 0  // build knob: 1 = request the density taps when the collision distance is drawn (measured slower: 59.8 -> 70.2 ms)
#endif

namespace avr {

// Compile-time shape of a tracking kernel.
//  EXACT   : restated CPU arithmetic (parity mode) vs. fast intrinsics + folded transforms
//  NCH     : 4 = SampledSpectrum, 1 = achromatic medium (sigma_a/sigma_s constant over lambda, so
//            all four channels carry bit-identical values and one is enough)
//  LAYOUT  : density layout (ignored when GENERAL)
//  GENERAL : any medium kind (rgbgrid, emissive uniformgrid) through sample_point_general; NCH = 4
template <bool EXACT_, int NCH_, int LAYOUT_, bool GENERAL_, bool ZSOBOL_ = false>
struct Traits {
    static constexpr bool ZSOBOL = ZSOBOL_;  // render kernels: the integrator's sampler is ZSobolSampler
    static constexpr bool EXACT = EXACT_;
    static constexpr int NCH = NCH_;
    static constexpr int LAYOUT = LAYOUT_;
    static constexpr bool GENERAL = GENERAL_;
    // FAST kernels of achromatic scalar media: with one channel every spectral-MIS weight of the path is an
    // exact ratio x / x = 1 (beta, r_u, r_l of the camera path, r_l of a shadow ray), so they are not carried
    // at all: analog delta tracking for the camera segments, plain ratio tracking for the shadow rays
    static constexpr bool LEAN = !EXACT_ && NCH_ == 1 && !GENERAL_;
    // One event op per real scattering event: the NEE shadow ray AND the next camera segment are set up together
    // (nothing the phase sampling draws or computes depends on the shadow ray's result: cpu/integrators.cpp:1041-1060
    // follows :1037 SampleLd unconditionally), the tracking loop walks the shadow ray, adds its contribution and
    // switches to the waiting camera ray by itself.  Halves the trips of a path through the pool's queues.
    static constexpr bool FUSED = LEAN;
};

template <int NCH>
struct PropsN {
    Spec<NCH> sigma_a, sigma_s;
    Spec<4> Le;  // only meaningful for GENERAL media
};

// Per-ray-batch constants: the medium's spectra sampled at the batch wavelengths.
template <int NCH>
struct LambdaCtx {
    float lambda[4];
    Spec<NCH> sig_a, sig_s, sigma_t;
};
template <int NCH>
AVR_DEV LambdaCtx<NCH> make_lambda_ctx(const MediumView &mv, const float lambda[4]) {
    LambdaCtx<NCH> c;
#pragma unroll
    for (int i = 0; i < 4; ++i) c.lambda[i] = lambda[i];
    if (mv.kind == KIND_RGB) {
        c.sig_a = sconst<NCH>(0.f);
        c.sig_s = sconst<NCH>(0.f);
        c.sigma_t = sconst<NCH>(1.f);  // RGBGridMedium::SampleRay: sigma_t(1), media.h:413
    } else {
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
            c.sig_a.v[i] = sample_dense1(mv.sigma_a_spec, lambda[i]);
            c.sig_s.v[i] = sample_dense1(mv.sigma_s_spec, lambda[i]);
        }
        c.sigma_t = c.sig_a + c.sig_s;
    }
    return c;
}

// Density-fetch helper bound to one ray: EXACT follows ray(t) -> ApplyInverse -> Offset -> p*n-.5;
// FAST folds the whole chain into s = so + sd*t.
template <typename TR>
struct Fetcher {
    V3 o, d;    // render-space ray (normalised d)
    V3 so, sd;  // FAST: sample-space origin / direction
    AVR_DEV void bind(const MediumView &mv, V3 o_, V3 d_) {
        o = o_;
        d = d_;
        if (!TR::EXACT) {
            const float *a = mv.gfr;
            so = v3(__fmaf_rn(a[0], o.x, __fmaf_rn(a[1], o.y, __fmaf_rn(a[2], o.z, a[3]))),
                    __fmaf_rn(a[4], o.x, __fmaf_rn(a[5], o.y, __fmaf_rn(a[6], o.z, a[7]))),
                    __fmaf_rn(a[8], o.x, __fmaf_rn(a[9], o.y, __fmaf_rn(a[10], o.z, a[11]))));
            sd = v3(__fmaf_rn(a[0], d.x, __fmaf_rn(a[1], d.y, a[2] * d.z)),
                    __fmaf_rn(a[4], d.x, __fmaf_rn(a[5], d.y, a[6] * d.z)),
                    __fmaf_rn(a[8], d.x, __fmaf_rn(a[9], d.y, a[10] * d.z)));
        }
    }
    AVR_DEV V3 point(float t) const { return v3(o.x + d.x * t, o.y + d.y * t, o.z + d.z * t); }
    // scalar density at ray(t) for a non-GENERAL medium
    AVR_DEV float density(const MediumView &mv, float t) const {
        if (TR::EXACT) {
            if (mv.kind == KIND_NVDB)
                return nv_density<TR::LAYOUT>(mv, nv_world_to_index(mv.nv_inv, mv.nv_vec, medium_point_exact(mv, point(t))));
            V3 p = grid_point_exact(mv, point(t));
            return lookup_density_s<true, TR::LAYOUT>(mv, p.x * mv.nx - .5f, p.y * mv.ny - .5f, p.z * mv.nz - .5f);
        }
        return lookup_density_s<false, TR::LAYOUT>(mv, __fmaf_rn(sd.x, t, so.x), __fmaf_rn(sd.y, t, so.y),
                                                   __fmaf_rn(sd.z, t, so.z));
    }
    // FAST octet layouts: the fetch split in two, so that the render kernel can request the 8 taps of a
    // tentative collision as soon as its distance is known and consume them one loop trip later.
    static constexpr bool kOctets = TR::LAYOUT == LAYOUT_OCTET_F32 || TR::LAYOUT == LAYOUT_OCTET_F16;
    static constexpr bool kSplitFetch = AVR_SPLIT_FETCH && !TR::EXACT && !TR::GENERAL && kOctets;
    AVR_DEV void request(const MediumView &mv, float t, float4 &a, float4 &b) const {
        float sx = __fmaf_rn(sd.x, t, so.x), sy = __fmaf_rn(sd.y, t, so.y), sz = __fmaf_rn(sd.z, t, so.z);
        int cx = (int)floorf(sx) + 1, cy = (int)floorf(sy) + 1, cz = (int)floorf(sz) + 1;
        a = b = make_float4(0.f, 0.f, 0.f, 0.f);
        if ((unsigned)cx > (unsigned)mv.nx || (unsigned)cy > (unsigned)mv.ny || (unsigned)cz > (unsigned)mv.nz) return;
        if (TR::LAYOUT == LAYOUT_OCTET_F32) {
            ldg256(mv.oct32 + 2 * size_t(octet_index(cx, cy, cz, mv.nbx, mv.nby)), a, b);
        } else {
            uint4 q = __ldg(mv.oct16 + octet_index(cx, cy, cz, mv.nbx, mv.nby));
            a = make_float4(__uint_as_float(q.x), __uint_as_float(q.y), __uint_as_float(q.z), __uint_as_float(q.w));
        }
    }
    // FAST octet layouts: fire-and-forget prefetch of the octet a tentative collision at t will read; the
    // render kernel issues it when the collision distance becomes known, up to walk_steps DDA steps
    // before op B consumes the octet.  No register is held across the gap.
    static constexpr int kPrefetch = (!TR::EXACT && !TR::GENERAL && kOctets) ? AVR_PREFETCH : 0;
    AVR_DEV void prefetch(const MediumView &mv, float t) const {
        float sx = __fmaf_rn(sd.x, t, so.x), sy = __fmaf_rn(sd.y, t, so.y), sz = __fmaf_rn(sd.z, t, so.z);
        int cx = (int)floorf(sx) + 1, cy = (int)floorf(sy) + 1, cz = (int)floorf(sz) + 1;
        if ((unsigned)cx > (unsigned)mv.nx || (unsigned)cy > (unsigned)mv.ny || (unsigned)cz > (unsigned)mv.nz) return;
        const void *p = TR::LAYOUT == LAYOUT_OCTET_F32
                            ? (const void *)(mv.oct32 + 2 * size_t(octet_index(cx, cy, cz, mv.nbx, mv.nby)))
                            : (const void *)(mv.oct16 + octet_index(cx, cy, cz, mv.nbx, mv.nby));
        if (kPrefetch == 1) asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
        else asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    }
    AVR_DEV float resolve(float t, const float4 &a, const float4 &b) const {
        float sx = __fmaf_rn(sd.x, t, so.x), sy = __fmaf_rn(sd.y, t, so.y), sz = __fmaf_rn(sd.z, t, so.z);
        float wx = sx - floorf(sx), wy = sy - floorf(sy), wz = sz - floorf(sz);
        if (TR::LAYOUT == LAYOUT_OCTET_F32) return trilerp8<false>(wx, wy, wz, a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w);
        float2 p = h2f(__float_as_uint(a.x)), q = h2f(__float_as_uint(a.y)), r = h2f(__float_as_uint(a.z)),
               u = h2f(__float_as_uint(a.w));
        return trilerp8<false>(wx, wy, wz, p.x, p.y, q.x, q.y, r.x, r.y, u.x, u.y);
    }
    template <int NCH>
    AVR_DEV PropsN<NCH> props(const MediumView &mv, const LambdaCtx<NCH> &lc, float t) const {
        PropsN<NCH> mp;
        if (TR::GENERAL) {
            // NCH == 4 by construction
            Spec<4> a4, s4;
#pragma unroll
            for (int i = 0; i < 4; ++i) { a4.v[i] = lc.sig_a.v[i % NCH]; s4.v[i] = lc.sig_s.v[i % NCH]; }
            // FAST general kernels of grid / RGB media: sample coordinates from the folded affine map, fast lookups
            // (the illuminant-weighted Le of an RGB grid multiplies per tap in the reference: linear, so once here)
            Props4 g = (!TR::EXACT && mv.kind != KIND_NVDB)
                           ? sample_point_general_fast(mv, __fmaf_rn(sd.x, t, so.x), __fmaf_rn(sd.y, t, so.y),
                                                       __fmaf_rn(sd.z, t, so.z), lc.lambda, a4, s4)
                           : sample_point_general(mv, point(t), lc.lambda, a4, s4);
#pragma unroll
            for (int i = 0; i < NCH; ++i) { mp.sigma_a.v[i] = g.sigma_a.v[i]; mp.sigma_s.v[i] = g.sigma_s.v[i]; }
            mp.Le = g.Le;
        } else {
            float dens = density(mv, t);
            mp.sigma_a = lc.sig_a * dens;
            mp.sigma_s = lc.sig_s * dens;
            mp.Le = sconst<4>(0.f);
        }
        return mp;
    }
};

template <bool EXACT, int N>
AVR_DEV Spec<N> exp_neg(float dt, Spec<N> sigma) {  // FastExp(-dt * sigma_maj)
    Spec<N> r;
#pragma unroll
    for (int i = 0; i < N; ++i) r.v[i] = fast_exp<EXACT>(-dt * sigma.v[i]);
    return r;
}

struct TmajCount {
    int n_fetch, n_seg;
};

// callback(t, mp, sigma_maj, T_maj) -> bool (continue)
template <typename TR, typename F>
AVR_DEV Spec<TR::NCH> sample_tmaj(const MediumView &mv, const float *maj, const LambdaCtx<TR::NCH> &lc, V3 o, V3 d,
                                  float tMax, float u, RNG &rng, F callback, TmajCount *cnt) {
    constexpr int N = TR::NCH;
    float len = sqrtf(sqr(d.x) + sqr(d.y) + sqr(d.z));
    tMax *= len;
    d = d / len;
    MRay mr = ray_to_medium(mv, o, d, tMax);
    DDA dda;
    if (mr.hit) dda.init<false>(mv, mr.o, mr.d, mr.tMin, mr.tMax); else dda.invalidate();
    Fetcher<TR> fe;
    fe.bind(mv, o, d);
    Spec<N> T_maj = sconst<N>(1.f);
    bool done = false;
    while (!done) {
        float sMin, sMax;
        int vox;
        if (!dda.next(mv, &sMin, &sMax, &vox)) return T_maj;
        cnt->n_seg++;
        Spec<N> sigma_maj = lc.sigma_t * maj[vox];
        if (sigma_maj[0] == 0) {
            float dt = sMax - sMin;
            if (isinf(dt)) dt = kFltMax;
            T_maj = T_maj * exp_neg<TR::EXACT>(dt, sigma_maj);
            continue;
        }
        float tMin = sMin;
        while (true) {
            float t = tMin + sample_exponential<TR::EXACT>(u, sigma_maj[0]);
            u = rng.uniform();
            if (t < sMax) {
                T_maj = T_maj * exp_neg<TR::EXACT>(t - tMin, sigma_maj);
                PropsN<N> mp = fe.template props<N>(mv, lc, t);
                cnt->n_fetch++;
                if (!callback(t, mp, sigma_maj, T_maj)) {
                    done = true;
                    break;
                }
                T_maj = sconst<N>(1.f);
                tMin = t;
            } else {
                float dt = sMax - tMin;
                if (isinf(dt)) dt = kFltMax;
                T_maj = T_maj * exp_neg<TR::EXACT>(dt, sigma_maj);
                break;
            }
        }
    }
    return sconst<N>(1.f);
}

}  // namespace avr
