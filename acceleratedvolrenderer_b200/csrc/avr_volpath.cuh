// avr_volpath.cuh — kernel 1b: persistent volumetric path tracer over a CTA-resident path pool.
//
// Replaces (for the scene class of include/avr.h) ImageTileIntegrator::Render + EvaluatePixelSample
// (cpu/integrators.cpp:72-298), VolPathIntegrator::Li volume branch (:962-1106) and SampleLd
// (:1282-1399), with SampleT_maj (media.h:741-806) unrolled into per-path ops; see avr_volpath.cu.
#pragma once
#include "avr_sampler.cuh"
#include "avr_tmaj.cuh"

namespace avr {

struct RenderParams {
    int width, height;
    int x0, y0, x1, y1;
    int s0, s1;
    int max_depth, seed;
    SamplerCfg smp;           // the integrator's sampler (independent | zsobol)
    float filter_rx, filter_ry;
    // pixel filter sampled through a table (FilterSampler; avr_filter_table) or null = box.  Packed floats:
    // f[ny*nx] | func = |f| [ny*nx] | conditional cdf [ny*(nx+1)] | conditional integrals [ny] | marginal func [ny] |
    // marginal cdf [ny+1] | marginal integral [1]
    const float *flt;
    int fnx, fny;
    // ... sampled for every pixel sample of the launch by a pre-pass (k_filter_prepass): {p.x, p.y, weight, 0} at
    // sample_slot(px, py, samp); the render kernels only load it (no binary search in their instruction stream)
    const float4 *fsamples;   // per pixel sample of the launch: {filter p.x, p.y, weight, zsobol wavelength draw}, or NULL
    int fweighted;            // the film adds weight * rgb (a tabulated filter); 0: weight 1
    float cfr[16];   // cameraFromRaster (projective)
    float rfc[12];   // renderFromCamera (affine rows)
    float light_dir[3];
    float scene_radius;
    int boundary_type;
    float boundary[6];
    const float *light_spec;  // device, kNSpec
    const float *sensor;      // device, 3*kNSpec
    float imaging_ratio, max_component;
    int rank, world;
    int tiles_x;
    long long local_items;    // work items owned by this rank (multiple of 32)
    long long total_items;    // padded global item count
    int dump_samples;         // 1: write L/lambda/pdf per sample instead of the film
    // CTA path pool and its scheduling knobs (avr_volpath.cu); they order the work, never change a result
    int pool;                 // paths resident per CTA (shared-memory slots)
    int maj_in_smem;          // 1: the majorant grid is staged in shared memory
    int walk_steps;           // DDA steps per tracking trip
    int collide_min;          // lanes inside a segment that make the collision op worth running
    int refill_lanes;         // idle lanes of a tracking warp that trigger a refill from Q_TRACK
    int track_min;            // Q_TRACK fill level at which a warp prefers tracking over partial event batches
    int yield_lanes;          // a dry tracking warp with <= this many paths gives them back if any event waits
    int yield_lanes_full;     // ... or with <= this many paths if a full event batch waits
    int event_smsp_mask;      // bit s set: warps on sub-partition s (warp id & 3) serve events first; the others track first
};

// stats slots in the counters array
enum { CNT_WORK = 0, CNT_RAYS = 1, CNT_CAM_FETCH = 2, CNT_SH_FETCH = 3, CNT_SEG = 4, CNT_SCATTER = 5 };

AVR_DEV bool boundary_hit(const RenderParams &rp, V3 o, V3 d, float *t0, float *t1) {
    if (rp.boundary_type == 0) {
        double ox = double(o.x) - rp.boundary[0], oy = double(o.y) - rp.boundary[1], oz = double(o.z) - rp.boundary[2];
        double r = rp.boundary[3];
        double a = double(d.x) * d.x + double(d.y) * d.y + double(d.z) * d.z;
        double b = 2 * (ox * d.x + oy * d.y + oz * d.z);
        double c = ox * ox + oy * oy + oz * oz - r * r;
        double disc = b * b - 4 * a * c;
        if (!(disc >= 0) || a == 0) return false;
        double s = sqrt(disc);
        double q = (b < 0) ? -0.5 * (b - s) : -0.5 * (b + s);
        double ta = q / a, tb = (q != 0) ? c / q : ta;
        if (ta > tb) { double t = ta; ta = tb; tb = t; }
        *t0 = (float)ta;
        *t1 = (float)tb;
        return true;
    }
    float a0 = __int_as_float(0xff800000), a1 = __int_as_float(0x7f800000);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float inv = 1 / comp(d, i);
        float tn = (rp.boundary[i] - comp(o, i)) * inv, tf = (rp.boundary[3 + i] - comp(o, i)) * inv;
        if (tn > tf) { float t = tn; tn = tf; tf = t; }
        a0 = tn > a0 ? tn : a0;
        a1 = tf < a1 ? tf : a1;
        if (a0 > a1) return false;
    }
    *t0 = a0;
    *t1 = a1;
    return true;
}

// FAST variant: float arithmetic and approximate reciprocals (statistically equivalent, not bit-compatible)
AVR_DEV bool boundary_hit_fast(const RenderParams &rp, V3 o, V3 d, float *t0, float *t1) {
    if (rp.boundary_type == 0) {
        float ox = o.x - rp.boundary[0], oy = o.y - rp.boundary[1], oz = o.z - rp.boundary[2];
        float r = rp.boundary[3];
        float a = __fmaf_rn(d.x, d.x, __fmaf_rn(d.y, d.y, d.z * d.z));
        float hb = __fmaf_rn(ox, d.x, __fmaf_rn(oy, d.y, oz * d.z));  // b / 2
        // |o - (o.d/a) d|^2 form of the discriminant: stable for origins far from / on the sphere
        float ia = ftz_rcp(a);
        float k = hb * ia;
        float px = __fmaf_rn(-k, d.x, ox), py = __fmaf_rn(-k, d.y, oy), pz = __fmaf_rn(-k, d.z, oz);
        float disc = r * r - __fmaf_rn(px, px, __fmaf_rn(py, py, pz * pz));  // = (b^2/4 - a c) / a
        if (!(disc >= 0) || a == 0) return false;
        float s = sqrtf(disc * ia);
        *t0 = -k - s;
        *t1 = -k + s;
        return true;
    }
    float a0 = __int_as_float(0xff800000), a1 = __int_as_float(0x7f800000);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float inv = ftz_rcp(comp(d, i));
        float tn = (rp.boundary[i] - comp(o, i)) * inv, tf = (rp.boundary[3 + i] - comp(o, i)) * inv;
        a0 = fmaxf(a0, fminf(tn, tf));
        a1 = fminf(a1, fmaxf(tn, tf));
    }
    *t0 = a0;
    *t1 = a1;
    return a0 <= a1;
}

// FindInterval (util/math.h) over a CDF of `sz` entries: the last index with cdf[index] <= u, clamped to [0, sz - 2]
AVR_DEV int find_interval_cdf(const float *cdf, int sz, float u) {
    int size = sz - 2, first = 1;
    while (size > 0) {
        int half = size >> 1, middle = first + half;
        bool pr = cdf[middle] <= u;
        first = pr ? middle + 1 : first;
        size = pr ? size - (half + 1) : half;
    }
    return min(max(first - 1, 0), sz - 2);
}
// PiecewiseConstant1D::Sample, util/sampling.h:657-675
AVR_DEV float pc1d_sample(const float *func, const float *cdf, float func_int, int n, float lo, float hi, float u,
                          float *pdf, int *offset) {
    int o = find_interval_cdf(cdf, n + 1, u);
    *offset = o;
    float du = u - cdf[o];
    if (cdf[o + 1] - cdf[o] > 0) du /= cdf[o + 1] - cdf[o];
    *pdf = func_int > 0 ? func[o] / func_int : 0.f;
    return lerp_exact((float(o) + du) / float(n), lo, hi);
}
// SampleTent (util/sampling.h:196-201): SampleDiscrete({.5, .5}, u) picks the half and remaps u (:79-113, the weights
// sum to 1 exactly), SampleLinear (:122-128) draws from the ramp of that half
AVR_DEV float sample_linear(float u, float a, float b) {
    if (u == 0 && a == 0) return 0;
    float x = u * (a + b) / (a + sqrtf((1 - u) * (a * a) + u * (b * b)));
    return fminf(x, kOneMinusEps);
}
AVR_DEV float sample_tent(float u, float r) {
    float up = u * 1.f;
    if (up == 1.f) up = kOneMinusEps;
    if (0.f + 0.5f <= up) {
        float ur = fminf((up - 0.5f) / 0.5f, kOneMinusEps);
        return r * sample_linear(ur, 1, 0);
    }
    float ur = fminf((up - 0.f) / 0.5f, kOneMinusEps);
    return -r + r * sample_linear(ur, 0, 1);
}
// FilterSampler::Sample (filters.h:33-38) over PiecewiseConstant2D::Sample (util/sampling.h:760-769): the marginal
// in y first, then the row's conditional in x; weight = f[pi] / (pdf_x * pdf_y)
AVR_DEV void filter_sample(const float *flt, int nx, int ny, float rx, float ry, float u0, float u1, float *px, float *py,
                           float *weight) {
    if (nx < 0) {  // TriangleFilter::Sample, filters.h:224-227
        *px = sample_tent(u0, rx);
        *py = sample_tent(u1, ry);
        *weight = 1.f;
        return;
    }
    const float *f = flt, *func = f + nx * ny, *ccdf = func + nx * ny, *cint = ccdf + ny * (nx + 1);
    const float *mfunc = cint + ny, *mcdf = mfunc + ny, *mint = mcdf + ny + 1;
    float pdf0, pdf1;
    int ox, oy;
    float d1 = pc1d_sample(mfunc, mcdf, mint[0], ny, -ry, ry, u1, &pdf1, &oy);
    float d0 = pc1d_sample(func + oy * nx, ccdf + oy * (nx + 1), cint[oy], nx, -rx, rx, u0, &pdf0, &ox);
    float pdf = pdf0 * pdf1;
    *px = d0;
    *py = d1;
    *weight = f[oy * nx + ox] / pdf;
}

// index of pixel sample (px, py, samp) in per-sample buffers of a launch (sample dump, filter samples)
AVR_DEV long long sample_slot(const RenderParams &rp, int px, int py, int samp) {
    long long nsamp = rp.s1 - rp.s0;
    return ((long long)(py - rp.y0) * (rp.x1 - rp.x0) + (px - rp.x0)) * nsamp + (samp - rp.s0);
}

template <typename TR>
struct PathLane {
    static constexpr int N = TR::NCH;
    // work / sampler
    int px, py, samp;
    RNG srng;          // IndependentSampler stream
    float lambda[4];
    // radiance state
    float L[4];
    Spec<N> beta, r_u, r_l;
    int depth;
    // tracking state (camera segment or shadow ray)
    int state;         // ST_* of avr_volpath.cu
    bool shadow;
    Fetcher<TR> fe;
    DDA dda;
    float cur_tMin, seg_tMax, u;
    float mj;          // majorant of the voxel the DDA stands in, requested one step ahead (registers only)
    float4 tapA, tapB; // density taps of the pending tentative collision, requested when its distance was drawn
    Spec<N> sigma_maj, T_maj;
    RNG trng;
    // shadow-ray state
    Spec<N> T_ray, s_rl, s_ru;
    float f_hat;
    V3 wo;
    // Traits::FUSED: the camera ray set up together with the shadow ray waits in the slot's second ray block
    bool cam_wait;
    // Traits::LEAN: L[i] = Lsum * Le_light(lambda_i) (one channel: every NEE term has the same scalar weight), and
    // the shadow ray's r_u is T_ray / 4^rr_k (both take the same ratios; T_ray also takes the roulette's x 4)
    float Lsum;
    int rr_k;
};

}  // namespace avr
