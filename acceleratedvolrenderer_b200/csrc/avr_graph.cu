// avr_graph.cu — kernel 3 and the graph-side kernels: radiance propagation (SpMV), graph lookup
// (ConnectToGraph), graph rendering and the light-vector estimator.
// Reference: src/graph/lighting_calculator.cpp:23-155, src/graph/graph_integrator.cpp:24-280,
// src/graph/util.h:344-366,814-839.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>
#include <numeric>
#include <string>

#include <unistd.h>

#include "avr_comm.h"
#include "avr_render_host.h"

using namespace avr;

namespace avr {

// ---- kernel 3: one propagation iteration -------------------------------------------------------------
// cur_out = T * cur_in over rows [row_begin,row_end); optionally total += cur_in (the previous
// iteration's result, valid because its NaN/Inf flag is final when this launch starts).
// Layout: CSR (int64 rowptr, int32 col, fp32 val).  A CTA owns kRowsPerCta consecutive rows; their nnz
// range is streamed through shared memory in coalesced 128-bit-friendly chunks as PRODUCTS
// val[k]*x[col[k]], then each thread sums its own row's products in column order starting from 0 —
// the exact summation order of Eigen's column-major SparseMatrix * SparseVector, so results are
// bit-identical to the reference while every global access is coalesced.
constexpr int kRowsPerCta = 256;
constexpr int kChunk = 2048;

// flags[0]: set to 1 when any NaN/Inf was produced; flags[1]: first bad iteration (atomicMin; starts at INT_MAX).
__global__ void __launch_bounds__(kRowsPerCta)
k_spmv(const int64_t *__restrict__ rowptr, const int32_t *__restrict__ col, const float *__restrict__ val,
       const float *__restrict__ x, float *__restrict__ y, float *__restrict__ total, int row_begin, int row_end,
       int *flags, int iteration, int add_prev) {
    __shared__ float s_prod[kChunk];
    // An EARLIER iteration was invalid: the reference has stopped.  Gated on the first bad iteration, not on
    // the raised flag: a CTA of this same launch may raise the flag while others have yet to start, and those
    // must still do their fused `total += x` for the valid previous iteration.
    if (flags[1] < iteration) return;
    int r0 = row_begin + blockIdx.x * kRowsPerCta;
    int r1 = min(r0 + kRowsPerCta, row_end);
    int r = r0 + threadIdx.x;
    bool have = r < r1;
    int64_t rs = 0, re = 0;
    if (have) { rs = rowptr[r]; re = rowptr[r + 1]; }
    int64_t k0 = rowptr[r0], k1 = rowptr[r1];
    float acc = 0;
    for (int64_t c0 = k0; c0 < k1; c0 += kChunk) {
        int64_t c1 = min(c0 + (int64_t)kChunk, k1);
        for (int64_t k = c0 + threadIdx.x; k < c1; k += kRowsPerCta) s_prod[k - c0] = val[k] * __ldg(x + col[k]);
        __syncthreads();
        if (have) {
            int64_t a = max(rs, c0), b = min(re, c1);
            for (int64_t k = a; k < b; ++k) acc += s_prod[k - c0];
        }
        __syncthreads();
    }
    if (have) {
        y[r] = acc;
        if (isnan(acc) || isinf(acc)) {
            atomicExch(&flags[0], 1);
            atomicMin(&flags[1], iteration);
        }
        if (add_prev) total[r] += x[r];
    }
}

// total[r] += cur[r] unless the iteration that produced cur was invalid.
__global__ void k_axpy(const float *__restrict__ cur, float *__restrict__ total, int row_begin, int row_end,
                       const int *flags) {
    if (flags[0]) return;
    int r = row_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (r < row_end) total[r] += cur[r];
}

// ---- ConnectToGraph over a uniform hash grid ------------------------------------------------------------
struct GraphView {
    const float4 *vert;   // x,y,z,lightScalar sorted by cell
    const float *range2;  // renderSearchRange^2, same order (or null)
    const int *cell_start;
    int gres[3];
    float gmin[3], inv_cell;
    float sq_radius, perc99, max_range;
    int use_range, n;
};

AVR_DEV bool graph_search(const GraphView &g, V3 p, float r2, bool filter, float *avg, float *wsum) {
    float r = sqrtf(r2);
    int lo[3], hi[3];
    float pp[3] = {p.x, p.y, p.z};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float flo = floorf((pp[i] - r - g.gmin[i]) * g.inv_cell), fhi = floorf((pp[i] + r - g.gmin[i]) * g.inv_cell);
        // one cell wider on both sides than the rounded range: the exact d2 < r2 test below filters the extra
        // candidates, and a vertex within an ulp of a cell face at distance ~r is never skipped
        lo[i] = (int)fmaxf(flo - 1.f, 0.f);
        hi[i] = (int)fminf(fhi + 1.f, float(g.gres[i] - 1));
        if (!(flo <= float(g.gres[i] - 1)) || !(fhi >= 0.f)) lo[i] = hi[i] + 1;  // outside the grid (or NaN)
    }
    float a = 0, w = 0;
    bool any = false;
    *avg = 0;
    *wsum = 0;
    if (lo[0] > hi[0] || lo[1] > hi[1] || lo[2] > hi[2]) return false;
    for (int z = lo[2]; z <= hi[2]; ++z)
        for (int y = lo[1]; y <= hi[1]; ++y) {
            int c0 = (z * g.gres[1] + y) * g.gres[0];
            int s = g.cell_start[c0 + lo[0]], e = g.cell_start[c0 + hi[0] + 1];
            for (int k = s; k < e; ++k) {
                float4 v = __ldg(g.vert + k);
                float dx = p.x - v.x, dy = p.y - v.y, dz = p.z - v.z;
                float d2 = dx * dx + dy * dy + dz * dz;
                if (d2 < r2) {
                    if (filter && d2 > __ldg(g.range2 + k)) continue;
                    float wt = 1.f / d2;
                    a += v.w * wt;
                    w += wt;
                    any = true;
                }
            }
        }
    *avg = a;
    *wsum = w;
    return any;
}
// graph_integrator.cpp:249-280 (summation order differs from nanoflann's distance-sorted list:
// parity is to 1e-5 relative, stated in the tests)
AVR_DEV float connect_to_graph(const GraphView &g, V3 p) {
    float a, w;
    bool any = graph_search(g, p, g.sq_radius, false, &a, &w);
    if (g.use_range && !any) {
        any = graph_search(g, p, g.perc99, true, &a, &w);
        if (!any) any = graph_search(g, p, g.max_range, true, &a, &w);
    }
    if (!any) return 0.f;
    if (w == 0) return 0.f;
    return a / w;
}

__global__ void k_graph_connect(GraphView g, const float *pts, int64_t n, float *out) {
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n) return;
    out[i] = connect_to_graph(g, v3(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]));
}

// ---- light-vector estimator: ComputeRaysToSphere + SampleTransmittance ------------------------------
// util::ComputeRaysToSphere (src/graph/util.h:814-839, rayInSphere absent) over one ray: `iterations`
// ratio-tracking estimates (util::SampleTransmittance, :344-366) with IndependentSampler streams
// StartPixelSample({px, py}, i) and tMax_i = start_scatter_t + dist_in_sphere * Get1D(); returns the mean.
template <typename TR>
AVR_DEV float light_ray_estimate(const MediumView &mv, const LambdaCtx<TR::NCH> &lc, V3 o, V3 d, int px, int py,
                                 float start_scatter_t, float dist_in_sphere, int iterations, int seed) {
    constexpr int N = TR::NCH;
    float average = 0, totalWeight = 0;
    for (int it = 0; it < iterations; ++it) {
        RNG srng;
        srng.set_sequence(hash_3u32((uint32_t)px, (uint32_t)py, (uint32_t)seed));
        srng.advance((long long)it * 65536ll);
        float curDist = dist_in_sphere * srng.uniform();
        float tMax = start_scatter_t + curDist;
        // unsequenced draws in the reference (util.h:346); GCC order: first draw -> offset
        uint64_t h1 = hash_float(srng.uniform());
        uint64_t h0 = hash_float(srng.uniform());
        RNG rng;
        rng.set_sequence(h0, h1);
        float u = srng.uniform();
        float Tr = 1;
        TmajCount cnt = {0, 0};
        sample_tmaj<TR>(mv, mv.maj, lc, o, d, tMax, u, rng,
                        [&](float, const PropsN<N> &mp, Spec<N> sigma_maj, Spec<N>) {
                            Spec<N> sigma_n = clamp_zero(sigma_maj - mp.sigma_a - mp.sigma_s);
                            Tr *= sigma_n[0] / sigma_maj[0];
                            return Tr != 0;
                        }, &cnt);
        average += Tr * 1.f;
        totalWeight += 1.f;
    }
    return (iterations == 0 || totalWeight == 0) ? 0.f : average / totalWeight;
}

template <typename TR>
__global__ void __launch_bounds__(128)
k_light_rays(MediumView mv, const float *rays, const int32_t *pixel, const float *ts, int64_t n, int iterations,
             int seed, float l0, float l1, float l2, float l3, float *out) {
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n) return;
    float lambda[4] = {l0, l1, l2, l3};
    LambdaCtx<TR::NCH> lc = make_lambda_ctx<TR::NCH>(mv, lambda);
    const float *r = rays + 6 * i;
    out[i] = light_ray_estimate<TR>(mv, lc, v3(r[0], r[1], r[2]), v3(r[3], r[4], r[5]), pixel[2 * i], pixel[2 * i + 1],
                                    ts[2 * i], ts[2 * i + 1], iterations, seed);
}

// ---- LightingCalculator::GetLightVector (src/graph/lighting_calculator.cpp:84-155) -------------------------
// Per vertex: a disk of rays along the light direction through the vertex sphere (util::GetDiskPoints,
// src/graph/util.h:179-204), each clipped against the medium's bounding shape and the vertex sphere
// (GetHits / GetStartEndT, :419-503), estimated by ComputeRaysToSphere, averaged over the rays that qualify,
// times Inv4Pi.  One thread per (vertex, lattice slot of the (2n+1)^2 disk lattice); the slot's index among the
// points INSIDE the disk (the reference's pointIndex, which selects the sampler pixel) is recounted by the
// thread itself.  The analytic sphere / box hits stand in for Primitive::Intersect + SkipIntersection: hit
// distances agree with the reference's to rounding, so parity of L0 is by tolerance (tests).
struct LightVecParams {
    float in_dir[3];       // direction the light travels (= -wi)
    int boundary_type;
    float boundary[6];
    float max_dist_to_center, vertex_radius;
    int iterations, points_on_radius, seed, resolution;
    float lambda[4];
};

AVR_DEV bool sphere_hit2(V3 c, float r, V3 o, V3 d, float *t0, float *t1) {
    double ox = double(o.x) - c.x, oy = double(o.y) - c.y, oz = double(o.z) - c.z;
    double a = double(d.x) * d.x + double(d.y) * d.y + double(d.z) * d.z;
    double b = 2 * (ox * d.x + oy * d.y + oz * d.z);
    double cc = ox * ox + oy * oy + oz * oz - double(r) * r;
    double disc = b * b - 4 * a * cc;
    if (!(disc > 0) || a == 0) return false;  // tangent rays have one hit: not "OutsideTwoHits"
    double s = sqrt(disc);
    double q = (b < 0) ? -0.5 * (b - s) : -0.5 * (b + s);
    double ta = q / a, tb = (q != 0) ? cc / q : ta;
    if (ta > tb) { double t = ta; ta = tb; tb = t; }
    *t0 = (float)ta;
    *t1 = (float)tb;
    return ta > 0;  // the origin lies outside, in front of the shape
}

template <typename TR>
__global__ void __launch_bounds__(128)
k_light_vector(MediumView mv, LightVecParams lp, const float *pos, const int32_t *list_index, int n, float *ray_val,
               unsigned char *ray_ok) {
    const int side = 2 * lp.points_on_radius + 1, slots = side * side;
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= (long long)n * slots) return;
    int v = int(i / slots), slot = int(i % slots);
    ray_ok[i] = 0;
    V3 dir = v3(lp.in_dir[0], lp.in_dir[1], lp.in_dir[2]);
    V3 vp = v3(pos[3 * v], pos[3 * v + 1], pos[3 * v + 2]);
    // origin = vertex.point - inDirection * maxDistToCenter * 2
    V3 center = vp - (dir * lp.max_dist_to_center) * 2.f;
    // GetDiskPoints: CoordinateSystem(direction) (vecmath.h:1007-1013), step = radius / (n + 1)
    float sign = copysignf(1.f, dir.z);
    float a = -1 / (sign + dir.z);
    float b = dir.x * dir.y * a;
    V3 xv = v3(1 + sign * sqr(dir.x) * a, sign * b, -sign * dir.x);
    V3 yv = v3(b, sign + sqr(dir.y) * a, -dir.y);
    int point_index = 0, count = 0;
    V3 disk_point = center;
    bool inside = lp.points_on_radius == 0 ? slot == 0 : false;
    if (lp.points_on_radius > 0) {
        float step = lp.vertex_radius / float(lp.points_on_radius + 1);
        xv = xv * step;
        yv = yv * step;
        for (int s = 0; s < slots; ++s) {
            int x = s / side - lp.points_on_radius, y = s % side - lp.points_on_radius;
            V3 p = (center + xv * float(x)) + yv * float(y);
            V3 dd = p - center;
            bool in = sqrtf(sqr(dd.x) + sqr(dd.y) + sqr(dd.z)) <= lp.vertex_radius;
            if (s == slot) { inside = in; disk_point = p; point_index = count; }
            count += in ? 1 : 0;
        }
    } else
        count = 1;
    if (!inside) { ray_ok[i] = 2; return; }  // not a disk point at all
    // GetHits on the medium's bounding shape and on the vertex sphere: both must be entered and left
    float m0, m1, s0, s1;
    bool mh;
    if (lp.boundary_type == 0)
        mh = sphere_hit2(v3(lp.boundary[0], lp.boundary[1], lp.boundary[2]), lp.boundary[3], disk_point, dir, &m0, &m1);
    else {
        float t0 = __int_as_float(0xff800000), t1 = __int_as_float(0x7f800000);
        mh = true;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float inv = 1 / comp(dir, k);
            float tn = (lp.boundary[k] - comp(disk_point, k)) * inv, tf = (lp.boundary[3 + k] - comp(disk_point, k)) * inv;
            if (tn > tf) { float t = tn; tn = tf; tf = t; }
            t0 = tn > t0 ? tn : t0;
            t1 = tf < t1 ? tf : t1;
        }
        mh = t0 < t1 && t0 > 0;
        m0 = t0; m1 = t1;
    }
    if (!mh || !sphere_hit2(vp, lp.vertex_radius, disk_point, dir, &s0, &s1)) return;
    // GetStartEndT, then SkipIntersection to the medium entry and StartEndT::SkipForward(startT)
    float startT = m0, endT = fminf(m1, s1), startScatterT = fmaxf(m0, s0), endScatterT = endT;
    if (endT < startScatterT || endScatterT < startT) return;
    V3 o = v3(disk_point.x + dir.x * startT, disk_point.y + dir.y * startT, disk_point.z + dir.z * startT);
    startScatterT -= startT;
    endScatterT -= startT;
    LambdaCtx<TR::NCH> lc = make_lambda_ctx<TR::NCH>(mv, lp.lambda);
    long long li = list_index ? list_index[v] : v;
    unsigned long long cur = (unsigned long long)li * (unsigned)count + (unsigned)point_index;
    int yCoor = int(cur / (unsigned)lp.resolution);
    int xCoor = int(cur - (unsigned long long)yCoor * (unsigned)lp.resolution);
    ray_val[i] = light_ray_estimate<TR>(mv, lc, o, dir, xCoor, yCoor, startScatterT, endScatterT - startScatterT,
                                        lp.iterations, lp.seed);
    ray_ok[i] = 1;
}

// util::Averager over the rays of a vertex in pointIndex order, then * Inv4Pi (lighting_calculator.cpp:147-152)
__global__ void k_light_vector_mean(const float *ray_val, const unsigned char *ray_ok, int n, int slots, float *out) {
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    float average = 0, total = 0;
    for (int s = 0; s < slots; ++s)
        if (ray_ok[(size_t)v * slots + s] == 1) { average += ray_val[(size_t)v * slots + s] * 1.f; total += 1.f; }
    out[v] = (total == 0 ? 0.f : average / total) * kInv4Pi;
}

// ---- GraphIntegrator::Li, free-graph branch (graph_integrator.cpp:180-244) ---------------------------
struct XfRows {
    float a[12];
};
template <typename TR>
__global__ void __launch_bounds__(128)
k_render_graph(MediumView mv, RenderParams rp, GraphView g, XfRows wfr, double *film) {
    constexpr int N = TR::NCH;
    long long npix = (long long)(rp.x1 - rp.x0) * (rp.y1 - rp.y0);
    long long total = npix * (rp.s1 - rp.s0);
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= total) return;
    if ((i >> 5) % rp.world != rp.rank) return;
    int s = rp.s0 + int(i / npix);
    int pi = int(i % npix);
    int px = rp.x0 + pi % (rp.x1 - rp.x0), py = rp.y0 + pi / (rp.x1 - rp.x0);
    PathLane<TR> ln;  // reuse start-of-path arithmetic
    ln.px = px; ln.py = py; ln.samp = s;
    // the scene's sampler (a run-time switch: this kernel is not worth a second set of instantiations)
    const bool zs = rp.smp.kind == AVR_SAMPLER_ZSOBOL;
    auto draw1 = [&]() -> float {
        if (zs) return PathSamplerRef<true>{rp.smp, ln.srng, px, py, s}.get1d();
        return ln.srng.uniform();
    };
    if (zs) PathSamplerRef<true>{rp.smp, ln.srng, px, py, s}.start_pixel_sample();
    else PathSamplerRef<false>{rp.smp, ln.srng, px, py, s}.start_pixel_sample();
    float lu = draw1();
    float pdf[4];
    sample_visible_wavelengths(lu, ln.lambda, pdf);
    LambdaCtx<N> lc = make_lambda_ctx<N>(mv, ln.lambda);
    float f0, f1;
    if (zs) PathSamplerRef<true>{rp.smp, ln.srng, px, py, s}.get2d(&f0, &f1);
    else PathSamplerRef<false>{rp.smp, ln.srng, px, py, s}.get2d(&f0, &f1);
    float fx = lerp_exact(f0, -rp.filter_rx, rp.filter_rx), fy = lerp_exact(f1, -rp.filter_ry, rp.filter_ry);
    float fw = 1.f;
    if (rp.fsamples) {  // tabulated filter: offset and weight from the pre-pass (same two draws), as in k_volpath
        const float4 fs = rp.fsamples[sample_slot(rp, px, py, s)];
        fx = fs.x; fy = fs.y; fw = fs.z;
    }
    float pfx = (float(px) + fx) + 0.5f, pfy = (float(py) + fy) + 0.5f;
    if (zs) PathSamplerRef<true>{rp.smp, ln.srng, px, py, s}.skip(3);  // time, lens
    else PathSamplerRef<false>{rp.smp, ln.srng, px, py, s}.skip(3);
    const float *c = rp.cfr;
    float xp = c[0] * pfx + c[1] * pfy + c[2] * 0.f + c[3];
    float yp = c[4] * pfx + c[5] * pfy + c[6] * 0.f + c[7];
    float zp = c[8] * pfx + c[9] * pfy + c[10] * 0.f + c[11];
    float wp = c[12] * pfx + c[13] * pfy + c[14] * 0.f + c[15];
    V3 pc = v3(xp, yp, zp);
    if (wp != 1) pc = pc / wp;
    V3 dir = pc / sqrtf(sqr(pc.x) + sqr(pc.y) + sqr(pc.z));
    V3 ro, rd;
    xform_ray_exact(rp.rfc, true, v3(0, 0, 0), dir, nullptr, &ro, &rd);
    float Lg = 0;
    float t0, t1;
    if (boundary_hit(rp, ro, rd, &t0, &t1) && t1 > 0) {
        if (t0 > 0) ro = v3(ro.x + rd.x * t0, ro.y + rd.y * t0, ro.z + rd.z * t0);
        if (boundary_hit(rp, ro, rd, &t0, &t1) && t1 > 0) {
            uint64_t hash0 = hash_float(draw1());
            uint64_t hash1 = hash_float(draw1());
            RNG rng;
            rng.set_sequence(hash0, hash1);
            float u = draw1();
            TmajCount cnt = {0, 0};
            float len = sqrtf(sqr(rd.x) + sqr(rd.y) + sqr(rd.z));
            V3 dn = rd / len;
            sample_tmaj<TR>(mv, mv.maj, lc, ro, rd, t1, u, rng,
                            [&](float t, const PropsN<N> &mp, Spec<N> sigma_maj, Spec<N>) {
                                float pAbsorb = mp.sigma_a[0] / sigma_maj[0];
                                float pScatter = mp.sigma_s[0] / sigma_maj[0];
                                float pNull = fmaxf(0.f, 1 - pAbsorb - pScatter);
                                float um = rng.uniform();
                                int md = sample_discrete3(pAbsorb, pScatter, pNull, um);
                                if (md == 0) return false;
                                if (md == 1) {
                                    V3 p = v3(ro.x + dn.x * t, ro.y + dn.y * t, ro.z + dn.z * t);
                                    // worldFromRender(p): Transform::operator()(Point3f)
                                    const float *a = wfr.a;
                                    V3 pw = v3(a[0] * p.x + a[1] * p.y + a[2] * p.z + a[3],
                                               a[4] * p.x + a[5] * p.y + a[6] * p.z + a[7],
                                               a[8] * p.x + a[9] * p.y + a[10] * p.z + a[11]);
                                    Lg += connect_to_graph(g, pw);
                                    return false;
                                }
                                return true;
                            }, &cnt);
        }
    }
    // spectrum * L (graph_integrator.cpp:243), then the film path of avr_render_volpath
#pragma unroll
    for (int k = 0; k < 4; ++k) ln.L[k] = sample_dense1(rp.light_spec, ln.lambda[k]) * Lg;
    RenderParams rpl = rp;
    rpl.dump_samples = 0;
    // inline copy of finish_path's film branch
    float rgb[3];
    float Ld[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        bool bad = isnan(ln.L[k]) || isinf(ln.L[k]);
        Ld[k] = (pdf[k] != 0 && !bad) ? ln.L[k] / pdf[k] : 0.f;
    }
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        float sacc = sample_dense1(rp.sensor + ch * kNSpec, ln.lambda[0]) * Ld[0];
#pragma unroll
        for (int k = 1; k < 4; ++k) sacc += sample_dense1(rp.sensor + ch * kNSpec, ln.lambda[k]) * Ld[k];
        rgb[ch] = rp.imaging_ratio * (sacc / 4);
    }
    float m = fmaxf(rgb[0], fmaxf(rgb[1], rgb[2]));
    if (m > rp.max_component) { float k = rp.max_component / m; rgb[0] *= k; rgb[1] *= k; rgb[2] *= k; }
    double *pxl = film + 4 * ((size_t)py * rp.width + px);
    // RGBFilm::AddSample: rgbSum += weight * rgb, weightSum += weight (weight 1 for the box filter)
    atomicAdd(pxl + 0, (double)(fw * rgb[0]));
    atomicAdd(pxl + 1, (double)(fw * rgb[1]));
    atomicAdd(pxl + 2, (double)(fw * rgb[2]));
    atomicAdd(pxl + 3, (double)fw);
}

}  // namespace avr

#include "avr_graph_host.inc"
#include "avr_graph_builder.inc"

// ---- avr_graph_propagate_noalloc: ComputeFinalLight with its scratch in a static device array ---------
// No counterpart in the reference; added for callers that
// cannot call cudaMalloc (CUDA graph capture, tight serving loops).  The ping-pong vectors and the
// NaN/Inf flags of the propagation live in ONE file-scope array that every GPU loading this library
// holds until exit: 2 * AVR_NOALLOC_MAX_VERTICES floats + one 256-byte flag line = 8 MiB + 256 B.
// Every call shares it, so calls that overlap on one GPU race; contents survive from call to call and
// the init kernel below rewrites what must start clean (flags, cur[0,n)); nxt is written before read.
namespace {

constexpr size_t kNoallocFloats = 2 * size_t(AVR_NOALLOC_MAX_VERTICES) + 64;
static __device__ __align__(256) float g_noalloc_scratch[kNoallocFloats];

// cur = total = L0 and flags = {0, INT_MAX}, in one launch instead of two copies and a host upload
__global__ void k_noalloc_init(const float *__restrict__ L0, float *__restrict__ cur, float *__restrict__ total,
                               int *flags, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { flags[0] = 0; flags[1] = 0x7fffffff; }
    if (i < n) { float v = L0[i]; cur[i] = v; total[i] = v; }
}

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address the
// kernels get comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int kMaxDevices = 64;
float *s_noalloc_ptr[kMaxDevices] = {};
std::mutex s_noalloc_mutex;

avr_status noalloc_scratch(avr_ctx *ctx, float **out) {
    if (ctx->device < 0 || ctx->device >= kMaxDevices)
        return set_err(ctx, AVR_ERR_UNSUPPORTED, "avr_graph_propagate_noalloc: device ordinal beyond 63");
    std::lock_guard<std::mutex> lock(s_noalloc_mutex);
    if (!s_noalloc_ptr[ctx->device]) {
        int cur_dev = -1;
        AVR_CUDA(ctx, cudaGetDevice(&cur_dev));
        if (cur_dev != ctx->device)
            return set_err(ctx, AVR_ERR_CUDA, "avr_graph_propagate_noalloc: the context's device is not current");
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return set_cuda_err(ctx, e, "cudaGetSymbolAddress(g_noalloc_scratch)");
        if (!p) return set_err(ctx, AVR_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch) returned NULL");
        s_noalloc_ptr[ctx->device] = static_cast<float *>(p);
    }
    *out = s_noalloc_ptr[ctx->device];
    return AVR_OK;
}

}  // namespace

extern "C" avr_status avr_graph_propagate_noalloc(avr_graph *g, const float *L0_dev, int32_t bounces, float *total_dev,
                                                  int32_t *iters_done, void *stream_) {
    if (!g || !L0_dev || !total_dev)
        return set_err(g ? g->ctx : nullptr, AVR_ERR_BAD_ARG, "avr_graph_propagate_noalloc: NULL argument");
    avr_ctx *ctx = g->ctx;
    if (bounces < 0) return set_err(ctx, AVR_ERR_BAD_ARG, "avr_graph_propagate_noalloc: bounces must be >= 0");
    if (g->n > AVR_NOALLOC_MAX_VERTICES)
        return set_err(ctx, AVR_ERR_UNSUPPORTED,
                       "avr_graph_propagate_noalloc: " + std::to_string(g->n) + " vertices exceed the static scratch of " +
                           std::to_string(AVR_NOALLOC_MAX_VERTICES) + " (AVR_NOALLOC_MAX_VERTICES); use avr_graph_propagate_device");
    AVR_CUDA(ctx, cudaSetDevice(ctx->device));
    float *scratch = nullptr;
    avr_status st = noalloc_scratch(ctx, &scratch);
    if (st != AVR_OK) return st;
    float *cur = scratch, *nxt = scratch + AVR_NOALLOC_MAX_VERTICES;
    int *flags = reinterpret_cast<int *>(scratch + 2 * size_t(AVR_NOALLOC_MAX_VERTICES));
    cudaStream_t s = stream_ ? (cudaStream_t)stream_ : ctx->stream;
    k_noalloc_init<<<blocks_for(g->n, 256), 256, 0, s>>>(L0_dev, cur, total_dev, flags, g->n);
    int nblk = blocks_for(g->n, kRowsPerCta);
    for (int it = 0; it < bounces; ++it) {
        k_spmv<<<nblk, kRowsPerCta, 0, s>>>(g->d_rowptr, g->d_col, g->d_val, cur, nxt, total_dev, 0, g->n, flags, it,
                                             it > 0 ? 1 : 0);
        std::swap(cur, nxt);
    }
    if (bounces > 0) k_axpy<<<blocks_for(g->n, 256), 256, 0, s>>>(cur, total_dev, 0, g->n, flags);
    AVR_CUDA(ctx, cudaGetLastError());
    if (iters_done) {
        int32_t fl[2];
        AVR_CUDA(ctx, cudaMemcpyAsync(fl, flags, sizeof(fl), cudaMemcpyDeviceToHost, s));
        AVR_CUDA(ctx, cudaStreamSynchronize(s));
        *iters_done = fl[0] ? fl[1] : bounces;
    }
    return AVR_OK;
}

// ---- avr_graph_propagate_multi: ComputeFinalLight with the rows of T split over the ranks ----------------
// 1-D vertex-range partition, equal chunks of ceil(n / world) rows; rank r multiplies rows
// [r * chunk, (r + 1) * chunk) and writes them straight into its block of the gathered vector, ONE in-place
// ncclAllGather per bounce makes that vector the next `cur` on every rank (rank-major blocks of equal size ARE
// the row order, so nothing is repacked).  The reference's NaN/Inf stop (lighting_calculator.cpp:36-48) needs no
// extra collective: after the gather every rank holds the whole vector and tests it itself (k_flag_nonfinite),
// so all ranks reach the same verdict.  The bounce sequence is captured once in a CUDA graph: no host
// synchronisation between bounces.
namespace {

__global__ void k_flag_nonfinite(const float *__restrict__ x, int n, int *flags, int iteration) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool bad = false;
    for (; i < n; i += gridDim.x * blockDim.x) {
        float v = x[i];
        bad = bad || isnan(v) || isinf(v);
    }
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) {
        atomicExch(&flags[0], 1);
        atomicMin(&flags[1], iteration);
    }
}

// cur = total = L0 (cur padded with zeros up to `padded`), flags = {0, INT_MAX}
__global__ void k_multi_init(const float *__restrict__ L0, float *__restrict__ cur, float *__restrict__ total, int *flags,
                             int n, int padded, int r0, int r1) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { flags[0] = 0; flags[1] = 0x7fffffff; }
    if (i < padded) {
        float v = i < n ? L0[i] : 0.f;
        cur[i] = v;
        if (i >= r0 && i < r1) total[i] = v;
    }
}


// ---- SpMV fused with its exchange: the new vector is written straight into every peer's copy ------------------
// Every rank owns ONE block {vector[2][padded] floats, SYNC_WORDS ints} that all peers have mapped (NVLink /
// NVSwitch peer memory).  A bounce is k_spmv_p2p (rows of this rank; each row's result is stored into the `next`
// vector of EVERY rank, so the transfer rides along with the arithmetic, 128 bytes per warp store) followed by
// k_p2p_barrier (a release/acquire handshake over the peers' sync words).  No NCCL call, no staging copy, no
// separate finiteness pass: a rank that produces a NaN/Inf posts the iteration into slot [rank] of every peer's
// `bad` words, which all ranks read after the barrier — the same verdict everywhere, as the reference's stop needs.
struct PeerTable {
    char *base[AVR_P2P_MAX_WORLD];
    int world, rank;
    unsigned long long padded;
};
enum { SYNC_ARRIVE = 0, SYNC_BAD = AVR_P2P_MAX_WORLD, SYNC_EPOCH = 2 * AVR_P2P_MAX_WORLD, SYNC_TIMEOUT = SYNC_EPOCH + 1,
       SYNC_WORDS = 64 };
AVR_DEV float *p2p_vec(const PeerTable &t, int p, int which) {
    return reinterpret_cast<float *>(t.base[p]) + size_t(which) * t.padded;
}
AVR_DEV int *p2p_sync(const PeerTable &t, int p) { return reinterpret_cast<int *>(reinterpret_cast<float *>(t.base[p]) + 2 * t.padded); }
AVR_DEV int p2p_first_bad(const PeerTable &t) {
    const volatile int *s = p2p_sync(t, t.rank) + SYNC_BAD;
    int m = 0x7fffffff;
    for (int p = 0; p < t.world; ++p) m = min(m, s[p]);
    return m;
}
AVR_DEV void st_release_sys(int *p, int v) { asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
AVR_DEV int ld_acquire_sys(const int *p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
AVR_DEV unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// One warp.  Everything this GPU wrote before (earlier kernels of the stream included) is released to the peers with
// the arrival word; passing means every peer has arrived, i.e. all their stores into this GPU's block are visible.
// A peer that never arrives (dead process) trips the 10 s timeout instead of hanging the GPU.
__global__ void k_p2p_barrier(PeerTable t) {
    const int lane = threadIdx.x;
    int *mine = p2p_sync(t, t.rank);
    const int e = *(volatile int *)(mine + SYNC_EPOCH) + 1;
    __threadfence_system();
    if (lane < t.world) st_release_sys(p2p_sync(t, lane) + SYNC_ARRIVE + t.rank, e);
    if (lane < t.world) {
        const unsigned long long t0 = global_timer_ns();
        while (ld_acquire_sys(mine + SYNC_ARRIVE + lane) - e < 0) {
            if (global_timer_ns() - t0 > 10000000000ull) {
                mine[SYNC_TIMEOUT] = 1;
                break;
            }
            __nanosleep(100);
        }
    }
    __syncwarp();
    if (lane == 0) *(volatile int *)(mine + SYNC_EPOCH) = e;
}

// total rows = L0 (bounce 0 reads L0 itself: no copy into the exchanged vectors), and this rank's `bad` slot reset on
// every peer (the writer of a slot resets it: a later post of the same call follows it on the same path)
__global__ void k_p2p_init(PeerTable t, const float *__restrict__ L0, float *__restrict__ total, int r0, int r1) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < t.world) {
        p2p_sync(t, i)[SYNC_BAD + t.rank] = 0x7fffffff;
        __threadfence_system();
    }
    if (r0 + i < r1) total[r0 + i] = L0[r0 + i];
}

__global__ void __launch_bounds__(kRowsPerCta)
k_spmv_p2p(PeerTable t, const float *__restrict__ L0, const int64_t *__restrict__ rowptr, const int32_t *__restrict__ col,
           const float *__restrict__ val, float *__restrict__ total, int row_begin, int row_end, int iteration, int add_prev) {
    __shared__ float s_prod[kChunk];
    if (p2p_first_bad(t) < iteration) return;  // an EARLIER iteration was invalid: the reference has stopped
    const float *__restrict__ x = iteration == 0 ? L0 : p2p_vec(t, t.rank, iteration & 1);
    int r0 = row_begin + blockIdx.x * kRowsPerCta;
    int r1 = min(r0 + kRowsPerCta, row_end);
    int r = r0 + threadIdx.x;
    bool have = r < r1;
    int64_t rs = 0, re = 0;
    if (have) { rs = rowptr[r]; re = rowptr[r + 1]; }
    int64_t k0 = rowptr[r0], k1 = rowptr[r1];
    float acc = 0;
    for (int64_t c0 = k0; c0 < k1; c0 += kChunk) {
        int64_t c1 = min(c0 + (int64_t)kChunk, k1);
        for (int64_t k = c0 + threadIdx.x; k < c1; k += kRowsPerCta) s_prod[k - c0] = val[k] * x[col[k]];
        __syncthreads();
        if (have) {
            int64_t a = max(rs, c0), b = min(re, c1);
            for (int64_t k = a; k < b; ++k) acc += s_prod[k - c0];
        }
        __syncthreads();
    }
    if (have) {
        const int nxt = (iteration + 1) & 1;
        for (int p = 0; p < t.world; ++p) p2p_vec(t, p, nxt)[r] = acc;  // the exchange: one coalesced store per peer
        if (isnan(acc) || isinf(acc))
            for (int p = 0; p < t.world; ++p) atomicMin_system(p2p_sync(t, p) + SYNC_BAD + t.rank, iteration);
        if (add_prev) total[r] += x[r];
    }
}

__global__ void k_axpy_p2p(PeerTable t, int which, float *__restrict__ total, int row_begin, int row_end) {
    if (p2p_first_bad(t) != 0x7fffffff) return;
    int r = row_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (r < row_end) total[r] += p2p_vec(t, t.rank, which)[r];
}

}  // namespace

namespace {

struct P2pInfo {  // what a rank tells the others about its block
    cudaIpcMemHandle_t handle;
    long long pid;
    void *ptr;
    int device, ok;
    char pad[128 - sizeof(cudaIpcMemHandle_t) - sizeof(long long) - sizeof(void *) - 2 * sizeof(int)];
};
static_assert(sizeof(P2pInfo) == 128, "P2pInfo is exchanged as 128 bytes");

void p2p_release(avr_comm *cm) {
    for (int p = 0; p < AVR_P2P_MAX_WORLD; ++p) {
        if (cm->p2p_base[p] && p != cm->rank && cm->p2p_ipc[p]) cudaIpcCloseMemHandle(cm->p2p_base[p]);
        if (p != cm->rank) cm->p2p_base[p] = nullptr;
        cm->p2p_ipc[p] = false;
    }
    if (cm->rank < AVR_P2P_MAX_WORLD) {
        cudaFree(cm->p2p_base[cm->rank]);
        cm->p2p_base[cm->rank] = nullptr;
    }
    cm->p2p_padded = 0;
}

// Allocate this rank's block, exchange IPC handles through the communicator, map every peer.  Collective: every rank
// calls it with the same `padded`, and all ranks reach the same verdict (p2p_state 1 or -1).
avr_status p2p_setup(avr_comm *cm, size_t padded, cudaStream_t s) {
    avr_ctx *ctx = cm->ctx;
    const int world = cm->world;
    if (cm->p2p_state == 1 && cm->p2p_padded >= padded) return AVR_OK;
    if (cm->p2p_state == -1) return AVR_OK;
    if (cm->exec) { cudaGraphExecDestroy(cm->exec); cm->exec = nullptr; }
    p2p_release(cm);
    int ok = (ctx->knobs.graph_p2p && world > 1 && world <= AVR_P2P_MAX_WORLD) ? 1 : 0;
    const size_t bytes = 2 * padded * sizeof(float) + SYNC_WORDS * sizeof(int);
    char *own = nullptr;
    P2pInfo mine{};
    if (ok) {
        if (cudaMalloc(&own, bytes) != cudaSuccess) { cudaGetLastError(); ok = 0; own = nullptr; }
        else {
            cudaMemsetAsync(own, 0, bytes, s);
            if (cudaIpcGetMemHandle(&mine.handle, own) != cudaSuccess) { cudaGetLastError(); ok = 0; }
        }
    }
    mine.pid = (long long)getpid();
    mine.ptr = own;
    mine.device = ctx->device;
    mine.ok = ok;
    // exchange (device staging: NCCL moves device memory)
    std::vector<P2pInfo> all(world);
    char *d_info = nullptr;
    AVR_CUDA(ctx, cudaMalloc(&d_info, sizeof(P2pInfo) * world));
    cudaError_t e = cudaMemcpyAsync(d_info + sizeof(P2pInfo) * cm->rank, &mine, sizeof(P2pInfo), cudaMemcpyHostToDevice, s);
    ncclResult_t nr = ncclSuccess;
    if (e == cudaSuccess) nr = cm->api->AllGather(d_info + sizeof(P2pInfo) * cm->rank, d_info, sizeof(P2pInfo), ncclChar, cm->comm, s);
    if (e == cudaSuccess && nr == ncclSuccess) e = cudaMemcpyAsync(all.data(), d_info, sizeof(P2pInfo) * world, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess || nr != ncclSuccess) {
        cudaFree(d_info); cudaFree(own);
        if (nr != ncclSuccess) return set_err(ctx, AVR_ERR_NCCL, std::string("ncclAllGather (peer handles): ") + cm->api->GetErrorString(nr));
        return set_cuda_err(ctx, e, "peer handle exchange");
    }
    for (int p = 0; p < world && ok; ++p) ok = ok && all[p].ok;
    if (ok) {
        cm->p2p_base[cm->rank] = own;
        for (int p = 0; p < world && ok; ++p) {
            if (p == cm->rank) continue;
            if (all[p].pid == mine.pid) {  // a thread of this process (avr_comm_init_all): direct peer access
                int can = 0;
                cudaDeviceCanAccessPeer(&can, ctx->device, all[p].device);
                cudaError_t pe = can ? cudaDeviceEnablePeerAccess(all[p].device, 0) : cudaErrorInvalidDevice;
                if (pe == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); pe = cudaSuccess; }
                if (pe != cudaSuccess) { cudaGetLastError(); ok = 0; break; }
                cm->p2p_base[p] = (char *)all[p].ptr;
            } else {
                void *mapped = nullptr;
                if (cudaIpcOpenMemHandle(&mapped, all[p].handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                    cudaGetLastError(); ok = 0; break;
                }
                cm->p2p_base[p] = (char *)mapped;
                cm->p2p_ipc[p] = true;
            }
        }
    }
    // all ranks must take the same path: the minimum of the local verdicts
    int *d_ok = reinterpret_cast<int *>(d_info);
    e = cudaMemcpyAsync(d_ok, &ok, sizeof(int), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) nr = cm->api->AllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, cm->comm, s);
    if (e == cudaSuccess && nr == ncclSuccess) e = cudaMemcpyAsync(&ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(d_info);
    if (nr != ncclSuccess) return set_err(ctx, AVR_ERR_NCCL, std::string("ncclAllReduce (peer verdict): ") + cm->api->GetErrorString(nr));
    if (e != cudaSuccess) return set_cuda_err(ctx, e, "peer verdict");
    if (!ok) {
        if (!cm->p2p_base[cm->rank]) cudaFree(own);
        p2p_release(cm);
        cm->p2p_state = -1;
        return AVR_OK;
    }
    cm->p2p_padded = padded;
    cm->p2p_state = 1;
    return AVR_OK;
}

}  // namespace

extern "C" avr_status avr_graph_propagate_multi(avr_comm *cm, avr_graph *g, const float *L0_dev, int32_t bounces,
                                                float *total_dev, int32_t *iters_done, void *stream_) {
    if (!cm || !g || !L0_dev || !total_dev)
        return set_err(cm ? cm->ctx : nullptr, AVR_ERR_BAD_ARG, "avr_graph_propagate_multi: NULL argument");
    avr_ctx *ctx = cm->ctx;
    if (g->ctx != ctx) return set_err(ctx, AVR_ERR_BAD_ARG, "avr_graph_propagate_multi: the graph belongs to another context");
    if (bounces < 0) return set_err(ctx, AVR_ERR_BAD_ARG, "avr_graph_propagate_multi: bounces must be >= 0");
    AVR_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = stream_ ? (cudaStream_t)stream_ : ctx->stream;
    const int n = g->n, world = cm->world;
    const int chunk = (n + world - 1) / world;
    const size_t padded = size_t(chunk) * world;
    const int r0 = std::min(cm->rank * chunk, n), r1 = std::min((cm->rank + 1) * chunk, n);
    // the fused path: SpMV writes the new vector into every peer's block (p2p_setup is collective; all ranks agree)
    if (world > 1 && cm->p2p_state >= 0) {
        avr_status pst = p2p_setup(cm, padded, s);
        if (pst != AVR_OK) return pst;
    }
    if (cm->p2p_state == 1 && world > 1) {
        PeerTable t{};
        for (int p = 0; p < world; ++p) t.base[p] = cm->p2p_base[p];
        t.world = world; t.rank = cm->rank; t.padded = cm->p2p_padded;
        const void *key[5] = {g->d_val, L0_dev, total_dev, (const void *)(intptr_t)(bounces + (1 << 20)), (const void *)s};
        if (!cm->exec || std::memcmp(key, cm->exec_key, sizeof(key)) != 0) {
            if (cm->exec) { cudaGraphExecDestroy(cm->exec); cm->exec = nullptr; }
            AVR_CUDA(ctx, cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
            // nobody may still read or write the previous call's vectors once this one starts
            k_p2p_barrier<<<1, 32, 0, s>>>(t);
            k_p2p_init<<<blocks_for(std::max(r1 - r0, world), 256), 256, 0, s>>>(t, L0_dev, total_dev, r0, r1);
            for (int it = 0; it < bounces; ++it) {
                if (r1 > r0)
                    k_spmv_p2p<<<blocks_for(r1 - r0, kRowsPerCta), kRowsPerCta, 0, s>>>(t, L0_dev, g->d_rowptr, g->d_col, g->d_val, total_dev,
                                                                                        r0, r1, it, it > 0 ? 1 : 0);
                k_p2p_barrier<<<1, 32, 0, s>>>(t);
            }
            if (bounces > 0 && r1 > r0) k_axpy_p2p<<<blocks_for(r1 - r0, 256), 256, 0, s>>>(t, bounces & 1, total_dev, r0, r1);
            cudaGraph_t graph = nullptr;
            cudaError_t e = cudaStreamEndCapture(s, &graph);
            avr_status st = AVR_OK;
            if (e != cudaSuccess) st = set_cuda_err(ctx, e, "cudaStreamEndCapture");
            if (st == AVR_OK) {
                e = cudaGraphInstantiate(&cm->exec, graph, 0);
                if (e != cudaSuccess) st = set_cuda_err(ctx, e, "cudaGraphInstantiate");
            }
            if (graph) cudaGraphDestroy(graph);
            if (st != AVR_OK) { cm->exec = nullptr; return st; }
            std::memcpy(cm->exec_key, key, sizeof(key));
        }
        AVR_CUDA(ctx, cudaGraphLaunch(cm->exec, s));
        if (iters_done) {
            int32_t sync[SYNC_WORDS];
            AVR_CUDA(ctx, cudaMemcpyAsync(sync, cm->p2p_base[cm->rank] + 2 * cm->p2p_padded * sizeof(float), sizeof(sync),
                                          cudaMemcpyDeviceToHost, s));
            AVR_CUDA(ctx, cudaStreamSynchronize(s));
            if (sync[SYNC_TIMEOUT])
                return set_err(ctx, AVR_ERR_NCCL, "avr_graph_propagate_multi: a peer GPU did not reach the exchange barrier within 10 s");
            int first_bad = 0x7fffffff;
            for (int p = 0; p < world; ++p) first_bad = std::min(first_bad, sync[SYNC_BAD + p]);
            *iters_done = first_bad == 0x7fffffff ? bounces : first_bad;
        }
        return AVR_OK;
    }
    if (cm->gather_floats < padded) {
        if (cm->exec) { cudaGraphExecDestroy(cm->exec); cm->exec = nullptr; }
        cudaFree(cm->d_gather[0]); cudaFree(cm->d_gather[1]);
        cm->d_gather[0] = cm->d_gather[1] = nullptr;
        cm->gather_floats = 0;
        AVR_CUDA(ctx, cudaMalloc(&cm->d_gather[0], padded * 4));
        AVR_CUDA(ctx, cudaMalloc(&cm->d_gather[1], padded * 4));
        cm->gather_floats = padded;
    }
    if (!cm->d_flags) AVR_CUDA(ctx, cudaMalloc(&cm->d_flags, 2 * sizeof(int32_t)));
    // one captured graph per (CSR, vectors, bounce count); anything else re-captures
    const void *key[5] = {g->d_val, L0_dev, total_dev, (const void *)(intptr_t)bounces, (const void *)s};
    if (!cm->exec || std::memcmp(key, cm->exec_key, sizeof(key)) != 0) {
        if (cm->exec) { cudaGraphExecDestroy(cm->exec); cm->exec = nullptr; }
        AVR_CUDA(ctx, cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        avr_status st = AVR_OK;
        ncclResult_t nr = ncclSuccess;
        float *cur = cm->d_gather[0], *nxt = cm->d_gather[1];
        k_multi_init<<<blocks_for((int64_t)padded, 256), 256, 0, s>>>(L0_dev, cur, total_dev, cm->d_flags, n, (int)padded, r0, r1);
        for (int it = 0; it < bounces && nr == ncclSuccess; ++it) {
            // rows [r0, r1) of T * cur into this rank's block of nxt; fused: total += cur for the previous bounce
            if (r1 > r0)
                k_spmv<<<blocks_for(r1 - r0, kRowsPerCta), kRowsPerCta, 0, s>>>(g->d_rowptr, g->d_col, g->d_val, cur, nxt, total_dev,
                                                                                r0, r1, cm->d_flags, it, it > 0 ? 1 : 0);
            if (world > 1)
                nr = cm->api->AllGather(nxt + size_t(cm->rank) * chunk, nxt, (size_t)chunk, ncclFloat, cm->comm, s);
            // the verdict on bounce `it`, from the whole vector, identical on every rank
            k_flag_nonfinite<<<std::min(blocks_for(n, 256), 1184), 256, 0, s>>>(nxt, n, cm->d_flags, it);
            std::swap(cur, nxt);
        }
        if (bounces > 0 && r1 > r0) k_axpy<<<blocks_for(r1 - r0, 256), 256, 0, s>>>(cur, total_dev, r0, r1, cm->d_flags);
        cudaGraph_t graph = nullptr;
        cudaError_t e = cudaStreamEndCapture(s, &graph);
        if (nr != ncclSuccess) st = set_err(ctx, AVR_ERR_NCCL, std::string("ncclAllGather: ") + cm->api->GetErrorString(nr));
        else if (e != cudaSuccess) st = set_cuda_err(ctx, e, "cudaStreamEndCapture");
        if (st == AVR_OK) {
            e = cudaGraphInstantiate(&cm->exec, graph, 0);
            if (e != cudaSuccess) st = set_cuda_err(ctx, e, "cudaGraphInstantiate");
        }
        if (graph) cudaGraphDestroy(graph);
        if (st != AVR_OK) { cm->exec = nullptr; return st; }
        std::memcpy(cm->exec_key, key, sizeof(key));
    }
    AVR_CUDA(ctx, cudaGraphLaunch(cm->exec, s));
    if (iters_done) {
        int32_t fl[2];
        AVR_CUDA(ctx, cudaMemcpyAsync(fl, cm->d_flags, sizeof(fl), cudaMemcpyDeviceToHost, s));
        AVR_CUDA(ctx, cudaStreamSynchronize(s));
        *iters_done = fl[0] ? fl[1] : bounces;
    }
    return AVR_OK;
}
