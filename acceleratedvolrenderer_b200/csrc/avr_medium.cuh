// avr_medium.cuh — device view of a grid medium, its HBM layouts, the DDA majorant iterator and the
// SampleT_maj driver (reference: src/pbrt/media.h:105-214,265-427,741-806; util/containers.h:765-869;
// util/transform.h:391-433 + transform.cpp:263-303).
#pragma once
#include "avr_math.cuh"

namespace avr {

enum { LAYOUT_DENSE_F32 = 0, LAYOUT_OCTET_F32 = 1, LAYOUT_OCTET_F16 = 2, LAYOUT_BRICK_F16 = 3 };
enum { KIND_GRID = 0, KIND_RGB = 1, KIND_NVDB = 2 };

// Plain-old-data view handed to kernels by value (lives in the kernel parameter bank).
struct MediumView {
    int kind, layout;
    int nx, ny, nz;
    // scalar density grid (KIND_GRID)
    const float *dense;       // [nz][ny][nx] fp32 — reference layout (released after the build for LAYOUT_BRICK_F16)
    const float4 *oct32;      // 2 x float4 per interpolation cell, 4x4x4-cell bricks (or null)
    const uint4 *oct16;       // 8 x fp16 per interpolation cell (or null)
    int nbx, nby, nbz;        // bricks per axis over the (nx+1)(ny+1)(nz+1) cell lattice
    // LAYOUT_BRICK_F16: sparse pool of 8x8x8-cell bricks behind an indirection grid (see brick16_taps)
    const uint2 *brick16;     // pool: 576 x 8 bytes per non-empty brick
    const int *brick_tab;     // [sbz][sby][sbx] pool slot of a brick, -1 = every tap of the brick is 0
    int sbx, sby, sbz;
    const float *temperature; // null or [nz][ny][nx]
    const float *lescale;     // [lnz][lny][lnx] (1x1x1 when the scene gave no "Lescale")
    int lnx, lny, lnz;
    const float *sigma_a_spec, *sigma_s_spec, *le_spec;  // kNSpec each; sigma_* already * "scale"
    int emissive, has_temperature, achromatic;
    float temp_offset, temp_scale;
    // RGB grids (KIND_RGB): {scale,c0,c1,c2} per voxel
    const float4 *rgb_a, *rgb_s, *rgb_le;
    const float *le_illum;
    float sigma_scale, le_scale;
    // common
    float g;
    float bmin[3], bmax[3];
    float m[12], minv[12];    // renderFromMedium / mediumFromRender, rows 0..2 (affine; row 3 = 0 0 0 1)
    const float *maj;         // majorant grid, x fastest
    int mres[3];
    // FAST render kernels: the same grid with every EMPTY cell (majorant 0) holding -d, d >= 1 the Chebyshev
    // distance in cells to the nearest non-empty cell (cells closer than d are all empty: the walk leaps to
    // the face of that cube, DDA::leap); non-empty cells hold their majorant
    const float *maj_fast;
    // The FAST grid may also be COARSER than the reference's (any grid of upper bounds of the density is a valid
    // majorant: the estimator stays unbiased, only the split between DDA steps and null collisions moves): cell
    // (i, j, k) of the FAST grid is the max over a c x c x c block of reference cells.  Its own resolution, cell
    // size and non-empty box:
    int fmres[3];
    float fcell[3], finv_cell[3];
    int fnz_lo[3], fnz_hi[3];
    float ach_sig_a, ach_sig_s;  // achromatic media: the constant sigma_a / sigma_s (already * "scale")
    // FAST path: sample coordinates straight from render space, s = gfr * p + gfr[.][3]  (s = p_grid*n - .5)
    float gfr[12];
    float cell[3], inv_cell[3];  // majorant cell size in medium space and its reciprocal
    int nz_lo[3], nz_hi[3];      // bounding box (inclusive voxel range) of the majorant cells != 0
    // KIND_NVDB (NanoVDBMedium): `dense`/octets hold the density grid flattened over its index box
    // (nx,ny,nz = box dims); index-space sampling, see nv_sample.  Temperature grid: own box and map.
    int nv_imin[3];
    float nv_inv[9], nv_vec[3];  // Map::mInvMatF / mVecF of the density grid
    int nv_tdim[3], nv_timin[3];
    float nv_tinv[9], nv_tvec[3];
};

// ---- Transform::ApplyInverse(Point3f) + Bounds3::Offset: render point -> grid [0,1]^3 ------------
AVR_DEV V3 grid_point_exact(const MediumView &mv, V3 p) {
    const float *a = mv.minv;
    float xp = (a[0] * p.x + a[1] * p.y) + (a[2] * p.z + a[3]);
    float yp = (a[4] * p.x + a[5] * p.y) + (a[6] * p.z + a[7]);
    float zp = (a[8] * p.x + a[9] * p.y) + (a[10] * p.z + a[11]);
    // wp == 1 exactly for an affine transform: (0*x + 0*y) + (0*z + 1)
    V3 o = v3(xp - mv.bmin[0], yp - mv.bmin[1], zp - mv.bmin[2]);
    if (mv.bmax[0] > mv.bmin[0]) o.x /= mv.bmax[0] - mv.bmin[0];
    if (mv.bmax[1] > mv.bmin[1]) o.y /= mv.bmax[1] - mv.bmin[1];
    if (mv.bmax[2] > mv.bmin[2]) o.z /= mv.bmax[2] - mv.bmin[2];
    return o;
}

// ---- SampledGrid<Float>::Lookup, util/containers.h:804-835 ----------------------------------------
AVR_DEV float dense_at(const float *v, int nx, int ny, int nz, int x, int y, int z) {
    if ((unsigned)x >= (unsigned)nx || (unsigned)y >= (unsigned)ny || (unsigned)z >= (unsigned)nz) return 0.f;
    return __ldg(v + (size_t(z) * ny + y) * nx + x);
}
template <bool EXACT>
AVR_DEV float trilerp8(float dx, float dy, float dz, float v000, float v100, float v010, float v110, float v001,
                       float v101, float v011, float v111) {
    float d00 = lerpf<EXACT>(dx, v000, v100);
    float d10 = lerpf<EXACT>(dx, v010, v110);
    float d01 = lerpf<EXACT>(dx, v001, v101);
    float d11 = lerpf<EXACT>(dx, v011, v111);
    return lerpf<EXACT>(dz, lerpf<EXACT>(dy, d00, d10), lerpf<EXACT>(dy, d01, d11));
}
// s = p*n - .5 already formed
template <bool EXACT>
AVR_DEV float lookup_dense_s(const float *v, int nx, int ny, int nz, float sx, float sy, float sz) {
    float fx = floorf(sx), fy = floorf(sy), fz = floorf(sz);
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    float dx = sx - fx, dy = sy - fy, dz = sz - fz;
    return trilerp8<EXACT>(dx, dy, dz, dense_at(v, nx, ny, nz, ix, iy, iz), dense_at(v, nx, ny, nz, ix + 1, iy, iz),
                           dense_at(v, nx, ny, nz, ix, iy + 1, iz), dense_at(v, nx, ny, nz, ix + 1, iy + 1, iz),
                           dense_at(v, nx, ny, nz, ix, iy, iz + 1), dense_at(v, nx, ny, nz, ix + 1, iy, iz + 1),
                           dense_at(v, nx, ny, nz, ix, iy + 1, iz + 1), dense_at(v, nx, ny, nz, ix + 1, iy + 1, iz + 1));
}
template <bool EXACT>
AVR_DEV float lookup_dense(const float *v, int nx, int ny, int nz, V3 p) {
    return lookup_dense_s<EXACT>(v, nx, ny, nz, p.x * nx - .5f, p.y * ny - .5f, p.z * nz - .5f);
}

// Octet layouts: cell (ix,iy,iz), ix in [-1, nx-1], stores its 8 corner taps contiguously
// {v000,v100,v010,v110 | v001,v101,v011,v111} (out-of-range taps baked as 0); cells are grouped
// in 4x4x4 bricks, x fastest inside the brick, so one fetch = one 32-byte sector (fp32) or one
// 16-byte vector (fp16) and the 4 x-neighbours of a cell share a 128-byte line.
// 32-bit cell index (medium creation rejects octet layouts with 2^32 or more cells)
AVR_HD uint32_t octet_index(int cx, int cy, int cz, int nbx, int nby) {
    uint32_t brick = (uint32_t(cz >> 2) * uint32_t(nby) + uint32_t(cy >> 2)) * uint32_t(nbx) + uint32_t(cx >> 2);
    return brick * 64u + uint32_t((cz & 3) * 16 + (cy & 3) * 4 + (cx & 3));
}
AVR_DEV void ldg256(const float4 *p, float4 &a, float4 &b) {
#if AVR_USE_LDG256 == 1
    // one 256-bit load; the L2::256B hint also pulls the neighbouring cells' sectors into L2
    asm volatile("ld.global.nc.L2::256B.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p));
#elif AVR_USE_LDG256 == 3
    // as 1, and the octet does not allocate in L1: a random 32-byte fetch is never reused, while the (small, hot)
    // majorant grid is served from the ~28 KB of L1 the path pool leaves
    asm volatile("ld.global.nc.L1::no_allocate.L2::256B.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p));
#elif AVR_USE_LDG256 == 2
    // one 256-bit load, no prefetch hint: DRAM traffic = the 32-byte sector of the cell
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p));
#else
    a = __ldg(p);
    b = __ldg(p + 1);
#endif
}
template <bool EXACT>
AVR_DEV float lookup_oct32_s(const MediumView &mv, float sx, float sy, float sz) {
    float fx = floorf(sx), fy = floorf(sy), fz = floorf(sz);
    int cx = (int)fx + 1, cy = (int)fy + 1, cz = (int)fz + 1;
    if ((unsigned)cx > (unsigned)mv.nx || (unsigned)cy > (unsigned)mv.ny || (unsigned)cz > (unsigned)mv.nz) return 0.f;
    float4 a, b;
    ldg256(mv.oct32 + 2 * size_t(octet_index(cx, cy, cz, mv.nbx, mv.nby)), a, b);
    return trilerp8<EXACT>(sx - fx, sy - fy, sz - fz, a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w);
}
AVR_DEV float2 h2f(uint32_t u) {
    __half2 h = *reinterpret_cast<__half2 *>(&u);
    return __half22float2(h);
}
template <bool EXACT>
AVR_DEV float lookup_oct16_s(const MediumView &mv, float sx, float sy, float sz) {
    float fx = floorf(sx), fy = floorf(sy), fz = floorf(sz);
    int cx = (int)fx + 1, cy = (int)fy + 1, cz = (int)fz + 1;
    if ((unsigned)cx > (unsigned)mv.nx || (unsigned)cy > (unsigned)mv.ny || (unsigned)cz > (unsigned)mv.nz) return 0.f;
    uint4 q = __ldg(mv.oct16 + octet_index(cx, cy, cz, mv.nbx, mv.nby));
    float2 a = h2f(q.x), b = h2f(q.y), c = h2f(q.z), d = h2f(q.w);
    return trilerp8<EXACT>(sx - fx, sy - fy, sz - fz, a.x, a.y, b.x, b.y, c.x, c.y, d.x, d.y);
}
// Sparse fp16 brick pool (LAYOUT_BRICK_F16).  The (nx+1)(ny+1)(nz+1) cell lattice is cut into 8x8x8-cell bricks;
// a brick with any non-zero tap owns 576 entries of 8 bytes in the pool, entry [ly][lx][r] (r = 0..8 fastest) =
// the four taps {v(x-1,y-1), v(x,y-1) | v(x-1,y), v(x,y)} of cell column (lx, ly) at voxel z = 8 bz - 1 + r, as
// fp16.  A trilinear fetch = one 4-byte load of the indirection grid + two ADJACENT 8-byte loads (rows r = lz,
// lz + 1); bricks are self-contained (their own apron), 2.25x the voxel count at 2 bytes = 1.1x the fp32 grid
// for a fully occupied volume, and empty bricks cost nothing.  Brick slots are 32-bit: no 2^32-cell limit.
constexpr int kBrickEntries = 8 * 8 * 9;
AVR_DEV bool brick16_taps(const MediumView &mv, int cx, int cy, int cz, float t[8]) {
    const int slot = __ldg(mv.brick_tab + (size_t(cz >> 3) * mv.sby + (cy >> 3)) * mv.sbx + (cx >> 3));
    if (slot < 0) return false;
    const uint2 *e = mv.brick16 + size_t(slot) * kBrickEntries + (((cy & 7) * 8 + (cx & 7)) * 9 + (cz & 7));
    const uint2 lo = __ldg(e), hi = __ldg(e + 1);
    float2 a = h2f(lo.x), b = h2f(lo.y), c = h2f(hi.x), d = h2f(hi.y);
    t[0] = a.x; t[1] = a.y; t[2] = b.x; t[3] = b.y; t[4] = c.x; t[5] = c.y; t[6] = d.x; t[7] = d.y;
    return true;
}
template <bool EXACT>
AVR_DEV float lookup_brick16_s(const MediumView &mv, float sx, float sy, float sz) {
    float fx = floorf(sx), fy = floorf(sy), fz = floorf(sz);
    int cx = (int)fx + 1, cy = (int)fy + 1, cz = (int)fz + 1;
    if ((unsigned)cx > (unsigned)mv.nx || (unsigned)cy > (unsigned)mv.ny || (unsigned)cz > (unsigned)mv.nz) return 0.f;
    float t[8];
    if (!brick16_taps(mv, cx, cy, cz, t)) return 0.f;
    return trilerp8<EXACT>(sx - fx, sy - fy, sz - fz, t[0], t[1], t[2], t[3], t[4], t[5], t[6], t[7]);
}
// density at grid-normalised p, any layout (runtime switch; the hot kernels are templated on LAYOUT)
template <bool EXACT, int LAYOUT>
AVR_DEV float lookup_density_s(const MediumView &mv, float sx, float sy, float sz) {
    if (LAYOUT == LAYOUT_OCTET_F32) return lookup_oct32_s<EXACT>(mv, sx, sy, sz);
    if (LAYOUT == LAYOUT_OCTET_F16) return lookup_oct16_s<EXACT>(mv, sx, sy, sz);
    if (LAYOUT == LAYOUT_BRICK_F16) return lookup_brick16_s<EXACT>(mv, sx, sy, sz);
    return lookup_dense_s<EXACT>(mv.dense, mv.nx, mv.ny, mv.nz, sx, sy, sz);
}
template <bool EXACT>
AVR_DEV float lookup_density_rt(const MediumView &mv, V3 p) {
    float sx = p.x * mv.nx - .5f, sy = p.y * mv.ny - .5f, sz = p.z * mv.nz - .5f;
    if (mv.layout == LAYOUT_OCTET_F32) return lookup_oct32_s<EXACT>(mv, sx, sy, sz);
    if (mv.layout == LAYOUT_OCTET_F16) return lookup_oct16_s<EXACT>(mv, sx, sy, sz);
    if (mv.layout == LAYOUT_BRICK_F16) return lookup_brick16_s<EXACT>(mv, sx, sy, sz);
    return lookup_dense_s<EXACT>(mv.dense, mv.nx, mv.ny, mv.nz, sx, sy, sz);
}

// ---- NanoVDBMedium sampling (media.h:624-638): worldToIndexF + SampleFromVoxels<Tree,1,false> ------
// NanoVDB.h:2028-2031 applyInverseMapF, :1409-1414 matMult (fmaf chain, innermost term a plain product)
AVR_DEV V3 nv_world_to_index(const float *inv, const float *vec, V3 p) {
    float x = p.x - vec[0], y = p.y - vec[1], z = p.z - vec[2];
    return v3(__fmaf_rn(x, inv[0], __fmaf_rn(y, inv[1], __fmul_rn(z, inv[2]))),
              __fmaf_rn(x, inv[3], __fmaf_rn(y, inv[4], __fmul_rn(z, inv[5]))),
              __fmaf_rn(x, inv[6], __fmaf_rn(y, inv[7], __fmul_rn(z, inv[8]))));
}
// util/SampleFromVoxels.h:419-461: lerp(a,b,w) = a + w*(b-a), along z, then y, then x.
// Tap naming as trilerp8: v<dx><dy><dz>.
AVR_DEV float nv_lerp(float a, float b, float w) { return __fadd_rn(a, __fmul_rn(w, __fsub_rn(b, a))); }
AVR_DEV float nv_trilerp8(float wx, float wy, float wz, float v000, float v100, float v010, float v110, float v001,
                          float v101, float v011, float v111) {
    float vx = nv_lerp(nv_lerp(v000, v001, wz), nv_lerp(v010, v011, wz), wy);
    float vx1 = nv_lerp(nv_lerp(v100, v101, wz), nv_lerp(v110, v111, wz), wy);
    return nv_lerp(vx, vx1, wx);
}
// xyz in index space; grid flattened over [imin, imin+n-1], background 0 (dense layout)
AVR_DEV float nv_sample_dense(const float *v, int nx, int ny, int nz, const int imin[3], V3 xyz) {
    float fx = floorf(xyz.x), fy = floorf(xyz.y), fz = floorf(xyz.z);
    int ix = (int)fx - imin[0], iy = (int)fy - imin[1], iz = (int)fz - imin[2];
    return nv_trilerp8(xyz.x - fx, xyz.y - fy, xyz.z - fz, dense_at(v, nx, ny, nz, ix, iy, iz),
                       dense_at(v, nx, ny, nz, ix + 1, iy, iz), dense_at(v, nx, ny, nz, ix, iy + 1, iz),
                       dense_at(v, nx, ny, nz, ix + 1, iy + 1, iz), dense_at(v, nx, ny, nz, ix, iy, iz + 1),
                       dense_at(v, nx, ny, nz, ix + 1, iy, iz + 1), dense_at(v, nx, ny, nz, ix, iy + 1, iz + 1),
                       dense_at(v, nx, ny, nz, ix + 1, iy + 1, iz + 1));
}
// density grid of a KIND_NVDB medium, any layout (the octet cell (c) holds taps c-1, c: same store as
// the SampledGrid convention, addressed by floor(index) - imin + 1)
template <int LAYOUT>
AVR_DEV float nv_density(const MediumView &mv, V3 xyz) {
    if (LAYOUT == LAYOUT_DENSE_F32) return nv_sample_dense(mv.dense, mv.nx, mv.ny, mv.nz, mv.nv_imin, xyz);
    float fx = floorf(xyz.x), fy = floorf(xyz.y), fz = floorf(xyz.z);
    int cx = (int)fx - mv.nv_imin[0] + 1, cy = (int)fy - mv.nv_imin[1] + 1, cz = (int)fz - mv.nv_imin[2] + 1;
    if ((unsigned)cx > (unsigned)mv.nx || (unsigned)cy > (unsigned)mv.ny || (unsigned)cz > (unsigned)mv.nz) return 0.f;
    float wx = xyz.x - fx, wy = xyz.y - fy, wz = xyz.z - fz;
    if (LAYOUT == LAYOUT_OCTET_F32) {
        float4 a, b;
        ldg256(mv.oct32 + 2 * size_t(octet_index(cx, cy, cz, mv.nbx, mv.nby)), a, b);
        return nv_trilerp8(wx, wy, wz, a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w);
    }
    if (LAYOUT == LAYOUT_BRICK_F16) {
        float t[8];
        if (!brick16_taps(mv, cx, cy, cz, t)) return 0.f;
        return nv_trilerp8(wx, wy, wz, t[0], t[1], t[2], t[3], t[4], t[5], t[6], t[7]);
    }
    uint4 q = __ldg(mv.oct16 + octet_index(cx, cy, cz, mv.nbx, mv.nby));
    float2 a = h2f(q.x), b = h2f(q.y), c = h2f(q.z), d = h2f(q.w);
    return nv_trilerp8(wx, wy, wz, a.x, a.y, b.x, b.y, c.x, c.y, d.x, d.y);
}
AVR_DEV float nv_density_rt(const MediumView &mv, V3 xyz) {
    if (mv.layout == LAYOUT_OCTET_F32) return nv_density<LAYOUT_OCTET_F32>(mv, xyz);
    if (mv.layout == LAYOUT_OCTET_F16) return nv_density<LAYOUT_OCTET_F16>(mv, xyz);
    if (mv.layout == LAYOUT_BRICK_F16) return nv_density<LAYOUT_BRICK_F16>(mv, xyz);
    return nv_density<LAYOUT_DENSE_F32>(mv, xyz);
}
// Transform::ApplyInverse(Point3f) alone (render -> medium space), util/transform.h:391-402
AVR_DEV V3 medium_point_exact(const MediumView &mv, V3 p) {
    const float *a = mv.minv;
    return v3((a[0] * p.x + a[1] * p.y) + (a[2] * p.z + a[3]), (a[4] * p.x + a[5] * p.y) + (a[6] * p.z + a[7]),
              (a[8] * p.x + a[9] * p.y) + (a[10] * p.z + a[11]));
}

// ---- general SamplePoint: GridMedium media.h:287-319, RGBGridMedium media.h:377-401 -------------
struct Props4 {
    Spec<4> sigma_a, sigma_s, Le;
};
AVR_DEV Spec<4> sample_dense4(const float *tab, const float lambda[4]) {
    Spec<4> s;
#pragma unroll
    for (int i = 0; i < 4; ++i) s.v[i] = sample_dense1(tab, lambda[i]);
    return s;
}
AVR_DEV Spec<4> rgb_tap(const float4 *gr, const MediumView &mv, int x, int y, int z, const float lambda[4],
                        const float *illum) {
    Spec<4> s = sconst<4>(0.f);
    bool oob = (unsigned)x >= (unsigned)mv.nx || (unsigned)y >= (unsigned)mv.ny || (unsigned)z >= (unsigned)mv.nz;
    if (oob) {
        // default-constructed RGBUnboundedSpectrum: scale 0, rsp(0,0,0) -> 0 * s(0) = 0 ; illuminant: 0
        return s;
    }
    float4 v = __ldg(gr + (size_t(z) * mv.ny + y) * mv.nx + x);
#pragma unroll
    for (int i = 0; i < 4; ++i) s.v[i] = v.x * rsp_eval(v.y, v.z, v.w, lambda[i]);
    if (illum) s = s * sample_dense4(illum, lambda);
    return s;
}
AVR_DEV Spec<4> lerp4(float t, Spec<4> a, Spec<4> b) { return a * (1 - t) + b * t; }
AVR_DEV Spec<4> rgb_lookup(const float4 *gr, const MediumView &mv, V3 p, const float lambda[4], const float *illum) {
    float sx = p.x * mv.nx - .5f, sy = p.y * mv.ny - .5f, sz = p.z * mv.nz - .5f;
    float fx = floorf(sx), fy = floorf(sy), fz = floorf(sz);
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    float dx = sx - fx, dy = sy - fy, dz = sz - fz;
    Spec<4> d00 = lerp4(dx, rgb_tap(gr, mv, ix, iy, iz, lambda, illum), rgb_tap(gr, mv, ix + 1, iy, iz, lambda, illum));
    Spec<4> d10 = lerp4(dx, rgb_tap(gr, mv, ix, iy + 1, iz, lambda, illum), rgb_tap(gr, mv, ix + 1, iy + 1, iz, lambda, illum));
    Spec<4> d01 = lerp4(dx, rgb_tap(gr, mv, ix, iy, iz + 1, lambda, illum), rgb_tap(gr, mv, ix + 1, iy, iz + 1, lambda, illum));
    Spec<4> d11 = lerp4(dx, rgb_tap(gr, mv, ix, iy + 1, iz + 1, lambda, illum), rgb_tap(gr, mv, ix + 1, iy + 1, iz + 1, lambda, illum));
    return lerp4(dz, lerp4(dy, d00, d10), lerp4(dy, d01, d11));
}
// FAST RGB lookup at sample coordinates (s = p * n - .5): the same taps and trilinear weights, the sigmoid
// through rsp_eval_fast, products and lerps as FMAs.  Statistically pinned against the exact lookup.
AVR_DEV Spec<4> rgb_tap_fast(const float4 *gr, const MediumView &mv, int x, int y, int z, const float lambda[4]) {
    Spec<4> s = sconst<4>(0.f);
    if ((unsigned)x >= (unsigned)mv.nx || (unsigned)y >= (unsigned)mv.ny || (unsigned)z >= (unsigned)mv.nz) return s;
    float4 v = __ldg(gr + (size_t(z) * mv.ny + y) * mv.nx + x);
#pragma unroll
    for (int i = 0; i < 4; ++i) s.v[i] = v.x * rsp_eval_fast(v.y, v.z, v.w, lambda[i]);
    return s;
}
AVR_DEV Spec<4> lerp4_fast(float t, Spec<4> a, Spec<4> b) {
    Spec<4> r;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.v[i] = __fmaf_rn(t, b.v[i] - a.v[i], a.v[i]);
    return r;
}
AVR_DEV Spec<4> rgb_lookup_fast(const float4 *gr, const MediumView &mv, float sx, float sy, float sz, const float lambda[4]) {
    float fx = floorf(sx), fy = floorf(sy), fz = floorf(sz);
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    float dx = sx - fx, dy = sy - fy, dz = sz - fz;
    Spec<4> acc[2];
#pragma unroll 1
    for (int k = 0; k < 2; ++k) {  // one z slice at a time: half the live taps
        Spec<4> r0 = lerp4_fast(dx, rgb_tap_fast(gr, mv, ix, iy, iz + k, lambda), rgb_tap_fast(gr, mv, ix + 1, iy, iz + k, lambda));
        Spec<4> r1 = lerp4_fast(dx, rgb_tap_fast(gr, mv, ix, iy + 1, iz + k, lambda), rgb_tap_fast(gr, mv, ix + 1, iy + 1, iz + k, lambda));
        acc[k] = lerp4_fast(dy, r0, r1);
    }
    return lerp4_fast(dz, acc[0], acc[1]);
}
// Emission part of GridMedium::SamplePoint (media.h:299-316); p grid-normalised
AVR_DEV Spec<4> grid_emission(const MediumView &mv, V3 p, const float lambda[4]) {
    Spec<4> Le = sconst<4>(0.f);
    float scale = lookup_dense<true>(mv.lescale, mv.lnx, mv.lny, mv.lnz, p);
    if (scale > 0) {
        if (mv.has_temperature) {
            float temp = lookup_dense<true>(mv.temperature, mv.nx, mv.ny, mv.nz, p);
            temp = (temp - mv.temp_offset) * mv.temp_scale;
            if (temp > 100.f) {
                float lambdaMax = 2.8977721e-3f / temp;
                float norm = 1 / blackbody(lambdaMax * 1e9f, temp);
#pragma unroll
                for (int i = 0; i < 4; ++i) Le.v[i] = blackbody(lambda[i], temp) * norm;
                Le = Le * scale;
            }
        } else
            Le = sample_dense4(mv.le_spec, lambda) * scale;
    }
    return Le;
}
// Full SamplePoint, exact arithmetic, any medium kind/layout (used by the general kernels and the
// avr_medium_sample_point entry point).  sig_a/sig_s: the medium's spectra sampled at lambda.
AVR_DEV Props4 sample_point_general(const MediumView &mv, V3 p_render, const float lambda[4], Spec<4> sig_a,
                                    Spec<4> sig_s) {
    Props4 mp;
    mp.Le = sconst<4>(0.f);
    if (mv.kind == KIND_NVDB) {  // NanoVDBMedium::SamplePoint + Le, media.h:624-638, :661-672
        V3 pm = medium_point_exact(mv, p_render);
        float d = nv_density_rt(mv, nv_world_to_index(mv.nv_inv, mv.nv_vec, pm));
        mp.sigma_a = sig_a * d;
        mp.sigma_s = sig_s * d;
        if (mv.has_temperature) {
            float temp = nv_sample_dense(mv.temperature, mv.nv_tdim[0], mv.nv_tdim[1], mv.nv_tdim[2], mv.nv_timin,
                                         nv_world_to_index(mv.nv_tinv, mv.nv_tvec, pm));
            temp = (temp - mv.temp_offset) * mv.temp_scale;
            if (!(temp <= 100.f)) {
                float lambdaMax = 2.8977721e-3f / temp;
                float norm = 1 / blackbody(lambdaMax * 1e9f, temp);
#pragma unroll
                for (int i = 0; i < 4; ++i) mp.Le.v[i] = blackbody(lambda[i], temp) * norm;
                mp.Le = mp.Le * mv.le_scale;
            }
        }
        return mp;
    }
    V3 p = grid_point_exact(mv, p_render);
    if (mv.kind == KIND_RGB) {
        mp.sigma_a = (mv.rgb_a ? rgb_lookup(mv.rgb_a, mv, p, lambda, nullptr) : sconst<4>(1.f)) * mv.sigma_scale;
        mp.sigma_s = (mv.rgb_s ? rgb_lookup(mv.rgb_s, mv, p, lambda, nullptr) : sconst<4>(1.f)) * mv.sigma_scale;
        if (mv.rgb_le && mv.le_scale > 0) mp.Le = rgb_lookup(mv.rgb_le, mv, p, lambda, mv.le_illum) * mv.le_scale;
        return mp;
    }
    float d = lookup_density_rt<true>(mv, p);
    mp.sigma_a = sig_a * d;
    mp.sigma_s = sig_s * d;
    if (mv.emissive) mp.Le = grid_emission(mv, p, lambda);
    return mp;
}

// FAST SamplePoint of the general kernels for KIND_GRID / KIND_RGB media, at sample coordinates s = p * n - .5
// (the folded render -> sample-space map of the FAST Fetcher): same estimator inputs, fast arithmetic.
AVR_DEV Props4 sample_point_general_fast(const MediumView &mv, float sx, float sy, float sz, const float lambda[4],
                                         Spec<4> sig_a, Spec<4> sig_s) {
    Props4 mp;
    mp.Le = sconst<4>(0.f);
    if (mv.kind == KIND_RGB) {
        mp.sigma_a = (mv.rgb_a ? rgb_lookup_fast(mv.rgb_a, mv, sx, sy, sz, lambda) : sconst<4>(1.f)) * mv.sigma_scale;
        mp.sigma_s = (mv.rgb_s ? rgb_lookup_fast(mv.rgb_s, mv, sx, sy, sz, lambda) : sconst<4>(1.f)) * mv.sigma_scale;
        if (mv.rgb_le && mv.le_scale > 0)
            mp.Le = rgb_lookup_fast(mv.rgb_le, mv, sx, sy, sz, lambda) * sample_dense4(mv.le_illum, lambda) * mv.le_scale;
        return mp;
    }
    float d;
    if (mv.layout == LAYOUT_OCTET_F32) d = lookup_oct32_s<false>(mv, sx, sy, sz);
    else if (mv.layout == LAYOUT_OCTET_F16) d = lookup_oct16_s<false>(mv, sx, sy, sz);
    else if (mv.layout == LAYOUT_BRICK_F16) d = lookup_brick16_s<false>(mv, sx, sy, sz);
    else d = lookup_dense_s<false>(mv.dense, mv.nx, mv.ny, mv.nz, sx, sy, sz);
    mp.sigma_a = sig_a * d;
    mp.sigma_s = sig_s * d;
    if (mv.emissive) {
        V3 p = v3((sx + .5f) * ftz_rcp((float)mv.nx), (sy + .5f) * ftz_rcp((float)mv.ny), (sz + .5f) * ftz_rcp((float)mv.nz));
        mp.Le = grid_emission(mv, p, lambda);
    }
    return mp;
}

// ---- medium-space ray: Transform::ApplyInverse(Ray,&tMax) + Bounds3::IntersectP ------------------
struct MRay {
    V3 o, d;
    float tMin, tMax;
    bool hit;
};
AVR_DEV void ivl_from(float v, float err, float &lo, float &hi) {  // Interval::FromValueAndError
    if (err == 0) { lo = hi = v; return; }
    lo = next_down(v - err);
    hi = next_up(v + err);
}
AVR_DEV void ivl_add_scalar(float &lo, float &hi, float f) {  // Interval += Float (math.h:873,921)
    float l = next_down(lo + f), h = next_up(hi + f);
    lo = fminf(l, h);
    hi = fmaxf(l, h);
}
// a: 3x4 affine rows of the matrix being applied; trans_err: forward (operator()) includes |a[i][3]|
AVR_DEV void xform_ray_exact(const float *a, bool trans_err, V3 o, V3 d, float *tMax, V3 *oo, V3 *od) {
    float xp = (a[0] * o.x + a[1] * o.y) + (a[2] * o.z + a[3]);
    float yp = (a[4] * o.x + a[5] * o.y) + (a[6] * o.z + a[7]);
    float zp = (a[8] * o.x + a[9] * o.y) + (a[10] * o.z + a[11]);
    float e[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float s = fabsf(a[4 * i] * o.x) + fabsf(a[4 * i + 1] * o.y) + fabsf(a[4 * i + 2] * o.z);
        if (trans_err) s = s + fabsf(a[4 * i + 3]);
        e[i] = kGamma3 * s;
    }
    float xl, xh, yl, yh, zl, zh;
    ivl_from(xp, e[0], xl, xh);
    ivl_from(yp, e[1], yl, yh);
    ivl_from(zp, e[2], zl, zh);
    V3 dd = v3(a[0] * d.x + a[1] * d.y + a[2] * d.z, a[4] * d.x + a[5] * d.y + a[6] * d.z,
               a[8] * d.x + a[9] * d.y + a[10] * d.z);
    float lengthSquared = sqr(dd.x) + sqr(dd.y) + sqr(dd.z);
    if (lengthSquared > 0) {
        V3 oErr = v3((xh - xl) / 2, (yh - yl) / 2, (zh - zl) / 2);
        float dt = dot(v3(fabsf(dd.x), fabsf(dd.y), fabsf(dd.z)), oErr) / lengthSquared;
        V3 off = dd * dt;
        ivl_add_scalar(xl, xh, off.x);
        ivl_add_scalar(yl, yh, off.y);
        ivl_add_scalar(zl, zh, off.z);
        if (tMax) *tMax -= dt;
    }
    *oo = v3((xl + xh) / 2, (yl + yh) / 2, (zl + zh) / 2);
    *od = dd;
}
AVR_DEV bool slab_test(const float bmin[3], const float bmax[3], V3 o, V3 d, float tMax, float *h0, float *h1) {
    float t0 = 0, t1 = tMax;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float inv = 1 / comp(d, i);
        float tNear = (bmin[i] - comp(o, i)) * inv;
        float tFar = (bmax[i] - comp(o, i)) * inv;
        if (tNear > tFar) { float t = tNear; tNear = tFar; tFar = t; }
        tFar *= 1 + 2 * kGamma3;
        t0 = tNear > t0 ? tNear : t0;
        t1 = tFar < t1 ? tFar : t1;
        if (t0 > t1) return false;
    }
    *h0 = t0;
    *h1 = t1;
    return true;
}
AVR_DEV MRay ray_to_medium(const MediumView &mv, V3 o, V3 d, float raytMax) {
    MRay r;
    xform_ray_exact(mv.minv, false, o, d, &raytMax, &r.o, &r.d);
    r.tMin = r.tMax = 0;
    r.hit = slab_test(mv.bmin, mv.bmax, r.o, r.d, raytMax, &r.tMin, &r.tMax);
    return r;
}

// ---- DDAMajorantIterator, media.h:136-214 (bit-exact contract) --------------------------------------
// State kept per ray: the reference's tMin/tMax/nextCrossingT/deltaT, plus the voxel as a linear index
// and, per axis, the number of steps left before the walk leaves the grid (10-bit fields of `rem`).
// The step direction of an axis is the sign of dTs (|dTs| = deltaT).  With TRIM the walk also ends
// when it leaves the bounding box of the non-zero majorant cells (mv.nz_lo/nz_hi): every voxel beyond
// has sigma_maj = 0, whose segments multiply T_maj by FastExp(-0) = 1 exactly, so results are
// unchanged and only the trailing empty segments are not enumerated.
struct DDA {
    float tMin, tMax;
    float nct[3], dTs[3];
    int lin;
    unsigned rem;

    AVR_DEV void invalidate() { tMin = __int_as_float(0x7f800000); tMax = __int_as_float(0xff800000); }
    // finish init from per-axis voxel / direction sign; returns false if no voxel can be visited
    // F: the walk runs over the FAST majorant grid (MediumView::maj_fast, fmres ...) instead of the reference's
    template <bool TRIM, bool F = false>
    AVR_DEV bool set_voxels(const MediumView &mv, const int vx[3], const bool pos[3]) {
        const int *mres = F ? mv.fmres : mv.mres, *nz_hi = F ? mv.fnz_hi : mv.nz_hi, *nz_lo = F ? mv.fnz_lo : mv.nz_lo;
        lin = vx[0] + mres[0] * (vx[1] + mres[1] * vx[2]);
        rem = 0;
        bool ok = true;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            int hi = TRIM ? nz_hi[a] + 1 : mres[a];  // leaving towards + when voxel == hi
            int lo = TRIM ? nz_lo[a] - 1 : -1;       // leaving towards - when voxel == lo
            int r = pos[a] ? hi - vx[a] : vx[a] - lo;
            ok = ok && r > 0;
            rem |= (unsigned)(r & 1023) << (10 * a);
        }
        return ok;
    }
    template <bool TRIM>
    AVR_DEV void init(const MediumView &mv, V3 o, V3 d, float tMin_, float tMax_) {
        tMin = tMin_;
        tMax = tMax_;
        V3 diag = v3(mv.bmax[0] - mv.bmin[0], mv.bmax[1] - mv.bmin[1], mv.bmax[2] - mv.bmin[2]);
        // bounds.Offset(ray.o)
        V3 ro = v3(o.x - mv.bmin[0], o.y - mv.bmin[1], o.z - mv.bmin[2]);
        if (mv.bmax[0] > mv.bmin[0]) ro.x /= diag.x;
        if (mv.bmax[1] > mv.bmin[1]) ro.y /= diag.y;
        if (mv.bmax[2] > mv.bmin[2]) ro.z /= diag.z;
        V3 rd = v3(d.x / diag.x, d.y / diag.y, d.z / diag.z);
        V3 gi = v3(ro.x + rd.x * tMin, ro.y + rd.y * tMin, ro.z + rd.z * tMin);
        int vx[3];
        bool pos[3];
#pragma unroll
        for (int axis = 0; axis < 3; ++axis) {
            int r = mv.mres[axis];
            float g = comp(gi, axis), dd = comp(rd, axis);
            float gv = g * r;
            int v;
            if (gv < 0) v = 0;
            else if (gv > float(r - 1)) v = r - 1;
            else v = (int)gv;
            vx[axis] = v;
            float dT = 1 / (fabsf(dd) * r);
            if (dd == -0.f) dd = 0.f;
            pos[axis] = dd >= 0;
            float nextVoxelPos = float(pos[axis] ? v + 1 : v) / r;
            nct[axis] = tMin + (nextVoxelPos - g) / dd;
            dTs[axis] = pos[axis] ? dT : -dT;
        }
        if (!set_voxels<TRIM>(mv, vx, pos)) invalidate();
    }
    // FAST variant: medium-space arithmetic with reciprocals (not bit-compatible with the reference)
    template <bool TRIM>
    AVR_DEV void init_fast(const MediumView &mv, V3 o, V3 inv_d, V3 d, float tMin_, float tMax_) {
        tMin = tMin_;
        tMax = tMax_;
        int vx[3];
        bool pos[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            int r = mv.fmres[a];
            float p = __fmaf_rn(comp(d, a), tMin, comp(o, a)) - mv.bmin[a];
            int v = (int)(p * mv.finv_cell[a]);
            v = v < 0 ? 0 : (v > r - 1 ? r - 1 : v);
            vx[a] = v;
            float id = comp(inv_d, a);
            pos[a] = !(comp(d, a) < 0);
            float edge = mv.bmin[a] + float(pos[a] ? v + 1 : v) * mv.fcell[a];
            nct[a] = (edge - comp(o, a)) * id;
            if (comp(d, a) == 0) nct[a] = __int_as_float(0x7f800000);
            float dT = fminf(fabsf(id) * mv.fcell[a], 1e30f);  // finite: leap() multiplies it by a cell count
            dTs[a] = pos[a] ? dT : -dT;
        }
        if (!set_voxels<TRIM, true>(mv, vx, pos)) invalidate();
    }
    // FAST variant only.  The current voxel and every voxel at Chebyshev distance < d (d >= 2) are empty
    // (sigma_maj = 0: nothing can happen there): advance straight to the first crossing that leaves that
    // cube, or to the end of the walk.  Per axis, k = the cells that may still be entered inside the cube
    // (and inside the trimmed box); the exit axis takes k + 1 crossings, the others the crossings before
    // the exit time.  No segment is emitted.
    AVR_DEV void leap(const MediumView &mv, int d) {
        int r[3] = {int(rem & 1023u), int((rem >> 10) & 1023u), int(rem >> 20)};
        float a[3], tb[3];
        int k[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            k[i] = min(d, r[i]) - 1;
            a[i] = fabsf(dTs[i]);
            tb[i] = __fmaf_rn(float(k[i]), a[i], nct[i]);
        }
        float T = fminf(fminf(tb[0], tb[1]), tb[2]);
        if (!(T < tMax)) { tMin = tMax; return; }  // the ray ends inside the cube
        bool e0 = tb[0] <= tb[1] && tb[0] <= tb[2], e1 = !e0 && tb[1] <= tb[2];
        bool ex[3] = {e0, e1, !e0 && !e1};
        int stride[3] = {1, mv.fmres[0], mv.fmres[0] * mv.fmres[1]};
        bool out = false;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            int n = k[i] + 1;
            if (!ex[i]) {
                int c = nct[i] < T ? (int)((T - nct[i]) * ftz_rcp(a[i])) + 1 : 0;
                n = min(c, k[i]);
            }
            nct[i] = __fmaf_rn(float(n), a[i], nct[i]);
            lin += (dTs[i] >= 0 ? n : -n) * stride[i];
            rem -= unsigned(n) << (10 * i);
            out = out || (ex[i] && n == r[i]);
        }
        tMin = out ? tMax : T;  // left the (trimmed) grid: every voxel beyond is empty
    }
    // Returns false when the iterator is exhausted.  *vox = linear voxel index of the segment.
    template <bool F = false>
    AVR_DEV bool next(const MediumView &mv, float *segMin, float *segMax, int *vox) {
        const int *mres = F ? mv.fmres : mv.mres;
        if (tMin >= tMax) return false;
        // stepAxis = cmpToAxis[bits], cmpToAxis = {2,1,2,1,2,2,0,0}: 0 iff c01 & c02, 1 iff !c01 & c12
        bool c01 = nct[0] < nct[1], c02 = nct[0] < nct[2], c12 = nct[1] < nct[2];
        bool p0 = c01 && c02, p1 = !c01 && c12;
        float n = p0 ? nct[0] : (p1 ? nct[1] : nct[2]);
        float ds = p0 ? dTs[0] : (p1 ? dTs[1] : dTs[2]);
        float tVoxelExit = n < tMax ? n : tMax;  // std::min(tMax, nextCrossingT[stepAxis])
        *vox = lin;
        *segMin = tMin;
        *segMax = tVoxelExit;
        tMin = tVoxelExit;
        int stride = p0 ? 1 : (p1 ? mres[0] : mres[0] * mres[1]);
        lin += ds >= 0 ? stride : -stride;
        unsigned sh = p0 ? 0u : (p1 ? 10u : 20u);
        rem -= 1u << sh;
        if (((rem >> sh) & 1023u) == 0) tMin = tMax;  // voxel[stepAxis] == voxelLimit[stepAxis]
        float nn = n + fabsf(ds);
        if (p0) nct[0] = nn;
        else if (p1) nct[1] = nn;
        else nct[2] = nn;
        return true;
    }
};

}  // namespace avr
