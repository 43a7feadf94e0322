// oracle/restate_medium.hpp — TEST INFRASTRUCTURE ONLY (see restate.hpp header).
// Grid media, majorant grid, DDA iterator and SampleT_maj, restated from src/pbrt/media.h,
// src/pbrt/media.cpp, src/pbrt/util/containers.h and src/pbrt/util/transform.{h,cpp}.
#pragma once
#include "restate.hpp"

namespace avro {

// ---- Interval (util/math.h:818-1035, CPU branch: NextFloatUp/Down of the rounded result) -----
struct Ivl {
    float lo, hi;
    float width() const { return hi - lo; }
    float mid() const { return (lo + hi) / 2; }
};
inline Ivl ivl_from_value_and_error(float v, float err) {  // math.h:828-837
    if (err == 0) return {v, v};
    return {next_down(v - err), next_up(v + err)};
}
inline Ivl ivl_add(Ivl a, Ivl b) {  // math.h:873-875, ctor orders min/max
    float l = next_down(a.lo + b.lo), h = next_up(a.hi + b.hi);
    return {std::min(l, h), std::max(l, h)};
}

struct Xform {
    float m[4][4], mInv[4][4];
};

// Transform::ApplyInverse(Point3f), util/transform.h:391-402
inline V3 apply_inverse_point(const Xform &t, V3 p) {
    float x = p.x, y = p.y, z = p.z;
    float xp = (t.mInv[0][0] * x + t.mInv[0][1] * y) + (t.mInv[0][2] * z + t.mInv[0][3]);
    float yp = (t.mInv[1][0] * x + t.mInv[1][1] * y) + (t.mInv[1][2] * z + t.mInv[1][3]);
    float zp = (t.mInv[2][0] * x + t.mInv[2][1] * y) + (t.mInv[2][2] * z + t.mInv[2][3]);
    float wp = (t.mInv[3][0] * x + t.mInv[3][1] * y) + (t.mInv[3][2] * z + t.mInv[3][3]);
    if (wp == 1) return {xp, yp, zp};
    return V3{xp, yp, zp} / wp;
}
// Transform::ApplyInverse(Vector3f), util/transform.h:404-410
inline V3 apply_inverse_vector(const Xform &t, V3 v) {
    return {t.mInv[0][0] * v.x + t.mInv[0][1] * v.y + t.mInv[0][2] * v.z,
            t.mInv[1][0] * v.x + t.mInv[1][1] * v.y + t.mInv[1][2] * v.z,
            t.mInv[2][0] * v.x + t.mInv[2][1] * v.y + t.mInv[2][2] * v.z};
}
// Transform::operator()(Point3f) / (Vector3f), util/transform.h:300-330
inline V3 apply_point(const Xform &t, V3 p) {
    float x = p.x, y = p.y, z = p.z;
    float xp = t.m[0][0] * x + t.m[0][1] * y + t.m[0][2] * z + t.m[0][3];
    float yp = t.m[1][0] * x + t.m[1][1] * y + t.m[1][2] * z + t.m[1][3];
    float zp = t.m[2][0] * x + t.m[2][1] * y + t.m[2][2] * z + t.m[2][3];
    float wp = t.m[3][0] * x + t.m[3][1] * y + t.m[3][2] * z + t.m[3][3];
    if (wp == 1) return {xp, yp, zp};
    return V3{xp, yp, zp} / wp;
}
inline V3 apply_vector(const Xform &t, V3 v) {
    return {t.m[0][0] * v.x + t.m[0][1] * v.y + t.m[0][2] * v.z,
            t.m[1][0] * v.x + t.m[1][1] * v.y + t.m[1][2] * v.z,
            t.m[2][0] * v.x + t.m[2][1] * v.y + t.m[2][2] * v.z};
}

// Shared tail of Transform::operator()(Ray) / ApplyInverse(Ray): offset the origin to the edge of
// its error box (util/transform.h:340-352, :420-433).  `a` is m (forward) or mInv (inverse);
// with_translation_err: the forward version's exact-point error bound includes |a[i][3]|
// (transform.h:148-153), the inverse version's does not (transform.cpp:274-279).
inline void xform_ray_impl(const float a[4][4], bool with_translation_err, V3 o, V3 d, float *tMax,
                           V3 *oOut, V3 *dOut) {
    float x = o.x, y = o.y, z = o.z;
    float xp = (a[0][0] * x + a[0][1] * y) + (a[0][2] * z + a[0][3]);
    float yp = (a[1][0] * x + a[1][1] * y) + (a[1][2] * z + a[1][3]);
    float zp = (a[2][0] * x + a[2][1] * y) + (a[2][2] * z + a[2][3]);
    float wp = (a[3][0] * x + a[3][1] * y) + (a[3][2] * z + a[3][3]);
    float e[3];
    for (int i = 0; i < 3; ++i) {
        float s = std::abs(a[i][0] * x) + std::abs(a[i][1] * y) + std::abs(a[i][2] * z);
        if (with_translation_err) s = s + std::abs(a[i][3]);
        e[i] = gamma_n(3) * s;
    }
    Ivl ox = ivl_from_value_and_error(xp, e[0]);
    Ivl oy = ivl_from_value_and_error(yp, e[1]);
    Ivl oz = ivl_from_value_and_error(zp, e[2]);
    // wp != 1 (projective) is outside the scene class; the reference would divide intervals.
    (void)wp;
    V3 dd = {a[0][0] * d.x + a[0][1] * d.y + a[0][2] * d.z,
             a[1][0] * d.x + a[1][1] * d.y + a[1][2] * d.z,
             a[2][0] * d.x + a[2][1] * d.y + a[2][2] * d.z};
    float lengthSquared = length_squared(dd);
    if (lengthSquared > 0) {
        V3 oError = {ox.width() / 2, oy.width() / 2, oz.width() / 2};
        float dt = dot(vabs(dd), oError) / lengthSquared;
        V3 off = dd * dt;
        ox = ivl_add(ox, {off.x, off.x});
        oy = ivl_add(oy, {off.y, off.y});
        oz = ivl_add(oz, {off.z, off.z});
        if (tMax) *tMax -= dt;
    }
    *oOut = {ox.mid(), oy.mid(), oz.mid()};
    *dOut = dd;
}
inline void apply_inverse_ray(const Xform &t, V3 o, V3 d, float *tMax, V3 *oOut, V3 *dOut) {
    xform_ray_impl(t.mInv, false, o, d, tMax, oOut, dOut);
}
inline void apply_ray(const Xform &t, V3 o, V3 d, float *tMax, V3 *oOut, V3 *dOut) {
    xform_ray_impl(t.m, true, o, d, tMax, oOut, dOut);
}

struct Bounds {
    V3 pMin, pMax;
    V3 offset(V3 p) const {  // vecmath.h:1323-1332
        V3 o = p - pMin;
        if (pMax.x > pMin.x) o.x /= pMax.x - pMin.x;
        if (pMax.y > pMin.y) o.y /= pMax.y - pMin.y;
        if (pMax.z > pMin.z) o.z /= pMax.z - pMin.z;
        return o;
    }
    // vecmath.h:1547-1570
    bool intersect_p(V3 o, V3 d, float tMax, float *hitt0, float *hitt1) const {
        float t0 = 0, t1 = tMax;
        for (int i = 0; i < 3; ++i) {
            float invRayDir = 1 / d[i];
            float tNear = (pMin[i] - o[i]) * invRayDir;
            float tFar = (pMax[i] - o[i]) * invRayDir;
            if (tNear > tFar) std::swap(tNear, tFar);
            tFar *= 1 + 2 * gamma_n(3);
            t0 = tNear > t0 ? tNear : t0;
            t1 = tFar < t1 ? tFar : t1;
            if (t0 > t1) return false;
        }
        *hitt0 = t0;
        *hitt1 = t1;
        return true;
    }
};

// ---- SampledGrid<Float>, util/containers.h:765-869 -----------------------------------------------
struct Grid {
    const float *v = nullptr;
    int nx = 0, ny = 0, nz = 0;
    float at(int x, int y, int z) const {  // Lookup(Point3i), :830-835
        if (x < 0 || y < 0 || z < 0 || x >= nx || y >= ny || z >= nz) return 0.f;
        // the reference indexes with int (util/containers.h SampledGrid): identical below 2^31 voxels, and UB
        // above; 64-bit here so that the checker stays defined for the device's larger grids
        return v[(size_t(z) * ny + y) * nx + x];
    }
    float lookup(V3 p) const {  // :804-819
        float sx = p.x * nx - .5f, sy = p.y * ny - .5f, sz = p.z * nz - .5f;
        int ix = (int)std::floor(sx), iy = (int)std::floor(sy), iz = (int)std::floor(sz);
        float dx = sx - (float)ix, dy = sy - (float)iy, dz = sz - (float)iz;
        float d00 = lerp(dx, at(ix, iy, iz), at(ix + 1, iy, iz));
        float d10 = lerp(dx, at(ix, iy + 1, iz), at(ix + 1, iy + 1, iz));
        float d01 = lerp(dx, at(ix, iy, iz + 1), at(ix + 1, iy, iz + 1));
        float d11 = lerp(dx, at(ix, iy + 1, iz + 1), at(ix + 1, iy + 1, iz + 1));
        return lerp(dz, lerp(dy, d00, d10), lerp(dy, d01, d11));
    }
    // MaxValue(bounds), :838-858
    template <typename F>
    float max_value(V3 bmin, V3 bmax, F conv) const {
        float ps0[3] = {bmin.x * nx - .5f, bmin.y * ny - .5f, bmin.z * nz - .5f};
        float ps1[3] = {bmax.x * nx - .5f, bmax.y * ny - .5f, bmax.z * nz - .5f};
        int n[3] = {nx, ny, nz}, lo[3], hi[3];
        for (int i = 0; i < 3; ++i) {
            lo[i] = std::max((int)std::floor(ps0[i]), 0);
            hi[i] = std::min((int)std::floor(ps1[i]) + 1, n[i] - 1);
        }
        float mv = conv(lo[0], lo[1], lo[2]);
        for (int z = lo[2]; z <= hi[2]; ++z)
            for (int y = lo[1]; y <= hi[1]; ++y)
                for (int x = lo[0]; x <= hi[0]; ++x) mv = std::max(mv, conv(x, y, z));
        return mv;
    }
};

// ---- NanoVDB FloatGrid as NanoVDBMedium reads it (media.h:624-672, media.cpp:511-616) --------------
// Arithmetic restated from the vendored headers the reference compiles against
// (src/ext/openvdb/nanovdb/nanovdb, submodule 414bed84): Map::applyInverseMapF (NanoVDB.h:2028-2031),
// matMult (NanoVDB.h:1409-1414), Floor (util/SampleFromVoxels.h:58-65), SampleFromVoxels<Tree,1,false>
// (util/SampleFromVoxels.h:419-461).  The tree is a dense flattening of the index bounding box
// (cmd/nanovdb2pbrt.cpp:100-110); voxels outside it hold the background value 0.
struct NvGrid {
    bool present = false;
    std::vector<float> v;
    int imin[3] = {0, 0, 0}, dims[3] = {0, 0, 0};
    float inv[9], vec[3];
    double wbox[6];
    float get(int i, int j, int k) const {  // ReadAccessor::getValue
        i -= imin[0]; j -= imin[1]; k -= imin[2];
        if (i < 0 || j < 0 || k < 0 || i >= dims[0] || j >= dims[1] || k >= dims[2]) return 0.f;
        return v[(size_t(k) * dims[1] + j) * dims[0] + i];
    }
    // worldToIndexF(Vec3<float>)
    V3 world_to_index_f(V3 p) const {
        float x = p.x - vec[0], y = p.y - vec[1], z = p.z - vec[2];
        return {std::fmaf(x, inv[0], std::fmaf(y, inv[1], z * inv[2])),
                std::fmaf(x, inv[3], std::fmaf(y, inv[4], z * inv[5])),
                std::fmaf(x, inv[6], std::fmaf(y, inv[7], z * inv[8]))};
    }
    // worldToIndexF(Vec3R): the subtraction and the innermost product run in double, every fmaf
    // argument is narrowed to float (matMult(const float*, Vec3<double>) calls fmaf)
    void world_to_index_f_from_double(const double p[3], double out[3]) const {
        double x = p[0] - vec[0], y = p[1] - vec[1], z = p[2] - vec[2];
        for (int r = 0; r < 3; ++r)
            out[r] = std::fmaf(float(x), inv[3 * r], std::fmaf(float(y), inv[3 * r + 1], float(z * inv[3 * r + 2])));
    }
    float sample(V3 xyz) const {  // SampleFromVoxels<TreeT,1,false>::operator()(Vec3<float>)
        float fi = std::floor(xyz.x), fj = std::floor(xyz.y), fk = std::floor(xyz.z);
        float wx = xyz.x - fi, wy = xyz.y - fj, wz = xyz.z - fk;
        int i = int32_t(fi), j = int32_t(fj), k = int32_t(fk);
        auto lerpn = [](float a, float b, float w) { return a + w * (b - a); };
        float vy = lerpn(get(i, j, k), get(i, j, k + 1), wz);
        float vy1 = lerpn(get(i, j + 1, k), get(i, j + 1, k + 1), wz);
        float vx = lerpn(vy, vy1, wy);
        vy1 = lerpn(get(i + 1, j + 1, k), get(i + 1, j + 1, k + 1), wz);
        vy = lerpn(get(i + 1, j, k), get(i + 1, j, k + 1), wz);
        float vx1 = lerpn(vy, vy1, wy);
        return lerpn(vx, vx1, wx);
    }
};

// ---- MajorantGrid, media.h:105-133 ------------------------------------------------------------------
struct MajGrid {
    Bounds bounds;
    int res[3];
    std::vector<float> vox;
    float lookup(int x, int y, int z) const { return vox[x + res[0] * (y + res[1] * z)]; }
    void voxel_bounds(int x, int y, int z, V3 *p0, V3 *p1) const {
        *p0 = {float(x) / res[0], float(y) / res[1], float(z) / res[2]};
        *p1 = {float(x + 1) / res[0], float(y + 1) / res[1], float(z + 1) / res[2]};
    }
};

struct Segment {
    float tMin, tMax;
    Spec4 sigma_maj;
    int voxel;  // x + rx*(y + ry*z) — not part of RayMajorantSegment; exposed for the DDA contract
};

// ---- DDAMajorantIterator, media.h:136-214 -----------------------------------------------------------
struct DDA {
    Spec4 sigma_t;
    float tMin = kInf, tMax = -kInf;
    const MajGrid *grid = nullptr;
    float nextCrossingT[3], deltaT[3];
    int step[3], voxelLimit[3], voxel[3];
    DDA() = default;
    DDA(V3 o, V3 d, float tMin_, float tMax_, const MajGrid *g, Spec4 st)
        : sigma_t(st), tMin(tMin_), tMax(tMax_), grid(g) {
        V3 diag = g->bounds.pMax - g->bounds.pMin;
        V3 ro = g->bounds.offset(o);
        V3 rd = {d.x / diag.x, d.y / diag.y, d.z / diag.z};
        V3 gi = ro + rd * tMin;  // Ray::operator()(t) = o + d * t
        for (int axis = 0; axis < 3; ++axis) {
            int r = g->res[axis];
            // Clamp(float, int, int) (util/math.h:144) compares in float and casts the float to int
            float gv = gi[axis] * r;
            if (gv < 0) voxel[axis] = 0;
            else if (gv > float(r - 1)) voxel[axis] = r - 1;
            else voxel[axis] = (int)gv;
            deltaT[axis] = 1 / (std::abs(rd[axis]) * r);
            if (rd[axis] == -0.f) rd[axis] = 0.f;
            if (rd[axis] >= 0) {
                float nextVoxelPos = float(voxel[axis] + 1) / r;
                nextCrossingT[axis] = tMin + (nextVoxelPos - gi[axis]) / rd[axis];
                step[axis] = 1;
                voxelLimit[axis] = r;
            } else {
                float nextVoxelPos = float(voxel[axis]) / r;
                nextCrossingT[axis] = tMin + (nextVoxelPos - gi[axis]) / rd[axis];
                step[axis] = -1;
                voxelLimit[axis] = -1;
            }
        }
    }
    bool next(Segment *seg) {
        if (tMin >= tMax) return false;
        int bits = ((nextCrossingT[0] < nextCrossingT[1]) << 2) + ((nextCrossingT[0] < nextCrossingT[2]) << 1) +
                   ((nextCrossingT[1] < nextCrossingT[2]));
        const int cmpToAxis[8] = {2, 1, 2, 1, 2, 2, 0, 0};
        int stepAxis = cmpToAxis[bits];
        float tVoxelExit = std::min(tMax, nextCrossingT[stepAxis]);
        seg->sigma_maj = sigma_t * grid->lookup(voxel[0], voxel[1], voxel[2]);
        seg->tMin = tMin;
        seg->tMax = tVoxelExit;
        seg->voxel = voxel[0] + grid->res[0] * (voxel[1] + grid->res[1] * voxel[2]);
        tMin = tVoxelExit;
        if (nextCrossingT[stepAxis] > tMax) tMin = tMax;
        voxel[stepAxis] += step[stepAxis];
        if (voxel[stepAxis] == voxelLimit[stepAxis]) tMin = tMax;
        nextCrossingT[stepAxis] += deltaT[stepAxis];
        return true;
    }
};

// ---- spectra -------------------------------------------------------------------------------------------
struct Lambda4 {
    float lambda[4], pdf[4];
};
// util/sampling.h:163-171, util/spectrum.h:334-346 SampleVisible
inline Lambda4 sample_visible_wavelengths(float u) {
    Lambda4 s;
    for (int i = 0; i < 4; ++i) {
        float up = u + float(i) / 4;
        if (up > 1) up -= 1;
        s.lambda[i] = 538 - 138.888889f * std::atanh(0.85691062f - 1.82750197f * up);
        float l = s.lambda[i];
        s.pdf[i] = (l < 360 || l > 830) ? 0.f : 0.0039398042f / sqr(std::cosh(0.0072f * (l - 538)));
    }
    return s;
}
// DenselySampledSpectrum::Sample, util/spectrum.h:390-399
inline Spec4 sample_dense(const float *tab, const float lambda[4]) {
    Spec4 s;
    for (int i = 0; i < 4; ++i) {
        int offset = (int)std::lround(lambda[i]) - 360;
        s.v[i] = (offset < 0 || offset >= NSPEC) ? 0.f : tab[offset];
    }
    return s;
}
// util/spectrum.h:69-80 Blackbody, :500-523 BlackbodySpectrum
inline float blackbody(float lambda, float T) {
    if (T <= 0) return 0;
    const float c = 299792458.f, h = 6.62606957e-34f, kb = 1.3806488e-23f;
    float l = lambda * 1e-9f;
    float l2 = l * l;            // Pow<5>(l): n2 = Pow<2>(l) = l*l ; n2*n2*Pow<1>(l)
    float l5 = l2 * l2 * l;
    return (2 * h * c * c) / (l5 * (fast_exp((h * c) / (l * kb * T)) - 1));
}
inline Spec4 blackbody_sample(float T, const float lambda[4]) {
    float lambdaMax = 2.8977721e-3f / T;
    float norm = 1 / blackbody(lambdaMax * 1e9f, T);
    Spec4 s;
    for (int i = 0; i < 4; ++i) s.v[i] = blackbody(lambda[i], T) * norm;
    return s;
}
// util/color.h:340-360 RGBSigmoidPolynomial, util/spectrum.h:560-588 RGBUnboundedSpectrum.
// voxel = {scale, c0, c1, c2}
inline float sigmoid_s(float x) {
    if (std::isinf(x)) return x > 0 ? 1 : 0;
    return .5f + x / (2 * std::sqrt(1 + sqr(x)));
}
inline float rsp_eval(const float *vx, float lambda) {  // EvaluatePolynomial(lambda, c2, c1, c0)
    return sigmoid_s(std::fma(lambda, std::fma(lambda, vx[1], vx[2]), vx[3]));
}
inline float rgb_unbounded_eval(const float *vx, float lambda) { return vx[0] * rsp_eval(vx, lambda); }
inline float rgb_unbounded_max(const float *vx) {
    float result = std::max(rsp_eval(vx, 360), rsp_eval(vx, 830));
    float lambda = -vx[2] / (2 * vx[1]);
    if (lambda >= 360 && lambda <= 830) result = std::max(result, rsp_eval(vx, lambda));
    return vx[0] * result;
}

struct MediumProps {
    Spec4 sigma_a, sigma_s, Le;
};

}  // namespace avro
