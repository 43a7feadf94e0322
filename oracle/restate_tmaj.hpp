// oracle/restate_tmaj.hpp — TEST INFRASTRUCTURE ONLY (see restate.hpp header).
// GridMedium / RGBGridMedium (media.h:265-427, media.cpp:212-377) and SampleT_maj (media.h:741-806).
#pragma once
#include "restate_medium.hpp"

namespace avro {

struct Medium {
    // common
    bool is_rgb = false;
    Bounds bounds;
    Xform xf;  // renderFromMedium
    float g = 0;
    MajGrid maj;
    // GridMedium
    std::vector<float> density_v, temperature_v, lescale_v;
    Grid density, temperature, lescale;
    bool has_temperature = false, emissive = false;
    std::vector<float> sigma_a_spec, sigma_s_spec, le_spec;  // sigma specs already * sigmaScale
    float temp_scale = 1, temp_offset = 0;
    // RGBGridMedium: voxels {scale,c0,c1,c2}
    std::vector<float> rgb_a, rgb_s, rgb_le, le_illum;
    int nx = 0, ny = 0, nz = 0;
    float sigma_scale = 1, le_scale = 0;
    // NanoVDBMedium
    bool is_nvdb = false;
    NvGrid nv_density, nv_temperature;

    // GridMedium ctor, media.cpp:212-247
    void init_grid(int nx_, int ny_, int nz_, const float *dens, const float *temp, const float *lesc,
                   float le_norm, const float p0[3], const float p1[3], const float m[16],
                   const float minv[16], const float *sa, const float *ss, const float *le,
                   float sigScale, float g_, float tOff, float tScale, const int mres[3]) {
        is_rgb = false;
        nx = nx_; ny = ny_; nz = nz_;
        size_t n = size_t(nx) * ny * nz;
        density_v.assign(dens, dens + n);
        density = {density_v.data(), nx, ny, nz};
        has_temperature = temp != nullptr;
        if (temp) { temperature_v.assign(temp, temp + n); temperature = {temperature_v.data(), nx, ny, nz}; }
        if (lesc) {  // media.cpp:295-306: LeScale[i] *= LeNorm
            lescale_v.resize(n);
            for (size_t i = 0; i < n; ++i) lescale_v[i] = lesc[i] * le_norm;
            lescale = {lescale_v.data(), nx, ny, nz};
        } else {
            lescale_v.assign(1, le_norm);
            lescale = {lescale_v.data(), 1, 1, 1};
        }
        bounds = {{p0[0], p0[1], p0[2]}, {p1[0], p1[1], p1[2]}};
        std::memcpy(xf.m, m, 64);
        std::memcpy(xf.mInv, minv, 64);
        sigma_a_spec.assign(sa, sa + NSPEC);
        sigma_s_spec.assign(ss, ss + NSPEC);
        for (float &v : sigma_a_spec) v *= sigScale;  // DenselySampledSpectrum::Scale
        for (float &v : sigma_s_spec) v *= sigScale;
        if (le) le_spec.assign(le, le + NSPEC); else le_spec.assign(NSPEC, 0.f);
        g = g_; temp_offset = tOff; temp_scale = tScale;
        float leMax = *std::max_element(le_spec.begin(), le_spec.end());
        emissive = has_temperature ? true : (leMax > 0);
        maj.bounds = bounds;
        for (int i = 0; i < 3; ++i) maj.res[i] = mres[i] > 0 ? mres[i] : 16;
        maj.vox.resize(size_t(maj.res[0]) * maj.res[1] * maj.res[2]);
        for (int z = 0; z < maj.res[2]; ++z)
            for (int y = 0; y < maj.res[1]; ++y)
                for (int x = 0; x < maj.res[0]; ++x) {
                    V3 b0, b1;
                    maj.voxel_bounds(x, y, z, &b0, &b1);
                    maj.vox[x + maj.res[0] * (y + maj.res[1] * z)] =
                        density.max_value(b0, b1, [&](int a, int b, int c) { return density.at(a, b, c); });
                }
    }

    // RGBGridMedium ctor, media.cpp:339-377
    void init_rgb(int nx_, int ny_, int nz_, const float *sa, const float *ss, const float *le,
                  const float *illum, float leScale, const float p0[3], const float p1[3],
                  const float m[16], const float minv[16], float sigScale, float g_, const int mres[3]) {
        is_rgb = true;
        nx = nx_; ny = ny_; nz = nz_;
        size_t n = size_t(nx) * ny * nz * 4;
        if (sa) rgb_a.assign(sa, sa + n);
        if (ss) rgb_s.assign(ss, ss + n);
        if (le) { rgb_le.assign(le, le + n); le_illum.assign(illum, illum + NSPEC); }
        le_scale = leScale; sigma_scale = sigScale; g = g_;
        bounds = {{p0[0], p0[1], p0[2]}, {p1[0], p1[1], p1[2]}};
        std::memcpy(xf.m, m, 64);
        std::memcpy(xf.mInv, minv, 64);
        emissive = !rgb_le.empty() && le_scale > 0;
        maj.bounds = bounds;
        for (int i = 0; i < 3; ++i) maj.res[i] = mres[i] > 0 ? mres[i] : 16;
        maj.vox.resize(size_t(maj.res[0]) * maj.res[1] * maj.res[2]);
        Grid shape{nullptr, nx, ny, nz};
        auto vmax = [&](const std::vector<float> &gr) {
            return [&gr, this](int x, int y, int z) {
                // Lookup(Point3i, convert): out of bounds -> convert(T{}) = default spectrum = 0
                if (x < 0 || y < 0 || z < 0 || x >= nx || y >= ny || z >= nz) return 0.f;
                return rgb_unbounded_max(&gr[4 * ((size_t(z) * ny + y) * nx + x)]);
            };
        };
        for (int z = 0; z < maj.res[2]; ++z)
            for (int y = 0; y < maj.res[1]; ++y)
                for (int x = 0; x < maj.res[0]; ++x) {
                    V3 b0, b1;
                    maj.voxel_bounds(x, y, z, &b0, &b1);
                    float maxSigma_t = (rgb_a.empty() ? 1 : shape.max_value(b0, b1, vmax(rgb_a))) +
                                       (rgb_s.empty() ? 1 : shape.max_value(b0, b1, vmax(rgb_s)));
                    maj.vox[x + maj.res[0] * (y + maj.res[1] * z)] = sigma_scale * maxSigma_t;
                }
    }

    // NanoVDBMedium ctor, media.cpp:511-616.  dg/tg carry what the reference reads from the two
    // GridHandles (flattened values, float inverse map, index and world bounding boxes).
    void init_nanovdb(const NvGrid &dg, const NvGrid &tg, const float m[16], const float minv[16], const float *sa,
                      const float *ss, float sigScale, float g_, float leScale, float tOff, float tScale,
                      const int mres[3]) {
        is_nvdb = true;
        nv_density = dg;
        nv_temperature = tg;
        std::memcpy(xf.m, m, 64);
        std::memcpy(xf.mInv, minv, 64);
        sigma_a_spec.assign(sa, sa + NSPEC);
        sigma_s_spec.assign(ss, ss + NSPEC);
        for (float &v : sigma_a_spec) v *= sigScale;
        for (float &v : sigma_s_spec) v *= sigScale;
        g = g_; le_scale = leScale; temp_offset = tOff; temp_scale = tScale;
        has_temperature = tg.present;
        emissive = tg.present && leScale > 0;  // media.h:621
        // bounds = Bounds3f(worldBBox) [double -> float], Union with the temperature grid's (:536-551)
        auto fbox = [](const double w[6]) {
            return Bounds{{float(w[0]), float(w[1]), float(w[2])}, {float(w[3]), float(w[4]), float(w[5])}};
        };
        bounds = fbox(dg.wbox);
        if (tg.present) {
            Bounds b2 = fbox(tg.wbox);
            for (int i = 0; i < 3; ++i) {
                bounds.pMin[i] = std::min(bounds.pMin[i], b2.pMin[i]);
                bounds.pMax[i] = std::max(bounds.pMax[i], b2.pMax[i]);
            }
        }
        maj.bounds = bounds;
        for (int i = 0; i < 3; ++i) maj.res[i] = mres[i] > 0 ? mres[i] : 64;
        maj.vox.resize(size_t(maj.res[0]) * maj.res[1] * maj.res[2]);
        auto blerp = [&](float tx, float ty, float tz, double out[3]) {  // Bounds3::Lerp -> Vec3R
            out[0] = lerp(tx, bounds.pMin.x, bounds.pMax.x);
            out[1] = lerp(ty, bounds.pMin.y, bounds.pMax.y);
            out[2] = lerp(tz, bounds.pMin.z, bounds.pMax.z);
        };
        for (int z = 0; z < maj.res[2]; ++z)
            for (int y = 0; y < maj.res[1]; ++y)
                for (int x = 0; x < maj.res[0]; ++x) {
                    double w0[3], w1[3], i0[3], i1[3];
                    blerp(float(x) / maj.res[0], float(y) / maj.res[1], float(z) / maj.res[2], w0);
                    blerp(float(x + 1) / maj.res[0], float(y + 1) / maj.res[1], float(z + 1) / maj.res[2], w1);
                    dg.world_to_index_f_from_double(w0, i0);
                    dg.world_to_index_f_from_double(w1, i1);
                    float delta = 1.f;
                    int lo[3], hi[3];
                    for (int a = 0; a < 3; ++a) {
                        lo[a] = std::max(int(i0[a] - delta), dg.imin[a]);
                        hi[a] = std::min(int(i1[a] + delta), dg.imin[a] + dg.dims[a] - 1);
                    }
                    float maxValue = 0;
                    for (int nz_ = lo[2]; nz_ <= hi[2]; ++nz_)
                        for (int ny_ = lo[1]; ny_ <= hi[1]; ++ny_)
                            for (int nx_ = lo[0]; nx_ <= hi[0]; ++nx_)
                                maxValue = std::max(maxValue, dg.get(nx_, ny_, nz_));
                    maj.vox[x + maj.res[0] * (y + maj.res[1] * z)] = maxValue;
                }
    }

    // SampledGrid<RGB*Spectrum>::Lookup(p, convert): convert-then-lerp, containers.h:785-801
    Spec4 rgb_lookup(const std::vector<float> &gr, V3 p, const float lambda[4], const float *illum) const {
        float sx = p.x * nx - .5f, sy = p.y * ny - .5f, sz = p.z * nz - .5f;
        int ix = (int)std::floor(sx), iy = (int)std::floor(sy), iz = (int)std::floor(sz);
        float dx = sx - (float)ix, dy = sy - (float)iy, dz = sz - (float)iz;
        auto tap = [&](int x, int y, int z) {
            Spec4 s;
            bool oob = (x < 0 || y < 0 || z < 0 || x >= nx || y >= ny || z >= nz);
            static const float zero_vx[4] = {0, 0, 0, 0};  // RGBUnboundedSpectrum(): rsp(0,0,0), scale 0
            const float *vx = oob ? zero_vx : &gr[4 * ((size_t(z) * ny + y) * nx + x)];
            if (illum) {
                // RGBIlluminantSpectrum::Sample (spectrum.h:614-621); default-constructed -> 0
                if (oob) return spec(0.f);
                for (int i = 0; i < 4; ++i) s.v[i] = vx[0] * rsp_eval(vx, lambda[i]);
                return s * sample_dense(illum, lambda);
            }
            for (int i = 0; i < 4; ++i) s.v[i] = vx[0] * rsp_eval(vx, lambda[i]);
            return s;
        };
        auto lerps = [](float t, Spec4 a, Spec4 b) { return (1 - t) * a + t * b; };
        Spec4 d00 = lerps(dx, tap(ix, iy, iz), tap(ix + 1, iy, iz));
        Spec4 d10 = lerps(dx, tap(ix, iy + 1, iz), tap(ix + 1, iy + 1, iz));
        Spec4 d01 = lerps(dx, tap(ix, iy, iz + 1), tap(ix + 1, iy, iz + 1));
        Spec4 d11 = lerps(dx, tap(ix, iy + 1, iz + 1), tap(ix + 1, iy + 1, iz + 1));
        return lerps(dz, lerps(dy, d00, d10), lerps(dy, d01, d11));
    }

    // GridMedium::SamplePoint media.h:287-319 ; RGBGridMedium::SamplePoint media.h:377-401
    MediumProps sample_point(V3 p, const float lambda[4]) const {
        MediumProps mp;
        p = apply_inverse_point(xf, p);
        mp.Le = spec(0.f);
        if (is_nvdb) {  // NanoVDBMedium::SamplePoint + Le, media.h:624-638, :661-672
            mp.sigma_a = sample_dense(sigma_a_spec.data(), lambda);
            mp.sigma_s = sample_dense(sigma_s_spec.data(), lambda);
            float d = nv_density.sample(nv_density.world_to_index_f(p));
            mp.sigma_a = mp.sigma_a * d;
            mp.sigma_s = mp.sigma_s * d;
            if (nv_temperature.present) {
                float temp = nv_temperature.sample(nv_temperature.world_to_index_f(p));
                temp = (temp - temp_offset) * temp_scale;
                if (!(temp <= 100.f)) mp.Le = le_scale * blackbody_sample(temp, lambda);
            }
            return mp;
        }
        p = bounds.offset(p);
        if (is_rgb) {
            mp.sigma_a = sigma_scale * (rgb_a.empty() ? spec(1.f) : rgb_lookup(rgb_a, p, lambda, nullptr));
            mp.sigma_s = sigma_scale * (rgb_s.empty() ? spec(1.f) : rgb_lookup(rgb_s, p, lambda, nullptr));
            if (!rgb_le.empty() && le_scale > 0) mp.Le = le_scale * rgb_lookup(rgb_le, p, lambda, le_illum.data());
            return mp;
        }
        mp.sigma_a = sample_dense(sigma_a_spec.data(), lambda);
        mp.sigma_s = sample_dense(sigma_s_spec.data(), lambda);
        float d = density.lookup(p);
        mp.sigma_a = mp.sigma_a * d;
        mp.sigma_s = mp.sigma_s * d;
        if (emissive) {
            float scale = lescale.lookup(p);
            if (scale > 0) {
                if (has_temperature) {
                    float temp = temperature.lookup(p);
                    temp = (temp - temp_offset) * temp_scale;
                    if (temp > 100.f) mp.Le = scale * blackbody_sample(temp, lambda);
                } else
                    mp.Le = scale * sample_dense(le_spec.data(), lambda);
            }
        }
        return mp;
    }

    // {Grid,RGBGrid}Medium::SampleRay prologue, media.h:322-337 / :404-415
    bool ray_to_medium(V3 o, V3 d, float raytMax, V3 *mo, V3 *md, float *tMin, float *tMax) const {
        apply_inverse_ray(xf, o, d, &raytMax, mo, md);
        return bounds.intersect_p(*mo, *md, raytMax, tMin, tMax);
    }
    DDA sample_ray(V3 o, V3 d, float raytMax, const float lambda[4]) const {
        V3 mo, md;
        float tMin, tMax;
        if (!ray_to_medium(o, d, raytMax, &mo, &md, &tMin, &tMax)) return DDA();
        Spec4 sigma_t = is_rgb ? spec(1.f)
                               : sample_dense(sigma_a_spec.data(), lambda) + sample_dense(sigma_s_spec.data(), lambda);
        return DDA(mo, md, tMin, tMax, &maj, sigma_t);
    }
};

struct TmajCounters {
    int n_fetch = 0, n_seg = 0;
};

// SampleT_maj<ConcreteMedium,F>, media.h:741-806.  callback(p, mp, sigma_maj, T_maj) -> bool.
template <typename F>
inline Spec4 sample_tmaj(const Medium &medium, V3 o, V3 d, float tMax, float u, RNG &rng,
                         const float lambda[4], F callback, TmajCounters *cnt = nullptr) {
    tMax *= length(d);
    d = normalize(d);
    DDA iter = medium.sample_ray(o, d, tMax, lambda);
    Spec4 T_maj = spec(1.f);
    bool done = false;
    while (!done) {
        Segment seg;
        if (!iter.next(&seg)) return T_maj;
        if (cnt) cnt->n_seg++;
        if (seg.sigma_maj[0] == 0) {
            float dt = seg.tMax - seg.tMin;
            if (std::isinf(dt)) dt = std::numeric_limits<float>::max();
            T_maj = T_maj * fast_exp(-dt * seg.sigma_maj);
            continue;
        }
        float tMin = seg.tMin;
        while (true) {
            float t = tMin + sample_exponential(u, seg.sigma_maj[0]);
            u = rng.uniform();
            if (t < seg.tMax) {
                T_maj = T_maj * fast_exp(-(t - tMin) * seg.sigma_maj);
                V3 p = o + d * t;
                MediumProps mp = medium.sample_point(p, lambda);
                if (cnt) cnt->n_fetch++;
                if (!callback(p, t, mp, seg.sigma_maj, T_maj)) {
                    done = true;
                    break;
                }
                T_maj = spec(1.f);
                tMin = t;
            } else {
                float dt = seg.tMax - tMin;
                if (std::isinf(dt)) dt = std::numeric_limits<float>::max();
                T_maj = T_maj * fast_exp(-dt * seg.sigma_maj);
                break;
            }
        }
    }
    return spec(1.f);
}

struct TmajResult {  // mirrors avr_tmaj_result (include/avr.h)
    float T_maj[4], Tr[4], r_u[4], r_l[4];
    float t_event;
    int32_t event, n_fetch, n_seg;
};

// mode 0: util::SampleTransmittance src/graph/util.h:344-366
// mode 1: util::SampleScatter src/graph/util.h:368-403 / GraphIntegrator::Li callback
// mode 2: VolPathIntegrator::SampleLd shadow-ray callback, cpu/integrators.cpp:1352-1380
inline TmajResult run_tmaj(const Medium &m, int mode, V3 o, V3 d, float tMax, uint64_t seq, uint64_t seed,
                           float u, const float lambda[4]) {
    RNG rng(seq, seed);
    TmajResult r;
    TmajCounters cnt;
    Spec4 T_ray = spec(1.f), r_l = spec(1.f), r_u = spec(1.f);
    float Tr = 1;
    r.t_event = -1;
    r.event = 2;
    Spec4 T;
    if (mode == 0) {
        T = sample_tmaj(m, o, d, tMax, u, rng, lambda,
                        [&](V3, float, const MediumProps &mp, Spec4 sigma_maj, Spec4) {
                            Spec4 sigma_n = clamp_zero(sigma_maj - mp.sigma_a - mp.sigma_s);
                            Tr *= sigma_n[0] / sigma_maj[0];
                            return Tr != 0;
                        }, &cnt);
        T_ray = spec(1.f);
        T_ray[0] = Tr;
    } else if (mode == 1) {
        T = sample_tmaj(m, o, d, tMax, u, rng, lambda,
                        [&](V3, float t, const MediumProps &mp, Spec4 sigma_maj, Spec4) {
                            float pAbsorb = mp.sigma_a[0] / sigma_maj[0];
                            float pScatter = mp.sigma_s[0] / sigma_maj[0];
                            float pNull = std::max<float>(0, 1 - pAbsorb - pScatter);
                            float um = rng.uniform();
                            float w[3] = {pAbsorb, pScatter, pNull};
                            int md = sample_discrete(w, 3, um);
                            if (md == 0) { r.event = 0; r.t_event = t; return false; }
                            if (md == 1) { r.event = 1; r.t_event = t; return false; }
                            return true;
                        }, &cnt);
    } else {
        T = sample_tmaj(m, o, d, tMax, u, rng, lambda,
                        [&](V3, float, const MediumProps &mp, Spec4 sigma_maj, Spec4 T_maj) {
                            Spec4 sigma_n = clamp_zero(sigma_maj - mp.sigma_a - mp.sigma_s);
                            float pdf = T_maj[0] * sigma_maj[0];
                            T_ray = T_ray * (T_maj * sigma_n / pdf);
                            r_l = r_l * (T_maj * sigma_maj / pdf);
                            r_u = r_u * (T_maj * sigma_n / pdf);
                            Spec4 Trr = T_ray / average(r_l + r_u);
                            if (max_component(Trr) < 0.05f) {
                                float q = 0.75f;
                                if (rng.uniform() < q) T_ray = spec(0.f);
                                else T_ray = T_ray / (1 - q);
                            }
                            return nonzero(T_ray);
                        }, &cnt);
        // cpu/integrators.cpp:1382-1385
        T_ray = T_ray * (T / T[0]);
        r_l = r_l * (T / T[0]);
        r_u = r_u * (T / T[0]);
    }
    for (int i = 0; i < 4; ++i) { r.T_maj[i] = T[i]; r.Tr[i] = T_ray[i]; r.r_u[i] = r_u[i]; r.r_l[i] = r_l[i]; }
    r.n_fetch = cnt.n_fetch;
    r.n_seg = cnt.n_seg;
    return r;
}

}  // namespace avro
