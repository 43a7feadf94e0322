// oracle/restate_volpath.hpp — TEST INFRASTRUCTURE ONLY (see restate.hpp header).
// Volume-only subset of VolPathIntegrator (cpu/integrators.cpp:962-1106, :1282-1399),
// RayIntegrator::EvaluatePixelSample (:235-298), GetCameraSample (samplers.h:797-814),
// PerspectiveCamera::GenerateRayDifferential (cameras.cpp:429-460), PixelSensor::ToSensorRGB
// (film.h:95-100) and RGBFilm::AddSample (film.h:239-255) for the scene class of include/avr.h
// (one grid medium inside one convex "interface" boundary, one distant light).
//
// PARITY of Li/SampleLd as a whole: pinned BY TOLERANCE against the reference's own VolPathIntegrator
// (oracle/refprobe_volpath.cpp compiles cpu/integrators.cpp and runs Li over the same camera rays and
// sampler streams): tests/test_oracle_golden.py::test_oracle_volpath_li_* (>= 99.5 % of pixel samples
// within 1e-3 where every random number comes from the sampler; converged images within the reference's
// own noise where next-event estimation is on).  Bit-exactness is not available: see the two choices
// below.  The primitives Li is made of are pinned bit-exact individually, and so are the camera rays
// (against the real PerspectiveCamera) and the film arithmetic (against the reference's inline film code).
// Deliberate restatement choices where the reference goes through its BVH / shape code:
//  * boundary hits are analytic (sphere: quadratic in double; box: slab test) and the ray is
//    re-spawned at o + d*tHit without pbrt's error-bound offset (interaction.cpp:91-97);
//  * "IsInf(L.y(lambda))" (cpu/integrators.cpp:277) is evaluated as "any component of L is inf".
#pragma once
#include "restate_filter.hpp"
#include "restate_graph.hpp"
#include "restate_sampler.hpp"
#include "restate_tmaj.hpp"

namespace avro {

struct Scene {
    int width = 0, height = 0, spp = 1, max_depth = 5, sampler_seed = 0;
    int sampler_kind = SAMPLER_INDEPENDENT, sampler_randomize = RANDOMIZE_FAST_OWEN;  // "independent" | "zsobol"
    float filter_radius[2] = {0.5f, 0.5f};
    FilterTable filter;  // active: a FilterSampler-tabulated pixel filter instead of the box
    Xform camera_from_raster, render_from_camera;  // only .m is used
    V3 light_dir;                                  // wi towards the light, normalised
    std::vector<float> light_spec;                 // NSPEC, includes scale
    float scene_radius = 1;
    int boundary_type = 0;
    float boundary[6];
    std::vector<float> sensor_rgb;  // 3*NSPEC
    float imaging_ratio = 1;
    float max_component = kInf;
};

struct RenderCounters {
    int64_t camera_rays = 0, camera_fetches = 0, shadow_fetches = 0, segments = 0, scatter_events = 0;
};

// Distance range [t0,t1] of the ray inside the boundary; false if missed.
inline bool boundary_hit(const Scene &sc, V3 o, V3 d, float *t0, float *t1) {
    if (sc.boundary_type == 0) {
        double ox = double(o.x) - sc.boundary[0], oy = double(o.y) - sc.boundary[1], oz = double(o.z) - sc.boundary[2];
        double r = sc.boundary[3];
        double a = double(d.x) * d.x + double(d.y) * d.y + double(d.z) * d.z;
        double b = 2 * (ox * d.x + oy * d.y + oz * d.z);
        double c = ox * ox + oy * oy + oz * oz - r * r;
        double disc = b * b - 4 * a * c;
        if (!(disc >= 0) || a == 0) return false;
        double s = std::sqrt(disc);
        double q = (b < 0) ? -0.5 * (b - s) : -0.5 * (b + s);
        double ta = q / a, tb = (q != 0) ? c / q : ta;
        if (ta > tb) std::swap(ta, tb);
        *t0 = (float)ta;
        *t1 = (float)tb;
        return true;
    }
    float a0 = -kInf, a1 = kInf;
    for (int i = 0; i < 3; ++i) {
        float inv = 1 / d[i];
        float tn = (sc.boundary[i] - o[i]) * inv, tf = (sc.boundary[3 + i] - o[i]) * inv;
        if (tn > tf) std::swap(tn, tf);
        a0 = tn > a0 ? tn : a0;
        a1 = tf < a1 ? tf : a1;
        if (a0 > a1) return false;
    }
    *t0 = a0;
    *t1 = a1;
    return true;
}

// VolPathIntegrator::SampleLd, medium-interaction branch, cpu/integrators.cpp:1282-1399
inline Spec4 sample_ld(const Scene &sc, const Medium &med, V3 p, V3 wo, const float lambda[4],
                       PathSampler &sampler, Spec4 beta, Spec4 r_p, RenderCounters *rc) {
    float u = sampler.get1d();  // light choice: single light, p = 1
    (void)u;
    float ul0, ul1;
    sampler.get2d(&ul0, &ul1);  // uLight
    // DistantLight::SampleLi, lights.h:284-291
    V3 wi = sc.light_dir;
    V3 pLight = p + wi * (2 * sc.scene_radius);
    Spec4 Ll = sample_dense(sc.light_spec.data(), lambda);
    if (!nonzero(Ll)) return spec(0.f);
    float p_l = 1.f * 1.f;
    float ph = henyey_greenstein(dot(wo, wi), med.g);  // HGPhaseFunction::p, media.h:51
    Spec4 f_hat = spec(ph);
    float scatterPDF = ph;
    (void)scatterPDF;
    if (!nonzero(f_hat)) return spec(0.f);
    // Interaction::SpawnRayTo(pLight): medium interaction has n = 0 and exact pi -> o = p, d = pLight - p
    V3 ro = p, rd = pLight - p;
    Spec4 T_ray = spec(1.f), r_l = spec(1.f), r_u = spec(1.f);
    RNG rng(hash_v3(ro), hash_v3(rd));
    if (!(rd.x == 0 && rd.y == 0 && rd.z == 0)) {
        float t0, t1;
        float tMax = 1 - kShadowEps;
        if (boundary_hit(sc, ro, rd, &t0, &t1) && t1 > 0 && t1 < tMax) tMax = t1;
        float uu = rng.uniform();
        TmajCounters cnt;
        Spec4 T_maj = sample_tmaj(med, ro, rd, tMax, uu, rng, lambda,
                                  [&](V3, float, const MediumProps &mp, Spec4 sigma_maj, Spec4 T_maj) {
                                      Spec4 sigma_n = clamp_zero(sigma_maj - mp.sigma_a - mp.sigma_s);
                                      float pdf = T_maj[0] * sigma_maj[0];
                                      T_ray = T_ray * (T_maj * sigma_n / pdf);
                                      r_l = r_l * (T_maj * sigma_maj / pdf);
                                      r_u = r_u * (T_maj * sigma_n / pdf);
                                      Spec4 Tr = T_ray / average(r_l + r_u);
                                      if (max_component(Tr) < 0.05f) {
                                          float q = 0.75f;
                                          if (rng.uniform() < q) T_ray = spec(0.f);
                                          else T_ray = T_ray / (1 - q);
                                      }
                                      return nonzero(T_ray);
                                  }, &cnt);
        if (rc) { rc->shadow_fetches += cnt.n_fetch; rc->segments += cnt.n_seg; }
        T_ray = T_ray * (T_maj / T_maj[0]);
        r_l = r_l * (T_maj / T_maj[0]);
        r_u = r_u * (T_maj / T_maj[0]);
        if (!nonzero(T_ray)) return spec(0.f);
    }
    r_l = r_l * (r_p * p_l);
    // delta light: beta * f_hat * T_ray * ls->L / r_l.Average()
    return beta * f_hat * T_ray * Ll / average(r_l);
}

// VolPathIntegrator::Li, cpu/integrators.cpp:962-1106, for a ray that starts outside the boundary.
inline Spec4 volpath_li(const Scene &sc, const Medium &med, V3 ro, V3 rd, const float lambda[4],
                        PathSampler &sampler, RenderCounters *rc) {
    Spec4 L = spec(0.f), beta = spec(1.f), r_u = spec(1.f), r_l = spec(1.f);
    int depth = 0;
    float t0, t1;
    if (!boundary_hit(sc, ro, rd, &t0, &t1) || t1 <= 0) return L;
    if (t0 > 0) ro = ro + rd * t0;  // SkipIntersection at the entry hit
    while (true) {
        float tMax = kInf;
        if (boundary_hit(sc, ro, rd, &t0, &t1) && t1 > 0) tMax = t1; else return L;
        bool scattered = false, terminated = false;
        uint64_t hash0 = hash_float(sampler.get1d());
        uint64_t hash1 = hash_float(sampler.get1d());
        RNG rng(hash0, hash1);
        float u0 = sampler.get1d();
        TmajCounters cnt;
        V3 nro = ro, nrd = rd;
        Spec4 T_maj = sample_tmaj(
            med, ro, rd, tMax, u0, rng, lambda,
            [&](V3 p, float, const MediumProps &mp, Spec4 sigma_maj, Spec4 T_maj) {
                if (!nonzero(beta)) { terminated = true; return false; }
                if (rc) rc->camera_fetches++;
                if (depth < sc.max_depth && nonzero(mp.Le)) {
                    float pdf = sigma_maj[0] * T_maj[0];
                    Spec4 betap = beta * T_maj / pdf;
                    Spec4 r_e = r_u * sigma_maj * T_maj / pdf;
                    if (nonzero(r_e)) L = L + betap * mp.sigma_a * mp.Le / average(r_e);
                }
                float pAbsorb = mp.sigma_a[0] / sigma_maj[0];
                float pScatter = mp.sigma_s[0] / sigma_maj[0];
                float pNull = std::max<float>(0, 1 - pAbsorb - pScatter);
                float um = rng.uniform();
                float w[3] = {pAbsorb, pScatter, pNull};
                int mode = sample_discrete(w, 3, um);
                if (mode == 0) { terminated = true; return false; }
                if (mode == 1) {
                    if (depth++ >= sc.max_depth) { terminated = true; return false; }
                    float pdf = T_maj[0] * mp.sigma_s[0];
                    beta = beta * (T_maj * mp.sigma_s / pdf);
                    r_u = r_u * (T_maj * mp.sigma_s / pdf);
                    if (nonzero(beta) && nonzero(r_u)) {
                        if (rc) rc->scatter_events++;
                        V3 wo = -rd;  // ray.d is the normalised direction inside SampleT_maj's caller
                        L = L + sample_ld(sc, med, p, wo, lambda, sampler, beta, r_u, rc);
                        float ua, ub;
                        sampler.get2d(&ua, &ub);
                        float pdfp;
                        V3 wi = sample_hg(wo, med.g, ua, ub, &pdfp);
                        if (pdfp == 0) terminated = true;
                        else {
                            beta = beta * (pdfp / pdfp);  // ps->p / ps->pdf, both = pdf
                            r_l = r_u / pdfp;
                            scattered = true;
                            nro = p;
                            nrd = wi;
                        }
                    }
                    return false;
                }
                Spec4 sigma_n = clamp_zero(sigma_maj - mp.sigma_a - mp.sigma_s);
                float pdf = T_maj[0] * sigma_n[0];
                beta = beta * (T_maj * sigma_n / pdf);
                if (pdf == 0) beta = spec(0.f);
                r_u = r_u * (T_maj * sigma_n / pdf);
                r_l = r_l * (T_maj * sigma_maj / pdf);
                return nonzero(beta) && nonzero(r_u);
            }, &cnt);
        if (rc) rc->segments += cnt.n_seg;
        if (terminated || !nonzero(beta) || !nonzero(r_u)) return L;
        if (scattered) { ro = nro; rd = nrd; continue; }
        beta = beta * (T_maj / T_maj[0]);
        r_u = r_u * (T_maj / T_maj[0]);
        r_l = r_l * (T_maj / T_maj[0]);
        // ray leaves the boundary: no infinite lights (a DistantLight is not one) -> done
        break;
    }
    return L;
}

struct PixelSampleOut {
    Spec4 L;
    Lambda4 lambda;
    float rgb[3];
    float weight;
};

// GraphIntegrator::Li, free-graph branch (graph/graph_integrator.cpp:180-246) for a camera ray that starts
// outside the boundary: skip the entry hit, one SampleT_maj through the medium whose first real collision
// either absorbs the path or connects it to the graph (ConnectToGraph of the WORLD-space point), then
// spectrum * L.  world_from_render: affine rows (12 floats).
inline Spec4 graph_li(const Scene &sc, const Medium &med, const GraphVerts &g, const float *wfr, V3 ro, V3 rd,
                      const float lambda[4], PathSampler &sampler) {
    float Lg = 0;
    float t0, t1;
    if (boundary_hit(sc, ro, rd, &t0, &t1) && t1 > 0) {
        if (t0 > 0) ro = ro + rd * t0;
        if (boundary_hit(sc, ro, rd, &t0, &t1) && t1 > 0) {
            uint64_t hash0 = hash_float(sampler.get1d());
            uint64_t hash1 = hash_float(sampler.get1d());
            RNG rng(hash0, hash1);
            float u0 = sampler.get1d();
            TmajCounters cnt;
            sample_tmaj(med, ro, rd, t1, u0, rng, lambda,
                        [&](V3 p, float, const MediumProps &mp, Spec4 sigma_maj, Spec4) {
                            float pAbsorb = mp.sigma_a[0] / sigma_maj[0];
                            float pScatter = mp.sigma_s[0] / sigma_maj[0];
                            float pNull = std::max<float>(0, 1 - pAbsorb - pScatter);
                            float um = rng.uniform();
                            float w[3] = {pAbsorb, pScatter, pNull};
                            int mode = sample_discrete(w, 3, um);
                            if (mode == 0) return false;
                            if (mode == 1) {
                                V3 pw = {wfr[0] * p.x + wfr[1] * p.y + wfr[2] * p.z + wfr[3],
                                         wfr[4] * p.x + wfr[5] * p.y + wfr[6] * p.z + wfr[7],
                                         wfr[8] * p.x + wfr[9] * p.y + wfr[10] * p.z + wfr[11]};
                                Lg += connect_to_graph(g, pw);
                                return false;
                            }
                            return true;
                        }, &cnt);
        }
    }
    Spec4 spectrum = sample_dense(sc.light_spec.data(), lambda);
    return spectrum * Lg;
}

// the scene's sampler: IndependentSampler(seed) or ZSobolSampler(spp, film resolution, randomize, seed)
inline PathSampler make_sampler(const Scene &sc) {
    PathSampler s;
    s.kind = sc.sampler_kind;
    s.ind.seed = sc.sampler_seed;
    if (s.kind == SAMPLER_ZSOBOL) s.zs.configure(sc.spp, sc.width, sc.height, sc.sampler_randomize, sc.sampler_seed);
    return s;
}

// Camera ray of one pixel sample (GetCameraSample + PerspectiveCamera::GenerateRay), as
// evaluate_pixel_sample below forms it; exported so that the reference probe traces the same rays.
inline void camera_ray(const Scene &sc, int px, int py, int sampleIndex, V3 *ro, V3 *rd, float *weight = nullptr) {
    PathSampler sampler = make_sampler(sc);
    sampler.start_pixel_sample(px, py, sampleIndex);
    sampler.get1d();
    float f0, f1;
    sampler.get2d(&f0, &f1);
    float fx, fy, fw = 1;
    if (sc.filter.active()) sc.filter.sample(f0, f1, &fx, &fy, &fw);
    else {
        fx = lerp(f0, -sc.filter_radius[0], sc.filter_radius[0]);
        fy = lerp(f1, -sc.filter_radius[1], sc.filter_radius[1]);
    }
    if (weight) *weight = fw;
    float pfx = (float(px) + fx) + 0.5f, pfy = (float(py) + fy) + 0.5f;
    V3 pCamera = apply_point(sc.camera_from_raster, {pfx, pfy, 0});
    apply_ray(sc.render_from_camera, {0, 0, 0}, normalize(pCamera), nullptr, ro, rd);
}

// PixelSensor::ToSensorRGB (film.h:95-100) followed by the clamp of RGBFilm::AddSample (film.h:244-247).
// Pinned bit for bit against the reference's inline film code (refprobe: ref_film_add_samples).
inline void sensor_rgb_clamped(const Scene &sc, Spec4 L, const float lambda[4], Spec4 pdf, float rgb[3]) {
    Spec4 Ld = safe_div(L, pdf);
    for (int c = 0; c < 3; ++c)
        rgb[c] = sc.imaging_ratio * average(sample_dense(sc.sensor_rgb.data() + c * NSPEC, lambda) * Ld);
    float m = std::max({rgb[0], rgb[1], rgb[2]});
    if (m > sc.max_component) for (int c = 0; c < 3; ++c) rgb[c] *= sc.max_component / m;
}

// RayIntegrator::EvaluatePixelSample, cpu/integrators.cpp:235-298
inline PixelSampleOut evaluate_pixel_sample(const Scene &sc, const Medium &med, int px, int py, int sampleIndex,
                                            RenderCounters *rc) {
    PathSampler sampler = make_sampler(sc);
    sampler.start_pixel_sample(px, py, sampleIndex);
    float lu = sampler.get1d();
    PixelSampleOut out;
    out.lambda = sample_visible_wavelengths(lu);
    // GetCameraSample (samplers.h:797-814) with BoxFilter::Sample (filters.h:67-70) or FilterSampler::Sample
    float f0, f1, l0, l1;
    sampler.get2d(&f0, &f1);  // GetPixel2D
    float fx, fy;
    out.weight = 1;
    if (sc.filter.active()) sc.filter.sample(f0, f1, &fx, &fy, &out.weight);
    else {
        fx = lerp(f0, -sc.filter_radius[0], sc.filter_radius[0]);
        fy = lerp(f1, -sc.filter_radius[1], sc.filter_radius[1]);
    }
    float pfx = (float(px) + fx) + 0.5f, pfy = (float(py) + fy) + 0.5f;
    sampler.get1d();          // time
    sampler.get2d(&l0, &l1);  // lens
    V3 pCamera = apply_point(sc.camera_from_raster, {pfx, pfy, 0});
    V3 dir = normalize(pCamera);
    V3 ro, rd;
    apply_ray(sc.render_from_camera, {0, 0, 0}, dir, nullptr, &ro, &rd);
    if (rc) rc->camera_rays++;
    Spec4 L = volpath_li(sc, med, ro, rd, out.lambda.lambda, sampler, rc);
    if (has_nans(L)) L = spec(0.f);
    else if (std::isinf(L[0]) || std::isinf(L[1]) || std::isinf(L[2]) || std::isinf(L[3])) L = spec(0.f);
    out.L = L;
    Spec4 pdf = {{out.lambda.pdf[0], out.lambda.pdf[1], out.lambda.pdf[2], out.lambda.pdf[3]}};
    sensor_rgb_clamped(sc, L, out.lambda.lambda, pdf, out.rgb);
    return out;
}

}  // namespace avro
