// oracle/oracle_capi.cpp — TEST INFRASTRUCTURE ONLY.
// Flat C entry points (ctypes) over the CPU restatement in restate*.hpp.  Built by oracle/Makefile
// into oracle/liboracle.so.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load it.  Signatures match oracle/refprobe.cpp (ref_* symbols), which
// runs the same calls through the reference's own templates.
#include "restate_graphfile.hpp"
#include "restate_tmaj.hpp"
#include "restate_volpath.hpp"

#include "../include/avr.h"

using namespace avro;

extern "C" {

const char *orc_kind() { return "port"; }

void orc_fast_exp(const float *x, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = fast_exp(x[i]);
}
// host libm exactly as the reference calls it (std::log / std::sin / std::cos on floats)
void orc_libm(int kind, const float *x, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = kind == 0 ? std::log(x[i]) : kind == 1 ? std::sin(x[i]) : std::cos(x[i]);
}
void orc_mix_bits(const uint64_t *x, int64_t n, uint64_t *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = mix_bits(x[i]);
}
void orc_hash_float(const float *x, int64_t n, uint64_t *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = hash_float(x[i]);
}
void orc_hash_point3(const float *x, int64_t n, uint64_t *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = hash_v3({x[3 * i], x[3 * i + 1], x[3 * i + 2]});
}
// RNG(seq, seed) (use_seed=1) or RNG(seq) (use_seed=0), Advance(adv), then n Uniform<uint32>/<float>
void orc_rng_stream(uint64_t seq, uint64_t seed, int use_seed, int64_t adv, int64_t n, uint32_t *out_u32,
                    float *out_f) {
    RNG r;
    if (use_seed) r.set_sequence(seq, seed); else r.set_sequence(seq);
    r.advance(adv);
    RNG r2 = r;
    for (int64_t i = 0; i < n; ++i) out_u32[i] = r.u32();
    for (int64_t i = 0; i < n; ++i) out_f[i] = r2.uniform();
}
// IndependentSampler(spp, seed).StartPixelSample({px,py}, sampleIndex, dim) then n Get1D()
void orc_independent_sampler(int px, int py, int seed, int sampleIndex, int dim, int64_t n, float *out) {
    IndependentSampler s;
    s.seed = seed;
    s.start_pixel_sample(px, py, sampleIndex, dim);
    for (int64_t i = 0; i < n; ++i) out[i] = s.get1d();
}
void orc_sample_discrete3(const float *w, const float *u, int64_t n, int32_t *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = sample_discrete(w + 3 * i, 3, u[i]);
}
void orc_sample_exponential(const float *u, const float *a, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = sample_exponential(u[i], a[i]);
}
void orc_hg_sample(const float *wo, float g, const float *u, int64_t n, float *wi, float *pdf) {
    for (int64_t i = 0; i < n; ++i) {
        V3 w = sample_hg({wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]}, g, u[2 * i], u[2 * i + 1], &pdf[i]);
        wi[3 * i] = w.x; wi[3 * i + 1] = w.y; wi[3 * i + 2] = w.z;
    }
}
void orc_hg_p(const float *cosTheta, float g, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = henyey_greenstein(cosTheta[i], g);
}
void orc_sample_wavelengths(const float *u, int64_t n, float *lambda_pdf) {
    for (int64_t i = 0; i < n; ++i) {
        Lambda4 l = sample_visible_wavelengths(u[i]);
        for (int k = 0; k < 4; ++k) { lambda_pdf[8 * i + k] = l.lambda[k]; lambda_pdf[8 * i + 4 + k] = l.pdf[k]; }
    }
}
void orc_blackbody(const float *T, int64_t n, const float lambda[4], float *out) {
    for (int64_t i = 0; i < n; ++i) {
        Spec4 s = blackbody_sample(T[i], lambda);
        for (int k = 0; k < 4; ++k) out[4 * i + k] = s[k];
    }
}
void orc_grid_lookup(const float *grid, int nx, int ny, int nz, const float *p, int64_t n, float *out) {
    Grid g{grid, nx, ny, nz};
    for (int64_t i = 0; i < n; ++i) out[i] = g.lookup({p[3 * i], p[3 * i + 1], p[3 * i + 2]});
}
// rays: 8 floats {o, d, tMin, tMax} medium space; bounds: 6 floats; see avr_dda_segments (avr.h)
void orc_dda(const float *bounds, const int32_t *res, const float *vox, const float *rays, int64_t n,
             float sigma_t, int32_t max_per_ray, int32_t *counts, float *seg_t, float *seg_sigma,
             int32_t *seg_voxel) {
    MajGrid g;
    g.bounds = {{bounds[0], bounds[1], bounds[2]}, {bounds[3], bounds[4], bounds[5]}};
    for (int i = 0; i < 3; ++i) g.res[i] = res[i];
    g.vox.assign(vox, vox + size_t(res[0]) * res[1] * res[2]);
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays + 8 * i;
        DDA it({r[0], r[1], r[2]}, {r[3], r[4], r[5]}, r[6], r[7], &g, spec(sigma_t));
        Segment s;
        int k = 0;
        while (it.next(&s)) {
            if (k < max_per_ray) {
                size_t j = size_t(i) * max_per_ray + k;
                seg_t[2 * j] = s.tMin; seg_t[2 * j + 1] = s.tMax;
                seg_sigma[j] = s.sigma_maj[0];
                seg_voxel[j] = s.voxel;
            }
            ++k;
            if (k > 100000) break;  // NaN-direction guard; the reference would spin as well
        }
        counts[i] = k;
    }
}
void orc_apply_inverse_ray(const float *m, const float *minv, const float *rays7, int64_t n, float *out7) {
    Xform t;
    std::memcpy(t.m, m, 64);
    std::memcpy(t.mInv, minv, 64);
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays7 + 7 * i;
        float tMax = r[6];
        V3 o, d;
        apply_inverse_ray(t, {r[0], r[1], r[2]}, {r[3], r[4], r[5]}, &tMax, &o, &d);
        float *q = out7 + 7 * i;
        q[0] = o.x; q[1] = o.y; q[2] = o.z; q[3] = d.x; q[4] = d.y; q[5] = d.z; q[6] = tMax;
    }
}
void orc_apply_ray(const float *m, const float *minv, const float *rays7, int64_t n, float *out7) {
    Xform t;
    std::memcpy(t.m, m, 64);
    std::memcpy(t.mInv, minv, 64);
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays7 + 7 * i;
        float tMax = r[6];
        V3 o, d;
        apply_ray(t, {r[0], r[1], r[2]}, {r[3], r[4], r[5]}, &tMax, &o, &d);
        float *q = out7 + 7 * i;
        q[0] = o.x; q[1] = o.y; q[2] = o.z; q[3] = d.x; q[4] = d.y; q[5] = d.z; q[6] = tMax;
    }
}

// ---- media --------------------------------------------------------------------------------------------
void *orc_medium_create_uniformgrid(const avr_grid_desc *d) {
    Medium *m = new Medium();
    m->init_grid(d->nx, d->ny, d->nz, d->density, d->temperature, d->lescale, d->le_norm, d->p0, d->p1,
                 d->render_from_medium, d->medium_from_render, d->sigma_a, d->sigma_s, d->le, d->sigma_scale,
                 d->g, d->temperature_offset, d->temperature_scale, d->majorant_res);
    return m;
}
void *orc_medium_create_rgbgrid(const avr_rgbgrid_desc *d) {
    Medium *m = new Medium();
    m->init_rgb(d->nx, d->ny, d->nz, d->sigma_a, d->sigma_s, d->le, d->le_illuminant, d->le_scale, d->p0, d->p1,
                d->render_from_medium, d->medium_from_render, d->sigma_scale, d->g, d->majorant_res);
    return m;
}
static NvGrid nv_from(const avr_nanovdb_grid &g) {
    NvGrid n;
    n.present = g.values != nullptr;
    if (!n.present) return n;
    for (int i = 0; i < 3; ++i) { n.imin[i] = g.index_min[i]; n.dims[i] = g.dims[i]; n.vec[i] = g.vec[i]; }
    for (int i = 0; i < 9; ++i) n.inv[i] = g.inv_mat[i];
    for (int i = 0; i < 6; ++i) n.wbox[i] = g.world_bbox[i];
    n.v.assign(g.values, g.values + size_t(g.dims[0]) * g.dims[1] * g.dims[2]);
    return n;
}
void *orc_medium_create_nanovdb(const avr_nanovdb_desc *d) {
    Medium *m = new Medium();
    m->init_nanovdb(nv_from(d->density), nv_from(d->temperature), d->render_from_medium, d->medium_from_render,
                    d->sigma_a, d->sigma_s, d->sigma_scale, d->g, d->le_scale, d->temperature_offset,
                    d->temperature_scale, d->majorant_res);
    return m;
}
void orc_medium_bounds(void *m, float out[6]) {
    const Bounds &b = ((Medium *)m)->bounds;
    for (int i = 0; i < 3; ++i) { out[i] = b.pMin[i]; out[3 + i] = b.pMax[i]; }
}
void orc_medium_destroy(void *m) { delete (Medium *)m; }
int orc_medium_is_emissive(void *m) { return ((Medium *)m)->emissive ? 1 : 0; }
void orc_medium_majorants(void *mp, float *out, int32_t res[3]) {
    Medium *m = (Medium *)mp;
    for (int i = 0; i < 3; ++i) res[i] = m->maj.res[i];
    std::memcpy(out, m->maj.vox.data(), m->maj.vox.size() * 4);
}
void orc_medium_sample_point(void *mp, const float *p, int64_t n, const float lambda[4], float *out) {
    Medium *m = (Medium *)mp;
    for (int64_t i = 0; i < n; ++i) {
        MediumProps q = m->sample_point({p[3 * i], p[3 * i + 1], p[3 * i + 2]}, lambda);
        for (int k = 0; k < 4; ++k) {
            out[12 * i + k] = q.sigma_a[k];
            out[12 * i + 4 + k] = q.sigma_s[k];
            out[12 * i + 8 + k] = q.Le[k];
        }
    }
}
void orc_medium_ray_to_medium(void *mp, const float *rays, int64_t n, float *out) {
    Medium *m = (Medium *)mp;
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays + 7 * i;
        V3 o, d;
        float t0 = 0, t1 = 0;
        bool hit = m->ray_to_medium({r[0], r[1], r[2]}, {r[3], r[4], r[5]}, r[6], &o, &d, &t0, &t1);
        float *q = out + 9 * i;
        q[0] = o.x; q[1] = o.y; q[2] = o.z; q[3] = d.x; q[4] = d.y; q[5] = d.z;
        q[6] = hit ? t0 : 0; q[7] = hit ? t1 : 0; q[8] = hit ? 1.f : 0.f;
    }
}
void orc_sample_tmaj(void *mp, int mode, const float *rays, int64_t n, const uint64_t *rng, const float *u,
                     const float lambda[4], avr_tmaj_result *out) {
    Medium *m = (Medium *)mp;
    static_assert(sizeof(TmajResult) == sizeof(avr_tmaj_result), "layout");
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays + 7 * i;
        TmajResult t = run_tmaj(*m, mode, {r[0], r[1], r[2]}, {r[3], r[4], r[5]}, r[6], rng[2 * i], rng[2 * i + 1],
                                u[i], lambda);
        std::memcpy(&out[i], &t, sizeof(t));
    }
}

}  // extern "C"

#include "oracle_capi_render.inc"
