// oracle/refprobe.cpp — TEST INFRASTRUCTURE ONLY.
//
// Thin probe over the REFERENCE'S OWN hot-path code, compiled from the sources where they lie under
// /root/reference (never copied): this file only #includes reference headers and calls into them;
// oracle/Makefile adds the reference's media.cpp, util/{pstd,transform,math,sampling,vecmath}.cpp.
// Output: oracle/_ref/libavr_refprobe.so (git-ignored, travels to the GPU box).
//
// It exports the same flat signatures as oracle/oracle_capi.cpp with a ref_ prefix, so that
// tests can check the restatement against the real GridMedium / RGBGridMedium / SampleT_maj<> /
// DDAMajorantIterator / SampledGrid / RNG / Hash / FastExp / SampleDiscrete / HG / Transform code.
// Built with -fno-access-control solely to fill RGBUnboundedSpectrum coefficients (the reference
// builds them from its generated rgb2spec tables, which need its cmake build).
#include <pbrt/media.h>
#include <pbrt/samplers.h>
#include <pbrt/util/hash.h>
#include <pbrt/util/rng.h>
#include <pbrt/util/sampling.h>
#include <pbrt/util/scattering.h>
#include <pbrt/util/spectrum.h>
#include <pbrt/util/transform.h>

#define NANOVDB_USE_ZIP
#include <nanovdb/util/GridBuilder.h>
#include <nanovdb/util/IO.h>
#include <fstream>

#include <Eigen/Sparse>
#include <graph/util.h>
#include <graph/graph.h>
#include <graph/lighting_calculator.h>

#include <fstream>

#include "../include/avr.h"

using namespace pbrt;

namespace {
Transform make_xf(const float *m, const float *minv) {
    SquareMatrix<4> M, MI;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            M[i][j] = m[4 * i + j];
            MI[i][j] = minv[4 * i + j];
        }
    return Transform(M, MI);
}
SampledWavelengths make_lambda(const float l[4]) {
    SampledWavelengths swl;
    for (int i = 0; i < 4; ++i) {
        swl.lambda[i] = l[i];
        swl.pdf[i] = 1;
    }
    return swl;
}
struct RefMedium {
    Medium medium;
    GridMedium *grid = nullptr;
    RGBGridMedium *rgb = nullptr;
    NanoVDBMedium *nvdb = nullptr;
    std::vector<DenselySampledSpectrum *> keep;
};
DenselySampledSpectrum *dense_from(const float *tab) {
    auto *d = new DenselySampledSpectrum(
        DenselySampledSpectrum::SampleFunction([&](Float l) { return tab[int(l) - 360]; }));
    return d;
}
}  // namespace

#pragma GCC visibility push(default)
extern "C" {

const char *ref_kind() { return "reference"; }

void ref_fast_exp(const float *x, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = FastExp(x[i]);
}
void ref_mix_bits(const uint64_t *x, int64_t n, uint64_t *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = MixBits(x[i]);
}
void ref_hash_float(const float *x, int64_t n, uint64_t *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = Hash(x[i]);
}
void ref_hash_point3(const float *x, int64_t n, uint64_t *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = Hash(Point3f(x[3 * i], x[3 * i + 1], x[3 * i + 2]));
}
void ref_rng_stream(uint64_t seq, uint64_t seed, int use_seed, int64_t adv, int64_t n, uint32_t *out_u32,
                    float *out_f) {
    RNG r;
    if (use_seed) r.SetSequence(seq, seed); else r.SetSequence(seq);
    r.Advance(adv);
    RNG r2 = r;
    for (int64_t i = 0; i < n; ++i) out_u32[i] = r.Uniform<uint32_t>();
    for (int64_t i = 0; i < n; ++i) out_f[i] = r2.Uniform<Float>();
}
void ref_independent_sampler(int px, int py, int seed, int sampleIndex, int dim, int64_t n, float *out) {
    IndependentSampler s(1, seed);
    s.StartPixelSample(Point2i(px, py), sampleIndex, dim);
    for (int64_t i = 0; i < n; ++i) out[i] = s.Get1D();
}
void ref_sample_discrete3(const float *w, const float *u, int64_t n, int32_t *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = SampleDiscrete({w[3 * i], w[3 * i + 1], w[3 * i + 2]}, u[i]);
}
void ref_sample_exponential(const float *u, const float *a, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = SampleExponential(u[i], a[i]);
}
void ref_hg_sample(const float *wo, float g, const float *u, int64_t n, float *wi, float *pdf) {
    HGPhaseFunction hg(g);
    for (int64_t i = 0; i < n; ++i) {
        auto ps = hg.Sample_p(Vector3f(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]), Point2f(u[2 * i], u[2 * i + 1]));
        wi[3 * i] = ps->wi.x; wi[3 * i + 1] = ps->wi.y; wi[3 * i + 2] = ps->wi.z;
        pdf[i] = ps->pdf;
    }
}
void ref_hg_p(const float *cosTheta, float g, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = HenyeyGreenstein(cosTheta[i], g);
}
void ref_sample_wavelengths(const float *u, int64_t n, float *lambda_pdf) {
    for (int64_t i = 0; i < n; ++i) {
        SampledWavelengths l = SampledWavelengths::SampleVisible(u[i]);
        for (int k = 0; k < 4; ++k) { lambda_pdf[8 * i + k] = l[k]; lambda_pdf[8 * i + 4 + k] = l.PDF()[k]; }
    }
}
void ref_blackbody(const float *T, int64_t n, const float lambda[4], float *out) {
    SampledWavelengths swl = make_lambda(lambda);
    for (int64_t i = 0; i < n; ++i) {
        SampledSpectrum s = BlackbodySpectrum(T[i]).Sample(swl);
        for (int k = 0; k < 4; ++k) out[4 * i + k] = s[k];
    }
}
void ref_grid_lookup(const float *grid, int nx, int ny, int nz, const float *p, int64_t n, float *out) {
    SampledGrid<Float> g(pstd::span<const Float>(grid, size_t(nx) * ny * nz), nx, ny, nz, Allocator());
    for (int64_t i = 0; i < n; ++i) out[i] = g.Lookup(Point3f(p[3 * i], p[3 * i + 1], p[3 * i + 2]));
}
void ref_dda(const float *bounds, const int32_t *res, const float *vox, const float *rays, int64_t n,
             float sigma_t, int32_t max_per_ray, int32_t *counts, float *seg_t, float *seg_sigma,
             int32_t *seg_voxel) {
    MajorantGrid g(Bounds3f(Point3f(bounds[0], bounds[1], bounds[2]), Point3f(bounds[3], bounds[4], bounds[5])),
                   Point3i(res[0], res[1], res[2]), Allocator());
    // a second grid whose values are the voxel's own linear index recovers the traversed voxel
    MajorantGrid gi = g;
    for (size_t i = 0; i < g.voxels.size(); ++i) { g.voxels[i] = vox[i]; gi.voxels[i] = Float(i); }
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays + 8 * i;
        Ray ray(Point3f(r[0], r[1], r[2]), Vector3f(r[3], r[4], r[5]));
        DDAMajorantIterator it(ray, r[6], r[7], &g, SampledSpectrum(sigma_t));
        DDAMajorantIterator iti(ray, r[6], r[7], &gi, SampledSpectrum(1.f));
        int k = 0;
        while (auto s = it.Next()) {
            auto si = iti.Next();
            if (k < max_per_ray) {
                size_t j = size_t(i) * max_per_ray + k;
                seg_t[2 * j] = s->tMin; seg_t[2 * j + 1] = s->tMax;
                seg_sigma[j] = s->sigma_maj[0];
                seg_voxel[j] = int(si->sigma_maj[0]);
            }
            ++k;
            if (k > 100000) break;
        }
        counts[i] = k;
    }
}
void ref_apply_inverse_ray(const float *m, const float *minv, const float *rays7, int64_t n, float *out7) {
    Transform t = make_xf(m, minv);
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays7 + 7 * i;
        Float tMax = r[6];
        Ray ray = t.ApplyInverse(Ray(Point3f(r[0], r[1], r[2]), Vector3f(r[3], r[4], r[5])), &tMax);
        float *q = out7 + 7 * i;
        q[0] = ray.o.x; q[1] = ray.o.y; q[2] = ray.o.z; q[3] = ray.d.x; q[4] = ray.d.y; q[5] = ray.d.z; q[6] = tMax;
    }
}
void ref_apply_ray(const float *m, const float *minv, const float *rays7, int64_t n, float *out7) {
    Transform t = make_xf(m, minv);
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays7 + 7 * i;
        Float tMax = r[6];
        Ray ray = t(Ray(Point3f(r[0], r[1], r[2]), Vector3f(r[3], r[4], r[5])), &tMax);
        float *q = out7 + 7 * i;
        q[0] = ray.o.x; q[1] = ray.o.y; q[2] = ray.o.z; q[3] = ray.d.x; q[4] = ray.d.y; q[5] = ray.d.z; q[6] = tMax;
    }
}

// ---- media: the reference's real constructors ----------------------------------------------------
void *ref_medium_create_uniformgrid(const avr_grid_desc *d) {
    Allocator alloc;
    RefMedium *rm = new RefMedium();
    size_t n = size_t(d->nx) * d->ny * d->nz;
    SampledGrid<Float> density(pstd::span<const Float>(d->density, n), d->nx, d->ny, d->nz, alloc);
    pstd::optional<SampledGrid<Float>> temperature;
    if (d->temperature)
        temperature = SampledGrid<Float>(pstd::span<const Float>(d->temperature, n), d->nx, d->ny, d->nz, alloc);
    // LeScale grid as GridMedium::Create builds it, media.cpp:293-306
    SampledGrid<Float> LeGrid(alloc);
    if (!d->lescale) {
        std::vector<Float> one = {d->le_norm};
        LeGrid = SampledGrid<Float>(one, 1, 1, 1, alloc);
    } else {
        std::vector<Float> ls(d->lescale, d->lescale + n);
        for (size_t i = 0; i < n; ++i) ls[i] *= d->le_norm;
        LeGrid = SampledGrid<Float>(ls, d->nx, d->ny, d->nz, alloc);
    }
    DenselySampledSpectrum *sa = dense_from(d->sigma_a), *ss = dense_from(d->sigma_s);
    rm->keep = {sa, ss};
    Spectrum Le;
    if (d->le) {
        DenselySampledSpectrum *le = dense_from(d->le);
        rm->keep.push_back(le);
        Le = le;
    } else
        Le = new ConstantSpectrum(0.f);
    Transform xf = make_xf(d->render_from_medium, d->medium_from_render);
    rm->grid = new GridMedium(Bounds3f(Point3f(d->p0[0], d->p0[1], d->p0[2]), Point3f(d->p1[0], d->p1[1], d->p1[2])),
                              xf, sa, ss, d->sigma_scale, d->g, std::move(density), std::move(temperature),
                              d->temperature_scale, d->temperature_offset, Le, std::move(LeGrid), alloc);
    rm->medium = rm->grid;
    return rm;
}
void *ref_medium_create_rgbgrid(const avr_rgbgrid_desc *d) {
    Allocator alloc;
    RefMedium *rm = new RefMedium();
    size_t n = size_t(d->nx) * d->ny * d->nz;
    auto mk = [&](const float *v) {
        std::vector<RGBUnboundedSpectrum> s(n);
        for (size_t i = 0; i < n; ++i) {
            s[i].scale = v[4 * i];
            s[i].rsp = RGBSigmoidPolynomial(v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
        return SampledGrid<RGBUnboundedSpectrum>(s, d->nx, d->ny, d->nz, alloc);
    };
    pstd::optional<SampledGrid<RGBUnboundedSpectrum>> ga, gs;
    pstd::optional<SampledGrid<RGBIlluminantSpectrum>> gle;
    if (d->sigma_a) ga = mk(d->sigma_a);
    if (d->sigma_s) gs = mk(d->sigma_s);
    if (d->le) {
        DenselySampledSpectrum *ill = dense_from(d->le_illuminant);
        rm->keep.push_back(ill);
        std::vector<RGBIlluminantSpectrum> s(n);
        for (size_t i = 0; i < n; ++i) {
            s[i].scale = d->le[4 * i];
            s[i].rsp = RGBSigmoidPolynomial(d->le[4 * i + 1], d->le[4 * i + 2], d->le[4 * i + 3]);
            s[i].illuminant = ill;
        }
        gle = SampledGrid<RGBIlluminantSpectrum>(s, d->nx, d->ny, d->nz, alloc);
    }
    Transform xf = make_xf(d->render_from_medium, d->medium_from_render);
    rm->rgb = new RGBGridMedium(Bounds3f(Point3f(d->p0[0], d->p0[1], d->p0[2]), Point3f(d->p1[0], d->p1[1], d->p1[2])),
                                xf, d->g, std::move(ga), std::move(gs), d->sigma_scale, std::move(gle), d->le_scale,
                                alloc);
    rm->medium = rm->rgb;
    return rm;
}
// ---- NanoVDB: a real nanovdb::FloatGrid built with the vendored GridBuilder, handed to the
// reference's own NanoVDBMedium constructor (media.cpp:511-616) -------------------------------------
using NvHandle = nanovdb::GridHandle<NanoVDBBuffer>;
void *ref_nanovdb_grid_build(const float *dense, const int32_t dims[3], const int32_t origin_ijk[3],
                             double voxel_size, const double origin_world[3]) {
    nanovdb::GridBuilder<float> builder(0.f, nanovdb::GridClass::FogVolume);
    nanovdb::CoordBBox box(nanovdb::Coord(origin_ijk[0], origin_ijk[1], origin_ijk[2]),
                           nanovdb::Coord(origin_ijk[0] + dims[0] - 1, origin_ijk[1] + dims[1] - 1,
                                          origin_ijk[2] + dims[2] - 1));
    builder([&](const nanovdb::Coord &c) -> float {
        int i = c[0] - origin_ijk[0], j = c[1] - origin_ijk[1], k = c[2] - origin_ijk[2];
        return dense[(size_t(k) * dims[1] + j) * dims[0] + i];
    }, box);
    NvHandle h = builder.getHandle<nanovdb::AbsDiff, NanoVDBBuffer>(
        voxel_size, nanovdb::Vec3d(origin_world[0], origin_world[1], origin_world[2]), "density", nanovdb::AbsDiff(),
        NanoVDBBuffer(Allocator()));
    return new NvHandle(std::move(h));
}
// the same with a grid name ("density" / "temperature": what NanoVDBMedium::Create looks up, media.cpp:636-645)
void *ref_nanovdb_grid_build_named(const float *dense, const int32_t dims[3], const int32_t origin_ijk[3],
                                   double voxel_size, const double origin_world[3], const char *name) {
    nanovdb::GridBuilder<float> builder(0.f, nanovdb::GridClass::FogVolume);
    nanovdb::CoordBBox box(nanovdb::Coord(origin_ijk[0], origin_ijk[1], origin_ijk[2]),
                           nanovdb::Coord(origin_ijk[0] + dims[0] - 1, origin_ijk[1] + dims[1] - 1,
                                          origin_ijk[2] + dims[2] - 1));
    builder([&](const nanovdb::Coord &c) -> float {
        int i = c[0] - origin_ijk[0], j = c[1] - origin_ijk[1], k = c[2] - origin_ijk[2];
        return dense[(size_t(k) * dims[1] + j) * dims[0] + i];
    }, box);
    NvHandle h = builder.getHandle<nanovdb::AbsDiff, NanoVDBBuffer>(
        voxel_size, nanovdb::Vec3d(origin_world[0], origin_world[1], origin_world[2]), name, nanovdb::AbsDiff(),
        NanoVDBBuffer(Allocator()));
    return new NvHandle(std::move(h));
}
// nanovdb::io::writeGrids of the reference's vendored NanoVDB: the .nvdb files the product's reader is tested on.
// codec 0 = NONE, 1 = ZIP.  Returns 0 on success.
int32_t ref_nanovdb_write_file(const char *path, void **handles, int32_t n, int32_t codec) {
    try {
        std::ofstream os(path, std::ios::out | std::ios::binary | std::ios::trunc);
        if (!os.is_open()) return 1;
        // writeGrids wants a vector of handles; write segment by segment instead (one grid each), which the file
        // format allows (readGrid scans all segments)
        for (int32_t i = 0; i < n; ++i)
            nanovdb::io::writeGrid(os, *(NvHandle *)handles[i], codec == 1 ? nanovdb::io::Codec::ZIP : nanovdb::io::Codec::NONE);
        return os ? 0 : 2;
    } catch (const std::exception &e) {
        std::fprintf(stderr, "refprobe: nanovdb write: %s\n", e.what());
        return 3;
    }
}
void ref_nanovdb_grid_free(void *h) { delete (NvHandle *)h; }
// what the host adapter reads off the grid to fill avr_nanovdb_grid
void ref_nanovdb_grid_info(void *hp, int32_t index_min[3], int32_t dims[3], float inv_mat[9], float vec[3],
                           double world_bbox[6]) {
    const nanovdb::FloatGrid *g = ((NvHandle *)hp)->grid<float>();
    auto ib = g->indexBBox();
    auto wb = g->worldBBox();
    for (int i = 0; i < 3; ++i) {
        index_min[i] = ib.min()[i];
        dims[i] = ib.max()[i] - ib.min()[i] + 1;
        vec[i] = g->map().mVecF[i];
        world_bbox[i] = wb.min()[i];
        world_bbox[3 + i] = wb.max()[i];
    }
    for (int i = 0; i < 9; ++i) inv_mat[i] = g->map().mInvMatF[i];
}
// cmd/nanovdb2pbrt.cpp:100-110 style flattening over the inclusive index box
void ref_nanovdb_grid_flatten(void *hp, float *out) {
    const nanovdb::FloatGrid *g = ((NvHandle *)hp)->grid<float>();
    auto ib = g->indexBBox();
    size_t n = 0;
    for (int z = ib.min()[2]; z <= ib.max()[2]; ++z)
        for (int y = ib.min()[1]; y <= ib.max()[1]; ++y)
            for (int x = ib.min()[0]; x <= ib.max()[0]; ++x) out[n++] = g->tree().getValue({x, y, z});
}
// consumes (moves from) the two grid handles; temperature may be NULL
void *ref_medium_create_nanovdb(const avr_nanovdb_desc *d, void *density_h, void *temperature_h) {
    Allocator alloc;
    RefMedium *rm = new RefMedium();
    DenselySampledSpectrum *sa = dense_from(d->sigma_a), *ss = dense_from(d->sigma_s);
    rm->keep = {sa, ss};
    Transform xf = make_xf(d->render_from_medium, d->medium_from_render);
    NvHandle dg = std::move(*(NvHandle *)density_h);
    NvHandle tg;
    if (temperature_h) tg = std::move(*(NvHandle *)temperature_h);
    rm->nvdb = new NanoVDBMedium(xf, sa, ss, d->sigma_scale, d->g, std::move(dg), std::move(tg), d->le_scale,
                                 d->temperature_offset, d->temperature_scale, alloc);
    rm->medium = rm->nvdb;
    return rm;
}
void ref_medium_bounds(void *m, float out[6]) {
    RefMedium *rm = (RefMedium *)m;
    const Bounds3f &b = rm->grid ? rm->grid->bounds : rm->rgb ? rm->rgb->bounds : rm->nvdb->bounds;
    for (int i = 0; i < 3; ++i) { out[i] = b.pMin[i]; out[3 + i] = b.pMax[i]; }
}
void ref_medium_destroy(void *m) { delete (RefMedium *)m; }
int ref_medium_is_emissive(void *m) { return ((RefMedium *)m)->medium.IsEmissive() ? 1 : 0; }
void ref_medium_majorants(void *mp, float *out, int32_t res[3]) {
    RefMedium *rm = (RefMedium *)mp;
    const MajorantGrid &g = rm->grid ? rm->grid->majorantGrid : rm->rgb ? rm->rgb->majorantGrid : rm->nvdb->majorantGrid;
    for (int i = 0; i < 3; ++i) res[i] = g.res[i];
    for (size_t i = 0; i < g.voxels.size(); ++i) out[i] = g.voxels[i];
}
void ref_medium_sample_point(void *mp, const float *p, int64_t n, const float lambda[4], float *out) {
    RefMedium *rm = (RefMedium *)mp;
    SampledWavelengths swl = make_lambda(lambda);
    for (int64_t i = 0; i < n; ++i) {
        MediumProperties q = rm->medium.SamplePoint(Point3f(p[3 * i], p[3 * i + 1], p[3 * i + 2]), swl);
        for (int k = 0; k < 4; ++k) {
            out[12 * i + k] = q.sigma_a[k];
            out[12 * i + 4 + k] = q.sigma_s[k];
            out[12 * i + 8 + k] = q.Le[k];
        }
    }
}
// SampleRay prologue exactly as media.h:322-337 writes it (ApplyInverse + IntersectP)
void ref_medium_ray_to_medium(void *mp, const float *rays, int64_t n, float *out) {
    RefMedium *rm = (RefMedium *)mp;
    const Transform &xf = rm->grid ? rm->grid->renderFromMedium : rm->rgb ? rm->rgb->renderFromMedium : rm->nvdb->renderFromMedium;
    const Bounds3f &b = rm->grid ? rm->grid->bounds : rm->rgb ? rm->rgb->bounds : rm->nvdb->bounds;
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays + 7 * i;
        Float raytMax = r[6];
        Ray ray = xf.ApplyInverse(Ray(Point3f(r[0], r[1], r[2]), Vector3f(r[3], r[4], r[5])), &raytMax);
        Float t0 = 0, t1 = 0;
        bool hit = b.IntersectP(ray.o, ray.d, raytMax, &t0, &t1);
        float *q = out + 9 * i;
        q[0] = ray.o.x; q[1] = ray.o.y; q[2] = ray.o.z; q[3] = ray.d.x; q[4] = ray.d.y; q[5] = ray.d.z;
        q[6] = hit ? t0 : 0; q[7] = hit ? t1 : 0; q[8] = hit ? 1.f : 0.f;
    }
}

// SampleT_maj (media.h:730-806) driven with the three callbacks the fork uses (bodies follow
// src/graph/util.h:344-403 and cpu/integrators.cpp:1352-1380 statement by statement).
void ref_sample_tmaj(void *mp, int mode, const float *rays, int64_t n, const uint64_t *rngs, const float *u,
                     const float lambda[4], avr_tmaj_result *out) {
    RefMedium *rm = (RefMedium *)mp;
    SampledWavelengths swl = make_lambda(lambda);
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays + 7 * i;
        Ray ray(Point3f(r[0], r[1], r[2]), Vector3f(r[3], r[4], r[5]), 0.f, rm->medium);
        RNG rng(rngs[2 * i], rngs[2 * i + 1]);
        avr_tmaj_result res;
        res.t_event = -1;
        res.event = 2;
        res.n_fetch = 0;
        res.n_seg = -1;  // the reference iterator does not expose a segment count
        SampledSpectrum T_ray(1.f), r_l(1.f), r_u(1.f), T;
        Vector3f dn = Normalize(ray.d);
        if (mode == 0) {
            float Tr = 1;
            T = SampleT_maj(ray, r[6], u[i], rng, swl,
                            [&](Point3f p, MediumProperties mp, SampledSpectrum sigma_maj, SampledSpectrum T_maj) {
                                res.n_fetch++;
                                SampledSpectrum sigma_n = ClampZero(sigma_maj - mp.sigma_a - mp.sigma_s);
                                Tr *= sigma_n[0] / sigma_maj[0];
                                if (Tr == 0) return false;
                                return true;
                            });
            T_ray[0] = Tr;
        } else if (mode == 1) {
            T = SampleT_maj(ray, r[6], u[i], rng, swl,
                            [&](Point3f p, MediumProperties mp, SampledSpectrum sigma_maj, SampledSpectrum T_maj) {
                                res.n_fetch++;
                                Float pAbsorb = mp.sigma_a[0] / sigma_maj[0];
                                Float pScatter = mp.sigma_s[0] / sigma_maj[0];
                                Float pNull = std::max<Float>(0, 1 - pAbsorb - pScatter);
                                Float um = rng.Uniform<Float>();
                                int md = SampleDiscrete({pAbsorb, pScatter, pNull}, um);
                                if (md == 0 || md == 1) {
                                    res.event = md;
                                    // recover t from p = o + dn*t along the dominant axis
                                    int ax = std::abs(dn.x) > std::abs(dn.y) ? (std::abs(dn.x) > std::abs(dn.z) ? 0 : 2)
                                                                             : (std::abs(dn.y) > std::abs(dn.z) ? 1 : 2);
                                    res.t_event = (p[ax] - ray.o[ax]) / dn[ax];
                                    return false;
                                }
                                return true;
                            });
        } else {
            T = SampleT_maj(ray, r[6], u[i], rng, swl,
                            [&](Point3f p, MediumProperties mp, SampledSpectrum sigma_maj, SampledSpectrum T_maj) {
                                res.n_fetch++;
                                SampledSpectrum sigma_n = ClampZero(sigma_maj - mp.sigma_a - mp.sigma_s);
                                Float pdf = T_maj[0] * sigma_maj[0];
                                T_ray *= T_maj * sigma_n / pdf;
                                r_l *= T_maj * sigma_maj / pdf;
                                r_u *= T_maj * sigma_n / pdf;
                                SampledSpectrum Tr = T_ray / (r_l + r_u).Average();
                                if (Tr.MaxComponentValue() < 0.05f) {
                                    Float q = 0.75f;
                                    if (rng.Uniform<Float>() < q) T_ray = SampledSpectrum(0.);
                                    else T_ray /= 1 - q;
                                }
                                if (!T_ray) return false;
                                return true;
                            });
            T_ray *= T / T[0];
            r_l *= T / T[0];
            r_u *= T / T[0];
        }
        for (int k = 0; k < 4; ++k) { res.T_maj[k] = T[k]; res.Tr[k] = T_ray[k]; res.r_u[k] = r_u[k]; res.r_l[k] = r_l[k]; }
        out[i] = res;
    }
}

// Eigen SpMV exactly as LightingCalculator::ComputeFinalLight drives it
// (src/graph/lighting_calculator.cpp:23-59, types from lighting_calculator.h).
int ref_graph_propagate(int32_t n, int64_t nnz, const int64_t *rowptr, const int32_t *col, const float *val,
                        const float *L0, int32_t bounces, float *total) {
    using SparseMat = Eigen::SparseMatrix<float>;
    using SparseVec = Eigen::SparseVector<float>;
    std::vector<Eigen::Triplet<float>> entries;
    for (int r = 0; r < n; ++r)
        for (int64_t k = rowptr[r]; k < rowptr[r + 1]; ++k) entries.emplace_back(r, col[k], val[k]);
    SparseMat T(n, n);
    T.setFromTriplets(entries.begin(), entries.end());
    SparseVec lightVec(n);
    for (int i = 0; i < n; ++i) lightVec.coeffRef(i) = L0[i];
    SparseVec totalLight(lightVec);
    int curIteration = 0;
    if (bounces > 0) {
        SparseVec curLight = totalLight;
        for (; curIteration < bounces; ++curIteration) {
            curLight = T * curLight;
            bool invalidNumbers = false;
            for (int i = 0; i < n; ++i)
                if (IsNaN(curLight.coeff(i)) || IsInf(curLight.coeff(i))) { invalidNumbers = true; break; }
            if (invalidNumbers) break;
            totalLight += curLight;
        }
    }
    for (int i = 0; i < n; ++i) total[i] = totalLight.coeff(i);
    return curIteration;
}


// LightingCalculator::GetLightVector's per-ray estimator: the fork's own graph::ComputeRaysToSphere
// (src/graph/util.h:814-839) -> util::SampleTransmittance (:344-366) -> SampleT_maj, unmodified.
// samplingIndex = y * 65536 + x with Options->graph.samplingResolution = (65536, 65536).
void ref_graph_light_rays(void *mp, const float *rays, const int32_t *pixel, const float *ts, int64_t n,
                          int32_t iterations, int32_t seed, const float lambda[4], float *out) {
    RefMedium *rm = (RefMedium *)mp;
    static PBRTOptions opts;
    opts.graph.samplingResolution = Point2i(65536, 65536);
    Options = &opts;
    util::MediumData md;
    md.defaultLambda = make_lambda(lambda);
    md.medium = rm->medium;
    IndependentSampler is(iterations, seed);
    Sampler sampler(&is);
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays + 6 * i;
        RayDifferential ray(Point3f(r[0], r[1], r[2]), Vector3f(r[3], r[4], r[5]), 0.f, rm->medium);
        util::StartEndT se{};
        se.startScatterT = ts[2 * i];
        se.endScatterT = ts[2 * i] + ts[2 * i + 1];
        uint64_t idx = uint64_t(pixel[2 * i + 1]) * 65536ull + uint64_t(pixel[2 * i]);
        out[i] = graph::ComputeRaysToSphere(ray, std::nullopt, se, md, sampler, iterations, idx);
    }
}


// ---- src/graph: the fork's own FreeGraph (graph.cpp) and LightingCalculator (lighting_calculator.cpp) ----
struct RefGraph {
    graph::FreeGraph g;
    explicit RefGraph(float r) : g(r) {}
    RefGraph() = default;
};
// Build a FreeGraph the way FreeGraphBuilder / ReadFromStream do: AddVertex(id, p, data), AddEdge(id, from, to,
// data) (graph.cpp:202-235: drops edges with unknown endpoints, merges duplicates), AddPath + AddVertexToPath.
void *ref_graph_build(float vertex_radius, int32_t nv, const int32_t *vid, const float *pos, const int32_t *vtype,
                      const float *vlight, const int32_t *vsamples, const float *vrange, const int32_t *vcoors,
                      int32_t ne, const int32_t *eid, const int32_t *efrom, const int32_t *eto, const int32_t *esamples,
                      int32_t np, const int32_t *pid, const int64_t *poff, const int32_t *pverts) {
    RefGraph *rg = new RefGraph(vertex_radius);
    for (int i = 0; i < nv; ++i) {
        graph::VertexData d{vtype ? graph::RayVertexType(vtype[i]) : graph::none, vlight ? vlight[i] : -1.f,
                            vsamples ? vsamples[i] : 0, vrange ? vrange[i] : -1.f};
        graph::Vertex &v = rg->g.AddVertex(vid[i], Point3f(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]), d);
        if (vcoors) v.coors = Point3i(vcoors[3 * i], vcoors[3 * i + 1], vcoors[3 * i + 2]);
    }
    for (int i = 0; i < ne; ++i) rg->g.AddEdge(eid[i], efrom[i], eto[i], graph::EdgeData{esamples ? esamples[i] : 0});
    for (int i = 0; i < np; ++i) {
        graph::Path &p = rg->g.AddPath(pid[i], graph::PathData{});
        for (int64_t j = poff[i]; j < poff[i + 1]; ++j) rg->g.AddVertexToPath(pverts[j], p.id);
    }
    return rg;
}
void ref_graph_free(void *h) { delete (RefGraph *)h; }
// Graph::WriteToDisk (graph.cpp:419-433) minus its "files/graphs/" path prefix
void ref_graph_write(void *h, const char *path, const char *desc, const int32_t flags[5], const int32_t opts[3]) {
    RefGraph *rg = (RefGraph *)h;
    rg->g.streamFlags = graph::StreamFlags{flags[0] != 0, flags[1] != 0, flags[2] != 0, flags[3] != 0, flags[4] != 0};
    rg->g.streamOptions = graph::StreamOptions{opts[0] != 0, opts[1] != 0, opts[2] != 0};
    std::ofstream file(path);
    file << desc << '\n';  // NEW, graph.cpp:11
    rg->g.WriteToStream(file);
    file.close();
}
// FreeGraph::ReadFromDisk (graph.cpp:657-663) minus the path prefix
void *ref_graph_read(const char *path) {
    RefGraph *rg = new RefGraph();
    std::ifstream file(path);
    rg->g.ReadFromStream(file);
    file.close();
    return rg;
}
void ref_graph_counts(void *h, int32_t out[11], float *vertex_radius) {
    RefGraph *rg = (RefGraph *)h;
    const auto &f = rg->g.streamFlags;
    out[0] = f && f->useCoors; out[1] = f && f->useSamples; out[2] = f && f->useRayVertexTypes;
    out[3] = f && f->useLighting; out[4] = f && f->useRenderSearchRange;
    out[5] = rg->g.GetCurVertexId(); out[6] = rg->g.GetCurEdgeId(); out[7] = rg->g.GetCurPathId();
    out[8] = int(rg->g.GetVerticesConst().size()); out[9] = int(rg->g.GetEdgesConst().size());
    out[10] = int(rg->g.GetPathsConst().size());
    *vertex_radius = rg->g.GetVertexRadius();
}
// vertices / edges in ascending id order
void ref_graph_get(void *h, int32_t *vid, float *pos, int32_t *vtype, float *vlight, int32_t *vsamples, float *vrange,
                   int32_t *vcoors, int32_t *eid, int32_t *efrom, int32_t *eto, int32_t *esamples) {
    RefGraph *rg = (RefGraph *)h;
    std::vector<int> ids;
    for (auto &kv : rg->g.GetVerticesConst()) ids.push_back(kv.first);
    std::sort(ids.begin(), ids.end());
    for (size_t i = 0; i < ids.size(); ++i) {
        const graph::Vertex &v = rg->g.GetVerticesConst().at(ids[i]);
        vid[i] = v.id;
        pos[3 * i] = v.point.x; pos[3 * i + 1] = v.point.y; pos[3 * i + 2] = v.point.z;
        vtype[i] = int(v.data.type); vlight[i] = v.data.lightScalar; vsamples[i] = v.data.samples;
        vrange[i] = v.data.renderSearchRange;
        for (int k = 0; k < 3; ++k) vcoors[3 * i + k] = v.coors ? (*v.coors)[k] : 0;
    }
    ids.clear();
    for (auto &kv : rg->g.GetEdgesConst()) ids.push_back(kv.first);
    std::sort(ids.begin(), ids.end());
    for (size_t i = 0; i < ids.size(); ++i) {
        const graph::Edge &e = rg->g.GetEdgesConst().at(ids[i]);
        eid[i] = e.id; efrom[i] = e.from; eto[i] = e.to; esamples[i] = e.data.samples;
    }
}
// LightingCalculator::GetTransportMatrix (lighting_calculator.cpp:61-82) and ComputeFinalLight(lightVec,
// transportMat, 0) (:23-59), the fork's own member functions.  CSR out: rows ascending, columns ascending.
// Returns nnz; total (n floats) = vertex.data.lightScalar after ComputeFinalLight when L0 != NULL.
int64_t ref_graph_transport(void *h, int64_t cap, int64_t *rowptr, int32_t *col, float *val, const float *L0,
                            int32_t bounces, float *total, int32_t *iters) {
    RefGraph *rg = (RefGraph *)h;
    util::MediumData md;
    IndependentSampler is(1, 0);
    graph::LightingCalculatorConfig cfg{1, 1, {bounces}, false};
    graph::LightingCalculator lc(rg->g, md, Vector3f(0, 0, 1), Sampler(&is), cfg, true);
    graph::SparseMat T = lc.GetTransportMatrix();
    int n = int(rg->g.GetVerticesConst().size());
    std::vector<std::vector<std::pair<int, float>>> rows(n);
    for (int c = 0; c < T.outerSize(); ++c)
        for (graph::SparseMat::InnerIterator it(T, c); it; ++it) rows[it.row()].emplace_back(int(it.col()), it.value());
    int64_t k = 0;
    for (int r = 0; r < n; ++r) {
        rowptr[r] = k;
        std::sort(rows[r].begin(), rows[r].end());
        for (auto &e : rows[r]) {
            if (k < cap) { col[k] = e.first; val[k] = e.second; }
            ++k;
        }
    }
    rowptr[n] = k;
    if (L0) {
        graph::SparseVec lightVec(n);
        for (int i = 0; i < n; ++i) lightVec.coeffRef(i) = L0[i];
        *iters = lc.ComputeFinalLight(lightVec, T, 0);
        for (int i = 0; i < n; ++i) total[i] = rg->g.GetVerticesConst().at(i).data.lightScalar;
    }
    return k;
}

}  // extern "C"
#pragma GCC visibility pop

// handle -> the reference Medium object (for refprobe_volpath.cpp)
extern "C" __attribute__((visibility("default"))) const pbrt::Medium *ref_medium_ptr(void *m) {
    return &((RefMedium *)m)->medium;
}
