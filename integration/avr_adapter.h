// integration/avr_adapter.h — the adapter translation unit a maintainer adds to the reference tree
// (tsvdh/AcceleratedVolRenderer, a pbrt-v4 fork) to put libavr_b200.so behind its Medium / Integrator switches
// (Medium::Create, src/pbrt/media.cpp:667-689; Integrator::Create, src/pbrt/cpu/integrators.cpp:3658-3700).
//
// It fills the C-ABI descriptors of include/avr.h from the reference's OWN constructed objects, so whatever the
// scene parser produced is what the device gets.  The members read here are private in the reference: in the
// reference tree add `friend struct avr_adapter::Access;` to GridMedium, NanoVDBMedium, HGPhaseFunction, GaussianFilter, MitchellFilter, LanczosSincFilter, FilterSampler,
// DistantLight, IndependentSampler, ZSobolSampler and PixelSensor (this repository's test compiles the header
// with -fno-access-control instead, oracle/Makefile).  No CUDA headers are needed on the pbrt side.
//
// This header IS compiled: oracle/refprobe_adapter.cpp instantiates every function against the reference's
// headers and tests/test_adapter.py checks the descriptors it produces, field by field, against the ones
// scenes.py builds for the same scene, and renders through them on the GPU.
#pragma once
#include <avr.h>

#include <pbrt/cameras.h>
#include <pbrt/film.h>
#include <pbrt/filters.h>
#include <pbrt/lights.h>
#include <pbrt/media.h>
#include <pbrt/samplers.h>

#include <algorithm>
#include <cstring>
#include <vector>

namespace avr_adapter {

struct Access {
    // Densely-sampled 360..830 nm table of a spectrum (avr.h: AVR_NSPEC = 471)
    static std::vector<float> Table(const pbrt::DenselySampledSpectrum &s) {
        std::vector<float> t(AVR_NSPEC);
        for (int i = 0; i < AVR_NSPEC; ++i) t[i] = s(pbrt::Float(360 + i));
        return t;
    }
    static void Matrix(const pbrt::SquareMatrix<4> &m, float out[16]) {
        for (int i = 0; i < 4; ++i)
            for (int j = 0; j < 4; ++j) out[4 * i + j] = m[i][j];
    }
    template <typename Desc>
    static void Transforms(const pbrt::Transform &renderFromMedium, Desc *d) {
        Matrix(renderFromMedium.GetMatrix(), d->render_from_medium);
        Matrix(renderFromMedium.GetInverseMatrix(), d->medium_from_render);
    }
    static std::vector<float> Grid(const pbrt::SampledGrid<pbrt::Float> &g) {
        return std::vector<float>(g.begin(), g.end());
    }

    // ---- "uniformgrid": GridMedium (media.h:265-352, constructed by GridMedium::Create, media.cpp:249-336) ----
    // The constructor has already applied "scale" to the spectra (media.cpp:229-230) and LeNorm to the LeScale grid
    // (:295-306): the tables go over as they are, with sigma_scale = le_norm = 1.
    struct GridDesc {
        avr_grid_desc d{};
        std::vector<float> density, temperature, lescale, sigma_a, sigma_s, le;
    };
    static void Describe(const pbrt::GridMedium &m, GridDesc *o) {
        avr_grid_desc &d = o->d;
        d.nx = m.densityGrid.XSize(); d.ny = m.densityGrid.YSize(); d.nz = m.densityGrid.ZSize();
        o->density = Grid(m.densityGrid);
        d.density = o->density.data();
        if (m.temperatureGrid) { o->temperature = Grid(*m.temperatureGrid); d.temperature = o->temperature.data(); }
        // a 1x1x1 LeScale grid is the "no Lescale parameter" case (media.cpp:293-294): its one value is LeNorm
        if (m.LeScale.XSize() * m.LeScale.YSize() * m.LeScale.ZSize() == 1) {
            d.le_norm = *m.LeScale.begin();
        } else {
            o->lescale = Grid(m.LeScale);
            d.lescale = o->lescale.data();
            d.le_norm = 1;
        }
        for (int i = 0; i < 3; ++i) { d.p0[i] = m.bounds.pMin[i]; d.p1[i] = m.bounds.pMax[i]; }
        Transforms(m.renderFromMedium, &d);
        o->sigma_a = Table(m.sigma_a_spec); o->sigma_s = Table(m.sigma_s_spec); o->le = Table(m.Le_spec);
        d.sigma_a = o->sigma_a.data(); d.sigma_s = o->sigma_s.data();
        bool hasLe = false;
        for (float v : o->le) hasLe = hasLe || v > 0;
        d.le = hasLe ? o->le.data() : nullptr;
        d.sigma_scale = 1;
        d.g = m.phase.g;
        d.temperature_offset = m.temperatureOffset; d.temperature_scale = m.temperatureScale;
        d.majorant_res[0] = m.majorantGrid.res.x; d.majorant_res[1] = m.majorantGrid.res.y; d.majorant_res[2] = m.majorantGrid.res.z;
    }

    // ---- "nanovdb": NanoVDBMedium (media.h:602-685, media.cpp:511-665) --------------------------------------
    // Each FloatGrid is flattened over its index box the way the fork's own cmd/nanovdb2pbrt.cpp:100-110 does it.
    struct NanoVDBDesc {
        avr_nanovdb_desc d{};
        std::vector<float> density, temperature, sigma_a, sigma_s;
        std::vector<int32_t> block_origin;   // density as 8^3 blocks (Describe(..., blocks = true))
    };
    static void FillGrid(const nanovdb::FloatGrid *fg, avr_nanovdb_grid *g, std::vector<float> *values) {
        auto ib = fg->indexBBox();  // inclusive on both ends
        auto acc = fg->getAccessor();
        for (int z = ib.min()[2]; z <= ib.max()[2]; ++z)
            for (int y = ib.min()[1]; y <= ib.max()[1]; ++y)
                for (int x = ib.min()[0]; x <= ib.max()[0]; ++x) values->push_back(acc.getValue(nanovdb::Coord(x, y, z)));
        g->values = values->data();
        for (int i = 0; i < 3; ++i) {
            g->index_min[i] = ib.min()[i];
            g->dims[i] = ib.max()[i] - ib.min()[i] + 1;
            g->vec[i] = fg->map().mVecF[i];
            g->world_bbox[i] = fg->worldBBox().min()[i];
            g->world_bbox[3 + i] = fg->worldBBox().max()[i];
        }
        for (int i = 0; i < 9; ++i) g->inv_mat[i] = fg->map().mInvMatF[i];
    }
    // The grid WITHOUT flattening (avr_nanovdb_grid.block_*): every leaf node that holds a non-zero value as it is
    // (transposed from NanoVDB's z-fastest order), and the non-zero value tiles of the internal nodes and of the root
    // cut into 8^3 blocks where they meet the index box.  For layout AVR_LAYOUT_BRICK_F16: host memory and upload are
    // proportional to the leaves, not to the box.
    static void FillBlocks(const nanovdb::FloatGrid *fg, avr_nanovdb_grid *g, std::vector<int32_t> *origins,
                           std::vector<float> *values) {
        using Tree = nanovdb::FloatGrid::TreeType;
        using Leaf = Tree::Node0;
        using Lower = Tree::Node1;
        using Upper = Tree::Node2;
        const Tree &tree = fg->tree();
        auto ib = fg->indexBBox();
        auto tile = [&](nanovdb::Coord o, int extent, float v) {
            if (v == 0) return;
            int a[3], b[3];
            for (int i = 0; i < 3; ++i) {
                a[i] = std::max(o[i], ib.min()[i]);
                b[i] = std::min(o[i] + extent - 1, ib.max()[i]);
                if (a[i] > b[i]) return;
            }
            for (int z = a[2] & ~7; z <= b[2]; z += 8)
                for (int y = a[1] & ~7; y <= b[1]; y += 8)
                    for (int x = a[0] & ~7; x <= b[0]; x += 8) {
                        origins->insert(origins->end(), {x, y, z});
                        values->insert(values->end(), 512, v);
                    }
        };
        const Leaf *leaf = tree.getFirstNode<Leaf>();
        for (uint32_t i = 0; leaf && i < tree.nodeCount<Leaf>(); ++i) {
            float tmp[512];
            bool any = false;
            for (int x = 0; x < 8; ++x)
                for (int y = 0; y < 8; ++y)
                    for (int z = 0; z < 8; ++z) {
                        float v = leaf[i].getValue(uint32_t((x << 6) | (y << 3) | z));
                        tmp[x + 8 * y + 64 * z] = v;
                        any |= v != 0;
                    }
            if (!any) continue;
            nanovdb::Coord o = leaf[i].origin();
            origins->insert(origins->end(), {o[0], o[1], o[2]});
            values->insert(values->end(), tmp, tmp + 512);
        }
        const Lower *lower = tree.getFirstNode<Lower>();
        for (uint32_t i = 0; lower && i < tree.nodeCount<Lower>(); ++i)
            for (uint32_t n = 0; n < 4096; ++n)
                if (!lower[i].childMask().isOn(n)) tile(lower[i].offsetToGlobalCoord(n), 8, lower[i].data()->mTable[n].value);
        const Upper *upper = tree.getFirstNode<Upper>();
        for (uint32_t i = 0; upper && i < tree.nodeCount<Upper>(); ++i)
            for (uint32_t n = 0; n < 32768; ++n)
                if (!upper[i].childMask().isOn(n)) tile(upper[i].offsetToGlobalCoord(n), 128, upper[i].data()->mTable[n].value);
        const auto &root = tree.root();
        for (uint32_t t = 0; t < root.tileCount(); ++t) {
            const auto *rt = root.data()->tile(t);
            if (!rt->isChild()) tile(rt->origin(), 4096, rt->value);
        }
        if (values->empty()) values->push_back(0.f);   // block_values != NULL marks the grid as present
        g->values = nullptr;
        g->n_blocks = int64_t(origins->size() / 3);
        g->block_origin = origins->data();
        g->block_values = values->data();
        for (int i = 0; i < 3; ++i) {
            g->index_min[i] = ib.min()[i];
            g->dims[i] = ib.max()[i] - ib.min()[i] + 1;
            g->vec[i] = fg->map().mVecF[i];
            g->world_bbox[i] = fg->worldBBox().min()[i];
            g->world_bbox[3 + i] = fg->worldBBox().max()[i];
        }
        for (int i = 0; i < 9; ++i) g->inv_mat[i] = fg->map().mInvMatF[i];
    }
    // blocks = true: the density grid as 8^3 blocks when the medium has no temperature grid (for
    // AVR_LAYOUT_BRICK_F16); the flattened box otherwise
    static void Describe(const pbrt::NanoVDBMedium &m, NanoVDBDesc *o, bool blocks = false) {
        avr_nanovdb_desc &d = o->d;
        if (blocks && !m.temperatureFloatGrid) FillBlocks(m.densityFloatGrid, &d.density, &o->block_origin, &o->density);
        else FillGrid(m.densityFloatGrid, &d.density, &o->density);
        if (m.temperatureFloatGrid) FillGrid(m.temperatureFloatGrid, &d.temperature, &o->temperature);
        Transforms(m.renderFromMedium, &d);
        o->sigma_a = Table(m.sigma_a_spec); o->sigma_s = Table(m.sigma_s_spec);
        d.sigma_a = o->sigma_a.data(); d.sigma_s = o->sigma_s.data();
        d.sigma_scale = 1;  // applied by the constructor (media.cpp:519-520)
        d.g = m.phase.g;
        d.le_scale = m.LeScale; d.temperature_offset = m.temperatureOffset; d.temperature_scale = m.temperatureScale;
        d.majorant_res[0] = m.majorantGrid.res.x; d.majorant_res[1] = m.majorantGrid.res.y; d.majorant_res[2] = m.majorantGrid.res.z;
    }

    // ---- camera, film, filter, sampler, light -> avr_render_desc -------------------------------------------------
    struct RenderDesc {
        avr_render_desc d{};
        std::vector<float> light, sensor;
        avr_filter_table filter{};       // referenced by d.filter when the pixel filter is a tabulated one
        std::vector<float> filter_values;
    };
    // PerspectiveCamera (cameras.h:343-370, cameras.cpp:404-460) over an RGBFilm (film.h:214-300)
    static void DescribeCamera(const pbrt::PerspectiveCamera &cam, RenderDesc *o) {
        avr_render_desc &d = o->d;
        pbrt::Film film = cam.GetFilm();
        pbrt::Point2i res = film.FullResolution();
        pbrt::Bounds2i pb = film.PixelBounds();
        d.width = res.x; d.height = res.y;
        d.pixel_bounds[0] = pb.pMin.x; d.pixel_bounds[1] = pb.pMin.y; d.pixel_bounds[2] = pb.pMax.x; d.pixel_bounds[3] = pb.pMax.y;
        Matrix(cam.cameraFromRaster.GetMatrix(), d.camera_from_raster);
        Matrix(cam.cameraTransform.RenderFromCamera().startTransform.GetMatrix(), d.render_from_camera);
    }
    // PixelFilter.  "box" (filters.h:48-82): analytic, weight 1.  "gaussian" (pbrt's default, scene.cpp:94), "mitchell",
    // "sinc": pbrt samples them through a FilterSampler (filters.h:26-45); its table `f` — the filter evaluated at the
    // cell centres of [-radius, radius]^2, filters.cpp:133-143 — goes over as it is and the library rebuilds the
    // sampling distribution from it.  "triangle" (filters.h:208-231): analytic tent sampling with weight 1, a `kind`
    // of the same descriptor.  Anything else: false = the caller keeps the stock integrator.
    static bool DescribeFilter(pbrt::Filter filter, RenderDesc *o) {
        pbrt::Vector2f r = filter.Radius();
        o->d.filter_radius[0] = r.x; o->d.filter_radius[1] = r.y;
        o->d.filter = nullptr;
        if (filter.Is<pbrt::BoxFilter>()) return true;
        o->filter = avr_filter_table{};
        if (filter.Is<pbrt::TriangleFilter>()) {
            o->filter.kind = AVR_FILTER_TRIANGLE;
            o->d.filter = &o->filter;
            return true;
        }
        const pbrt::FilterSampler *fs = nullptr;
        if (filter.Is<pbrt::GaussianFilter>()) fs = &filter.Cast<pbrt::GaussianFilter>()->sampler;
        else if (filter.Is<pbrt::MitchellFilter>()) fs = &filter.Cast<pbrt::MitchellFilter>()->sampler;
        else if (filter.Is<pbrt::LanczosSincFilter>()) fs = &filter.Cast<pbrt::LanczosSincFilter>()->sampler;
        if (!fs) return false;
        o->filter_values.assign(fs->f.begin(), fs->f.end());
        o->filter.nx = fs->f.XSize();
        o->filter.ny = fs->f.YSize();
        o->filter.f = o->filter_values.data();
        o->d.filter = &o->filter;
        return true;
    }
    // PixelSensor (film.h:60-130) + RGBFilm's outputRGBFromSensorRGB and maxComponentValue (film.h:239-275)
    static void DescribeSensor(const pbrt::PixelSensor &s, const pbrt::SquareMatrix<3> &outputRGBFromSensorRGB,
                               pbrt::Float maxComponentValue, RenderDesc *o) {
        std::vector<float> r = Table(s.r_bar), g = Table(s.g_bar), b = Table(s.b_bar);
        o->sensor = r;
        o->sensor.insert(o->sensor.end(), g.begin(), g.end());
        o->sensor.insert(o->sensor.end(), b.begin(), b.end());
        o->d.sensor_rgb = o->sensor.data();
        o->d.imaging_ratio = s.imagingRatio;
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) o->d.output_rgb_from_sensor[3 * i + j] = outputRGBFromSensorRGB[i][j];
        o->d.max_component_value = maxComponentValue;
    }
    // Sampler "independent" (samplers.h:442-476) or "zsobol" (:225-364); false = not on this path
    static bool DescribeSampler(pbrt::Sampler sampler, RenderDesc *o) {
        avr_render_desc &d = o->d;
        d.spp = sampler.SamplesPerPixel();
        if (sampler.Is<pbrt::IndependentSampler>()) {
            d.sampler = AVR_SAMPLER_INDEPENDENT;
            d.sampler_seed = sampler.Cast<pbrt::IndependentSampler>()->seed;
            return true;
        }
        if (sampler.Is<pbrt::ZSobolSampler>()) {
            const pbrt::ZSobolSampler *z = sampler.Cast<pbrt::ZSobolSampler>();
            d.sampler = AVR_SAMPLER_ZSOBOL;
            d.sampler_seed = z->seed;
            d.randomize = int(z->randomize);  // RandomizeStrategy and AVR_RANDOMIZE_* share their order
            return true;
        }
        return false;
    }
    // DistantLight after Preprocess(sceneBounds) (lights.h:262-304): direction towards the light, L = scale * Lemit
    static void DescribeLight(const pbrt::DistantLight &l, RenderDesc *o) {
        pbrt::Vector3f wi = pbrt::Normalize(l.renderFromLight(pbrt::Vector3f(0, 0, 1)));
        o->d.light_dir[0] = wi.x; o->d.light_dir[1] = wi.y; o->d.light_dir[2] = wi.z;
        o->light = Table(*l.Lemit);
        for (float &v : o->light) v *= l.scale;
        o->d.light_spectrum = o->light.data();
        o->d.scene_radius = l.sceneRadius;
    }
    // the "interface" shape around the medium: a Sphere (shapes.h) translated to its centre
    static void DescribeBoundarySphere(const pbrt::Sphere &s, RenderDesc *o) {
        pbrt::Point3f c = (*s.renderFromObject)(pbrt::Point3f(0, 0, 0));
        o->d.boundary_type = AVR_BOUNDARY_SPHERE;
        o->d.boundary[0] = c.x; o->d.boundary[1] = c.y; o->d.boundary[2] = c.z; o->d.boundary[3] = s.radius;
    }
};

// What Integrator::Create's new case "volpath-b200" runs instead of ImageTileIntegrator::Render
// (cpu/integrators.cpp:72-226): one C-ABI call; sums = width * height * {rgbSum[3], weightSum} (RGBFilm::Pixel).
// A non-zero status is the caller's cue to ErrorExit with avr_last_error (or to hand the job to the stock integrator
// when it is AVR_ERR_UNSUPPORTED).
inline avr_status RenderVolPath(avr_ctx *ctx, avr_medium *medium, const Access::RenderDesc &r, int maxDepth, bool exact,
                                std::vector<double> *sums, avr_render_stats *stats) {
    avr_render_desc d = r.d;
    d.max_depth = maxDepth;
    d.exact_math = exact ? 1 : 0;
    sums->assign(size_t(d.width) * d.height * 4, 0.0);
    return avr_render_volpath(ctx, medium, &d, sums->data(), stats);
}

}  // namespace avr_adapter
