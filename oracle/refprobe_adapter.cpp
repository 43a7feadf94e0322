// oracle/refprobe_adapter.cpp — TEST INFRASTRUCTURE ONLY.
// Compiles integration/avr_adapter.h against the reference's headers (in place, /root/reference) and runs it on
// the reference's REAL objects — the GridMedium / NanoVDBMedium the probe constructs (refprobe.cpp), a
// PerspectiveCamera over an RGBFilm, a DistantLight after Preprocess, IndependentSampler / ZSobolSampler, a Sphere —
// handing the descriptors the adapter fills back to the tests (tests/test_adapter.py).
#include "../integration/avr_adapter.h"

#include <pbrt/options.h>
#include <pbrt/shapes.h>

using namespace pbrt;
using A = avr_adapter::Access;

extern "C" const pbrt::Medium *ref_medium_ptr(void *m);  // refprobe.cpp

#pragma GCC visibility push(default)
extern "C" {

// kind: 0 grid, 2 nanovdb.  Returns an adapter descriptor holder (free with ref_adapter_free_medium); *desc points
// at the avr_grid_desc / avr_nanovdb_desc inside it.
void *ref_adapter_describe_medium(void *medium_h, int32_t *kind, const void **desc) {
    const Medium &m = *ref_medium_ptr(medium_h);
    if (m.Is<GridMedium>()) {
        auto *h = new A::GridDesc();
        A::Describe(*m.Cast<GridMedium>(), h);
        *kind = 0; *desc = &h->d;
        return h;
    }
    if (m.Is<NanoVDBMedium>()) {
        auto *h = new A::NanoVDBDesc();
        A::Describe(*m.Cast<NanoVDBMedium>(), h);
        *kind = 2; *desc = &h->d;
        return h;
    }
    *kind = -1; *desc = nullptr;
    return nullptr;
}
// the NanoVDBMedium with its density grid as 8^3 blocks (A::FillBlocks)
void *ref_adapter_describe_medium_blocks(void *medium_h, const void **desc) {
    const Medium &m = *ref_medium_ptr(medium_h);
    if (!m.Is<NanoVDBMedium>()) { *desc = nullptr; return nullptr; }
    auto *h = new A::NanoVDBDesc();
    A::Describe(*m.Cast<NanoVDBMedium>(), h, true);
    *desc = &h->d;
    return h;
}
void ref_adapter_free_medium(void *h, int32_t kind) {
    if (kind == 0) delete (A::GridDesc *)h;
    else if (kind == 2) delete (A::NanoVDBDesc *)h;
}

// Builds the reference objects of a scene the way the scene loader would and lets the adapter describe them.
// cfg = {width, height, spp, sampler (0 independent, 1 zsobol), randomize, seed, filter (0 box, 1 gaussian = pbrt's default, 2 triangle, 3 mitchell)}
void *ref_adapter_describe_render(const float world_from_camera[16], float fov, const float screen_window[4],
                                  const int32_t cfg[7], const float light_wi[3], const float *light_spec, float light_scale,
                                  const float sphere[4], const float *sensor_rgb, float imaging_ratio,
                                  const float output_rgb[9], const avr_render_desc **desc) {
    static PBRTOptions opts;
    opts.disableTextureFiltering = true;
    opts.renderingSpace = RenderingCoordinateSystem::World;
    Options = &opts;
    Allocator alloc;
    int width = cfg[0], height = cfg[1];
    Filter filter;
    if (cfg[6] == 0) filter = new BoxFilter(Vector2f(0.5f, 0.5f));
    else if (cfg[6] == 1) filter = new GaussianFilter(Vector2f(1.5f, 1.5f), 0.5f, alloc);  // filters.cpp:114 defaults
    else if (cfg[6] == 2) filter = new TriangleFilter(Vector2f(2.f, 2.f));
    else filter = new MitchellFilter(Vector2f(2.f, 2.f), 1.f / 3.f, 1.f / 3.f, alloc);
    // RGBFilm / PixelSensor want RGBColorSpace objects (rgb2spec tables: not buildable here): laid out in zeroed
    // storage with the members the adapter reads constructed, as oracle/refprobe_volpath.cpp does for the film tests
    RGBFilm *f = (RGBFilm *)::operator new(sizeof(RGBFilm));
    std::memset((void *)f, 0, sizeof(RGBFilm));
    f->fullResolution = Point2i(width, height);
    f->pixelBounds = Bounds2i(Point2i(0, 0), Point2i(width, height));
    f->filter = filter;
    PixelSensor *sensor = (PixelSensor *)::operator new(sizeof(PixelSensor));
    std::memset((void *)sensor, 0, sizeof(PixelSensor));
    auto dense = [&](const float *t) {
        return DenselySampledSpectrum(DenselySampledSpectrum::SampleFunction([t](Float l) { return t[int(l) - 360]; }));
    };
    new (&sensor->r_bar) DenselySampledSpectrum(dense(sensor_rgb));
    new (&sensor->g_bar) DenselySampledSpectrum(dense(sensor_rgb + AVR_NSPEC));
    new (&sensor->b_bar) DenselySampledSpectrum(dense(sensor_rgb + 2 * AVR_NSPEC));
    sensor->imagingRatio = imaging_ratio;
    const float *w = world_from_camera;
    Transform worldFromCamera(SquareMatrix<4>(w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7], w[8], w[9], w[10], w[11], w[12],
                                              w[13], w[14], w[15]));
    CameraBaseParameters base;
    base.cameraTransform = CameraTransform{AnimatedTransform(worldFromCamera, 0, worldFromCamera, 1)};
    base.film = Film(f);
    PerspectiveCamera *cam = new PerspectiveCamera(
        base, fov, Bounds2f(Point2f(screen_window[0], screen_window[1]), Point2f(screen_window[2], screen_window[3])), 0.f, 1e6f);
    // LightSource "distant": renderFromLight maps +z to the direction towards the light
    Vector3f wi = Normalize(Vector3f(light_wi[0], light_wi[1], light_wi[2]));
    Frame fr = Frame::FromZ(wi);
    Transform renderFromLight(SquareMatrix<4>(fr.x.x, fr.y.x, fr.z.x, 0, fr.x.y, fr.y.y, fr.z.y, 0, fr.x.z, fr.y.z, fr.z.z, 0, 0, 0, 0, 1));
    DenselySampledSpectrum *Lemit = new DenselySampledSpectrum(dense(light_spec));
    DistantLight *dl = new DistantLight(renderFromLight, Lemit, light_scale);
    Transform *rfo = new Transform(Translate(Vector3f(sphere[0], sphere[1], sphere[2])));
    Transform *ofr = new Transform(Inverse(*rfo));
    Sphere *sph = new Sphere(rfo, ofr, false, sphere[3], -sphere[3], sphere[3], 360.f);
    dl->Preprocess(sph->Bounds());   // the scene bounds are the boundary shape's (the only primitive)
    Sampler sampler;
    if (cfg[3] == 0) sampler = new IndependentSampler(cfg[2], cfg[5]);
    else sampler = new ZSobolSampler(cfg[2], Point2i(width, height), RandomizeStrategy(cfg[4]), cfg[5]);

    auto *h = new A::RenderDesc();
    A::DescribeCamera(*cam, h);
    bool ok = A::DescribeFilter(filter, h);
    SquareMatrix<3> out(output_rgb[0], output_rgb[1], output_rgb[2], output_rgb[3], output_rgb[4], output_rgb[5], output_rgb[6],
                        output_rgb[7], output_rgb[8]);
    A::DescribeSensor(*sensor, out, Infinity, h);
    ok = A::DescribeSampler(sampler, h) && ok;
    A::DescribeLight(*dl, h);
    A::DescribeBoundarySphere(*sph, h);
    h->d.max_depth = 5;
    h->d.exact_math = 1;
    *desc = ok ? &h->d : nullptr;
    return h;
}
void ref_adapter_free_render(void *h) { delete (A::RenderDesc *)h; }

}  // extern "C"
#pragma GCC visibility pop
