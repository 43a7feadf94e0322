// oracle/refprobe_volpath.cpp — TEST INFRASTRUCTURE ONLY.
//
// Probe over the reference's OWN VolPathIntegrator (src/pbrt/cpu/integrators.cpp:962-1106 Li, :1282-1399
// SampleLd), compiled from the sources where they lie under /root/reference together with its Sphere,
// GeometricPrimitive, DistantLight, BVHLightSampler and IndependentSampler.  Builds the scene class of
// include/avr.h in the reference's own objects (one grid medium inside a sphere with Material
// "interface", one distant light) and evaluates pixel samples the way
// RayIntegrator::EvaluatePixelSample does (:235-298): StartPixelSample, wavelengths from Get1D,
// GetCameraSample (samplers.h:797-814), then Li on the camera ray.  Camera rays are passed in (the film /
// camera classes need the reference's generated rgb2spec tables and OpenEXR and are not built here).
// Everything the integrator references but never reaches for this scene class (surface materials,
// textures, image IO, other shapes) is closed at link time by generated abort() stubs, see Makefile.
#include <cstring>

#include <pbrt/cpu/aggregates.h>
#include <pbrt/cpu/integrators.h>
#include <pbrt/cpu/primitive.h>
#include <pbrt/cameras.h>
#include <pbrt/film.h>
#include <pbrt/filters.h>
#include <pbrt/lights.h>
#include <pbrt/media.h>
#include <pbrt/samplers.h>
#include <pbrt/shapes.h>
#include <pbrt/util/buffercache.h>
#include <pbrt/util/mesh.h>
#include <pbrt/util/spectrum.h>

using namespace pbrt;

extern "C" const pbrt::Medium *ref_medium_ptr(void *m);  // refprobe.cpp

#pragma GCC visibility push(default)
extern "C" {

// samples: per sample {px, py, sampleIndex}; rays: per sample {o.xyz, d.xyz} (render space, as the oracle's
// camera generates them).  out: 12 floats per sample {L[4], lambda[4], pdf[4]} (lambda/pdf after Li, which
// may terminate secondary wavelengths).  info: {wi.xyz of the light as DistantLight::SampleLi computes it,
// sceneRadius after Preprocess}.
void ref_volpath_li(void *medium_h, const float sphere[4], const float light_dir[3], const float *light_spec,
                    int32_t max_depth, int32_t sampler_seed, int32_t spp, const int32_t *samples, const float *rays,
                    int64_t n, float *out, float *info, const int32_t sampler_cfg[5]) {
    // sampler_cfg = {0 independent | 1 zsobol, RandomizeStrategy, film width, film height,
    //                boundary: 0 = Shape "sphere" (sphere[0..3] = centre, radius),
    //                          1 = Shape "trianglemesh" box (sphere[0..5] = min xyz, max xyz), 12 outward-wound
    //                              triangles under the default BVH aggregate}
    // --disable-texture-filtering: SurfaceInteraction::ComputeDifferentials (interaction.cpp:41-47) then
    // skips the camera-based footprint estimate, the only use Li makes of the camera (none is built
    // here); it feeds texture lookups only, which the "interface" boundary does not have.
    static PBRTOptions opts;
    opts.disableTextureFiltering = true;
    Options = &opts;
    Allocator alloc;
    Medium medium = *ref_medium_ptr(medium_h);
    Primitive aggregate;
    if (sampler_cfg[4] == 0) {
        // Shape "sphere" at Translate(c): shapes.cpp Sphere::Create defaults zmin = -radius, zmax = radius, phimax = 360
        Transform *rfo = new Transform(Translate(Vector3f(sphere[0], sphere[1], sphere[2])));
        Transform *ofr = new Transform(Inverse(*rfo));
        Sphere *sph = new Sphere(rfo, ofr, false, sphere[3], -sphere[3], sphere[3], 360.f);
        // Material "interface" = null material; MediumInterface(inside = medium, outside = none)
        aggregate = new GeometricPrimitive(Shape(sph), Material(), Light(), MediumInterface(medium, nullptr));
    } else {
        // Shape "trianglemesh": the 8 corners, 12 triangles with outward geometric normals (SURVEY appendix A:
        // an inward-wound cube assigns the medium to the wrong side); BVHAggregate defaults of the scene
        // loader (maxnodeprims 4, SAH: cpu/aggregates.cpp BVHAggregate::Create)
        static bool init = false;
        if (!init) {
            InitBufferCaches();
            Triangle::Init(alloc);
            init = true;
        }
        const float *b = sphere;
        std::vector<Point3f> P;
        for (int i = 0; i < 8; ++i) P.push_back(Point3f(b[(i & 1) ? 3 : 0], b[(i & 2) ? 4 : 1], b[(i & 4) ? 5 : 2]));
        std::vector<int> idx = {0, 2, 1, 1, 2, 3,   // z = min (normal -z)
                                4, 5, 6, 5, 7, 6,   // z = max (+z)
                                0, 1, 4, 1, 5, 4,   // y = min (-y)
                                2, 6, 3, 3, 6, 7,   // y = max (+y)
                                0, 4, 2, 2, 4, 6,   // x = min (-x)
                                1, 3, 5, 3, 7, 5};  // x = max (+x)
        TriangleMesh *mesh = new TriangleMesh(Transform(), false, idx, P, {}, {}, {}, {}, alloc);
        pstd::vector<Shape> tris = Triangle::CreateTriangles(mesh, alloc);
        std::vector<Primitive> prims;
        for (Shape sh : tris)
            prims.push_back(new GeometricPrimitive(sh, Material(), Light(), MediumInterface(medium, nullptr)));
        aggregate = new BVHAggregate(prims, 4, BVHAggregate::SplitMethod::SAH);
    }
    // LightSource "distant": renderFromLight maps +z to the direction towards the light
    Vector3f w = Normalize(Vector3f(light_dir[0], light_dir[1], light_dir[2]));
    Frame f = Frame::FromZ(w);
    SquareMatrix<4> m(f.x.x, f.y.x, f.z.x, 0, f.x.y, f.y.y, f.z.y, 0, f.x.z, f.y.z, f.z.z, 0, 0, 0, 0, 1);
    Transform renderFromLight(m);
    DenselySampledSpectrum *Lemit = new DenselySampledSpectrum(
        DenselySampledSpectrum::SampleFunction([&](Float l) { return light_spec[int(l) - 360]; }));
    DistantLight *dl = new DistantLight(renderFromLight, Lemit, 1.f);
    std::vector<Light> lights = {Light(dl)};
    IndependentSampler proto(spp, sampler_seed);
    VolPathIntegrator integrator(max_depth, Camera(), Sampler(&proto), aggregate, lights, "bvh", false);
    Vector3f wi = Normalize(renderFromLight(Vector3f(0, 0, 1)));
    info[0] = wi.x; info[1] = wi.y; info[2] = wi.z;
    info[3] = dl->sceneRadius;
    Filter filter = new BoxFilter(Vector2f(0.5f, 0.5f));
    ScratchBuffer scratch(65536);
    for (int64_t i = 0; i < n; ++i) {
        IndependentSampler is(spp, sampler_seed);
        ZSobolSampler zs(spp, Point2i(sampler_cfg[2], sampler_cfg[3]), RandomizeStrategy(sampler_cfg[1]), sampler_seed);
        Sampler sampler = sampler_cfg[0] == 1 ? Sampler(&zs) : Sampler(&is);
        Point2i pPixel(samples[3 * i], samples[3 * i + 1]);
        sampler.StartPixelSample(pPixel, samples[3 * i + 2]);
        Float lu = sampler.Get1D();
        SampledWavelengths lambda = SampledWavelengths::SampleVisible(lu);  // RGBFilm::SampleWavelengths, film.h
        CameraSample cs = GetCameraSample(sampler, pPixel, filter);
        (void)cs;
        const float *r = rays + 6 * i;
        RayDifferential ray(Point3f(r[0], r[1], r[2]), Vector3f(r[3], r[4], r[5]), 0.f, Medium());
        SampledSpectrum L = integrator.Li(ray, lambda, sampler, scratch, nullptr);
        scratch.Reset();
        for (int k = 0; k < 4; ++k) {
            out[12 * i + k] = L[k];
            out[12 * i + 4 + k] = lambda[k];
            out[12 * i + 8 + k] = lambda.PDF()[k];
        }
    }
}

// The reference's inline film code: PixelSensor::ToSensorRGB (film.h:95-100), RGBFilm::AddSample (:239-255) and
// RGBFilm::GetPixelRGB (:257-275).  Their constructors need RGBColorSpace objects (generated rgb2spec tables,
// not buildable here) that none of the three methods reads: sensor and film are laid out in zeroed storage
// and only the members those methods use are constructed (the three response curves, imagingRatio,
// maxComponentValue, filterIntegral = 1 for the splat term, outputRGBFromSensorRGB, the pixel array).
// samples: pixels {px, py}, values {L[4], lambda[4], pdf[4]}, weight 1, in the order given.
// film: width*height*4 doubles {rgbSum[3], weightSum}; rgb: width*height*3 floats from GetPixelRGB.
void ref_film_add_samples(int32_t width, int32_t height, const float *sensor_rgb, float imaging_ratio,
                          float max_component, const float out_from_sensor[9], const int32_t *pixels, const float *vals,
                          int64_t n, double *film, float *rgb_out) {
    Allocator alloc;
    PixelSensor *sensor = (PixelSensor *)::operator new(sizeof(PixelSensor));
    std::memset((void *)sensor, 0, sizeof(PixelSensor));
    auto curve = [&](int c) {
        return DenselySampledSpectrum::SampleFunction([&](Float l) { return sensor_rgb[c * 471 + (int(l) - 360)]; });
    };
    new (&sensor->r_bar) DenselySampledSpectrum(curve(0));
    new (&sensor->g_bar) DenselySampledSpectrum(curve(1));
    new (&sensor->b_bar) DenselySampledSpectrum(curve(2));
    sensor->imagingRatio = imaging_ratio;
    RGBFilm *f = (RGBFilm *)::operator new(sizeof(RGBFilm));
    std::memset((void *)f, 0, sizeof(RGBFilm));
    f->sensor = sensor;
    f->fullResolution = Point2i(width, height);
    f->pixelBounds = Bounds2i(Point2i(0, 0), Point2i(width, height));
    f->maxComponentValue = max_component;
    f->filterIntegral = 1;
    const float *M = out_from_sensor;
    f->outputRGBFromSensorRGB = SquareMatrix<3>(M[0], M[1], M[2], M[3], M[4], M[5], M[6], M[7], M[8]);
    new (&f->pixels) Array2D<RGBFilm::Pixel>(f->pixelBounds, alloc);
    for (int64_t i = 0; i < n; ++i) {
        const float *v = vals + 12 * i;
        SampledSpectrum L(0.f);
        SampledWavelengths lambda;
        for (int k = 0; k < 4; ++k) { L[k] = v[k]; lambda.lambda[k] = v[4 + k]; lambda.pdf[k] = v[8 + k]; }
        f->AddSample(Point2i(pixels[2 * i], pixels[2 * i + 1]), L, lambda, nullptr, 1.f);
    }
    for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
            const RGBFilm::Pixel &p = f->pixels[Point2i(x, y)];
            double *o = film + 4 * (size_t(y) * width + x);
            o[0] = p.rgbSum[0]; o[1] = p.rgbSum[1]; o[2] = p.rgbSum[2]; o[3] = p.weightSum;
            RGB c = f->GetPixelRGB(Point2i(x, y));
            float *q = rgb_out + 3 * (size_t(y) * width + x);
            q[0] = c.r; q[1] = c.g; q[2] = c.b;
        }
}

// The reference's camera: the REAL PerspectiveCamera constructor (cameras.h:343-370 over ProjectiveCamera and
// CameraBase; world-space rendering so that render space == world space as include/avr.h assumes) over a film
// that only has to answer FullResolution() and GetFilter() (an RGBFilm laid out in zeroed storage, see
// ref_film_add_samples), then, per pixel sample, the prologue of RayIntegrator::EvaluatePixelSample
// (cpu/integrators.cpp:235-262): StartPixelSample, the wavelength draw, GetCameraSample (samplers.h:797-814),
// camera.GenerateRayDifferential (cameras.cpp:428-490).
// world_from_camera: row-major 4x4; screen_window = {xmin, ymin, xmax, ymax}.  rays: {o, d} per sample;
// matrices: cameraFromRaster then renderFromCamera, row-major 4x4 each.
// filter_cfg (may be NULL = box 0.5): {kind, radius.x, radius.y, p0, p1} as make_filter below; weights (may be NULL):
// CameraSample::filterWeight per sample.
static Filter make_filter(const float *cfg) {
    if (!cfg) return new BoxFilter(Vector2f(0.5f, 0.5f));
    Vector2f radius(cfg[1], cfg[2]);
    switch (int(cfg[0])) {
    case 1: return new GaussianFilter(radius, cfg[3]);         // "gaussian": sigma (filters.cpp:114, default 0.5)
    case 2: return new MitchellFilter(radius, cfg[3], cfg[4]);  // "mitchell": B, C
    case 3: return new LanczosSincFilter(radius, cfg[3]);      // "sinc": tau
    case 4: return new TriangleFilter(radius);                 // "triangle"
    default: return new BoxFilter(radius);
    }
}

// Filter::Sample(u) of the reference's own filter objects: out = {p.x, p.y, weight} per point.
void ref_filter_sample(const float *filter_cfg, const float *u, int64_t n, float *out) {
    Filter filter = make_filter(filter_cfg);
    for (int64_t i = 0; i < n; ++i) {
        FilterSample fs = filter.Sample(Point2f(u[2 * i], u[2 * i + 1]));
        out[3 * i] = fs.p.x; out[3 * i + 1] = fs.p.y; out[3 * i + 2] = fs.weight;
    }
}

// The table FilterSampler's constructor builds (filters.cpp:133-143): nxny = {int(32 rx), int(32 ry)}; f may be NULL
// to query the size first.
void ref_filter_table(const float *filter_cfg, int32_t nxny[2], float *f) {
    Filter filter = make_filter(filter_cfg);
    Vector2f r = filter.Radius();
    Bounds2f domain(Point2f(-r.x, -r.y), Point2f(r.x, r.y));
    int nx = int(32 * r.x), ny = int(32 * r.y);
    nxny[0] = nx; nxny[1] = ny;
    if (!f) return;
    for (int y = 0; y < ny; ++y)
        for (int x = 0; x < nx; ++x) {
            Point2f p = domain.Lerp(Point2f((x + 0.5f) / nx, (y + 0.5f) / ny));
            f[y * nx + x] = filter.Evaluate(p);
        }
}

void ref_camera_rays(const float world_from_camera[16], float fov, const float screen_window[4], int32_t width,
                     int32_t height, int32_t sampler_seed, int32_t spp, const int32_t *samples, int64_t n, float *rays,
                     float *matrices, const float *filter_cfg, float *weights) {
    static PBRTOptions opts;
    opts.disableTextureFiltering = true;
    opts.renderingSpace = RenderingCoordinateSystem::World;
    Options = &opts;
    Allocator alloc;
    Filter filter = make_filter(filter_cfg);
    RGBFilm *f = (RGBFilm *)::operator new(sizeof(RGBFilm));
    std::memset((void *)f, 0, sizeof(RGBFilm));
    f->fullResolution = Point2i(width, height);
    f->pixelBounds = Bounds2i(Point2i(0, 0), Point2i(width, height));
    f->filter = filter;
    const float *w = world_from_camera;
    Transform worldFromCamera(SquareMatrix<4>(w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7], w[8], w[9], w[10], w[11], w[12],
                                              w[13], w[14], w[15]));
    CameraTransform cameraTransform{AnimatedTransform(worldFromCamera, 0, worldFromCamera, 1)};
    CameraBaseParameters base;
    base.cameraTransform = cameraTransform;
    base.film = Film(f);
    PerspectiveCamera *cam = new PerspectiveCamera(
        base, fov, Bounds2f(Point2f(screen_window[0], screen_window[1]), Point2f(screen_window[2], screen_window[3])), 0.f,
        1e6f);
    Camera camera(cam);
    const SquareMatrix<4> &a = cam->cameraFromRaster.GetMatrix();
    const SquareMatrix<4> &b = cam->cameraTransform.RenderFromCamera().startTransform.GetMatrix();
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) { matrices[4 * i + j] = a[i][j]; matrices[16 + 4 * i + j] = b[i][j]; }
    for (int64_t i = 0; i < n; ++i) {
        IndependentSampler is(spp, sampler_seed);
        Sampler sampler(&is);
        Point2i pPixel(samples[3 * i], samples[3 * i + 1]);
        sampler.StartPixelSample(pPixel, samples[3 * i + 2]);
        Float lu = sampler.Get1D();
        SampledWavelengths lambda = SampledWavelengths::SampleVisible(lu);
        CameraSample cs = GetCameraSample(sampler, pPixel, filter);
        if (weights) weights[i] = cs.filterWeight;
        pstd::optional<CameraRayDifferential> cr = camera.GenerateRayDifferential(cs, lambda);
        float *q = rays + 6 * i;
        q[0] = cr->ray.o.x; q[1] = cr->ray.o.y; q[2] = cr->ray.o.z;
        q[3] = cr->ray.d.x; q[4] = cr->ray.d.y; q[5] = cr->ray.d.z;
    }
}

// ZSobolSampler streams (samplers.h:225-364): per sample {px, py, sampleIndex}, `rounds` times
// {Get1D, Get2D} after StartPixelSample -> 3 floats per round.
void ref_zsobol_samples(int32_t spp, int32_t resx, int32_t resy, int32_t randomize, int32_t seed, const int32_t *samples,
                        int64_t n, int32_t rounds, float *out) {
    ZSobolSampler zs(spp, Point2i(resx, resy), RandomizeStrategy(randomize), seed);
    for (int64_t i = 0; i < n; ++i) {
        zs.StartPixelSample(Point2i(samples[3 * i], samples[3 * i + 1]), samples[3 * i + 2], 0);
        for (int r = 0; r < rounds; ++r) {
            float *q = out + (i * rounds + r) * 3;
            q[0] = zs.Get1D();
            Point2f u = zs.Get2D();
            q[1] = u.x;
            q[2] = u.y;
        }
    }
}

// first two dimensions of util/sobolmatrices.cpp: 2 * SobolMatrixSize words
void ref_sobol_matrices(uint32_t *out) {
    for (int i = 0; i < 2 * SobolMatrixSize; ++i) out[i] = SobolMatrices32[i];
}

}  // extern "C"
#pragma GCC visibility pop
