// oracle/refprobe_graphbuild.cpp — TEST INFRASTRUCTURE ONLY.
//
// Probe over the reference's OWN graph pipeline, compiled from the sources where they lie under
// /root/reference and driven the way cmd/graph_maker.cpp:86-197 drives it:
//   FreeGraphBuilder::BuildGraph (src/graph/free/free_graph_builder.cpp:143-214: TracePath :19-141,
//   ReinforceSparseVertices off, ComputeSearchRanges :498-548)  ->  LightingCalculator::GetLightVector
//   (src/graph/lighting_calculator.cpp:84-155)  ->  GetTransportMatrix (:61-82)  ->  ComputeFinalLight (:23-59)
// on a scene of the class the fork accepts (src/graph/util.h:61-132): one grid medium inside ONE
// Shape "sphere" of Material "interface" under a BVHAggregate, one distant light.  The sampler is the
// IndependentSampler graph_maker would get from a scene that names it (pixelSamples =
// RoundUpPow2(lightIterations), cmd/graph_maker.cpp:95), seed 0.
#include <cmath>
#include <cstring>
#include <iostream>
#include <vector>

#include <graph/free/free_graph_builder.h>
#include <graph/lighting_calculator.h>
#include <pbrt/cpu/aggregates.h>
#include <pbrt/cpu/primitive.h>
#include <pbrt/lights.h>
#include <pbrt/options.h>
#include <pbrt/samplers.h>
#include <pbrt/shapes.h>

using namespace pbrt;

extern "C" const pbrt::Medium *ref_medium_ptr(void *m);  // refprobe.cpp

namespace {
struct Pipeline {
    int nv = 0, ne = 0, resolution = 0, pixel_samples = 0;
    float vertex_radius = 0, max_dist_to_center = 0, lambda[4] = {0, 0, 0, 0}, center[3] = {0, 0, 0};
    std::vector<float> pos, range, L0, totals;
    std::vector<int32_t> vsamples, efrom, eto, esamples, iters, list_index;
};
}  // namespace

#pragma GCC visibility push(default)
extern "C" {

// icfg = {dimensionSteps, iterationsPerStep, maxDepth, lightIterations, pointsOnRadiusLight,
//         renderSearchRange.active, renderSearchRange.neighboursToUse, sampler seed}; fcfg = {radiusModifier}
// light_wi: direction TOWARDS the light (what avr_render_desc.light_dir holds).
void *ref_graphpipe_run(void *medium_h, const float sphere[4], const float light_wi[3], const int32_t icfg[8],
                        const float fcfg[1], const int32_t *bounces, int32_t nb) {
    graph::Config cfg{};
    cfg.graphBuilder.radiusModifier = fcfg[0];
    cfg.graphBuilder.maxDepth = icfg[2];
    cfg.graphBuilder.renderSearchRangeConfig = {icfg[5] != 0, icfg[6], false};
    cfg.graphBuilder.dimensionSteps = icfg[0];
    cfg.graphBuilder.iterationsPerStep = icfg[1];
    cfg.graphBuilder.edgeReinforcement = {};
    cfg.graphBuilder.neighbourReinforcement = {};
    cfg.lightingCalculator.lightIterations = icfg[3];
    cfg.lightingCalculator.pointsOnRadiusLight = icfg[4];
    cfg.lightingCalculator.bounces.assign(bounces, bounces + nb);
    cfg.lightingCalculator.runInParallel = false;

    // cmd/graph_maker.cpp:86-108
    int maxDiskPoints = util::GetDiskPointsSize(cfg.lightingCalculator.pointsOnRadiusLight);
    int64_t maxNumVertices = int64_t(Sqr(cfg.graphBuilder.dimensionSteps)) * cfg.graphBuilder.iterationsPerStep * cfg.graphBuilder.maxDepth;
    int64_t maxNumRays = maxNumVertices * maxDiskPoints;
    int res = int(std::ceil(std::sqrt(double(maxNumRays))));
    static PBRTOptions opts;
    opts.disablePixelJitter = true;
    opts.disableWavelengthJitter = true;
    opts.disableTextureFiltering = true;
    opts.pixelSamples = RoundUpPow2(cfg.lightingCalculator.lightIterations);
    opts.graph.samplingResolution = Point2i(res, res);
    Options = &opts;

    Medium medium = *ref_medium_ptr(medium_h);
    Transform *rfo = new Transform(Translate(Vector3f(sphere[0], sphere[1], sphere[2])));
    Transform *ofr = new Transform(Inverse(*rfo));
    Sphere *sph = new Sphere(rfo, ofr, false, sphere[3], -sphere[3], sphere[3], 360.f);
    std::vector<Primitive> prims = {
        Primitive(new GeometricPrimitive(Shape(sph), Material(), Light(), MediumInterface(medium, nullptr)))};
    Primitive accel = new BVHAggregate(prims, 4, BVHAggregate::SplitMethod::SAH);

    // Film::SampleWavelengths(0) of the scene's film: SampleVisible (film.h SampleWavelengths)
    util::MediumData mediumData(accel, SampledWavelengths::SampleVisible(0));
    Vector3f wi = Normalize(Vector3f(light_wi[0], light_wi[1], light_wi[2]));
    Vector3f lightDir = -wi;  // graph_maker.cpp:128: -Normalize(renderFromLight(0, 0, 1))
    IndependentSampler proto(*opts.pixelSamples, icfg[7]);
    Sampler sampler(&proto);

    std::streambuf *old = std::cout.rdbuf(nullptr);
    graph::FreeGraphBuilder builder(mediumData, lightDir, sampler, cfg.graphBuilder, true, {});
    graph::FreeGraph g = builder.BuildGraph();
    graph::LightingCalculator lighting(g, mediumData, lightDir, sampler, cfg.lightingCalculator, true);
    // the order GetLightVector walks the vertex map in (lighting_calculator.cpp:91-94): it numbers the sampler pixels
    std::vector<int32_t> list_index(g.GetVertices().size(), -1);
    {
        int32_t k = 0;
        for (auto &[id, v] : g.GetVertices()) list_index[id] = k++;
    }
    graph::SparseVec lightVec = lighting.GetLightVector();
    graph::SparseMat transportMat = lighting.GetTransportMatrix();

    Pipeline *p = new Pipeline();
    p->list_index = list_index;
    p->nv = int(g.GetVertices().size());
    p->ne = int(g.GetEdges().size());
    p->resolution = res;
    p->pixel_samples = *opts.pixelSamples;
    p->vertex_radius = g.GetVertexRadius();
    p->max_dist_to_center = mediumData.primitiveData.maxDistToCenter;
    for (int i = 0; i < 4; ++i) p->lambda[i] = mediumData.defaultLambda[i];
    for (int i = 0; i < 3; ++i) p->center[i] = mediumData.primitiveData.boundsCenter[i];
    p->pos.resize(3 * size_t(p->nv)); p->range.resize(p->nv); p->L0.resize(p->nv); p->vsamples.resize(p->nv);
    for (auto &[id, v] : g.GetVertices()) {  // ids are sequential (CheckSequentialIds in the calculator's constructor)
        p->pos[3 * id] = v.point.x; p->pos[3 * id + 1] = v.point.y; p->pos[3 * id + 2] = v.point.z;
        p->range[id] = v.data.renderSearchRange;
        p->vsamples[id] = v.data.samples;
        p->L0[id] = lightVec.coeff(id);
    }
    for (auto &[id, e] : g.GetEdges()) {
        p->efrom.push_back(e.from); p->eto.push_back(e.to); p->esamples.push_back(e.data.samples);
    }
    p->totals.resize(size_t(nb) * p->nv);
    p->iters.resize(nb);
    for (int b = 0; b < nb; ++b) {
        p->iters[b] = lighting.ComputeFinalLight(lightVec, transportMat, b);
        for (auto &[id, v] : g.GetVertices()) p->totals[size_t(b) * p->nv + id] = v.data.lightScalar;
    }
    std::cout.rdbuf(old);
    std::cout.clear();
    return p;
}

// {vertices, edges, samplingResolution, pixelSamples}; info = {vertexRadius, maxDistToCenter, lambda[4], boundsCenter[3]}
void ref_graphpipe_info(void *h, int32_t counts[4], float info[9]) {
    Pipeline *p = (Pipeline *)h;
    counts[0] = p->nv; counts[1] = p->ne; counts[2] = p->resolution; counts[3] = p->pixel_samples;
    info[0] = p->vertex_radius; info[1] = p->max_dist_to_center;
    for (int i = 0; i < 4; ++i) info[2 + i] = p->lambda[i];
    for (int i = 0; i < 3; ++i) info[6 + i] = p->center[i];
}

void ref_graphpipe_get(void *h, float *pos, int32_t *vsamples, float *vrange, int32_t *efrom, int32_t *eto,
                       int32_t *esamples, float *L0, float *totals, int32_t *iters, int32_t *list_index) {
    Pipeline *p = (Pipeline *)h;
    std::memcpy(pos, p->pos.data(), p->pos.size() * 4);
    std::memcpy(vsamples, p->vsamples.data(), p->vsamples.size() * 4);
    std::memcpy(vrange, p->range.data(), p->range.size() * 4);
    std::memcpy(efrom, p->efrom.data(), p->efrom.size() * 4);
    std::memcpy(eto, p->eto.data(), p->eto.size() * 4);
    std::memcpy(esamples, p->esamples.data(), p->esamples.size() * 4);
    std::memcpy(L0, p->L0.data(), p->L0.size() * 4);
    std::memcpy(totals, p->totals.data(), p->totals.size() * 4);
    std::memcpy(iters, p->iters.data(), p->iters.size() * 4);
    std::memcpy(list_index, p->list_index.data(), p->list_index.size() * 4);
}

void ref_graphpipe_free(void *h) { delete (Pipeline *)h; }

}  // extern "C"
#pragma GCC visibility pop
