// oracle/refprobe_graphint.cpp — TEST INFRASTRUCTURE ONLY.
//
// Probe over the reference's OWN GraphIntegrator::Initialize (src/graph/graph_integrator.cpp:24-80: graph
// file -> search-range statistics -> nanoflann kd-tree) and GraphIntegrator::ConnectToGraph (:249-280),
// compiled from the source where it lies under /root/reference.  The integrator's constructor wants a
// camera with a film (rgb2spec tables, OpenEXR: not buildable here), which neither of the two methods
// uses; the probe therefore lays the object out in zeroed storage, constructs the one member the methods
// read without assigning first (worldFromRender = identity: render space == world space, the
// "cameraworld"-free case) and calls the two non-virtual methods.  The graph comes in through the
// reference's own reader from a file the reference's own writer produced (tests write it with
// ref_graph_write under <dir>/files/graphs/, where util::FileNameToPath looks).
#include <unistd.h>

#include <cstring>
#include <iostream>

#include <graph/graph_integrator.h>
#include <pbrt/options.h>

using namespace pbrt;

#pragma GCC visibility push(default)
extern "C" {

void *ref_graph_integrator_open(const char *dir, const char *name) {
    static PBRTOptions opts;
    opts.graph.dataFile = std::string(name);
    Options = &opts;
    char cwd[4096];
    if (!getcwd(cwd, sizeof(cwd)) || chdir(dir) != 0) return nullptr;
    void *mem = ::operator new(sizeof(graph::GraphIntegrator));
    std::memset(mem, 0, sizeof(graph::GraphIntegrator));
    graph::GraphIntegrator *gi = reinterpret_cast<graph::GraphIntegrator *>(mem);
    new (&gi->worldFromRender) Transform();
    std::streambuf *old = std::cout.rdbuf(nullptr);  // TaskTimeTracker progress lines
    gi->Initialize();
    std::cout.rdbuf(old);
    std::cout.clear();
    int rc = chdir(cwd);
    (void)rc;
    return gi;
}

// {squaredVertexRadius, perc99SquaredSearchRange, maxSquaredSearchRange, number of vertices}
void ref_graph_integrator_info(void *h, float out[4]) {
    graph::GraphIntegrator *gi = (graph::GraphIntegrator *)h;
    out[0] = gi->squaredVertexRadius;
    out[1] = gi->perc99SquaredSearchRange;
    out[2] = gi->maxSquaredSearchRange;
    out[3] = float(gi->freeGraph->GetVertices().size());
}

void ref_graph_connect(void *h, const float *pts, int64_t n, float *out) {
    graph::GraphIntegrator *gi = (graph::GraphIntegrator *)h;
    for (int64_t i = 0; i < n; ++i) out[i] = gi->ConnectToGraph(Point3f(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]));
}

void ref_graph_integrator_close(void *h) {
    graph::GraphIntegrator *gi = (graph::GraphIntegrator *)h;
    if (!gi) return;
    gi->searchTree.reset();
    gi->freeGraph.reset();
    ::operator delete(gi);  // vHolder's vector leaks with the probe object: test infrastructure
}

}  // extern "C"
#pragma GCC visibility pop
