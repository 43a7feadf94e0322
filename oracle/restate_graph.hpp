// oracle/restate_graph.hpp — TEST INFRASTRUCTURE ONLY (see restate.hpp header).
// src/graph radiance propagation and lookup: LightingCalculator::ComputeFinalLight
// (lighting_calculator.cpp:23-59), GetTransportMatrix (:61-82), GraphIntegrator::ConnectToGraph
// (graph_integrator.cpp:249-280) + Initialize search-range statistics (:55-73), util::Averager
// (util.h:529-568), ComputeRaysToSphere (util.h:814-839).
#pragma once
#include <utility>

#include "restate_tmaj.hpp"

namespace avro {

struct Csr {
    int n = 0;
    std::vector<int64_t> rowptr;
    std::vector<int32_t> col;
    std::vector<float> val;
};

// Eigen: SparseMatrix<float> (column-major) * SparseVector<float>; per output row the products are
// accumulated in increasing column order starting from 0 (checked against the vendored Eigen in
// oracle/refprobe.cpp).  CSR rows here hold columns in increasing order.
inline void spmv(const Csr &A, const float *x, float *y) {
    for (int r = 0; r < A.n; ++r) {
        float acc = 0;
        for (int64_t k = A.rowptr[r]; k < A.rowptr[r + 1]; ++k) acc += A.val[k] * x[A.col[k]];
        y[r] = acc;
    }
}

// ComputeFinalLight, lighting_calculator.cpp:23-59.  Returns curIteration.
inline int compute_final_light(const Csr &T, const float *L0, int bounces, float *total) {
    std::vector<float> cur(L0, L0 + T.n), nxt(T.n);
    for (int i = 0; i < T.n; ++i) total[i] = L0[i];
    int it = 0;
    for (; it < bounces; ++it) {
        spmv(T, cur.data(), nxt.data());
        bool invalid = false;
        for (int i = 0; i < T.n; ++i)
            if (std::isnan(nxt[i]) || std::isinf(nxt[i])) { invalid = true; break; }
        cur.swap(nxt);
        if (invalid) break;
        for (int i = 0; i < T.n; ++i) total[i] += cur[i];
    }
    return it;
}

struct GraphVerts {
    int n = 0;
    std::vector<float> pos;  // 3n, world space
    std::vector<float> light, range;
    bool use_range = false;
    float vertex_radius = 0;
    float sq_radius = 0, max_sq_range = -1, perc99_sq_range = 0;
    void finalize() {  // GraphIntegrator::Initialize, graph_integrator.cpp:53-73
        sq_radius = sqr(vertex_radius);
        if (use_range) {
            std::vector<float> s(n);
            for (int i = 0; i < n; ++i) s[i] = sqr(range[i]);
            std::sort(s.begin(), s.end());
            max_sq_range = s.back();
            perc99_sq_range = s[size_t(n) * 99 / 100];
        }
    }
};

// ConnectToGraph, graph_integrator.cpp:249-280.  nanoflann radiusSearch returns matches with
// squared L2 distance <= radius... (nanoflann RadiusResultSet::addPoint uses dist < radius) sorted
// by distance ascending; ties keep index order.
inline float connect_to_graph(const GraphVerts &g, V3 p) {
    std::vector<std::pair<float, int>> res;
    auto search = [&](float r2, bool filter) {
        res.clear();
        for (int i = 0; i < g.n; ++i) {
            float dx = p.x - g.pos[3 * i], dy = p.y - g.pos[3 * i + 1], dz = p.z - g.pos[3 * i + 2];
            float d2 = dx * dx + dy * dy + dz * dz;  // L2_Simple_Adaptor: sum of squared diffs in order
            if (d2 < r2) {
                if (filter && d2 > sqr(g.range[i])) continue;
                res.emplace_back(d2, i);
            }
        }
        std::stable_sort(res.begin(), res.end(), [](auto &a, auto &b) { return a.first < b.first; });
    };
    search(g.sq_radius, false);
    if (g.use_range && res.empty()) {
        search(g.perc99_sq_range, true);
        if (res.empty()) search(g.max_sq_range, true);
    }
    if (res.empty()) return 0;
    float average = 0, totalWeight = 0;
    for (auto &r : res) {
        float w = 1.f / r.first;
        average += g.light[r.second] * w;
        totalWeight += w;
    }
    if (totalWeight == 0) return 0;
    return average / totalWeight;
}

// ComputeRaysToSphere (rayInSphere == nullopt branch), util.h:814-839 + SampleTransmittance :344-366.
inline float light_ray_transmittance(const Medium &m, V3 o, V3 d, int px, int py, float startScatterT,
                                     float distInSphere, int iterations, int seed, const float lambda[4]) {
    float average = 0, totalWeight = 0;  // util::Averager
    for (int i = 0; i < iterations; ++i) {
        IndependentSampler s;
        s.seed = seed;
        s.start_pixel_sample(px, py, i);
        float curDistInSphere = distInSphere * s.get1d();
        float tMax = startScatterT + curDistInSphere;
        // util.h:346 `RNG rng(Hash(sampler.Get1D()), Hash(sampler.Get1D()))`: the two draws are
        // unsequenced in the reference; GCC (the reference's Linux compiler, and the refprobe's)
        // evaluates constructor arguments right to left, so the FIRST draw seeds `offset` and the
        // second the sequence index.  Pinned against refprobe.
        uint64_t h1 = hash_float(s.get1d());
        uint64_t h0 = hash_float(s.get1d());
        float u = s.get1d();
        TmajResult r = run_tmaj(m, 0, o, d, tMax, h0, h1, u, lambda);
        average += r.Tr[0] * 1.f;
        totalWeight += 1.f;
    }
    if (iterations == 0 || totalWeight == 0) return 0;
    return average / totalWeight;
}

}  // namespace avro
