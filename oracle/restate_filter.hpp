// oracle/restate_filter.hpp — TEST INFRASTRUCTURE ONLY (CPU restatement; the product never includes it).
//
// pbrt's tabulated pixel-filter sampling, restated: FilterSampler (src/pbrt/filters.h:26-45, filters.cpp:133-147) over
// PiecewiseConstant2D / PiecewiseConstant1D (src/pbrt/util/sampling.h:603-675, 696-790) and FindInterval
// (src/pbrt/util/math.h).  Pinned bit for bit against the reference's own GaussianFilter / MitchellFilter objects through
// the probe (refprobe_adapter.cpp: ref_filter_sample), tests/test_filter.py.
#pragma once
#include <cmath>
#include <vector>

#include "restate.hpp"

namespace avro {

struct Pc1d {  // PiecewiseConstant1D, util/sampling.h:625-649
    std::vector<float> func, cdf;
    float lo = 0, hi = 1, func_int = 0;
    void build(const float *f, int n, float lo_, float hi_) {
        lo = lo_; hi = hi_;
        func.assign(f, f + n);
        for (float &v : func) v = std::abs(v);
        cdf.assign(n + 1, 0.f);
        for (int i = 1; i < n + 1; ++i) cdf[i] = cdf[i - 1] + func[i - 1] * (hi - lo) / n;
        func_int = cdf[n];
        if (func_int == 0)
            for (int i = 1; i < n + 1; ++i) cdf[i] = float(i) / float(n);
        else
            for (int i = 1; i < n + 1; ++i) cdf[i] /= func_int;
    }
    // util/sampling.h:657-675 with FindInterval (util/math.h) written out
    float sample(float u, float *pdf, int *offset) const {
        int sz = (int)cdf.size(), size = sz - 2, first = 1;
        while (size > 0) {
            int half = size >> 1, middle = first + half;
            bool pr = cdf[middle] <= u;
            first = pr ? middle + 1 : first;
            size = pr ? size - (half + 1) : half;
        }
        int o = std::min(std::max(first - 1, 0), sz - 2);
        *offset = o;
        float du = u - cdf[o];
        if (cdf[o + 1] - cdf[o] > 0) du /= cdf[o + 1] - cdf[o];
        *pdf = func_int > 0 ? func[o] / func_int : 0.f;
        return lerp((o + du) / func.size(), lo, hi);
    }
};

// SampleLinear / SampleTent, util/sampling.h:122-128,196-201; SampleDiscrete({.5, .5}, u, nullptr, &u) (:79-113) is
// written out: the weights sum to 1, the first is passed over when .5 <= u
inline float sample_linear(float u, float a, float b) {
    if (u == 0 && a == 0) return 0;
    float x = u * (a + b) / (a + std::sqrt(lerp(u, a * a, b * b)));
    return std::min(x, kOneMinusEps);
}
inline float sample_tent(float u, float r) {
    float sum_w = 0.f;
    sum_w += 0.5f; sum_w += 0.5f;
    float up = u * sum_w;
    if (up == sum_w) up = std::nextafter(up, 0.f);
    int offset = 0;
    float sum = 0;
    const float w[2] = {0.5f, 0.5f};
    while (sum + w[offset] <= up) sum += w[offset++];
    float ur = std::min((up - sum) / w[offset], kOneMinusEps);
    if (offset == 0) return -r + r * sample_linear(ur, 0, 1);
    return r * sample_linear(ur, 1, 0);
}

struct FilterTable {  // FilterSampler: domain [-radius, radius]^2, f tabulated nx x ny, distrib = PiecewiseConstant2D(f)
    int nx = 0, ny = 0;
    bool triangle = false;  // TriangleFilter (filters.h:224-227): analytic, weight 1
    float trx = 0, try_ = 0;
    std::vector<float> f;
    std::vector<Pc1d> rows;
    Pc1d marginal;
    bool active() const { return nx > 0 || triangle; }
    void build_triangle(float rx, float ry) { triangle = true; trx = rx; try_ = ry; }
    void build(const float *values, int nx_, int ny_, float rx, float ry) {
        nx = nx_; ny = ny_;
        f.assign(values, values + size_t(nx) * ny);
        rows.resize(ny);
        std::vector<float> m(ny);
        for (int y = 0; y < ny; ++y) {
            rows[y].build(f.data() + size_t(y) * nx, nx, -rx, rx);
            m[y] = rows[y].func_int;
        }
        marginal.build(m.data(), ny, -ry, ry);
    }
    // FilterSampler::Sample (filters.h:33-38) = PiecewiseConstant2D::Sample (util/sampling.h:760-769) + f[pi] / pdf
    void sample(float u0, float u1, float *px, float *py, float *weight) const {
        if (triangle) {
            *px = sample_tent(u0, trx);
            *py = sample_tent(u1, try_);
            *weight = 1.f;
            return;
        }
        float pdf0, pdf1;
        int ox, oy;
        float d1 = marginal.sample(u1, &pdf1, &oy);
        float d0 = rows[oy].sample(u0, &pdf0, &ox);
        float pdf = pdf0 * pdf1;
        *px = d0;
        *py = d1;
        *weight = f[size_t(oy) * nx + ox] / pdf;
    }
};

}  // namespace avro
