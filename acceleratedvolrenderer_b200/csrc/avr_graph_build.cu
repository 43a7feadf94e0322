// avr_graph_build.cu — LightingCalculator::GetTransportMatrix (src/graph/lighting_calculator.cpp:61-82)
// on the device: edge list -> CSR transport matrix.
//
// The reference walks unordered_map<int,Vertex> and each vertex's std::map outEdges (other vertex id ->
// edge id) and hands (row, col, edgeSamples/vertexSamples) triplets to Eigen::setFromTriplets.  What
// reaches the SpMV is therefore: one entry per distinct (from, to), columns ascending inside a row,
// duplicates of (from, to) already merged by Graph::AddEdge (graph.cpp:202-235, samples added), edges
// with an endpoint that is not a vertex never created.  Here: 64-bit keys (from << 32 | to), one radix
// sort, one reduce-by-key, rowptr by binary search over the unique keys (no atomics, deterministic).
// The sort / segmented reduction primitives come from CUB (part of the CUDA toolkit); they are
// load-time plumbing, the propagation kernel itself is k_spmv (avr_graph.cu).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_reduce.cuh>

#include "avr_internal.h"

using namespace avr;

namespace {

constexpr unsigned long long kInvalidKey = ~0ull;

__global__ void k_edge_keys(const int32_t *__restrict__ from, const int32_t *__restrict__ to, int64_t n_edges,
                            int32_t n_vertices, unsigned long long *__restrict__ keys) {
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n_edges) return;
    int32_t f = from[i], t = to[i];
    bool ok = (unsigned)f < (unsigned)n_vertices && (unsigned)t < (unsigned)n_vertices;  // vertices.find(id) hits
    keys[i] = ok ? ((unsigned long long)(unsigned)f << 32) | (unsigned)t : kInvalidKey;
}

// entries [0, nnz): col = low word, val = float(edge samples) / float(vertex samples of the row)
__global__ void k_fill_csr(const unsigned long long *__restrict__ ukeys, const int32_t *__restrict__ sums,
                           const int32_t *__restrict__ vertex_samples, const int *__restrict__ num_runs,
                           int32_t *__restrict__ col, float *__restrict__ val) {
    int64_t nr = *num_runs;
    int64_t nnz = (nr > 0 && ukeys[nr - 1] == kInvalidKey) ? nr - 1 : nr;
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= nnz) return;
    unsigned long long k = ukeys[i];
    int32_t row = int32_t(k >> 32);
    col[i] = int32_t(k & 0xffffffffu);
    val[i] = __fdiv_rn(float(sums[i]), float(vertex_samples[row]));
}

// rowptr[r] = first entry whose row >= r, r in [0, n]
__global__ void k_rowptr(const unsigned long long *__restrict__ ukeys, const int *__restrict__ num_runs, int32_t n,
                         int64_t *__restrict__ rowptr) {
    int64_t nr = *num_runs;
    int64_t nnz = (nr > 0 && ukeys[nr - 1] == kInvalidKey) ? nr - 1 : nr;
    int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (r > n) return;
    unsigned long long target = (unsigned long long)r << 32;
    int64_t lo = 0, hi = nnz;
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        if (ukeys[mid] < target) lo = mid + 1; else hi = mid;
    }
    rowptr[r] = lo;
}

struct Scratch {
    std::vector<void *> ptrs;
    ~Scratch() { for (void *p : ptrs) cudaFree(p); }
    cudaError_t alloc(void **p, size_t bytes) {
        cudaError_t e = cudaMalloc(p, bytes ? bytes : 8);
        if (e == cudaSuccess) ptrs.push_back(*p);
        return e;
    }
};

inline int blocks_for(int64_t n, int b) { return (int)((n + b - 1) / b); }

}  // namespace

extern "C" avr_status avr_graph_create_from_edges(avr_ctx *ctx, int32_t n, const int32_t *vertex_samples, int64_t n_edges,
                                                  const int32_t *edge_from, const int32_t *edge_to,
                                                  const int32_t *edge_samples, avr_graph **out) {
    if (!ctx || !out || !vertex_samples || (n_edges > 0 && (!edge_from || !edge_to || !edge_samples)))
        return set_err(ctx, AVR_ERR_BAD_ARG, "avr_graph_create_from_edges: NULL argument");
    *out = nullptr;
    if (n < 1 || n_edges < 0 || n_edges > 0x7fffffff)
        return set_err(ctx, AVR_ERR_BAD_ARG, "avr_graph_create_from_edges: need n >= 1 and 0 <= n_edges < 2^31");
    AVR_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    Scratch sc;
    int32_t *d_from = nullptr, *d_to = nullptr, *d_smp = nullptr, *d_smp2 = nullptr, *d_sums = nullptr, *d_vs = nullptr;
    unsigned long long *d_keys = nullptr, *d_keys2 = nullptr, *d_ukeys = nullptr;
    int *d_runs = nullptr;
    size_t E = size_t(n_edges);
    AVR_CUDA(ctx, sc.alloc((void **)&d_from, E * 4));
    AVR_CUDA(ctx, sc.alloc((void **)&d_to, E * 4));
    AVR_CUDA(ctx, sc.alloc((void **)&d_smp, E * 4));
    AVR_CUDA(ctx, sc.alloc((void **)&d_smp2, E * 4));
    AVR_CUDA(ctx, sc.alloc((void **)&d_sums, E * 4));
    AVR_CUDA(ctx, sc.alloc((void **)&d_vs, size_t(n) * 4));
    AVR_CUDA(ctx, sc.alloc((void **)&d_keys, E * 8));
    AVR_CUDA(ctx, sc.alloc((void **)&d_keys2, E * 8));
    AVR_CUDA(ctx, sc.alloc((void **)&d_ukeys, E * 8));
    AVR_CUDA(ctx, sc.alloc((void **)&d_runs, sizeof(int)));
    AVR_CUDA(ctx, cudaMemcpyAsync(d_vs, vertex_samples, size_t(n) * 4, cudaMemcpyHostToDevice, s));
    AVR_CUDA(ctx, cudaMemsetAsync(d_runs, 0, sizeof(int), s));
    int num_runs = 0;
    if (E) {
        AVR_CUDA(ctx, cudaMemcpyAsync(d_from, edge_from, E * 4, cudaMemcpyHostToDevice, s));
        AVR_CUDA(ctx, cudaMemcpyAsync(d_to, edge_to, E * 4, cudaMemcpyHostToDevice, s));
        AVR_CUDA(ctx, cudaMemcpyAsync(d_smp, edge_samples, E * 4, cudaMemcpyHostToDevice, s));
        k_edge_keys<<<blocks_for(n_edges, 256), 256, 0, s>>>(d_from, d_to, n_edges, n, d_keys);
        AVR_CUDA(ctx, cudaGetLastError());
        size_t tmp1 = 0, tmp2 = 0;
        AVR_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp1, d_keys, d_keys2, d_smp, d_smp2, (int)n_edges, 0, 64, s));
        AVR_CUDA(ctx, cub::DeviceReduce::ReduceByKey(nullptr, tmp2, d_keys2, d_ukeys, d_smp2, d_sums, d_runs, cub::Sum(),
                                                     (int)n_edges, s));
        void *d_tmp = nullptr;
        size_t tmp = tmp1 > tmp2 ? tmp1 : tmp2;
        AVR_CUDA(ctx, sc.alloc(&d_tmp, tmp));
        AVR_CUDA(ctx, cub::DeviceRadixSort::SortPairs(d_tmp, tmp, d_keys, d_keys2, d_smp, d_smp2, (int)n_edges, 0, 64, s));
        AVR_CUDA(ctx, cub::DeviceReduce::ReduceByKey(d_tmp, tmp, d_keys2, d_ukeys, d_smp2, d_sums, d_runs, cub::Sum(),
                                                     (int)n_edges, s));
        AVR_CUDA(ctx, cudaMemcpyAsync(&num_runs, d_runs, sizeof(int), cudaMemcpyDeviceToHost, s));
        AVR_CUDA(ctx, cudaStreamSynchronize(s));
    }
    int64_t nnz = num_runs;
    if (nnz > 0) {
        unsigned long long last = 0;
        AVR_CUDA(ctx, cudaMemcpy(&last, d_ukeys + (nnz - 1), 8, cudaMemcpyDeviceToHost));
        if (last == kInvalidKey) --nnz;
    }
    avr_graph *g = new avr_graph();
    g->ctx = ctx;
    g->n = n;
    g->nnz = nnz;
    cudaError_t e = cudaSuccess;
    auto A = [&](void **p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes ? bytes : 4); };
    A((void **)&g->d_rowptr, (size_t(n) + 1) * 8);
    A((void **)&g->d_col, size_t(nnz) * 4);
    A((void **)&g->d_val, size_t(nnz) * 4);
    A((void **)&g->d_cur, size_t(n) * 4);
    A((void **)&g->d_nxt, size_t(n) * 4);
    A((void **)&g->d_total, size_t(n) * 4);
    A((void **)&g->d_l0, size_t(n) * 4);
    A((void **)&g->d_flag, 2 * 4);
    if (e != cudaSuccess) { avr_graph_destroy(g); return set_cuda_err(ctx, e, "avr_graph_create_from_edges alloc"); }
    if (nnz) k_fill_csr<<<blocks_for(nnz, 256), 256, 0, s>>>(d_ukeys, d_sums, d_vs, d_runs, g->d_col, g->d_val);
    k_rowptr<<<blocks_for(int64_t(n) + 1, 256), 256, 0, s>>>(d_ukeys, d_runs, n, g->d_rowptr);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) { avr_graph_destroy(g); return set_cuda_err(ctx, e, "avr_graph_create_from_edges"); }
    *out = g;
    return AVR_OK;
}

extern "C" avr_status avr_graph_get_csr(avr_graph *g, int32_t *n_vertices, int64_t *nnz, int64_t *rowptr, int32_t *col,
                                        float *val) {
    if (!g) return set_err(nullptr, AVR_ERR_BAD_ARG, "avr_graph_get_csr: NULL graph");
    avr_ctx *ctx = g->ctx;
    AVR_CUDA(ctx, cudaSetDevice(ctx->device));
    if (n_vertices) *n_vertices = g->n;
    if (nnz) *nnz = g->nnz;
    if (rowptr) AVR_CUDA(ctx, cudaMemcpy(rowptr, g->d_rowptr, (size_t(g->n) + 1) * 8, cudaMemcpyDeviceToHost));
    if (col && g->nnz) AVR_CUDA(ctx, cudaMemcpy(col, g->d_col, size_t(g->nnz) * 4, cudaMemcpyDeviceToHost));
    if (val && g->nnz) AVR_CUDA(ctx, cudaMemcpy(val, g->d_val, size_t(g->nnz) * 4, cudaMemcpyDeviceToHost));
    return AVR_OK;
}
