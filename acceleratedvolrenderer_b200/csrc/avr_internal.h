// avr_internal.h — host-side objects behind the opaque handles of include/avr.h.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "../../include/avr.h"
#include "avr_medium.cuh"

// Scheduling knobs of the render kernels' path pool (avr_volpath.cu): they order the work, never change a
// result.  Read ONCE from the environment when the context is created (AVR_WALK_STEPS, AVR_COLLIDE_MIN,
// AVR_REFILL_LANES, AVR_TRACK_MIN, AVR_YIELD_LANES, AVR_YIELD_LANES_FULL, AVR_EVENT_SMSP_MASK, AVR_BLOCKS_PER_SM,
// AVR_THREADS, AVR_POOL); a render call reads no environment variable.
struct avr_knobs {
    int walk_steps = 0 /* 0: 3 for FAST, 4 for EXACT kernels */, collide_min = 1, refill_lanes = 16, track_min = 32, yield_lanes = 8, yield_lanes_full = 20;
    int event_smsp_mask = 12;
    int blocks_per_sm = 0, threads = 0, pool = 0;  // 0 = the kernel variant's default
    // FAST majorant grid of a medium created under this context: c x c x c reference cells per FAST cell
    // (AVR_FAST_MAJ_COARSEN; 1 = the reference's grid)
    int fast_maj_coarsen = 2;
    // avr_graph_propagate_multi: 1 = exchange the vector inside the SpMV kernel over peer memory when the GPUs can
    // map each other (AVR_GRAPH_P2P, default 1); 0 = always ncclAllGather
    int graph_p2p = 1;
};
// What launch_render settled on for one kernel variant (function pointer): shared-memory opt-in done, occupancy
// queried, pool sized — once per context, not per call.
struct avr_launch_plan {
    const void *fn = nullptr;
    int threads = 0, pool = 0, per_sm = 0;
    unsigned maj_bytes = 0;  // majorant grid staged in shared memory (0 = read through L1 / L2)
    size_t smem = 0;
};

struct avr_ctx {
    int device = 0;
    avr_knobs knobs;
    std::vector<avr_launch_plan> plans;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int sm_count = 0;
    size_t smem_per_sm = 0, smem_per_block_optin = 0;
    std::string err;
    // render scratch (grown on demand)
    unsigned long long *d_counters = nullptr;  // work counter + stats
    double *d_film = nullptr;
    size_t film_bytes = 0;
    double *h_film = nullptr;  // pinned staging for the film read-back
    size_t h_film_bytes = 0;
    cudaEvent_t ev_chunk[8] = {};
    float *d_tables = nullptr;  // light spectrum + sensor tables for the current render
    // tabulated pixel filter of the current render (avr_filter_table -> packed sampling tables, avr_volpath.cuh)
    float *d_filter = nullptr;
    size_t d_filter_floats = 0;
    std::vector<float> h_filter;
    float4 *d_fsamples = nullptr;  // {filter p.x, p.y, weight, zsobol wavelength draw} per pixel sample of the current launch
    bool ev0_recorded = false;     // fill_render_params started the kernel timer before its pre-pass
    size_t d_fsamples_count = 0;
};

struct avr_medium {
    avr_ctx *ctx = nullptr;
    avr::MediumView view{};
    std::vector<void *> allocs;
    int64_t bytes = 0;
    float *d_maj = nullptr;
    int mres[3] = {16, 16, 16};
    int reports_emissive = -1;  // IsEmissive() when it differs from view.emissive (kernel selection)
};

struct avr_graph {
    avr_ctx *ctx = nullptr;
    int32_t n = 0;
    int64_t nnz = 0;
    int64_t *d_rowptr = nullptr;
    int32_t *d_col = nullptr;
    float *d_val = nullptr;
    float *d_cur = nullptr, *d_nxt = nullptr, *d_total = nullptr, *d_l0 = nullptr;  // propagation vectors, allocated once
    int32_t *d_flag = nullptr;
    // vertices + uniform hash grid for ConnectToGraph
    float4 *d_vert = nullptr;     // x,y,z,lightScalar  (sorted by cell)
    float *d_range2 = nullptr;    // renderSearchRange^2 (sorted by cell) or null
    int32_t *d_vid = nullptr;     // original vertex id (sorted by cell)
    int32_t *d_cell_start = nullptr;
    int gres[3] = {0, 0, 0};
    float gmin[3] = {0, 0, 0}, cell = 0;
    float sq_radius = 0, perc99 = 0, max_range = -1;
    bool use_range = false;
};

namespace avr {
avr_status set_err(avr_ctx *ctx, avr_status code, const std::string &msg);
avr_status set_cuda_err(avr_ctx *ctx, cudaError_t e, const char *what);
}  // namespace avr

#define AVR_CUDA(ctx, call)                                                   \
    do {                                                                      \
        cudaError_t e__ = (call);                                             \
        if (e__ != cudaSuccess) return avr::set_cuda_err((ctx), e__, #call);  \
    } while (0)
