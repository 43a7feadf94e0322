// avr_comm.h — multi-GPU plumbing behind the C ABI (avr_comm_init; 8(e): film reduce,
// per-bounce all-gather).  NCCL is bound at run time (dlopen of libnccl.so.2: the copy already loaded in the
// process, e.g. torch's, or the system one), so libavr_b200.so itself only depends on libcudart.
#pragma once
#include <nccl.h>

#include "avr_internal.h"

#define AVR_P2P_MAX_WORLD 16

namespace avr {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId *);
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int);
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*Reduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
    const char *(*GetErrorString)(ncclResult_t);
    ncclResult_t (*GetVersion)(int *);
};
// The bound API, or nullptr with the reason in *why.
const NcclApi *nccl_api(std::string *why);

}  // namespace avr

struct avr_comm {
    avr_ctx *ctx = nullptr;
    int rank = 0, world = 1;
    ncclComm_t comm = nullptr;
    const avr::NcclApi *api = nullptr;
    // row-sharded propagation: gathered `cur` vector (chunk * world floats), twice (ping-pong), flags, and the
    // captured bounce sequence (re-captured when the graph, the pointers or the bounce count change)
    float *d_gather[2] = {nullptr, nullptr};
    size_t gather_floats = 0;
    int32_t *d_flags = nullptr;
    cudaGraphExec_t exec = nullptr;
    const void *exec_key[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    // Fused SpMV + exchange over peer memory (avr_graph.cu, k_spmv_p2p): ONE allocation per rank
    // {vector[2][padded], sync words}, mapped into every peer (cudaIpc* between processes, direct peer access
    // between the threads of one process).  p2p_state: 0 untried, 1 in use, -1 unavailable (NCCL all-gather path).
    int p2p_state = 0;
    char *p2p_base[AVR_P2P_MAX_WORLD] = {};   // [rank] = own block (cudaMalloc), others = mapped peers
    bool p2p_ipc[AVR_P2P_MAX_WORLD] = {};     // opened with cudaIpcOpenMemHandle (to be closed)
    size_t p2p_padded = 0;
};

#define AVR_NCCL(cm, call)                                                                              \
    do {                                                                                                \
        ncclResult_t r__ = (call);                                                                      \
        if (r__ != ncclSuccess)                                                                         \
            return avr::set_err((cm)->ctx, AVR_ERR_NCCL, std::string(#call) + ": " + (cm)->api->GetErrorString(r__)); \
    } while (0)
