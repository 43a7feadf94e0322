// avr_comm.cu — avr_comm_* and avr_render_volpath_multi: one rank per GPU (process or thread), NCCL over
// NVLink / NVSwitch.  Reference counterparts: the tile / wave loop of ImageTileIntegrator::Render
// (cpu/integrators.cpp:168-226), whose ParallelFor2D workers all add into ONE film — here every rank
// renders its share of the sample indices into its own device film and one ncclReduce (fp64 sum) lands the
// complete film on rank 0.
#include <dlfcn.h>

#include <cstring>
#include <mutex>
#include <string>

#include "avr_comm.h"
#include "avr_render_host.h"

using namespace avr;

namespace avr {

const NcclApi *nccl_api(std::string *why) {
    static NcclApi api;
    static bool tried = false, ok = false;
    static std::string err;
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    if (!tried) {
        tried = true;
        // RTLD_NOLOAD first: reuse the NCCL the host process already runs (two copies in one process would
        // each set up their own transports); else the library the loader finds under that SONAME
        void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) {
            err = std::string("libnccl.so.2 not found (") + dlerror() + ")";
        } else {
            bool all = true;
            auto sym = [&](const char *name) {
                void *p = dlsym(h, name);
                if (!p) { all = false; err = std::string("libnccl: missing symbol ") + name; }
                return p;
            };
            api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
            api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
            api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
            api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
            api.Reduce = (decltype(api.Reduce))sym("ncclReduce");
            api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
            api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
            api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
            api.GetVersion = (decltype(api.GetVersion))sym("ncclGetVersion");
            ok = all;
        }
    }
    if (!ok && why) *why = err;
    return ok ? &api : nullptr;
}

}  // namespace avr

static_assert(AVR_COMM_ID_BYTES == sizeof(ncclUniqueId), "avr.h: AVR_COMM_ID_BYTES must equal NCCL_UNIQUE_ID_BYTES");

extern "C" avr_status avr_comm_unique_id(uint8_t id[AVR_COMM_ID_BYTES]) {
    if (!id) return set_err(nullptr, AVR_ERR_BAD_ARG, "avr_comm_unique_id: NULL argument");
    std::string why;
    const NcclApi *api = nccl_api(&why);
    if (!api) return set_err(nullptr, AVR_ERR_NCCL, "avr_comm_unique_id: " + why);
    ncclUniqueId u;
    ncclResult_t r = api->GetUniqueId(&u);
    if (r != ncclSuccess) return set_err(nullptr, AVR_ERR_NCCL, std::string("ncclGetUniqueId: ") + api->GetErrorString(r));
    std::memcpy(id, &u, sizeof(u));
    return AVR_OK;
}

extern "C" avr_status avr_comm_init_rank(avr_ctx *ctx, const uint8_t id[AVR_COMM_ID_BYTES], int32_t rank, int32_t world,
                                         avr_comm **out) {
    if (!ctx || !id || !out) return set_err(ctx, AVR_ERR_BAD_ARG, "avr_comm_init_rank: NULL argument");
    *out = nullptr;
    if (world < 1 || rank < 0 || rank >= world) return set_err(ctx, AVR_ERR_BAD_ARG, "avr_comm_init_rank: rank outside [0, world)");
    std::string why;
    const NcclApi *api = nccl_api(&why);
    if (!api) return set_err(ctx, AVR_ERR_NCCL, "avr_comm_init_rank: " + why);
    AVR_CUDA(ctx, cudaSetDevice(ctx->device));
    avr_comm *cm = new avr_comm();
    cm->ctx = ctx; cm->rank = rank; cm->world = world; cm->api = api;
    ncclUniqueId u;
    std::memcpy(&u, id, sizeof(u));
    ncclResult_t r = api->CommInitRank(&cm->comm, world, u, rank);
    if (r != ncclSuccess) {
        avr_status st = set_err(ctx, AVR_ERR_NCCL, std::string("ncclCommInitRank: ") + api->GetErrorString(r));
        delete cm;
        return st;
    }
    *out = cm;
    return AVR_OK;
}

extern "C" avr_status avr_comm_init_all(avr_ctx **ctx, int32_t n, avr_comm **out) {
    if (!ctx || !out || n < 1) return set_err(nullptr, AVR_ERR_BAD_ARG, "avr_comm_init_all: NULL argument or n < 1");
    for (int i = 0; i < n; ++i) {
        out[i] = nullptr;
        if (!ctx[i]) return set_err(nullptr, AVR_ERR_BAD_ARG, "avr_comm_init_all: NULL context");
    }
    std::string why;
    const NcclApi *api = nccl_api(&why);
    if (!api) return set_err(ctx[0], AVR_ERR_NCCL, "avr_comm_init_all: " + why);
    std::vector<int> devs(n);
    std::vector<ncclComm_t> comms(n);
    for (int i = 0; i < n; ++i) devs[i] = ctx[i]->device;
    ncclResult_t r = api->CommInitAll(comms.data(), n, devs.data());
    if (r != ncclSuccess) return set_err(ctx[0], AVR_ERR_NCCL, std::string("ncclCommInitAll: ") + api->GetErrorString(r));
    for (int i = 0; i < n; ++i) {
        avr_comm *cm = new avr_comm();
        cm->ctx = ctx[i]; cm->rank = i; cm->world = n; cm->api = api; cm->comm = comms[i];
        out[i] = cm;
    }
    return AVR_OK;
}

extern "C" void avr_comm_destroy(avr_comm *cm) {
    if (!cm) return;
    cudaSetDevice(cm->ctx->device);
    if (cm->exec) cudaGraphExecDestroy(cm->exec);
    for (int p = 0; p < AVR_P2P_MAX_WORLD; ++p)
        if (cm->p2p_base[p] && p != cm->rank && cm->p2p_ipc[p]) cudaIpcCloseMemHandle(cm->p2p_base[p]);
    if (cm->rank < AVR_P2P_MAX_WORLD) cudaFree(cm->p2p_base[cm->rank]);
    cudaFree(cm->d_gather[0]);
    cudaFree(cm->d_gather[1]);
    cudaFree(cm->d_flags);
    if (cm->comm) cm->api->CommDestroy(cm->comm);
    delete cm;
}

extern "C" int32_t avr_comm_rank(const avr_comm *cm) { return cm ? cm->rank : -1; }
extern "C" int32_t avr_comm_world(const avr_comm *cm) { return cm ? cm->world : 0; }
extern "C" int32_t avr_comm_graph_exchange(const avr_comm *cm) { return cm ? cm->p2p_state : 0; }

extern "C" avr_status avr_render_volpath_multi(avr_comm *cm, avr_medium *m, const avr_render_desc *d, double *film,
                                               avr_render_stats *stats) {
    if (!cm || !m || !d) return set_err(cm ? cm->ctx : nullptr, AVR_ERR_BAD_ARG, "avr_render_volpath_multi: NULL argument");
    avr_ctx *ctx = cm->ctx;
    if (m->ctx != ctx) return set_err(ctx, AVR_ERR_BAD_ARG, "avr_render_volpath_multi: the medium belongs to another context");
    if (cm->rank == 0 && !film) return set_err(ctx, AVR_ERR_BAD_ARG, "avr_render_volpath_multi: rank 0 needs the host film");
    AVR_CUDA(ctx, cudaSetDevice(ctx->device));
    // this rank's share: work items (32-path chunks of the pixel x sample lattice) with item % world == rank
    avr_render_desc mine = *d;
    mine.rank = cm->rank;
    mine.world = cm->world;
    size_t bytes = size_t(d->width) * d->height * 4 * sizeof(double);
    avr_status st = film_scratch(ctx, bytes, ctx->stream);
    if (st != AVR_OK) return st;
    avr_render_stats local;
    st = avr_render_volpath_device(ctx, m, &mine, ctx->d_film, ctx->stream, stats ? stats : &local);
    if (st != AVR_OK) return st;
    if (cm->world > 1)
        AVR_NCCL(cm, cm->api->Reduce(ctx->d_film, ctx->d_film, bytes / sizeof(double), ncclDouble, ncclSum, 0, cm->comm, ctx->stream));
    if (cm->rank == 0) return film_add_to_host(ctx, bytes, film);
    AVR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return AVR_OK;
}
