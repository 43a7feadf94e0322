// avr_lib.cu — context, media (upload, majorant-grid build, octet packing) and the small query
// kernels behind include/avr.h.  sm_100a only; no CPU fallback: without a CUDA device every entry
// point fails with AVR_ERR_CUDA.
#include <cstdlib>
#include <cuda_fp16.h>

#include <cmath>
#include <cstdio>
#include <cstring>

#include "avr_internal.h"
#include "avr_tmaj.cuh"

using namespace avr;

#define AVR_STR2(x) #x
#define AVR_STR(x) AVR_STR2(x)

static thread_local std::string g_create_err;

namespace avr {
avr_status set_err(avr_ctx *ctx, avr_status code, const std::string &msg) {
    if (ctx) ctx->err = msg; else g_create_err = msg;
    return code;
}
avr_status set_cuda_err(avr_ctx *ctx, cudaError_t e, const char *what) {
    std::string m = std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what;
    cudaGetLastError();
    return set_err(ctx, e == cudaErrorMemoryAllocation ? AVR_ERR_OOM : AVR_ERR_CUDA, m);
}
}  // namespace avr

extern "C" const char *avr_version(void) { return "avr_b200 0.1 sm_100a (CUDA " AVR_STR(CUDART_VERSION) ")"; }

extern "C" const char *avr_last_error(const avr_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

extern "C" avr_status avr_ctx_create(int device, avr_ctx **out) {
    if (!out) return set_err(nullptr, AVR_ERR_BAD_ARG, "avr_ctx_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return set_err(nullptr, AVR_ERR_CUDA,
                       std::string("avr_ctx_create: no CUDA device (") + cudaGetErrorString(e) +
                           "); libavr_b200 has no CPU fallback");
    if (device < 0 || device >= count) return set_err(nullptr, AVR_ERR_BAD_ARG, "avr_ctx_create: bad device index");
    avr_ctx *ctx = new avr_ctx();
    ctx->device = device;
    {
        auto env_int = [](const char *name, int def, int lo, int hi) {
            const char *e = getenv(name);
            if (!e || !*e) return def;
            int v = atoi(e);
            return v < lo ? lo : (v > hi ? hi : v);
        };
        avr_knobs &k = ctx->knobs;
        k.walk_steps = env_int("AVR_WALK_STEPS", k.walk_steps, 0, 1 << 20);
        k.collide_min = env_int("AVR_COLLIDE_MIN", k.collide_min, 1, 32);
        k.refill_lanes = env_int("AVR_REFILL_LANES", k.refill_lanes, 1, 32);
        k.track_min = env_int("AVR_TRACK_MIN", k.track_min, 1, 1 << 20);
        k.yield_lanes = env_int("AVR_YIELD_LANES", k.yield_lanes, 0, 32);
        k.yield_lanes_full = env_int("AVR_YIELD_LANES_FULL", k.yield_lanes_full, 0, 32);
        k.event_smsp_mask = env_int("AVR_EVENT_SMSP_MASK", k.event_smsp_mask, 0, 15);
        k.blocks_per_sm = env_int("AVR_BLOCKS_PER_SM", 0, 0, 8);
        k.threads = env_int("AVR_THREADS", 0, 0, 1024);
        k.pool = env_int("AVR_POOL", 0, 0, 4096);
        k.fast_maj_coarsen = env_int("AVR_FAST_MAJ_COARSEN", k.fast_maj_coarsen, 1, 16);
        k.graph_p2p = env_int("AVR_GRAPH_P2P", k.graph_p2p, 0, 1);
    }
    AVR_CUDA(nullptr, cudaSetDevice(device));
    cudaDeviceProp prop;
    AVR_CUDA(nullptr, cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    ctx->smem_per_sm = prop.sharedMemPerMultiprocessor;
    ctx->smem_per_block_optin = prop.sharedMemPerBlockOptin;
    AVR_CUDA(nullptr, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    AVR_CUDA(nullptr, cudaEventCreate(&ctx->ev0));
    AVR_CUDA(nullptr, cudaEventCreate(&ctx->ev1));
    AVR_CUDA(nullptr, cudaMalloc(&ctx->d_counters, 64 * sizeof(unsigned long long)));
    AVR_CUDA(nullptr, cudaMalloc(&ctx->d_tables, 4 * kNSpec * sizeof(float)));
    *out = ctx;
    return AVR_OK;
}

extern "C" void avr_ctx_destroy(avr_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaFree(ctx->d_counters);
    cudaFree(ctx->d_film);
    if (ctx->h_film) cudaFreeHost(ctx->h_film);
    for (cudaEvent_t e : ctx->ev_chunk) if (e) cudaEventDestroy(e);
    cudaFree(ctx->d_tables);
    cudaFree(ctx->d_filter);
    cudaFree(ctx->d_fsamples);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

// ---- read-bandwidth probe: the L2 / HBM denominators of the density-fetch roofline ---------------------------
// "L2 peak to be measured with a 64 MiB-resident read kernel".  Every thread streams 128-bit
// loads over a buffer of `bytes` (grid-stride, 8 loads in flight), `iters` passes; a buffer that fits the 126 MB L2
// is served from it after the first pass, a larger one from HBM.
__global__ void __launch_bounds__(256) k_read_stream(const uint4 *__restrict__ buf, size_t n16, int iters, unsigned *sink) {
    unsigned acc = 0;
    const size_t stride = size_t(gridDim.x) * blockDim.x;
    for (int it = 0; it < iters; ++it) {
        size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
        for (; i + 7 * stride < n16; i += 8 * stride) {
            uint4 v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = __ldcg(buf + i + k * stride);  // L2 only: L1 must not serve the re-reads
#pragma unroll
            for (int k = 0; k < 8; ++k) acc ^= v[k].x ^ v[k].y ^ v[k].z ^ v[k].w;
        }
        for (; i < n16; i += stride) { uint4 v = __ldcg(buf + i); acc ^= v.x ^ v.y ^ v.z ^ v.w; }
    }
    if (acc == 0x9e3779b9u) *sink = acc;  // keeps the loads alive, practically never taken
}

extern "C" avr_status avr_measure_read_bandwidth(avr_ctx *ctx, int64_t bytes, int32_t iters, float *gb_per_s) {
    if (!ctx || !gb_per_s) return set_err(ctx, AVR_ERR_BAD_ARG, "avr_measure_read_bandwidth: NULL argument");
    if (bytes < 4096 || iters < 1) return set_err(ctx, AVR_ERR_BAD_ARG, "avr_measure_read_bandwidth: bytes >= 4096, iters >= 1");
    AVR_CUDA(ctx, cudaSetDevice(ctx->device));
    uint4 *buf = nullptr;
    unsigned *sink = nullptr;
    size_t n16 = size_t(bytes) / 16;
    AVR_CUDA(ctx, cudaMalloc(&buf, n16 * 16));
    cudaError_t e = cudaMalloc(&sink, 4);
    if (e == cudaSuccess) e = cudaMemsetAsync(buf, 1, n16 * 16, ctx->stream);
    float ms = 0;
    if (e == cudaSuccess) {
        int blocks = ctx->sm_count * 8;
        k_read_stream<<<blocks, 256, 0, ctx->stream>>>(buf, n16, 2, sink);  // warm-up: the buffer settles in L2
        cudaEventRecord(ctx->ev0, ctx->stream);
        k_read_stream<<<blocks, 256, 0, ctx->stream>>>(buf, n16, iters, sink);
        cudaEventRecord(ctx->ev1, ctx->stream);
        e = cudaStreamSynchronize(ctx->stream);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    }
    cudaFree(buf);
    cudaFree(sink);
    if (e != cudaSuccess) return set_cuda_err(ctx, e, "avr_measure_read_bandwidth");
    *gb_per_s = float(double(n16) * 16 * iters / (double(ms) * 1e6));
    return AVR_OK;
}

// ---- build kernels -----------------------------------------------------------------------------------
// MajorantGrid construction, media.cpp:240-246 (GridMedium) / :364-377 (RGBGridMedium):
// one CTA per majorant voxel, max over the density voxels SampledGrid::MaxValue(bounds) visits
// (util/containers.h:838-858).  Bit-exact: only IEEE divisions/multiplications feed the index box.
template <bool RGB>
__global__ void k_build_majorants(MediumView mv, float *out) {
    __shared__ float sm[32];
    int mx = blockIdx.x, my = blockIdx.y, mz = blockIdx.z;
    int r[3] = {mv.mres[0], mv.mres[1], mv.mres[2]};
    int n[3] = {mv.nx, mv.ny, mv.nz};
    int idx[3] = {mx, my, mz};
    int lo[3], hi[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float b0 = float(idx[i]) / r[i], b1 = float(idx[i] + 1) / r[i];  // MajorantGrid::VoxelBounds
        float p0 = b0 * n[i] - .5f, p1 = b1 * n[i] - .5f;
        lo[i] = max((int)floorf(p0), 0);
        hi[i] = min((int)floorf(p1) + 1, n[i] - 1);
    }
    auto value = [&](const float4 *gr, int x, int y, int z) -> float {
        if ((unsigned)x >= (unsigned)mv.nx || (unsigned)y >= (unsigned)mv.ny || (unsigned)z >= (unsigned)mv.nz) return 0.f;
        float4 v = gr[(size_t(z) * mv.ny + y) * mv.nx + x];
        // RGBUnboundedSpectrum::MaxValue = scale * RGBSigmoidPolynomial::MaxValue (util/color.h:346-352)
        float result = fmaxf(rsp_eval(v.y, v.z, v.w, 360.f), rsp_eval(v.y, v.z, v.w, 830.f));
        float lambda = -v.z / (2 * v.y);
        if (lambda >= 360 && lambda <= 830) result = fmaxf(result, rsp_eval(v.y, v.z, v.w, lambda));
        return v.x * result;
    };
    auto box_max = [&](const float *dense, const float4 *gr) -> float {
        float m = RGB ? value(gr, lo[0], lo[1], lo[2]) : dense_at(dense, mv.nx, mv.ny, mv.nz, lo[0], lo[1], lo[2]);
        int ex = hi[0] - lo[0] + 1, ey = hi[1] - lo[1] + 1, ez = hi[2] - lo[2] + 1;
        if (ex > 0 && ey > 0 && ez > 0) {
            int tot = ex * ey * ez;
            for (int i = threadIdx.x; i < tot; i += blockDim.x) {
                int x = lo[0] + i % ex, y = lo[1] + (i / ex) % ey, z = lo[2] + i / (ex * ey);
                float v = RGB ? value(gr, x, y, z) : dense[(size_t(z) * mv.ny + y) * mv.nx + x];
                m = fmaxf(m, v);
            }
        }
        // block max
        __syncthreads();
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
        __syncthreads();
        int nw = (blockDim.x + 31) >> 5;
        float r2 = sm[0];
        for (int w = 1; w < nw; ++w) r2 = fmaxf(r2, sm[w]);
        __syncthreads();
        return r2;
    };
    float res;
    if (RGB) {
        float ma = mv.rgb_a ? box_max(nullptr, mv.rgb_a) : 1.f;
        float ms = mv.rgb_s ? box_max(nullptr, mv.rgb_s) : 1.f;
        res = mv.sigma_scale * (ma + ms);
    } else
        res = box_max(mv.dense, nullptr);
    if (threadIdx.x == 0) out[mx + r[0] * (my + r[1] * mz)] = res;
}

// Where the build kernels of a scalar grid read voxel (x, y, z) of the index box from: the dense array, or — the
// leaf-direct .nvdb ingest — 8^3 blocks behind a table over block coordinates (avr_nanovdb_grid's sparse form).
struct VoxelSrc {
    const float *dense;
    const float *bvals;   // 512 floats per block, x fastest
    const int *btab;      // block index or -1, over tb[0] x tb[1] x tb[2] block coordinates
    int n[3], imin[3], bo[3], tb[3];
    __device__ float at(int x, int y, int z) const {
        if (x < 0 || y < 0 || z < 0 || x >= n[0] || y >= n[1] || z >= n[2]) return 0.f;
        if (!bvals) return dense[(size_t(z) * n[1] + y) * n[0] + x];
        const int X = x + imin[0], Y = y + imin[1], Z = z + imin[2];
        const int b = btab[(size_t((Z >> 3) - bo[2]) * tb[1] + ((Y >> 3) - bo[1])) * tb[0] + ((X >> 3) - bo[0])];
        return b < 0 ? 0.f : bvals[size_t(b) * 512 + (X & 7) + 8 * (Y & 7) + 64 * (Z & 7)];
    }
};
VoxelSrc dense_src(const MediumView &v) {
    VoxelSrc s{};
    s.dense = v.dense;
    s.n[0] = v.nx; s.n[1] = v.ny; s.n[2] = v.nz;
    return s;
}

// NanoVDBMedium majorants, media.cpp:560-616: per majorant cell, Bounds3::Lerp of the cell corners ->
// worldToIndexF(Vec3R) (subtraction and innermost product in double, fmaf arguments narrowed to float:
// NanoVDB.h:1409-1414 instantiated for Vec3<double>) -> int(i0 - 1) .. int(i1 + 1) clipped to the index
// box -> max of the voxel values, starting from 0.
__global__ void k_build_majorants_nvdb(MediumView mv, VoxelSrc src, float *out) {
    __shared__ float sm[32];
    int idx[3] = {(int)blockIdx.x, (int)blockIdx.y, (int)blockIdx.z};
    int n[3] = {mv.nx, mv.ny, mv.nz};
    double w0[3], w1[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float t0 = float(idx[i]) / mv.mres[i], t1 = float(idx[i] + 1) / mv.mres[i];
        w0[i] = double((1 - t0) * mv.bmin[i] + t0 * mv.bmax[i]);  // pbrt::Lerp(t, pMin, pMax), float
        w1[i] = double((1 - t1) * mv.bmin[i] + t1 * mv.bmax[i]);
    }
    auto to_index = [&](const double w[3], int r) -> double {
        double x = w[0] - double(mv.nv_vec[0]), y = w[1] - double(mv.nv_vec[1]), z = w[2] - double(mv.nv_vec[2]);
        float inner = float(z * double(mv.nv_inv[3 * r + 2]));
        return double(__fmaf_rn(float(x), mv.nv_inv[3 * r], __fmaf_rn(float(y), mv.nv_inv[3 * r + 1], inner)));
    };
    int lo[3], hi[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        lo[i] = max(__double2int_rz(to_index(w0, i) - 1.0), mv.nv_imin[i]) - mv.nv_imin[i];
        hi[i] = min(__double2int_rz(to_index(w1, i) + 1.0), mv.nv_imin[i] + n[i] - 1) - mv.nv_imin[i];
    }
    float m = 0.f;
    int ex = hi[0] - lo[0] + 1, ey = hi[1] - lo[1] + 1, ez = hi[2] - lo[2] + 1;
    if (ex > 0 && ey > 0 && ez > 0) {
        int tot = ex * ey * ez;
        for (int i = threadIdx.x; i < tot; i += blockDim.x) {
            int x = lo[0] + i % ex, y = lo[1] + (i / ex) % ey, z = lo[2] + i / (ex * ey);
            m = fmaxf(m, src.at(x, y, z));
        }
    }
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        int nw = (blockDim.x + 31) >> 5;
        float r2 = sm[0];
        for (int w = 1; w < nw; ++w) r2 = fmaxf(r2, sm[w]);
        out[idx[0] + mv.mres[0] * (idx[1] + mv.mres[1] * idx[2])] = r2;
    }
}

// Octet packing: one thread per interpolation cell (cx,cy,cz) in [0,nx]x[0,ny]x[0,nz]
// (cell c covers taps c-1 and c); bricks beyond the lattice edge are zero-filled by the memset.
template <bool HALF>
__global__ void k_pack_octets(MediumView mv, float4 *out32, uint4 *out16) {
    size_t ncx = mv.nx + 1, ncy = mv.ny + 1, ncz = mv.nz + 1;
    size_t total = ncx * ncy * ncz;
    for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
        int cx = int(i % ncx), cy = int((i / ncx) % ncy), cz = int(i / (ncx * ncy));
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k)
            v[k] = dense_at(mv.dense, mv.nx, mv.ny, mv.nz, cx - 1 + (k & 1), cy - 1 + ((k >> 1) & 1), cz - 1 + (k >> 2));
        size_t o = octet_index(cx, cy, cz, mv.nbx, mv.nby);
        if (HALF) {
            __half2 h0 = __floats2half2_rn(v[0], v[1]), h1 = __floats2half2_rn(v[2], v[3]);
            __half2 h2 = __floats2half2_rn(v[4], v[5]), h3 = __floats2half2_rn(v[6], v[7]);
            uint4 q;
            q.x = *reinterpret_cast<uint32_t *>(&h0);
            q.y = *reinterpret_cast<uint32_t *>(&h1);
            q.z = *reinterpret_cast<uint32_t *>(&h2);
            q.w = *reinterpret_cast<uint32_t *>(&h3);
            out16[o] = q;
        } else {
            out32[2 * o] = make_float4(v[0], v[1], v[2], v[3]);
            out32[2 * o + 1] = make_float4(v[4], v[5], v[6], v[7]);
        }
    }
}

// Sparse brick pool (LAYOUT_BRICK_F16, avr_medium.cuh brick16_taps).  One CTA per brick of 8x8x8 cells.
// flags: does any voxel the brick's cells touch (x in [8 bx - 1, 8 bx + 7], ...) hold a non-zero value?
__global__ void k_brick_flags(VoxelSrc src, int *flags) {
    const int bx = blockIdx.x, by = blockIdx.y, bz = blockIdx.z;
    int any = 0;
    for (int i = threadIdx.x; i < 729; i += blockDim.x) {
        int x = 8 * bx - 1 + i % 9, y = 8 * by - 1 + (i / 9) % 9, z = 8 * bz - 1 + i / 81;
        any |= src.at(x, y, z) != 0.f;
    }
    any = __syncthreads_or(any);
    if (threadIdx.x == 0) flags[(size_t(bz) * gridDim.y + by) * gridDim.x + bx] = any;
}
// fill: CTA b writes the 576 entries of pool slot b (bricks lists the non-empty bricks' linear ids)
__global__ void k_brick_fill(VoxelSrc src, const int *bricks, int sbx, int sby, uint2 *pool) {
    const int id = bricks[blockIdx.x];
    const int bx = id % sbx, by = (id / sbx) % sby, bz = id / (sbx * sby);
    for (int i = threadIdx.x; i < kBrickEntries; i += blockDim.x) {
        int r = i % 9, lx = (i / 9) & 7, ly = i / 72;
        int x = 8 * bx + lx, y = 8 * by + ly, z = 8 * bz - 1 + r;  // cell column (x, y): taps x-1, x / y-1, y
        __half2 h0 = __floats2half2_rn(src.at(x - 1, y - 1, z), src.at(x, y - 1, z));
        __half2 h1 = __floats2half2_rn(src.at(x - 1, y, z), src.at(x, y, z));
        uint2 q;
        q.x = *reinterpret_cast<uint32_t *>(&h0);
        q.y = *reinterpret_cast<uint32_t *>(&h1);
        pool[size_t(blockIdx.x) * kBrickEntries + i] = q;
    }
}

// ---- empty-space distance field of the majorant grid (FAST render kernels, MediumView::maj_fast) --------
// Chebyshev distance in cells from every EMPTY majorant cell to the nearest non-empty one, by three 1-D
// passes: d1 = distance along x; d2 = min over y' of max(|y - y'|, d1(y')); d3 the same along z.
// Distances are capped at kLeapCap; a thread scans outwards and stops once the offset reaches its best.
constexpr int kLeapCap = 255;
__global__ void k_empty_dist_x(const float *maj, int rx, int ry, int rz, unsigned char *out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rx * ry * rz) return;
    int x = i % rx;
    const float *row = maj + (i - x);
    int best = kLeapCap;
    for (int o = 0; o < best; ++o) {
        bool hit = (x - o >= 0 && row[x - o] != 0.f) || (x + o < rx && row[x + o] != 0.f);
        if (hit) { best = o; break; }
    }
    out[i] = (unsigned char)best;
}
// axis 1: stride rx, extent ry; axis 2: stride rx*ry, extent rz
__global__ void k_empty_dist_next(const unsigned char *in, int n_total, int stride, int extent, unsigned char *out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_total) return;
    int c = (i / stride) % extent;
    int best = in[i];
    for (int o = 1; o < best; ++o) {
        int a = c - o >= 0 ? in[i - o * stride] : kLeapCap, b = c + o < extent ? in[i + o * stride] : kLeapCap;
        int m = min(a, b);
        best = min(best, max(o, m));
    }
    out[i] = (unsigned char)best;
}
__global__ void k_maj_fast(const float *maj, const unsigned char *dist, int n, float *out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float m = maj[i];
    // a non-empty cell keeps its majorant; an empty one holds -distance (>= 1); a negative majorant (invalid
    // input) must not read as a distance: it becomes 0 = "empty, no leap"
    out[i] = m != 0.f ? (m > 0.f || m != m ? m : 0.f) : -float(max((int)dist[i], 1));
}

// ---- medium creation ---------------------------------------------------------------------------------
namespace {

template <typename T>
avr_status upload(avr_medium *m, const T *host, size_t count, const T **dev) {
    void *d = nullptr;
    AVR_CUDA(m->ctx, cudaMalloc(&d, count * sizeof(T)));
    m->allocs.push_back(d);
    m->bytes += int64_t(count * sizeof(T));
    AVR_CUDA(m->ctx, cudaMemcpyAsync(d, host, count * sizeof(T), cudaMemcpyHostToDevice, m->ctx->stream));
    *dev = (const T *)d;
    return AVR_OK;
}

bool affine_rows(const float m16[16], float out12[12]) {
    if (m16[12] != 0.f || m16[13] != 0.f || m16[14] != 0.f || m16[15] != 1.f) return false;
    std::memcpy(out12, m16, 12 * sizeof(float));
    return true;
}

// FAST-path folding: s_i = ((minv_i . p) - bmin_i) / (bmax_i - bmin_i) * n_i - .5
void fold_grid_from_render(MediumView &v) {
    int n[3] = {v.nx, v.ny, v.nz};
    for (int i = 0; i < 3; ++i) {
        double ext = double(v.bmax[i]) - double(v.bmin[i]);
        double s = (ext > 0 ? 1.0 / ext : 1.0) * n[i];
        for (int j = 0; j < 3; ++j) v.gfr[4 * i + j] = float(double(v.minv[4 * i + j]) * s);
        v.gfr[4 * i + 3] = float((double(v.minv[4 * i + 3]) - double(v.bmin[i])) * s - 0.5);
    }
}

// KIND_NVDB: s = invMat * (minv . p - vec) - index_min  (sample space = index space relative to the box)
void fold_index_from_render(MediumView &v) {
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) {
            double a = 0;
            for (int k = 0; k < 3; ++k) a += double(v.nv_inv[3 * i + k]) * double(v.minv[4 * k + j]);
            v.gfr[4 * i + j] = float(a);
        }
        double t = 0;
        for (int k = 0; k < 3; ++k) t += double(v.nv_inv[3 * i + k]) * (double(v.minv[4 * k + 3]) - double(v.nv_vec[k]));
        v.gfr[4 * i + 3] = float(t - double(v.nv_imin[i]));
    }
}

avr_status finish_common(avr_medium *m, const float p0[3], const float p1[3], const float rfm[16],
                         const float mfr[16], const int32_t mres_in[3], int mres_default = 16) {
    MediumView &v = m->view;
    for (int i = 0; i < 3; ++i) {
        v.bmin[i] = p0[i];
        v.bmax[i] = p1[i];
        m->mres[i] = v.mres[i] = mres_in[i] > 0 ? mres_in[i] : mres_default;
    }
    if (v.mres[0] > 1023 || v.mres[1] > 1023 || v.mres[2] > 1023)
        return set_err(m->ctx, AVR_ERR_UNSUPPORTED, "majorant grid resolution > 1023 per axis");
    for (int i = 0; i < 3; ++i) {
        v.cell[i] = (v.bmax[i] - v.bmin[i]) / float(v.mres[i]);
        v.inv_cell[i] = v.cell[i] > 0 ? 1.f / v.cell[i] : 0.f;
        v.nz_lo[i] = 0;
        v.nz_hi[i] = v.mres[i] - 1;
    }
    if (!affine_rows(rfm, v.m) || !affine_rows(mfr, v.minv))
        return set_err(m->ctx, AVR_ERR_UNSUPPORTED, "medium transform must be affine (last row 0 0 0 1)");
    if (v.kind == KIND_NVDB) fold_index_from_render(v); else fold_grid_from_render(v);
    size_t nm = size_t(v.mres[0]) * v.mres[1] * v.mres[2];
    AVR_CUDA(m->ctx, cudaMalloc(&m->d_maj, nm * sizeof(float)));
    m->allocs.push_back(m->d_maj);
    v.maj = m->d_maj;
    return AVR_OK;
}

// Bounding box of the non-zero majorant cells (the renderer's DDA stops at its faces, avr_medium.cuh).
avr_status finish_majorants(avr_medium *m) {
    MediumView &v = m->view;
    size_t nm = size_t(v.mres[0]) * v.mres[1] * v.mres[2];
    std::vector<float> h(nm);
    AVR_CUDA(m->ctx, cudaMemcpyAsync(h.data(), m->d_maj, nm * sizeof(float), cudaMemcpyDeviceToHost, m->ctx->stream));
    AVR_CUDA(m->ctx, cudaStreamSynchronize(m->ctx->stream));
    int lo[3] = {v.mres[0], v.mres[1], v.mres[2]}, hi[3] = {-1, -1, -1};
    for (int z = 0; z < v.mres[2]; ++z)
        for (int y = 0; y < v.mres[1]; ++y)
            for (int x = 0; x < v.mres[0]; ++x) {
                float f = h[(size_t(z) * v.mres[1] + y) * v.mres[0] + x];
                if (f == 0.f) continue;  // NaN / negative majorants stay inside the box
                int c[3] = {x, y, z};
                for (int i = 0; i < 3; ++i) {
                    lo[i] = c[i] < lo[i] ? c[i] : lo[i];
                    hi[i] = c[i] > hi[i] ? c[i] : hi[i];
                }
            }
    for (int i = 0; i < 3; ++i) {
        if (hi[i] < lo[i]) { lo[i] = 0; hi[i] = -1; }  // all-zero grid: every walk ends at once
        v.nz_lo[i] = lo[i];
        v.nz_hi[i] = hi[i];
    }
    // FAST kernels: their own majorant grid — c x c x c reference cells per cell (max), with the empty-space
    // distance field folded in
    const int c = m->ctx->knobs.fast_maj_coarsen;
    int fr[3];
    for (int i = 0; i < 3; ++i) {
        fr[i] = (v.mres[i] + c - 1) / c;
        v.fmres[i] = fr[i];
        v.fcell[i] = v.cell[i] * float(c);
        v.finv_cell[i] = v.fcell[i] > 0 ? 1.f / v.fcell[i] : 0.f;
        v.fnz_lo[i] = hi[i] < lo[i] ? 0 : lo[i] / c;
        v.fnz_hi[i] = hi[i] < lo[i] ? -1 : hi[i] / c;
    }
    const size_t nf = size_t(fr[0]) * fr[1] * fr[2];
    std::vector<float> hc(nf, 0.f);
    for (int z = 0; z < v.mres[2]; ++z)
        for (int y = 0; y < v.mres[1]; ++y)
            for (int x = 0; x < v.mres[0]; ++x) {
                float f = h[(size_t(z) * v.mres[1] + y) * v.mres[0] + x];
                float &dst = hc[(size_t(z / c) * fr[1] + y / c) * fr[0] + x / c];
                if (f != f) dst = f;                 // NaN stays visible
                else if (!(dst != dst) && f > dst) dst = f;
            }
    float *mf = nullptr, *mc = nullptr;
    unsigned char *d0 = nullptr, *d1 = nullptr;
    AVR_CUDA(m->ctx, cudaMalloc(&mf, nf * sizeof(float)));
    m->allocs.push_back(mf);
    m->bytes += int64_t(nf * sizeof(float));
    AVR_CUDA(m->ctx, cudaMalloc(&mc, nf * sizeof(float)));
    cudaError_t e = cudaMalloc(&d0, nf);
    if (e == cudaSuccess) e = cudaMalloc(&d1, nf);
    if (e != cudaSuccess) { cudaFree(mc); cudaFree(d0); return set_cuda_err(m->ctx, e, "cudaMalloc(distance field)"); }
    int nb = int((nf + 255) / 256), rx = fr[0], ry = fr[1], rz = fr[2];
    cudaStream_t s = m->ctx->stream;
    e = cudaMemcpyAsync(mc, hc.data(), nf * sizeof(float), cudaMemcpyHostToDevice, s);
    k_empty_dist_x<<<nb, 256, 0, s>>>(mc, rx, ry, rz, d0);
    k_empty_dist_next<<<nb, 256, 0, s>>>(d0, int(nf), rx, ry, d1);
    k_empty_dist_next<<<nb, 256, 0, s>>>(d1, int(nf), rx * ry, rz, d0);
    k_maj_fast<<<nb, 256, 0, s>>>(mc, d0, int(nf), mf);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(mc);
    cudaFree(d0);
    cudaFree(d1);
    if (e != cudaSuccess) return set_cuda_err(m->ctx, e, "empty-space distance field");
    v.maj_fast = mf;
    return AVR_OK;
}

// Octet store of the scalar density grid (layouts 1, 2): cell c holds taps c-1, c on every axis.
avr_status build_bricks(avr_medium *m, const VoxelSrc &src);
avr_status build_octets(avr_medium *m, int layout) {
    if (layout == LAYOUT_DENSE_F32) return AVR_OK;
    if (layout == LAYOUT_BRICK_F16) return build_bricks(m, dense_src(m->view));
    avr_ctx *ctx = m->ctx;
    MediumView &v = m->view;
    v.nbx = (v.nx + 1 + 3) / 4; v.nby = (v.ny + 1 + 3) / 4; v.nbz = (v.nz + 1 + 3) / 4;
    size_t cells = size_t(v.nbx) * v.nby * v.nbz * 64;
    if (cells >= (size_t(1) << 32)) return set_err(ctx, AVR_ERR_UNSUPPORTED, "octet layout: 2^32 or more cells");
    size_t bytes = cells * (layout == LAYOUT_OCTET_F32 ? 32 : 16);
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) return set_cuda_err(ctx, e, "cudaMalloc(octets)");
    m->allocs.push_back(p);
    m->bytes += int64_t(bytes);
    AVR_CUDA(ctx, cudaMemsetAsync(p, 0, bytes, ctx->stream));
    int blocks = ctx->sm_count * 8;
    if (layout == LAYOUT_OCTET_F32) {
        v.oct32 = (const float4 *)p;
        k_pack_octets<false><<<blocks, 256, 0, ctx->stream>>>(v, (float4 *)p, nullptr);
    } else {
        v.oct16 = (const uint4 *)p;
        k_pack_octets<true><<<blocks, 256, 0, ctx->stream>>>(v, nullptr, (uint4 *)p);
    }
    AVR_CUDA(ctx, cudaGetLastError());
    return AVR_OK;
}

// Sparse fp16 brick pool (layout 3): flags -> host prefix sum over the (small) brick grid -> pool fill; the dense
// fp32 grid is released afterwards (every lookup of this medium goes through the pool from then on).
avr_status build_bricks(avr_medium *m, const VoxelSrc &src) {
    avr_ctx *ctx = m->ctx;
    MediumView &v = m->view;
    v.sbx = (v.nx + 1 + 7) / 8; v.sby = (v.ny + 1 + 7) / 8; v.sbz = (v.nz + 1 + 7) / 8;
    const size_t nb = size_t(v.sbx) * v.sby * v.sbz;
    if (nb >= (size_t(1) << 31) || v.sby > 65535 || v.sbz > 65535)
        return set_err(ctx, AVR_ERR_UNSUPPORTED, "brick layout: too many bricks");
    int *d_flags = nullptr;
    AVR_CUDA(ctx, cudaMalloc(&d_flags, nb * sizeof(int)));
    k_brick_flags<<<dim3(v.sbx, v.sby, v.sbz), 128, 0, ctx->stream>>>(src, d_flags);
    std::vector<int> tab(nb), list;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(tab.data(), d_flags, nb * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { cudaFree(d_flags); return set_cuda_err(ctx, e, "brick flags"); }
    for (size_t i = 0; i < nb; ++i) {
        if (tab[i]) { tab[i] = (int)list.size(); list.push_back((int)i); }
        else tab[i] = -1;
    }
    int *d_tab = d_flags;  // reused as the indirection grid
    m->allocs.push_back(d_tab);
    m->bytes += int64_t(nb * sizeof(int));
    AVR_CUDA(ctx, cudaMemcpyAsync(d_tab, tab.data(), nb * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    v.brick_tab = d_tab;
    uint2 *pool = nullptr;
    const size_t slots = list.size() ? list.size() : 1;
    AVR_CUDA(ctx, cudaMalloc(&pool, slots * kBrickEntries * sizeof(uint2)));
    m->allocs.push_back(pool);
    m->bytes += int64_t(slots * kBrickEntries * sizeof(uint2));
    v.brick16 = pool;
    if (!list.empty()) {
        int *d_list = nullptr;
        AVR_CUDA(ctx, cudaMalloc(&d_list, list.size() * sizeof(int)));
        e = cudaMemcpyAsync(d_list, list.data(), list.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->stream);
        if (e == cudaSuccess) {
            k_brick_fill<<<(unsigned)list.size(), 192, 0, ctx->stream>>>(src, d_list, v.sbx, v.sby, pool);
            e = cudaGetLastError();
        }
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        cudaFree(d_list);
        if (e != cudaSuccess) return set_cuda_err(ctx, e, "brick fill");
    } else
        AVR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return AVR_OK;
}
// after every consumer of the dense grid has run (majorants, FAST majorants): give it back
void release_dense(avr_medium *m) {
    MediumView &v = m->view;
    if (v.layout != LAYOUT_BRICK_F16 || !v.dense) return;
    for (size_t i = 0; i < m->allocs.size(); ++i)
        if (m->allocs[i] == (const void *)v.dense) {
            cudaFree(m->allocs[i]);
            m->allocs.erase(m->allocs.begin() + i);
            m->bytes -= int64_t(size_t(v.nx) * v.ny * v.nz * sizeof(float));
            break;
        }
    v.dense = nullptr;
}

void free_medium(avr_medium *m) {
    if (!m) return;
    cudaSetDevice(m->ctx->device);
    for (void *p : m->allocs) cudaFree(p);
    delete m;
}

}  // namespace

extern "C" avr_status avr_medium_create_uniformgrid(avr_ctx *ctx, const avr_grid_desc *d, int layout,
                                                    avr_medium **out) {
    if (!ctx || !d || !out) return set_err(ctx, AVR_ERR_BAD_ARG, "avr_medium_create_uniformgrid: NULL argument");
    *out = nullptr;
    // reference error behaviour: GridMedium::Create ErrorExit()s on these (media.cpp:252-270)
    if (!d->density) return set_err(ctx, AVR_ERR_BAD_ARG, "No \"density\" value provided for grid medium.");
    if (d->nx < 1 || d->ny < 1 || d->nz < 1) return set_err(ctx, AVR_ERR_BAD_ARG, "grid medium: nx, ny, nz must be >= 1");
    if (!d->sigma_a || !d->sigma_s) return set_err(ctx, AVR_ERR_BAD_ARG, "grid medium: sigma_a / sigma_s tables missing");
    if (d->le && d->temperature)
        return set_err(ctx, AVR_ERR_BAD_ARG, "Both \"Le\" and \"temperature\" values were provided.");
    if (layout < 0 || layout > 3) return set_err(ctx, AVR_ERR_BAD_ARG, "unknown grid layout");
    AVR_CUDA(ctx, cudaSetDevice(ctx->device));
    avr_medium *m = new avr_medium();
    m->ctx = ctx;
    MediumView &v = m->view;
    v.kind = KIND_GRID;
    v.layout = layout;
    v.nx = d->nx; v.ny = d->ny; v.nz = d->nz;
    size_t n = size_t(d->nx) * d->ny * d->nz;
    avr_status st;
#define TRY(x) do { st = (x); if (st != AVR_OK) { free_medium(m); return st; } } while (0)
    TRY(upload(m, d->density, n, &v.dense));
    if (d->temperature) TRY(upload(m, d->temperature, n, &v.temperature));
    v.has_temperature = d->temperature != nullptr;
    if (d->lescale) {  // media.cpp:295-306: LeScale[i] *= LeNorm
        std::vector<float> ls(d->lescale, d->lescale + n);
        for (float &x : ls) x *= d->le_norm;
        TRY(upload(m, ls.data(), n, &v.lescale));
        AVR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        v.lnx = d->nx; v.lny = d->ny; v.lnz = d->nz;
    } else {
        float one = d->le_norm;
        TRY(upload(m, &one, 1, &v.lescale));
        AVR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        v.lnx = v.lny = v.lnz = 1;
    }
    std::vector<float> sa(d->sigma_a, d->sigma_a + kNSpec), ss(d->sigma_s, d->sigma_s + kNSpec), le(kNSpec, 0.f);
    for (float &x : sa) x *= d->sigma_scale;  // DenselySampledSpectrum::Scale, media.cpp:229-230
    for (float &x : ss) x *= d->sigma_scale;
    if (d->le) le.assign(d->le, d->le + kNSpec);
    float leMax = le[0];
    bool achro = true;
    for (int i = 0; i < kNSpec; ++i) {
        leMax = std::fmax(leMax, le[i]);
        if (sa[i] != sa[0] || ss[i] != ss[0]) achro = false;
    }
    v.emissive = v.has_temperature ? 1 : (leMax > 0 ? 1 : 0);  // media.cpp:238
    v.achromatic = achro ? 1 : 0;
    v.ach_sig_a = sa[0];
    v.ach_sig_s = ss[0];
    TRY(upload(m, sa.data(), (size_t)kNSpec, &v.sigma_a_spec));
    TRY(upload(m, ss.data(), (size_t)kNSpec, &v.sigma_s_spec));
    TRY(upload(m, le.data(), (size_t)kNSpec, &v.le_spec));
    AVR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    v.g = d->g;
    v.temp_offset = d->temperature_offset;
    v.temp_scale = d->temperature_scale;
    v.sigma_scale = 1.f;
    TRY(finish_common(m, d->p0, d->p1, d->render_from_medium, d->medium_from_render, d->majorant_res));
    k_build_majorants<false><<<dim3(v.mres[0], v.mres[1], v.mres[2]), 128, 0, ctx->stream>>>(v, m->d_maj);
    AVR_CUDA(ctx, cudaGetLastError());
    TRY(build_octets(m, layout));
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { free_medium(m); return set_cuda_err(ctx, e, "medium build"); }
    TRY(finish_majorants(m));
    release_dense(m);
    *out = m;
    return AVR_OK;
}

extern "C" avr_status avr_medium_create_nanovdb(avr_ctx *ctx, const avr_nanovdb_desc *d, int layout, avr_medium **out) {
    if (!ctx || !d || !out) return set_err(ctx, AVR_ERR_BAD_ARG, "avr_medium_create_nanovdb: NULL argument");
    *out = nullptr;
    // NanoVDBMedium::Create ErrorExit()s when the density grid is missing (media.cpp:636-637)
    const bool sparse = !d->density.values && d->density.block_values;
    if (!d->density.values && !sparse) return set_err(ctx, AVR_ERR_BAD_ARG, "nanovdb: didn't find \"density\" grid.");
    for (int i = 0; i < 3; ++i)
        if (d->density.dims[i] < 1 || (d->temperature.values && d->temperature.dims[i] < 1))
            return set_err(ctx, AVR_ERR_BAD_ARG, "nanovdb: empty index bounding box");
    if (!d->sigma_a || !d->sigma_s) return set_err(ctx, AVR_ERR_BAD_ARG, "nanovdb medium: sigma_a / sigma_s tables missing");
    if (layout < 0 || layout > 3) return set_err(ctx, AVR_ERR_BAD_ARG, "unknown grid layout");
    if (sparse && (layout != LAYOUT_BRICK_F16 || d->temperature.values || d->temperature.block_values))
        return set_err(ctx, AVR_ERR_UNSUPPORTED, "nanovdb: a grid given as 8^3 blocks needs layout brick16 and no temperature grid");
    if (sparse && (d->density.n_blocks < 0 || d->density.n_blocks >= (int64_t(1) << 31) || (d->density.n_blocks > 0 && !d->density.block_origin)))
        return set_err(ctx, AVR_ERR_BAD_ARG, "nanovdb: n_blocks / block_origin inconsistent");
    if (d->temperature.block_values && !d->temperature.values)
        return set_err(ctx, AVR_ERR_UNSUPPORTED, "nanovdb: the temperature grid must be given flattened");
    AVR_CUDA(ctx, cudaSetDevice(ctx->device));
    avr_medium *m = new avr_medium();
    m->ctx = ctx;
    MediumView &v = m->view;
    v.kind = KIND_NVDB;
    v.layout = layout;
    const avr_nanovdb_grid &dg = d->density, &tg = d->temperature;
    v.nx = dg.dims[0]; v.ny = dg.dims[1]; v.nz = dg.dims[2];
    avr_status st;
    VoxelSrc src{};
    float *d_bvals = nullptr;   // the blocks and their table live only until the pool and the majorants are built
    int *d_btab = nullptr;
    if (!sparse) {
        TRY(upload(m, dg.values, size_t(v.nx) * v.ny * v.nz, &v.dense));
        src = dense_src(v);
    } else {
        src.n[0] = v.nx; src.n[1] = v.ny; src.n[2] = v.nz;
        size_t cells = 1;
        for (int i = 0; i < 3; ++i) {
            src.imin[i] = dg.index_min[i];
            src.bo[i] = dg.index_min[i] >> 3;
            src.tb[i] = ((dg.index_min[i] + dg.dims[i] - 1) >> 3) - src.bo[i] + 1;
            cells *= size_t(src.tb[i]);
        }
        std::vector<int> tab(cells, -1);
        for (int64_t b = 0; b < dg.n_blocks; ++b) {
            const int32_t *o = dg.block_origin + 3 * b;
            int c[3];
            bool inside = true;
            for (int i = 0; i < 3; ++i) {
                if (o[i] & 7) { free_medium(m); return set_err(ctx, AVR_ERR_BAD_ARG, "nanovdb: block origins must be multiples of 8"); }
                c[i] = (o[i] >> 3) - src.bo[i];
                inside = inside && c[i] >= 0 && c[i] < src.tb[i];
            }
            if (inside) tab[(size_t(c[2]) * src.tb[1] + c[1]) * src.tb[0] + c[0]] = int(b);   // blocks off the index box are never read
        }
        const size_t nv = size_t(dg.n_blocks > 0 ? dg.n_blocks : 1) * 512;
        cudaError_t e = cudaMalloc(&d_bvals, nv * sizeof(float));
        if (e == cudaSuccess) e = cudaMalloc(&d_btab, cells * sizeof(int));
        if (e == cudaSuccess && dg.n_blocks > 0)
            e = cudaMemcpyAsync(d_bvals, dg.block_values, size_t(dg.n_blocks) * 512 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(d_btab, tab.data(), cells * sizeof(int), cudaMemcpyHostToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);   // `tab` leaves scope
        if (e != cudaSuccess) { cudaFree(d_bvals); cudaFree(d_btab); free_medium(m); return set_cuda_err(ctx, e, "nanovdb blocks"); }
        src.bvals = d_bvals;
        src.btab = d_btab;
    }
    struct SrcGuard {   // freed on every way out
        float *a; int *b;
        ~SrcGuard() { cudaFree(a); cudaFree(b); }
    } guard{d_bvals, d_btab};
    float p0[3], p1[3];
    for (int i = 0; i < 3; ++i) {
        v.nv_imin[i] = dg.index_min[i];
        v.nv_vec[i] = dg.vec[i];
        p0[i] = float(dg.world_bbox[i]);  // Bounds3f(Point3f(bbox.min()...)), media.cpp:536-538
        p1[i] = float(dg.world_bbox[3 + i]);
    }
    for (int i = 0; i < 9; ++i) v.nv_inv[i] = dg.inv_mat[i];
    v.has_temperature = tg.values != nullptr;
    if (v.has_temperature) {
        TRY(upload(m, tg.values, size_t(tg.dims[0]) * tg.dims[1] * tg.dims[2], &v.temperature));
        for (int i = 0; i < 3; ++i) {
            v.nv_tdim[i] = tg.dims[i];
            v.nv_timin[i] = tg.index_min[i];
            v.nv_tvec[i] = tg.vec[i];
            p0[i] = std::fmin(p0[i], float(tg.world_bbox[i]));  // Union, media.cpp:546-550
            p1[i] = std::fmax(p1[i], float(tg.world_bbox[3 + i]));
        }
        for (int i = 0; i < 9; ++i) v.nv_tinv[i] = tg.inv_mat[i];
    }
    std::vector<float> sa(d->sigma_a, d->sigma_a + kNSpec), ss(d->sigma_s, d->sigma_s + kNSpec), le(kNSpec, 0.f);
    bool achro = true;
    for (int i = 0; i < kNSpec; ++i) {
        sa[i] *= d->sigma_scale;
        ss[i] *= d->sigma_scale;
        if (sa[i] != sa[0] || ss[i] != ss[0]) achro = false;
    }
    v.achromatic = achro ? 1 : 0;
    v.ach_sig_a = sa[0];
    v.ach_sig_s = ss[0];
    // IsEmissive, media.h:621.  A temperature grid with Lescale <= 0 still goes through Le() in the
    // reference; the general kernels are used whenever the grid is present.
    v.emissive = v.has_temperature ? 1 : 0;
    m->reports_emissive = (v.has_temperature && d->le_scale > 0) ? 1 : 0;
    TRY(upload(m, sa.data(), (size_t)kNSpec, &v.sigma_a_spec));
    TRY(upload(m, ss.data(), (size_t)kNSpec, &v.sigma_s_spec));
    TRY(upload(m, le.data(), (size_t)kNSpec, &v.le_spec));
    AVR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    v.g = d->g;
    v.le_scale = d->le_scale;
    v.temp_offset = d->temperature_offset;
    v.temp_scale = d->temperature_scale;
    v.sigma_scale = 1.f;
    TRY(finish_common(m, p0, p1, d->render_from_medium, d->medium_from_render, d->majorant_res, 64));
    k_build_majorants_nvdb<<<dim3(v.mres[0], v.mres[1], v.mres[2]), 128, 0, ctx->stream>>>(v, src, m->d_maj);
    AVR_CUDA(ctx, cudaGetLastError());
    if (sparse) TRY(build_bricks(m, src));
    else TRY(build_octets(m, layout));
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { free_medium(m); return set_cuda_err(ctx, e, "nanovdb medium build"); }
    TRY(finish_majorants(m));
    release_dense(m);
    *out = m;
    return AVR_OK;
}

extern "C" avr_status avr_medium_create_rgbgrid(avr_ctx *ctx, const avr_rgbgrid_desc *d, avr_medium **out) {
    if (!ctx || !d || !out) return set_err(ctx, AVR_ERR_BAD_ARG, "avr_medium_create_rgbgrid: NULL argument");
    *out = nullptr;
    // RGBGridMedium::Create, media.cpp:386-407
    if (!d->sigma_a && !d->sigma_s)
        return set_err(ctx, AVR_ERR_BAD_ARG, "RGB grid requires \"sigma_a\" and/or \"sigma_s\" parameter values.");
    if (d->le && !d->sigma_a) return set_err(ctx, AVR_ERR_BAD_ARG, "RGB grid requires \"sigma_a\" if \"Le\" value provided.");
    if (d->le && !d->le_illuminant) return set_err(ctx, AVR_ERR_BAD_ARG, "rgbgrid: le_illuminant table missing");
    if (d->nx < 1 || d->ny < 1 || d->nz < 1) return set_err(ctx, AVR_ERR_BAD_ARG, "rgbgrid: nx, ny, nz must be >= 1");
    AVR_CUDA(ctx, cudaSetDevice(ctx->device));
    avr_medium *m = new avr_medium();
    m->ctx = ctx;
    MediumView &v = m->view;
    v.kind = KIND_RGB;
    v.layout = LAYOUT_DENSE_F32;
    v.nx = d->nx; v.ny = d->ny; v.nz = d->nz;
    size_t n = size_t(d->nx) * d->ny * d->nz;
    avr_status st;
    if (d->sigma_a) TRY(upload(m, (const float4 *)d->sigma_a, n, &v.rgb_a));
    if (d->sigma_s) TRY(upload(m, (const float4 *)d->sigma_s, n, &v.rgb_s));
    if (d->le) {
        TRY(upload(m, (const float4 *)d->le, n, &v.rgb_le));
        TRY(upload(m, d->le_illuminant, (size_t)kNSpec, &v.le_illum));
    }
    v.le_scale = d->le_scale;
    v.sigma_scale = d->sigma_scale;
    v.g = d->g;
    v.emissive = (d->le && d->le_scale > 0) ? 1 : 0;  // media.h:374
    v.achromatic = 0;
    TRY(finish_common(m, d->p0, d->p1, d->render_from_medium, d->medium_from_render, d->majorant_res));
    k_build_majorants<true><<<dim3(v.mres[0], v.mres[1], v.mres[2]), 128, 0, ctx->stream>>>(v, m->d_maj);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { free_medium(m); return set_cuda_err(ctx, e, "rgb medium build"); }
    TRY(finish_majorants(m));
    release_dense(m);
    *out = m;
    return AVR_OK;
#undef TRY
}

extern "C" void avr_medium_destroy(avr_medium *m) { free_medium(m); }
extern "C" int avr_medium_is_emissive(const avr_medium *m) {
    return m ? (m->reports_emissive >= 0 ? m->reports_emissive : m->view.emissive) : 0;
}
extern "C" avr_status avr_medium_bounds(const avr_medium *m, float out[6]) {
    if (!m || !out) return set_err(m ? m->ctx : nullptr, AVR_ERR_BAD_ARG, "avr_medium_bounds: NULL argument");
    for (int i = 0; i < 3; ++i) { out[i] = m->view.bmin[i]; out[3 + i] = m->view.bmax[i]; }
    return AVR_OK;
}
extern "C" int64_t avr_medium_bytes(const avr_medium *m) { return m ? m->bytes : 0; }

extern "C" avr_status avr_medium_majorants(avr_medium *m, float *out, int32_t res[3]) {
    if (!m || !out || !res) return set_err(m ? m->ctx : nullptr, AVR_ERR_BAD_ARG, "avr_medium_majorants: NULL argument");
    AVR_CUDA(m->ctx, cudaSetDevice(m->ctx->device));
    for (int i = 0; i < 3; ++i) res[i] = m->mres[i];
    size_t nm = size_t(m->mres[0]) * m->mres[1] * m->mres[2];
    AVR_CUDA(m->ctx, cudaMemcpy(out, m->d_maj, nm * sizeof(float), cudaMemcpyDeviceToHost));
    return AVR_OK;
}

#include "avr_queries.inc"
