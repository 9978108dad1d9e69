// Data-parallel gradient exchange and optimiser step as ONE kernel over NVLink peer memory (one process per GPU).
//
// The reference sums the replicas' gradients inside nn.DataParallel and runs torch.optim.Adam on the full model
// (scripts/train_rl_captioning_module.py:81-83,97-99).  The NCCL form of that is all-reduce (2 (N-1)/N x 208 MB over
// the links) followed by N identical Adam updates (1.3 GB of HBM traffic on every GPU).  Here the flat gradient buffer,
// the flat parameter buffer and the bf16 weight shadows of every rank live in cudaMalloc'd arenas that the other ranks
// map through CUDA IPC, at identical offsets, and a rank OWNS every N-th 8192-element chunk:
//
//   adam_peer_kernel   for an owned chunk: g = sum over ranks r = 0 .. N-1 of rank r's gradient (peer loads, fixed order:
//                      bit-reproducible, and every element is summed by exactly one rank, so the replicas cannot drift
//                      apart), Adam on the owner's moments, then the new parameter (fp32) and its bf16 tensor-core
//                      shadow are stored into EVERY rank's buffers (peer stores).
//                      = reduce-scatter + sharded optimiser + all-gather, one pass, Adam state traffic divided by N.
//   peer_signal_kernel / peer_wait_kernel   the cross-GPU ordering around it: "my gradients of this stage are final"
//                      before anyone reads them, "I am done reading / writing" at the end of the step.  A signal is a
//                      release store of the slot's step count into every rank's flag table, a wait spins (one warp, one
//                      CTA, never the big grid) on the local table.  Counts only grow: CUDA-graph replays need no reset.
//
// Nothing here allocates or synchronises on the host after set-up; every launch is capturable.
#include <cuda_bf16.h>
#include <string.h>
#include "common.cuh"
#include "../../include/bmhrl_b200.h"

namespace bmhrl {

constexpr int PEER_MAX = 8;
constexpr int PEER_SLOTS = 32;
constexpr int ADAM_PEER_CHUNK = 8192;   // = ADAM_CHUNK of misc.cu (checked by the caller through bmhrl_adam_chunk_elems)

struct PeerOff {
    long long b[PEER_MAX];   // BYTE offset from a local arena address to the same address in rank r's arena (0 for self)
};

// flags: (PEER_SLOTS, PEER_MAX) uint64 arrival counts + (PEER_SLOTS) uint64 local signal counts behind them
__global__ void peer_signal_kernel(unsigned long long* flags, PeerOff off, int n, int rank, int slot) {
    __shared__ unsigned long long epoch;
    if (threadIdx.x == 0) {
        unsigned long long* cnt = flags + PEER_SLOTS * PEER_MAX + slot;
        epoch = *cnt + 1ull;
        *cnt = epoch;
    }
    __syncthreads();
    if ((int)threadIdx.x < n) {
        __threadfence_system();   // everything this GPU wrote before the signal (previous kernels of the stream included)
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(flags) + off.b[threadIdx.x]) +
                                  slot * PEER_MAX + rank;
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(dst), "l"(epoch) : "memory");
    }
}

__global__ void peer_wait_kernel(const unsigned long long* flags, int n, int slot) {
    if ((int)threadIdx.x < n) {
        const unsigned long long want = flags[PEER_SLOTS * PEER_MAX + slot];   // what the preceding signal of this rank sent
        const unsigned long long* src = flags + slot * PEER_MAX + threadIdx.x;
        unsigned long long got;
        unsigned long long spins = 0;
        do {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(got) : "l"(src) : "memory");
            if (got >= want) break;
            __nanosleep(200);
            if (++spins > (1ull << 24)) __trap();   // ~4 s: a rank that never arrives must fault, not hang the GPU
        } while (true);
    }
    __threadfence_system();
}

__device__ __forceinline__ float4 ld_peer_f4(const float* p, long long byte_off) {
    return *reinterpret_cast<const float4*>(reinterpret_cast<const char*>(p) + byte_off);
}

// Same arithmetic as adam_multi_kernel (misc.cu) on the chunks this rank owns.
__device__ __forceinline__ void adam_peer_chunk(int chunk, float* const* __restrict__ params,
                                                        const float* const* __restrict__ grads,
                                                        float* const* __restrict__ exp_avg,
                                                        float* const* __restrict__ exp_avg_sq,
                                                        const long long* __restrict__ numels,
                                                        const int* __restrict__ chunk_tensor,
                                                        const int* __restrict__ chunk_start,
                                                        const float* __restrict__ step_ptr, float lr, float beta1, float beta2,
                                                        float eps, float grad_scale, void* const* __restrict__ shadow,
                                                        const int* __restrict__ shadow_cols, const int* __restrict__ shadow_ld,
                                                        const int* __restrict__ shadow_kind, PeerOff og, PeerOff op, PeerOff os,
                                                        int n_ranks, int owner_only_master) {
    const int t = chunk_tensor[chunk];
    const long long s0 = chunk_start[chunk];
    const long long n = numels[t];
    float* p = params[t];
    const float* g = grads[t];
    float* __restrict__ m = exp_avg[t];
    float* __restrict__ v = exp_avg_sq[t];
    const float step = step_ptr[t];
    const float bc1 = 1.f - powf(beta1, step);
    const float bc2 = 1.f - powf(beta2, step);
    const float step_size = lr / bc1;
    const float inv_sqrt_bc2 = rsqrtf(bc2);
    long long end = s0 + ADAM_PEER_CHUNK;
    if (end > n) end = n;
    void* sh = shadow ? shadow[t] : nullptr;
    const int sk = sh ? shadow_kind[t] : 0;
    const int scols = sh ? shadow_cols[t] : 1, sld = sh ? shadow_ld[t] : 1;
    auto upd = [&](float pi, float gi, float& mi, float& vi) {
        return adam_update(pi, gi, mi, vi, beta1, beta2, eps, grad_scale, step_size, inv_sqrt_bc2);
    };
    // owner_only_master: a weight that has a bf16 tensor-core shadow is READ through that shadow only (bf16 math mode), so
    // its fp32 master is needed where its Adam moments are -- on the owner.  Not broadcasting it removes 4 of the 6 bytes a
    // GEMM weight element sends to every peer.  The replicas' fp32 copies of such weights then go stale until
    // bmhrl_peer_bcast_params_f32 (checkpoints, math-mode switches).
    const bool bcast_master = !(owner_only_master && sh && sk == 0);
    const bool vec = ((((uintptr_t)p) | ((uintptr_t)g) | ((uintptr_t)m) | ((uintptr_t)v)) & 15u) == 0 &&
                     (!sh || (sk == 1 ? ((uintptr_t)sh & 15u) == 0
                                      : (((uintptr_t)sh & 7u) == 0 && (scols == sld || ((scols & 3) == 0 && (sld & 3) == 0)))));
    if (vec) {
        const long long e4 = s0 + ((end - s0) & ~3ll);
        for (long long i = s0 + 4 * threadIdx.x; i < e4; i += 4 * 256) {
            float4 g4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int r = 0; r < PEER_MAX; ++r)
                if (r < n_ranks) {
                    const float4 t4 = ld_peer_f4(g + i, og.b[r]);
                    g4.x += t4.x; g4.y += t4.y; g4.z += t4.z; g4.w += t4.w;
                }
            const float4 p4 = ld_f4(p + i);
            float4 m4 = ld_f4(m + i), v4 = ld_f4(v + i), o;
            o.x = upd(p4.x, g4.x, m4.x, v4.x);
            o.y = upd(p4.y, g4.y, m4.y, v4.y);
            o.z = upd(p4.z, g4.z, m4.z, v4.z);
            o.w = upd(p4.w, g4.w, m4.w, v4.w);
            st_f4(m + i, m4);
            st_f4(v + i, v4);
            uint2 w;
            const long long si = (sh && sk == 0 && scols != sld) ? (i / scols) * sld + (i % scols) : i;
            if (sh && sk == 0) {
                asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w.x) : "f"(o.y), "f"(o.x));
                asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w.y) : "f"(o.w), "f"(o.z));
            }
            if (!bcast_master) st_f4(p + i, o);
#pragma unroll
            for (int r = 0; r < PEER_MAX; ++r)
                if (r < n_ranks) {
                    if (bcast_master) *reinterpret_cast<float4*>(reinterpret_cast<char*>(p + i) + op.b[r]) = o;
                    if (sh) {
                        if (sk == 1)
                            *reinterpret_cast<float4*>(reinterpret_cast<char*>(reinterpret_cast<float*>(sh) + i) + os.b[r]) = o;
                        else   // 4 consecutive elements never straddle a row (scols % 4 == 0); padded pitch sld
                            *reinterpret_cast<uint2*>(reinterpret_cast<char*>(reinterpret_cast<unsigned short*>(sh) + si) + os.b[r]) = w;
                    }
                }
        }
        if (threadIdx.x >= (int)(end - e4)) return;
    }
    // scalar form: the tail of a tensor whose numel is not a multiple of 4, or a shadow with a padded row pitch
    const long long b0 = vec ? s0 + ((end - s0) & ~3ll) : s0;
    for (long long i = b0 + threadIdx.x; i < end; i += 256) {
        float gi = 0.f;
        for (int r = 0; r < n_ranks; ++r) gi += *reinterpret_cast<const float*>(reinterpret_cast<const char*>(g + i) + og.b[r]);
        float mi = m[i], vi = v[i];
        const float o = upd(p[i], gi, mi, vi);
        m[i] = mi;
        v[i] = vi;
        if (!bcast_master) p[i] = o;
        for (int r = 0; r < n_ranks; ++r) {
            if (bcast_master) *reinterpret_cast<float*>(reinterpret_cast<char*>(p + i) + op.b[r]) = o;
            if (sh) {
                if (sk == 1) {
                    *reinterpret_cast<float*>(reinterpret_cast<char*>(reinterpret_cast<float*>(sh) + i) + os.b[r]) = o;
                } else {
                    const long long row = i / scols, col = i - row * scols;
                    *reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(reinterpret_cast<__nv_bfloat16*>(sh) + row * sld + col) +
                                                      os.b[r]) = __float2bfloat16(o);
                }
            }
        }
    }
}

// One CTA per owned chunk, or a grid-stride loop under bmhrl_adam_set_grid_limit (see adam_multi_kernel, misc.cu).
__global__ void __launch_bounds__(256) adam_peer_kernel(float* const* __restrict__ params, const float* const* __restrict__ grads,
                                                        float* const* __restrict__ exp_avg, float* const* __restrict__ exp_avg_sq,
                                                        const long long* __restrict__ numels, const int* __restrict__ chunk_tensor,
                                                        const int* __restrict__ chunk_start, int n_chunks,
                                                        const float* __restrict__ step_ptr, float lr, float beta1, float beta2,
                                                        float eps, float grad_scale, void* const* __restrict__ shadow,
                                                        const int* __restrict__ shadow_cols, const int* __restrict__ shadow_ld,
                                                        const int* __restrict__ shadow_kind, PeerOff og, PeerOff op, PeerOff os,
                                                        int n_ranks, int owner_only_master) {
    for (int c = blockIdx.x; c < n_chunks; c += gridDim.x)
        adam_peer_chunk(c, params, grads, exp_avg, exp_avg_sq, numels, chunk_tensor, chunk_start, step_ptr, lr, beta1, beta2, eps,
                        grad_scale, shadow, shadow_cols, shadow_ld, shadow_kind, og, op, os, n_ranks, owner_only_master);
}

// The owner's fp32 values of the listed chunks into every rank's parameter arena (after steps with owner_only_master).
__global__ void __launch_bounds__(256) peer_bcast_params_kernel(float* const* __restrict__ params,
                                                                const long long* __restrict__ numels,
                                                                const int* __restrict__ chunk_tensor,
                                                                const int* __restrict__ chunk_start, PeerOff op, int n_ranks) {
    const int t = chunk_tensor[blockIdx.x];
    const long long s0 = chunk_start[blockIdx.x];
    const long long n = numels[t];
    const float* p = params[t];
    long long end = s0 + ADAM_PEER_CHUNK;
    if (end > n) end = n;
    for (long long i = s0 + threadIdx.x; i < end; i += 256) {
        const float o = p[i];
        for (int r = 0; r < n_ranks; ++r)
            if (op.b[r] != 0) *reinterpret_cast<float*>(reinterpret_cast<char*>(const_cast<float*>(p) + i) + op.b[r]) = o;
    }
}

static int fill_off(PeerOff& o, const long long* src, int n) {
    for (int r = 0; r < PEER_MAX; ++r) o.b[r] = (src && r < n) ? src[r] : 0;
    return 0;
}

}  // namespace bmhrl
using namespace bmhrl;

extern "C" int bmhrl_peer_max_ranks(void) { return PEER_MAX; }
extern "C" long long bmhrl_peer_flag_bytes(void) { return (long long)sizeof(unsigned long long) * (PEER_SLOTS * PEER_MAX + PEER_SLOTS); }
extern "C" int bmhrl_peer_slots(void) { return PEER_SLOTS; }

extern "C" int bmhrl_peer_alloc(long long bytes, void** out) {
    BMHRL_CHECK_ARG(bytes > 0 && out, "bmhrl_peer_alloc: bad arguments");
    void* p = nullptr;
    if (cudaMalloc(&p, (size_t)bytes) != cudaSuccess || cudaMemset(p, 0, (size_t)bytes) != cudaSuccess) {
        bmhrl_set_error(cudaGetErrorString(cudaGetLastError()));
        return BMHRL_ERR_CUDA;
    }
    *out = p;
    return BMHRL_OK;
}

extern "C" int bmhrl_peer_free(void* p) {
    if (p && cudaFree(p) != cudaSuccess) {
        bmhrl_set_error(cudaGetErrorString(cudaGetLastError()));
        return BMHRL_ERR_CUDA;
    }
    return BMHRL_OK;
}

extern "C" int bmhrl_peer_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }

extern "C" int bmhrl_peer_get_handle(const void* p, unsigned char* handle) {
    BMHRL_CHECK_ARG(p && handle, "bmhrl_peer_get_handle: null pointer");
    cudaIpcMemHandle_t h;
    if (cudaIpcGetMemHandle(&h, const_cast<void*>(p)) != cudaSuccess) {
        bmhrl_set_error(cudaGetErrorString(cudaGetLastError()));
        return BMHRL_ERR_CUDA;
    }
    memcpy(handle, &h, sizeof(h));
    return BMHRL_OK;
}

extern "C" int bmhrl_peer_open_handle(const unsigned char* handle, void** out) {
    BMHRL_CHECK_ARG(handle && out, "bmhrl_peer_open_handle: null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        bmhrl_set_error(cudaGetErrorString(cudaGetLastError()));
        return BMHRL_ERR_CUDA;
    }
    *out = p;
    return BMHRL_OK;
}

extern "C" int bmhrl_peer_close_handle(void* p) {
    if (p && cudaIpcCloseMemHandle(p) != cudaSuccess) {
        bmhrl_set_error(cudaGetErrorString(cudaGetLastError()));
        return BMHRL_ERR_CUDA;
    }
    return BMHRL_OK;
}

extern "C" int bmhrl_peer_signal(unsigned long long* flags, const long long* flag_off_bytes, int n_ranks, int rank, int slot,
                                 void* stream) {
    BMHRL_CHECK_ARG(flags && flag_off_bytes && n_ranks >= 1 && n_ranks <= PEER_MAX && rank >= 0 && rank < n_ranks &&
                        slot >= 0 && slot < PEER_SLOTS,
                    "bmhrl_peer_signal: bad arguments");
    PeerOff o;
    fill_off(o, flag_off_bytes, n_ranks);
    peer_signal_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(flags, o, n_ranks, rank, slot);
    BMHRL_POST_LAUNCH("peer_signal");
    return BMHRL_OK;
}

extern "C" int bmhrl_peer_wait(const unsigned long long* flags, int n_ranks, int slot, void* stream) {
    BMHRL_CHECK_ARG(flags && n_ranks >= 1 && n_ranks <= PEER_MAX && slot >= 0 && slot < PEER_SLOTS,
                    "bmhrl_peer_wait: bad arguments");
    peer_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(flags, n_ranks, slot);
    BMHRL_POST_LAUNCH("peer_wait");
    return BMHRL_OK;
}

extern "C" int bmhrl_adam_peer_f32(float* const* params, const float* const* grads, float* const* exp_avg,
                                   float* const* exp_avg_sq, const long long* numels, const int* chunk_tensor,
                                   const int* chunk_start, int n_chunks, const float* step_counts, float lr, float beta1,
                                   float beta2, float eps, float grad_scale, void* const* shadow, const int* shadow_cols,
                                   const int* shadow_ld, const int* shadow_kind, const long long* grad_off_bytes,
                                   const long long* param_off_bytes, const long long* shadow_off_bytes, int n_ranks,
                                   int owner_only_master, void* stream) {
    BMHRL_CHECK_ARG(params && grads && exp_avg && exp_avg_sq && numels && chunk_tensor && chunk_start && step_counts,
                    "bmhrl_adam_peer_f32: null pointer");
    BMHRL_CHECK_ARG(!shadow || (shadow_cols && shadow_ld && shadow_kind && shadow_off_bytes), "bmhrl_adam_peer_f32: shadow tables");
    BMHRL_CHECK_ARG(grad_off_bytes && param_off_bytes && n_ranks >= 1 && n_ranks <= PEER_MAX, "bmhrl_adam_peer_f32: peer offsets");
    if (n_chunks <= 0) return BMHRL_OK;
    PeerOff og, op, os;
    fill_off(og, grad_off_bytes, n_ranks);
    fill_off(op, param_off_bytes, n_ranks);
    fill_off(os, shadow_off_bytes, n_ranks);
    const int lim = bmhrl_adam_get_grid_limit();
    const int grid = (lim > 0 && n_chunks > lim) ? lim : n_chunks;
    adam_peer_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(params, grads, exp_avg, exp_avg_sq, numels, chunk_tensor,
                                                                 chunk_start, n_chunks, step_counts, lr, beta1, beta2, eps, grad_scale,
                                                                 shadow, shadow_cols, shadow_ld, shadow_kind, og, op, os,
                                                                 n_ranks, owner_only_master ? 1 : 0);
    BMHRL_POST_LAUNCH("adam_peer");
    return BMHRL_OK;
}

extern "C" int bmhrl_peer_bcast_params_f32(float* const* params, const long long* numels, const int* chunk_tensor,
                                           const int* chunk_start, int n_chunks, const long long* param_off_bytes, int n_ranks,
                                           void* stream) {
    BMHRL_CHECK_ARG(params && numels && chunk_tensor && chunk_start && param_off_bytes && n_ranks >= 1 && n_ranks <= PEER_MAX,
                    "bmhrl_peer_bcast_params_f32: bad arguments");
    if (n_chunks <= 0) return BMHRL_OK;
    PeerOff op;
    fill_off(op, param_off_bytes, n_ranks);
    peer_bcast_params_kernel<<<n_chunks, 256, 0, (cudaStream_t)stream>>>(params, numels, chunk_tensor, chunk_start, op, n_ranks);
    BMHRL_POST_LAUNCH("peer_bcast_params");
    return BMHRL_OK;
}
