// Shared device helpers for libbmhrl_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include <atomic>

#define BMHRL_OK 0
#define BMHRL_ERR_ARG (-1)
#define BMHRL_ERR_CUDA (-2)
#define BMHRL_ERR_UNSUPPORTED (-3)

extern "C" void bmhrl_set_error(const char* msg);
extern "C" void bmhrl_count_launch(int n);

#define BMHRL_CHECK_ARG(cond, msg)                 \
    do {                                           \
        if (!(cond)) {                             \
            bmhrl_set_error(msg);                  \
            return BMHRL_ERR_ARG;                  \
        }                                          \
    } while (0)

// Call after every kernel launch: surfaces launch-configuration errors without synchronising.
#define BMHRL_POST_LAUNCH(name)                                    \
    do {                                                           \
        cudaError_t _e = cudaGetLastError();                       \
        if (_e != cudaSuccess) {                                   \
            bmhrl_set_error(cudaGetErrorString(_e));               \
            return BMHRL_ERR_CUDA;                                 \
        }                                                          \
        bmhrl_count_launch(1);                                     \
    } while (0)

namespace bmhrl {

// ---------------------------------------------------------------- per-device one-time set-up (host)
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) and the SM count are PER DEVICE: a process that uses several
// devices (tests on cuda:1, single-process multi-GPU) must repeat the set-up on each.  The flags are atomics and the
// set-up calls idempotent, so two threads racing through the first call on a device are harmless.
constexpr int kMaxDevices = 64;
struct DeviceOnce {
    std::atomic<unsigned char> done[kMaxDevices];
};
inline int current_device_slot() {
    int d = 0;
    cudaGetDevice(&d);
    return (d < 0 || d >= kMaxDevices) ? 0 : d;
}
// true when `once` has not been completed on the current device yet; call mark_done(once) after a successful set-up
inline bool needs_setup(DeviceOnce& once) { return once.done[current_device_slot()].load(std::memory_order_acquire) == 0; }
inline void mark_done(DeviceOnce& once) { once.done[current_device_slot()].store(1, std::memory_order_release); }
inline int device_sm_count() {
    static std::atomic<int> n[kMaxDevices];
    const int slot = current_device_slot();
    int v = n[slot].load(std::memory_order_relaxed);
    if (v <= 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
        if (v <= 0) v = 148;
        n[slot].store(v, std::memory_order_relaxed);
    }
    return v;
}

// ---------------------------------------------------------------- programmatic dependent launch (PDL)
// The kernels of a residual block form a dependency chain on one stream (LayerNorm -> projection GEMM -> attention ->
// out-projection GEMM -> ...), each a few microseconds long.  Launched with the programmatic-stream-serialisation
// attribute, the NEXT kernel's CTAs become resident as soon as every CTA of the current one has passed its own
// `griddep_wait()` + `griddep_launch_dependents()`: its prologue (barrier init, TMEM allocation, tensor-map fetch) and
// the TMA loads of operands that no earlier launch of the stream produces (weights) overlap the current kernel's body.
// Rules every kernel launched through launch_pdl() follows: (1) no memory a predecessor may have written is read
// before griddep_wait(); (2) launch_dependents is issued only AFTER the own wait, so a dependent never runs ahead of its
// predecessor's predecessors.  Without the launch attribute both instructions are no-ops.
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

int pdl_enabled();   // api.cu: bmhrl_set_pdl(), BMHRL_PDL environment variable
// api.cu: bmhrl_set_launch_priority().  kNoPriority: launches carry no priority attribute (they take their stream's);
// otherwise the value becomes cudaLaunchAttributePriority of every launch_pdl / GEMM launch: inside a CUDA graph that was
// instantiated with cudaGraphInstantiateFlagUseNodePriority the pending CTAs of a higher-priority node are placed first.
constexpr int kNoPriority = 1 << 30;
int launch_priority();

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    int na = 0;
    if (pdl_enabled()) {
        at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    const int prio = launch_priority();
    if (prio != kNoPriority) {
        at[na].id = cudaLaunchAttributePriority;
        at[na].val.priority = prio;
        ++na;
    }
    if (na) {
        cfg.attrs = at;
        cfg.numAttrs = na;
    }
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#define BMHRL_LAUNCH_PDL(name, ...)                                \
    do {                                                           \
        cudaError_t _e = launch_pdl(__VA_ARGS__);                  \
        if (_e != cudaSuccess) {                                   \
            bmhrl_set_error(cudaGetErrorString(_e));               \
            return BMHRL_ERR_CUDA;                                 \
        }                                                          \
        bmhrl_count_launch(1);                                     \
    } while (0)

constexpr float kNegFill = -1e9f;  // reference replaces masked scores by -1e9 (model/multihead_attention.py:21-22)

// ---------------------------------------------------------------- Philox4x32-10 (counter based RNG)
struct Philox4 {
    uint32_t x, y, z, w;
};

__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                 uint32_t k1) {
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    return Philox4{c0, c1, c2, c3};
}

// rng_state: device pointer to {seed, step}.  salt identifies the call site; idx the element.
// Four consecutive elements share one Philox block.
__device__ __forceinline__ uint32_t philox_u32(const unsigned long long* rng_state, uint32_t salt, uint64_t idx) {
    unsigned long long seed = rng_state[0], step = rng_state[1];
    uint64_t blk = idx >> 2;
    Philox4 r = philox4x32_10((uint32_t)blk, (uint32_t)(blk >> 32), salt, (uint32_t)step, (uint32_t)seed,
                              (uint32_t)(seed >> 32) ^ (uint32_t)(step >> 32));
    uint32_t lane = (uint32_t)idx & 3u;
    return lane == 0 ? r.x : lane == 1 ? r.y : lane == 2 ? r.z : r.w;
}

__device__ __forceinline__ Philox4 philox_block(const unsigned long long* rng_state, uint32_t salt, uint64_t blk) {
    unsigned long long seed = rng_state[0], step = rng_state[1];
    return philox4x32_10((uint32_t)blk, (uint32_t)(blk >> 32), salt, (uint32_t)step, (uint32_t)seed,
                         (uint32_t)(seed >> 32) ^ (uint32_t)(step >> 32));
}

// uniform in [0,1): 24 random bits
__device__ __forceinline__ float u32_to_unit(uint32_t u) { return (float)(u >> 8) * (1.0f / 16777216.0f); }

// dropout keep decision shared by forward and backward of every op: keep iff u >= p
__device__ __forceinline__ bool dropout_keep(const unsigned long long* rng_state, uint32_t salt, uint64_t idx,
                                             float p) {
    return u32_to_unit(philox_u32(rng_state, salt, idx)) >= p;
}

// ---------------------------------------------------------------- reductions
// One Adam update (torch.optim.Adam, no weight decay / amsgrad) with every rounding spelled out: adam_multi_kernel
// (misc.cu) and adam_peer_kernel (peer.cu) must give the same bits for the same inputs whatever loop the compiler finds
// the expression in (left to the contraction heuristics the two differed by one ulp in a tail element).
__device__ __forceinline__ float adam_update(float pi, float gi, float& mi, float& vi, float beta1, float beta2, float eps,
                                             float grad_scale, float step_size, float inv_sqrt_bc2) {
    gi = __fmul_rn(gi, grad_scale);
    mi = __fmaf_rn(beta1, mi, __fmul_rn(1.f - beta1, gi));
    vi = __fmaf_rn(beta2, vi, __fmul_rn(__fmul_rn(1.f - beta2, gi), gi));
    const float denom = __fadd_rn(__fmul_rn(sqrtf(vi), inv_sqrt_bc2), eps);
    return __fsub_rn(pi, __fdiv_rn(__fmul_rn(step_size, mi), denom));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide sum for blockDim.x <= 1024 (multiple of 32).  `red` is >= 32 floats of shared memory.
__device__ __forceinline__ float block_sum(float v, float* red) {
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    float r = (lane < nw) ? red[lane] : 0.f;
    r = warp_sum(r);
    return r;
}
__device__ __forceinline__ float block_max(float v, float* red) {
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    float r = (lane < nw) ? red[lane] : -INFINITY;
    r = warp_max(r);
    return r;
}

__device__ __forceinline__ float4 ld_f4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st_f4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

__host__ __device__ __forceinline__ bool aligned16(const void* p) { return (((uintptr_t)p) & 15u) == 0; }

}  // namespace bmhrl
