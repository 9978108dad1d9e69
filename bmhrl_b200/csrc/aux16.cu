// bf16 math mode, the HBM-bound helpers around the tensor-core kernels (all 128-bit / 64-bit vectorised, all
// deterministic: no floating-point atomics anywhere):
//   layernorm_fwd16   LayerNorm of the fp32 residual stream with a bf16 output (the A operand of the projection that
//                     follows it) -- ResidualConnection model/blocks.py:135-144, normCA/normCV bm_hrl_agent.py:107-108
//   cast16 (+colsum)  fp32 -> bf16 with optional Philox dropout and the bias gradient db[n] += sum_m out[m,n]: the head of
//                     every linear layer's backward (dZ = dropout-mask o dY), fixed-order column sums
//   colsum16          bias gradients of dZ tensors that a GEMM / attention epilogue already emitted in bf16
//   to_bf16 strided   weight shadows at start-up / after load_state_dict (Adam keeps them fresh afterwards, misc.cu)
#include <cuda_bf16.h>
#include "common.cuh"
#include "../../include/bmhrl_b200.h"

namespace bmhrl {

__device__ __forceinline__ uint32_t pk2(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}

constexpr int LN16_WARPS = 4;

// y16 = bf16((x - mean) * rstd * gamma + beta); one warp per row, row cached in registers when D <= 1024
__global__ void __launch_bounds__(LN16_WARPS * 32) layernorm_fwd16_kernel(
    const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
    __nv_bfloat16* __restrict__ y, long long ldy, float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows,
    int D, float eps, int vec) {
    griddep_wait();
    griddep_launch_dependents();
    const int row = blockIdx.x * LN16_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float* xr = x + (long long)row * D;
    __nv_bfloat16* yr = y + (long long)row * ldy;
    if (vec) {  // D % 4 == 0, D <= 1024: up to 8 float4 per lane stay in registers
        const int D4 = D >> 2;
        float4 v[8];
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int i = lane + 32 * j;
            v[j] = i < D4 ? ld_f4(xr + 4 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
            s += (v[j].x + v[j].y) + (v[j].z + v[j].w);
        }
        const float mean = warp_sum(s) / (float)D;
        float q = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int i = lane + 32 * j;
            if (i < D4) {
                const float a = v[j].x - mean, b = v[j].y - mean, c = v[j].z - mean, d = v[j].w - mean;
                q += (a * a + b * b) + (c * c + d * d);
            }
        }
        const float rstd = rsqrtf(warp_sum(q) / (float)D + eps);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int i = lane + 32 * j;
            if (i < D4) {
                const float4 g = ld_f4(gamma + 4 * i), b = ld_f4(beta + 4 * i);
                uint2 o;
                o.x = pk2((v[j].x - mean) * rstd * g.x + b.x, (v[j].y - mean) * rstd * g.y + b.y);
                o.y = pk2((v[j].z - mean) * rstd * g.z + b.z, (v[j].w - mean) * rstd * g.w + b.w);
                *reinterpret_cast<uint2*>(yr + 4 * i) = o;
            }
        }
        if (lane == 0) { mean_out[row] = mean; rstd_out[row] = rstd; }
        return;
    }
    float s = 0.f;
    for (int i = lane; i < D; i += 32) s += xr[i];
    const float mean = warp_sum(s) / (float)D;
    float q = 0.f;
    for (int i = lane; i < D; i += 32) q += (xr[i] - mean) * (xr[i] - mean);
    const float rstd = rsqrtf(warp_sum(q) / (float)D + eps);
    for (int i = lane; i < D; i += 32) yr[i] = __float2bfloat16((xr[i] - mean) * rstd * gamma[i] + beta[i]);
    if (lane == 0) { mean_out[row] = mean; rstd_out[row] = rstd; }
}

// ---- fp32 -> bf16 (+ dropout), element-wise, for casts without a bias gradient
__global__ void cast16_kernel(const float* __restrict__ x, long long ldx, __nv_bfloat16* __restrict__ y, long long ldy,
                              int rows, int cols, float drop_p, const unsigned long long* __restrict__ rng,
                              unsigned int salt) {
    griddep_wait();
    griddep_launch_dependents();
    const int c4 = cols >> 2;
    const long long n4 = (long long)rows * c4;
    const float ks = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(i / c4), n = (int)(i - (long long)m * c4) * 4;
        float4 v = ld_f4(x + (long long)m * ldx + n);
        if (drop_p > 0.f) {
            const Philox4 rb = philox_block(rng, salt, ((unsigned long long)m * cols + n) >> 2);
            v.x = u32_to_unit(rb.x) >= drop_p ? v.x * ks : 0.f;
            v.y = u32_to_unit(rb.y) >= drop_p ? v.y * ks : 0.f;
            v.z = u32_to_unit(rb.z) >= drop_p ? v.z * ks : 0.f;
            v.w = u32_to_unit(rb.w) >= drop_p ? v.w * ks : 0.f;
        }
        *reinterpret_cast<uint2*>(y + (long long)m * ldy + n) = make_uint2(pk2(v.x, v.y), pk2(v.z, v.w));
    }
}

// ---- fp32 -> bf16 (+ dropout) AND the column sums (bias gradient).  A CTA owns a 64-column strip and ALL rows: its 16
// row lanes stride the rows, partial sums meet in shared memory and are added in a fixed order -> bit-reproducible.
__global__ void __launch_bounds__(256) cast16_colsum_kernel(const float* __restrict__ x, long long ldx,
                                                            __nv_bfloat16* __restrict__ y, long long ldy, int rows,
                                                            int cols, float drop_p,
                                                            const unsigned long long* __restrict__ rng, unsigned int salt,
                                                            float* __restrict__ db, int accumulate) {
    __shared__ float red[16][64 + 1];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int n = blockIdx.x * 64 + tx * 4;
    const float ks = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    if (n < cols) {
        for (int m = ty; m < rows; m += 16) {
            float4 v = ld_f4(x + (long long)m * ldx + n);
            if (drop_p > 0.f) {
                const Philox4 rb = philox_block(rng, salt, ((unsigned long long)m * cols + n) >> 2);
                v.x = u32_to_unit(rb.x) >= drop_p ? v.x * ks : 0.f;
                v.y = u32_to_unit(rb.y) >= drop_p ? v.y * ks : 0.f;
                v.z = u32_to_unit(rb.z) >= drop_p ? v.z * ks : 0.f;
                v.w = u32_to_unit(rb.w) >= drop_p ? v.w * ks : 0.f;
            }
            if (y) *reinterpret_cast<uint2*>(y + (long long)m * ldy + n) = make_uint2(pk2(v.x, v.y), pk2(v.z, v.w));
            a0 += v.x; a1 += v.y; a2 += v.z; a3 += v.w;
        }
    }
    red[ty][tx * 4 + 0] = a0; red[ty][tx * 4 + 1] = a1; red[ty][tx * 4 + 2] = a2; red[ty][tx * 4 + 3] = a3;
    __syncthreads();
    if (threadIdx.x < 64) {
        const int c = blockIdx.x * 64 + threadIdx.x;
        if (c < cols) {
            float t = 0.f;
#pragma unroll
            for (int i = 0; i < 16; ++i) t += red[i][threadIdx.x];
            db[c] = accumulate ? db[c] + t : t;
        }
    }
}

// ---- column sums of a bf16 matrix; column n goes to out[n / seg_cols][n % seg_cols] (Q|K|V bias gradients)
__global__ void __launch_bounds__(256) colsum16_kernel(const __nv_bfloat16* __restrict__ z, long long ldz, int rows, int cols,
                                                       float* __restrict__ o0, float* __restrict__ o1,
                                                       float* __restrict__ o2, int seg_cols, int accumulate) {
    griddep_wait();
    griddep_launch_dependents();
    __shared__ float red[16][64 + 1];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int n = blockIdx.x * 64 + tx * 4;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    if (n < cols) {
        int m = ty;
        for (; m + 48 < rows; m += 64) {  // four independent loads in flight per thread
            uint2 w[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) w[u] = *reinterpret_cast<const uint2*>(z + (long long)(m + 16 * u) * ldz + n);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                a0 += __uint_as_float(w[u].x << 16); a1 += __uint_as_float(w[u].x & 0xffff0000u);
                a2 += __uint_as_float(w[u].y << 16); a3 += __uint_as_float(w[u].y & 0xffff0000u);
            }
        }
        for (; m < rows; m += 16) {
            const uint2 w = *reinterpret_cast<const uint2*>(z + (long long)m * ldz + n);
            a0 += __uint_as_float(w.x << 16); a1 += __uint_as_float(w.x & 0xffff0000u);
            a2 += __uint_as_float(w.y << 16); a3 += __uint_as_float(w.y & 0xffff0000u);
        }
    }
    red[ty][tx * 4 + 0] = a0; red[ty][tx * 4 + 1] = a1; red[ty][tx * 4 + 2] = a2; red[ty][tx * 4 + 3] = a3;
    __syncthreads();
    if (threadIdx.x < 64) {
        const int c = blockIdx.x * 64 + threadIdx.x;
        if (c < cols) {
            float t = 0.f;
#pragma unroll
            for (int i = 0; i < 16; ++i) t += red[i][threadIdx.x];
            const int sg = seg_cols > 0 ? c / seg_cols : 0;
            float* o = (sg == 0 ? o0 : (sg == 1 ? o1 : o2)) + (c - sg * (seg_cols > 0 ? seg_cols : 0));
            *o = accumulate ? *o + t : t;
        }
    }
}

// ---- generic strided fp32 -> bf16 copy (any cols): weight shadows
__global__ void to_bf16_kernel(const float* __restrict__ x, long long ldx, __nv_bfloat16* __restrict__ y, long long ldy,
                               long long rows, int cols) {
    const long long n = rows * cols;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const long long m = i / cols;
        const int c = (int)(i - m * cols);
        y[m * ldy + c] = __float2bfloat16(x[m * ldx + c]);
    }
}

static inline int grid1d(long long n, int threads, int cap) {
    long long b = (n + threads - 1) / threads;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

}  // namespace bmhrl
using namespace bmhrl;

extern "C" int bmhrl_layernorm_fwd_bf16(const float* x, const float* gamma, const float* beta, void* y, long long ldy,
                                        float* mean, float* rstd, int rows, int D, float eps, void* stream) {
    BMHRL_CHECK_ARG(x && gamma && beta && y && mean && rstd, "bmhrl_layernorm_fwd_bf16: null pointer");
    BMHRL_CHECK_ARG(rows >= 0 && D > 0 && ldy >= D, "bmhrl_layernorm_fwd_bf16: bad sizes");
    if (rows == 0) return BMHRL_OK;
    const int vec = (D % 4 == 0) && D <= 1024 && aligned16(x) && aligned16(gamma) && aligned16(beta) &&
                    (((uintptr_t)y) & 7u) == 0 && (ldy & 3) == 0;
    BMHRL_LAUNCH_PDL("layernorm_fwd16", layernorm_fwd16_kernel, dim3((rows + LN16_WARPS - 1) / LN16_WARPS),
                     dim3(LN16_WARPS * 32), 0, (cudaStream_t)stream, x, gamma, beta, (__nv_bfloat16*)y, ldy, mean, rstd, rows, D,
                     eps, vec);
    return BMHRL_OK;
}

extern "C" int bmhrl_cast_bf16(const float* x, long long ldx, void* y, long long ldy, int rows, int cols, float drop_p,
                               const unsigned long long* rng_state, unsigned int salt, float* db, int accumulate_db,
                               void* stream) {
    BMHRL_CHECK_ARG(x && (y || db) && rows >= 0 && cols > 0 && ldx >= cols, "bmhrl_cast_bf16: bad arguments");
    BMHRL_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f && (drop_p == 0.f || rng_state), "bmhrl_cast_bf16: dropout");
    if (rows == 0) return BMHRL_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const bool vec = (cols % 4 == 0) && (ldx % 4 == 0) && aligned16(x) && (!y || ((((uintptr_t)y) & 7u) == 0 && (ldy & 3) == 0));
    if (!vec) {
        BMHRL_CHECK_ARG(drop_p == 0.f && !db && y, "bmhrl_cast_bf16: dropout / bias gradient need cols % 4 == 0 and aligned rows");
        to_bf16_kernel<<<grid1d((long long)rows * cols, 256, 148 * 8), 256, 0, st>>>(x, ldx, (__nv_bfloat16*)y, ldy, rows, cols);
        BMHRL_POST_LAUNCH("to_bf16");
        return BMHRL_OK;
    }
    if (db) {
        cast16_colsum_kernel<<<(cols + 63) / 64, 256, 0, st>>>(x, ldx, (__nv_bfloat16*)y, ldy, rows, cols, drop_p, rng_state,
                                                              salt, db, accumulate_db);
        BMHRL_POST_LAUNCH("cast16_colsum");
    } else {
        BMHRL_LAUNCH_PDL("cast16", cast16_kernel, dim3(grid1d((long long)rows * (cols / 4), 256, 148 * 4)), dim3(256), 0, st, x,
                         ldx, (__nv_bfloat16*)y, ldy, rows, cols, drop_p, rng_state, salt);
    }
    return BMHRL_OK;
}

extern "C" int bmhrl_colsum_bf16(const void* z, long long ldz, int rows, int cols, float* out0, float* out1, float* out2,
                                 int seg_cols, int accumulate, void* stream) {
    BMHRL_CHECK_ARG(z && out0 && rows >= 0 && cols > 0 && ldz >= cols, "bmhrl_colsum_bf16: bad arguments");
    BMHRL_CHECK_ARG((cols % 4 == 0) && (ldz % 4 == 0) && ((((uintptr_t)z) & 7u) == 0), "bmhrl_colsum_bf16: cols / pitch % 4");
    BMHRL_CHECK_ARG(seg_cols == 0 || (seg_cols > 0 && cols <= 3 * seg_cols && (cols <= seg_cols || out1) &&
                                      (cols <= 2 * seg_cols || out2)), "bmhrl_colsum_bf16: segments");
    BMHRL_LAUNCH_PDL("colsum16", colsum16_kernel, dim3((cols + 63) / 64), dim3(256), 0, (cudaStream_t)stream,
                     (const __nv_bfloat16*)z, ldz, rows, cols, out0, out1, out2, seg_cols, accumulate);
    return BMHRL_OK;
}
