// K2: fused residual + LayerNorm (fwd/bwd) and the sigmoid-gated audio/video mix of the fusion decoder.
// Reference: ResidualConnection model/blocks.py:128-144 (nn.LayerNorm eps 1e-5, biased variance),
// BMFusionLayer tail model/bm_hrl_agent.py:103-114.  HBM-bound: one warp per row, 128-bit accesses,
// two-pass mean/variance like ATen (re-reads hit L1).
#include <cuda_bf16.h>
#include "common.cuh"
#include "../../include/bmhrl_b200.h"

namespace bmhrl {

constexpr int LN_WARPS = 4;

// z = x (+ resid);  y = (z - mean) * rstd * gamma + beta.  Optionally stores z (when resid given).
__global__ void __launch_bounds__(LN_WARPS * 32) layernorm_fwd_kernel(
    const float* __restrict__ x, const float* __restrict__ resid, const float* __restrict__ gamma,
    const float* __restrict__ beta, float* __restrict__ z_out, float* __restrict__ y, float* __restrict__ mean_out,
    float* __restrict__ rstd_out, int rows, int D, float eps, int vec) {
    const int row = blockIdx.x * LN_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float* xr = x + (long long)row * D;
    const float* rr = resid ? resid + (long long)row * D : nullptr;
    float* zr = z_out ? z_out + (long long)row * D : nullptr;
    float* yr = y + (long long)row * D;
    const int D4 = vec ? (D >> 2) : 0;

    float s = 0.f;
    for (int i = lane; i < D4; i += 32) {
        float4 v = ld_f4(xr + 4 * i);
        if (rr) { float4 r = ld_f4(rr + 4 * i); v.x += r.x; v.y += r.y; v.z += r.z; v.w += r.w; }
        if (zr) st_f4(zr + 4 * i, v);
        s += (v.x + v.y) + (v.z + v.w);
    }
    for (int i = 4 * D4 + lane; i < D; i += 32) {
        float v = xr[i] + (rr ? rr[i] : 0.f);
        if (zr) zr[i] = v;
        s += v;
    }
    const float mean = warp_sum(s) / (float)D;
    // second pass re-reads z (or x + resid) -- served from L1/L2
    const float* src = zr ? zr : xr;
    const bool add_r = (rr != nullptr) && (zr == nullptr);
    __syncwarp();
    float q = 0.f;
    for (int i = lane; i < D4; i += 32) {
        float4 v = ld_f4(src + 4 * i);
        if (add_r) { float4 r = ld_f4(rr + 4 * i); v.x += r.x; v.y += r.y; v.z += r.z; v.w += r.w; }
        float a = v.x - mean, b = v.y - mean, c = v.z - mean, d = v.w - mean;
        q += (a * a + b * b) + (c * c + d * d);
    }
    for (int i = 4 * D4 + lane; i < D; i += 32) {
        float v = src[i] + (add_r ? rr[i] : 0.f);
        q += (v - mean) * (v - mean);
    }
    const float rstd = rsqrtf(warp_sum(q) / (float)D + eps);
    for (int i = lane; i < D4; i += 32) {
        float4 v = ld_f4(src + 4 * i);
        if (add_r) { float4 r = ld_f4(rr + 4 * i); v.x += r.x; v.y += r.y; v.z += r.z; v.w += r.w; }
        float4 g = ld_f4(gamma + 4 * i), b = ld_f4(beta + 4 * i);
        float4 o;
        o.x = (v.x - mean) * rstd * g.x + b.x;
        o.y = (v.y - mean) * rstd * g.y + b.y;
        o.z = (v.z - mean) * rstd * g.z + b.z;
        o.w = (v.w - mean) * rstd * g.w + b.w;
        st_f4(yr + 4 * i, o);
    }
    for (int i = 4 * D4 + lane; i < D; i += 32) {
        float v = src[i] + (add_r ? rr[i] : 0.f);
        yr[i] = (v - mean) * rstd * gamma[i] + beta[i];
    }
    if (lane == 0) {
        mean_out[row] = mean;
        rstd_out[row] = rstd;
    }
}

// dx = rstd * (g - mean(g) - xhat * mean(g * xhat)) (+ dres),  g = dy * gamma
__global__ void __launch_bounds__(LN_WARPS * 32) layernorm_bwd_dx_kernel(
    const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ gamma,
    const float* __restrict__ mean_in, const float* __restrict__ rstd_in, const float* __restrict__ dres,
    float* __restrict__ dx, int rows, int D, int vec) {
    const int row = blockIdx.x * LN_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float* dyr = dy + (long long)row * D;
    const float* xr = x + (long long)row * D;
    const float* dr = dres ? dres + (long long)row * D : nullptr;
    float* dxr = dx + (long long)row * D;
    const float mean = mean_in[row], rstd = rstd_in[row];
    const int D4 = vec ? (D >> 2) : 0;
    float s1 = 0.f, s2 = 0.f;
    for (int i = lane; i < D4; i += 32) {
        float4 g = ld_f4(dyr + 4 * i), w = ld_f4(gamma + 4 * i), v = ld_f4(xr + 4 * i);
        float g0 = g.x * w.x, g1 = g.y * w.y, g2 = g.z * w.z, g3 = g.w * w.w;
        s1 += (g0 + g1) + (g2 + g3);
        s2 += (g0 * (v.x - mean) + g1 * (v.y - mean)) + (g2 * (v.z - mean) + g3 * (v.w - mean));
    }
    for (int i = 4 * D4 + lane; i < D; i += 32) {
        float g = dyr[i] * gamma[i];
        s1 += g;
        s2 += g * (xr[i] - mean);
    }
    s1 = warp_sum(s1) / (float)D;
    s2 = warp_sum(s2) * rstd / (float)D;  // mean(g * xhat)
    for (int i = lane; i < D4; i += 32) {
        float4 g = ld_f4(dyr + 4 * i), w = ld_f4(gamma + 4 * i), v = ld_f4(xr + 4 * i);
        float4 o;
        o.x = rstd * (g.x * w.x - s1 - (v.x - mean) * rstd * s2);
        o.y = rstd * (g.y * w.y - s1 - (v.y - mean) * rstd * s2);
        o.z = rstd * (g.z * w.z - s1 - (v.z - mean) * rstd * s2);
        o.w = rstd * (g.w * w.w - s1 - (v.w - mean) * rstd * s2);
        if (dr) { float4 r = ld_f4(dr + 4 * i); o.x += r.x; o.y += r.y; o.z += r.z; o.w += r.w; }
        st_f4(dxr + 4 * i, o);
    }
    for (int i = 4 * D4 + lane; i < D; i += 32) {
        float o = rstd * (dyr[i] * gamma[i] - s1 - (xr[i] - mean) * rstd * s2);
        if (dr) o += dr[i];
        dxr[i] = o;
    }
}

// dgamma[c] = sum_r dy[r,c] * xhat[r,c];  dbeta[c] = sum_r dy[r,c].  One block per 32 columns, 8 warps stride
// the rows (128-byte coalesced row segments), deterministic shared-memory tree at the end.
__global__ void __launch_bounds__(256) layernorm_bwd_affine_kernel(
    const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ mean_in,
    const float* __restrict__ rstd_in, float* __restrict__ dgamma, float* __restrict__ dbeta, int rows, int D,
    int accumulate) {
    __shared__ float sg[8][33], sb[8][33];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + lane;
    float ag = 0.f, ab = 0.f;
    if (c < D) {
        for (int r = w; r < rows; r += 8) {
            float d = dy[(long long)r * D + c];
            float xh = (x[(long long)r * D + c] - mean_in[r]) * rstd_in[r];
            ag += d * xh;
            ab += d;
        }
    }
    sg[w][lane] = ag;
    sb[w][lane] = ab;
    __syncthreads();
    if (w == 0 && c < D) {
        float tg = 0.f, tb = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) { tg += sg[i][lane]; tb += sb[i][lane]; }
        dgamma[c] = accumulate ? dgamma[c] + tg : tg;
        dbeta[c] = accumulate ? dbeta[c] + tb : tb;
    }
}

// Fused backward: dx (as layernorm_bwd_dx_kernel) AND the affine gradients in one pass over dy / x.  A block owns
// `rows_per_block` consecutive rows, its warps stride them; every lane keeps the dgamma / dbeta partial sums of its
// NV float4 column groups in registers, the 4 warps are combined in shared memory and each block issues one RED per
// column into dgamma / dbeta (the parameter's slice of the flat gradient buffer, or a zeroed temporary).
template <int NV>
__global__ void __launch_bounds__(LN_WARPS * 32) layernorm_bwd_fused_kernel(
    const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ gamma,
    const float* __restrict__ mean_in, const float* __restrict__ rstd_in, const float* __restrict__ dres,
    float* __restrict__ dx, float* __restrict__ dgamma, float* __restrict__ dbeta, int rows, int D,
    int rows_per_block) {
    extern __shared__ float sred[];  // [LN_WARPS][2][D]
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int D4 = D >> 2;
    const int r_begin = blockIdx.x * rows_per_block;
    const int r_end = min(rows, r_begin + rows_per_block);
    float4 ag[NV], ab[NV], gw[NV];
#pragma unroll
    for (int j = 0; j < NV; ++j) {
        ag[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        ab[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        const int i = lane + 32 * j;
        gw[j] = i < D4 ? ld_f4(gamma + 4 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (int row = r_begin + w; row < r_end; row += LN_WARPS) {
        const float* dyr = dy + (long long)row * D;
        const float* xr = x + (long long)row * D;
        const float mean = mean_in[row], rstd = rstd_in[row];
        float4 g[NV], v[NV];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            const int i = lane + 32 * j;
            if (i < D4) {
                g[j] = ld_f4(dyr + 4 * i);
                v[j] = ld_f4(xr + 4 * i);
                v[j].x = (v[j].x - mean) * rstd; v[j].y = (v[j].y - mean) * rstd;
                v[j].z = (v[j].z - mean) * rstd; v[j].w = (v[j].w - mean) * rstd;
                ab[j].x += g[j].x; ab[j].y += g[j].y; ab[j].z += g[j].z; ab[j].w += g[j].w;
                ag[j].x += g[j].x * v[j].x; ag[j].y += g[j].y * v[j].y;
                ag[j].z += g[j].z * v[j].z; ag[j].w += g[j].w * v[j].w;
                g[j].x *= gw[j].x; g[j].y *= gw[j].y; g[j].z *= gw[j].z; g[j].w *= gw[j].w;
                s1 += (g[j].x + g[j].y) + (g[j].z + g[j].w);
                s2 += (g[j].x * v[j].x + g[j].y * v[j].y) + (g[j].z * v[j].z + g[j].w * v[j].w);
            }
        }
        if (dx) {
            s1 = warp_sum(s1) / (float)D;
            s2 = warp_sum(s2) / (float)D;  // mean(g * xhat)
            float* dxr = dx + (long long)row * D;
            const float* dr = dres ? dres + (long long)row * D : nullptr;
#pragma unroll
            for (int j = 0; j < NV; ++j) {
                const int i = lane + 32 * j;
                if (i < D4) {
                    float4 o;
                    o.x = rstd * (g[j].x - s1 - v[j].x * s2);
                    o.y = rstd * (g[j].y - s1 - v[j].y * s2);
                    o.z = rstd * (g[j].z - s1 - v[j].z * s2);
                    o.w = rstd * (g[j].w - s1 - v[j].w * s2);
                    if (dr) { float4 r = ld_f4(dr + 4 * i); o.x += r.x; o.y += r.y; o.z += r.z; o.w += r.w; }
                    st_f4(dxr + 4 * i, o);
                }
            }
        }
    }
    if (!dgamma) return;
    float* mg = sred + (size_t)w * 2 * D;
    float* mb = mg + D;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
        const int i = lane + 32 * j;
        if (i < D4) {
            st_f4(mg + 4 * i, ag[j]);
            st_f4(mb + 4 * i, ab[j]);
        }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < D; c += LN_WARPS * 32) {
        float tg = 0.f, tb = 0.f;
#pragma unroll
        for (int i = 0; i < LN_WARPS; ++i) {
            tg += sred[(size_t)i * 2 * D + c];
            tb += sred[(size_t)i * 2 * D + D + c];
        }
        atomicAdd(dgamma + c, tg);
        atomicAdd(dbeta + c, tb);
    }
}

// Deterministic backward in ONE launch: blocks [0, row_blocks) compute dx (one warp per row, two passes over dy / x),
// blocks [row_blocks, row_blocks + ceil(D / 32)) each own a 32-column strip and ALL rows and produce dgamma / dbeta with a
// fixed summation order (8 warps stride the rows, shared-memory combine).  No atomics: gradients are bit-reproducible
// run to run.  The strips re-read dy / x from L2 (2 x rows x D x 4 B, a few MB).
__global__ void __launch_bounds__(256) layernorm_bwd_det_kernel(
    const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ gamma,
    const float* __restrict__ mean_in, const float* __restrict__ rstd_in, const float* __restrict__ dres,
    float* __restrict__ dx, float* __restrict__ dgamma, float* __restrict__ dbeta, int rows, int D, int vec,
    int row_blocks, int accumulate, __nv_bfloat16* __restrict__ dx16, long long ld16, float drop_p,
    const unsigned long long* __restrict__ rng, unsigned int salt) {
    griddep_wait();
    griddep_launch_dependents();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const float ks = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
    if ((int)blockIdx.x < row_blocks) {
        const int row = blockIdx.x * 8 + w;
        if (row >= rows) return;
        const float* dyr = dy + (long long)row * D;
        const float* xr = x + (long long)row * D;
        const float* dr = dres ? dres + (long long)row * D : nullptr;
        float* dxr = dx + (long long)row * D;
        const float mean = mean_in[row], rstd = rstd_in[row];
        const int D4 = vec ? (D >> 2) : 0;
        float s1 = 0.f, s2 = 0.f;
        for (int i = lane; i < D4; i += 32) {
            float4 g = ld_f4(dyr + 4 * i), ww = ld_f4(gamma + 4 * i), v = ld_f4(xr + 4 * i);
            float g0 = g.x * ww.x, g1 = g.y * ww.y, g2 = g.z * ww.z, g3 = g.w * ww.w;
            s1 += (g0 + g1) + (g2 + g3);
            s2 += (g0 * (v.x - mean) + g1 * (v.y - mean)) + (g2 * (v.z - mean) + g3 * (v.w - mean));
        }
        for (int i = 4 * D4 + lane; i < D; i += 32) {
            float g = dyr[i] * gamma[i];
            s1 += g;
            s2 += g * (xr[i] - mean);
        }
        s1 = warp_sum(s1) / (float)D;
        s2 = warp_sum(s2) * rstd / (float)D;  // mean(g * xhat)
        for (int i = lane; i < D4; i += 32) {
            float4 g = ld_f4(dyr + 4 * i), ww = ld_f4(gamma + 4 * i), v = ld_f4(xr + 4 * i);
            float4 o;
            o.x = rstd * (g.x * ww.x - s1 - (v.x - mean) * rstd * s2);
            o.y = rstd * (g.y * ww.y - s1 - (v.y - mean) * rstd * s2);
            o.z = rstd * (g.z * ww.z - s1 - (v.z - mean) * rstd * s2);
            o.w = rstd * (g.w * ww.w - s1 - (v.w - mean) * rstd * s2);
            if (dr) { float4 r = ld_f4(dr + 4 * i); o.x += r.x; o.y += r.y; o.z += r.z; o.w += r.w; }
            st_f4(dxr + 4 * i, o);
            if (dx16) {   // bf16(dropout-mask o dx): what the backward of the block that PRODUCED x starts from (cast16_kernel)
                if (drop_p > 0.f) {
                    const Philox4 rb = philox_block(rng, salt, ((unsigned long long)row * D + 4 * i) >> 2);
                    o.x = u32_to_unit(rb.x) >= drop_p ? o.x * ks : 0.f;
                    o.y = u32_to_unit(rb.y) >= drop_p ? o.y * ks : 0.f;
                    o.z = u32_to_unit(rb.z) >= drop_p ? o.z * ks : 0.f;
                    o.w = u32_to_unit(rb.w) >= drop_p ? o.w * ks : 0.f;
                }
                uint2 pk;
                asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(pk.x) : "f"(o.y), "f"(o.x));
                asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(pk.y) : "f"(o.w), "f"(o.z));
                *reinterpret_cast<uint2*>(dx16 + (long long)row * ld16 + 4 * i) = pk;
            }
        }
        for (int i = 4 * D4 + lane; i < D; i += 32) {
            float o = rstd * (dyr[i] * gamma[i] - s1 - (xr[i] - mean) * rstd * s2);
            if (dr) o += dr[i];
            dxr[i] = o;
        }
        return;
    }
    __shared__ float sg[8][33], sb[8][33];
    const int c = ((int)blockIdx.x - row_blocks) * 32 + lane;
    float ag = 0.f, ab = 0.f;
    if (c < D) {
        int r = w;
        for (; r + 24 < rows; r += 32) {  // four independent row loads in flight per lane (fixed order of the adds)
            float d[4], xv[4], mu[4], rs[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                d[u] = dy[(long long)(r + 8 * u) * D + c];
                xv[u] = x[(long long)(r + 8 * u) * D + c];
                mu[u] = mean_in[r + 8 * u];
                rs[u] = rstd_in[r + 8 * u];
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                ag += d[u] * ((xv[u] - mu[u]) * rs[u]);
                ab += d[u];
            }
        }
        for (; r < rows; r += 8) {
            const float d = dy[(long long)r * D + c];
            const float xh = (x[(long long)r * D + c] - mean_in[r]) * rstd_in[r];
            ag += d * xh;
            ab += d;
        }
    }
    sg[w][lane] = ag;
    sb[w][lane] = ab;
    __syncthreads();
    if (w == 0 && c < D) {
        float tg = 0.f, tb = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) { tg += sg[i][lane]; tb += sb[i][lane]; }
        dgamma[c] = accumulate ? dgamma[c] + tg : tg;
        dbeta[c] = accumulate ? dbeta[c] + tb : tb;
    }
}

// ---------------------------------------------------------------- gated mix
__device__ __forceinline__ float gate_of(float c) {
    float cc = fminf(fmaxf(c, -2.f), 2.f);
    return 1.f / (1.f + expf(-cc));
}

// out = s * v + (1 - s) * a,  s = sigmoid(clamp(c, -2, 2))
__global__ void gated_mix_fwd_kernel(const float* __restrict__ v, const float* __restrict__ a,
                                     const float* __restrict__ cptr, float* __restrict__ out, long long n, int vec) {
    const float s = gate_of(cptr[0]);
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    if (vec) {
        const long long n4 = n >> 2;
        for (long long j = i; j < n4; j += stride) {
            float4 x = ld_f4(v + 4 * j), y = ld_f4(a + 4 * j), o;
            o.x = s * x.x + (1.f - s) * y.x; o.y = s * x.y + (1.f - s) * y.y;
            o.z = s * x.z + (1.f - s) * y.z; o.w = s * x.w + (1.f - s) * y.w;
            st_f4(out + 4 * j, o);
        }
        for (long long j = 4 * n4 + i; j < n; j += stride) out[j] = s * v[j] + (1.f - s) * a[j];
    } else {
        for (long long j = i; j < n; j += stride) out[j] = s * v[j] + (1.f - s) * a[j];
    }
}

// dv = s*g, da = (1-s)*g, partial[b] = sum_block g*(v-a)
__global__ void __launch_bounds__(256) gated_mix_bwd_kernel(const float* __restrict__ g, const float* __restrict__ v,
                                                            const float* __restrict__ a,
                                                            const float* __restrict__ cptr, float* __restrict__ dv,
                                                            float* __restrict__ da, float* __restrict__ partial,
                                                            long long n) {
    __shared__ float red[32];
    const float s = gate_of(cptr[0]);
    float acc = 0.f;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride) {
        float gg = g[j];
        dv[j] = s * gg;
        da[j] = (1.f - s) * gg;
        acc += gg * (v[j] - a[j]);
    }
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

// dc = (sum partial) * s(1-s) * [ -2 <= c <= 2 ]
__global__ void gated_mix_bwd_final_kernel(const float* __restrict__ partial, int nparts,
                                           const float* __restrict__ cptr, float* __restrict__ dc, int accumulate) {
    __shared__ float red[32];
    float acc = 0.f;
    for (int i = threadIdx.x; i < nparts; i += blockDim.x) acc += partial[i];
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) {
        float c = cptr[0];
        float s = gate_of(c);
        const float d = (c >= -2.f && c <= 2.f) ? acc * s * (1.f - s) : 0.f;
        dc[0] = accumulate ? dc[0] + d : d;
    }
}

}  // namespace bmhrl
using namespace bmhrl;

extern "C" int bmhrl_layernorm_fwd_f32(const float* x, const float* resid, const float* gamma, const float* beta,
                                       float* z_out, float* y, float* mean, float* rstd, int rows, int D, float eps,
                                       void* stream) {
    BMHRL_CHECK_ARG(x && gamma && beta && y && mean && rstd, "bmhrl_layernorm_fwd_f32: null pointer");
    BMHRL_CHECK_ARG(rows >= 0 && D > 0, "bmhrl_layernorm_fwd_f32: bad sizes");
    if (rows == 0) return BMHRL_OK;
    int vec = (D % 4 == 0) && aligned16(x) && aligned16(y) && aligned16(gamma) && aligned16(beta) &&
              (!resid || aligned16(resid)) && (!z_out || aligned16(z_out));
    int blocks = (rows + LN_WARPS - 1) / LN_WARPS;
    layernorm_fwd_kernel<<<blocks, LN_WARPS * 32, 0, (cudaStream_t)stream>>>(x, resid, gamma, beta, z_out, y, mean,
                                                                            rstd, rows, D, eps, vec);
    BMHRL_POST_LAUNCH("layernorm_fwd");
    return BMHRL_OK;
}

extern "C" int bmhrl_layernorm_bwd_f32(const float* dy, const float* x, const float* gamma, const float* mean,
                                       const float* rstd, const float* dres, float* dx, float* dgamma, float* dbeta,
                                       int rows, int D, int accumulate_affine, void* stream) {
    BMHRL_CHECK_ARG(dy && x && gamma && mean && rstd, "bmhrl_layernorm_bwd_f32: null pointer");
    BMHRL_CHECK_ARG(rows >= 0 && D > 0, "bmhrl_layernorm_bwd_f32: bad sizes");
    cudaStream_t st = (cudaStream_t)stream;
    const bool want_affine = dgamma && dbeta;
    if (rows == 0 || (!dx && !want_affine)) return BMHRL_OK;
    const int vec = (D % 4 == 0) && aligned16(x) && aligned16(dy) && aligned16(gamma) && (!dx || aligned16(dx)) &&
                    (!dres || aligned16(dres));
    const int row_blocks = dx ? (rows + 7) / 8 : 0;
    const int strips = want_affine ? (D + 31) / 32 : 0;
    BMHRL_LAUNCH_PDL("layernorm_bwd_det", layernorm_bwd_det_kernel, dim3(row_blocks + strips), dim3(256), 0, st, dy, x, gamma,
                     mean, rstd, dres, dx, dgamma, dbeta, rows, D, vec, row_blocks, accumulate_affine,
                     (__nv_bfloat16*)nullptr, 0ll, 0.f, (const unsigned long long*)nullptr, 0u);
    return BMHRL_OK;
}

extern "C" int bmhrl_layernorm_bwd_twin_f32(const float* dy, const float* x, const float* gamma, const float* mean,
                                            const float* rstd, const float* dres, float* dx, void* dx16, long long lddx16,
                                            float drop_p, const unsigned long long* rng_state, unsigned int salt, int rows,
                                            int D, void* stream) {
    BMHRL_CHECK_ARG(dy && x && gamma && mean && rstd && dx && dx16, "bmhrl_layernorm_bwd_twin_f32: null pointer");
    BMHRL_CHECK_ARG(rows >= 0 && D > 0 && D % 4 == 0 && lddx16 >= D && lddx16 % 4 == 0, "bmhrl_layernorm_bwd_twin_f32: D and the bf16 pitch must be multiples of 4");
    BMHRL_CHECK_ARG(aligned16(x) && aligned16(dy) && aligned16(gamma) && aligned16(dx) && (!dres || aligned16(dres)) &&
                        ((uintptr_t)dx16 & 7u) == 0, "bmhrl_layernorm_bwd_twin_f32: alignment");
    BMHRL_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f && (drop_p == 0.f || rng_state), "bmhrl_layernorm_bwd_twin_f32: dropout");
    if (rows == 0) return BMHRL_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int row_blocks = (rows + 7) / 8;
    BMHRL_LAUNCH_PDL("layernorm_bwd_det", layernorm_bwd_det_kernel, dim3(row_blocks), dim3(256), 0, st, dy, x, gamma, mean, rstd,
                     dres, dx, (float*)nullptr, (float*)nullptr, rows, D, 1, row_blocks, 0, (__nv_bfloat16*)dx16, lddx16,
                     drop_p, rng_state, salt);
    return BMHRL_OK;
}

extern "C" int bmhrl_gated_mix_fwd_f32(const float* v, const float* a, const float* gate_param, float* out,
                                       long long n, void* stream) {
    BMHRL_CHECK_ARG(v && a && gate_param && out, "bmhrl_gated_mix_fwd_f32: null pointer");
    if (n <= 0) return BMHRL_OK;
    int vec = aligned16(v) && aligned16(a) && aligned16(out);
    long long work = vec ? (n + 3) / 4 : n;
    int blocks = (int)((work + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    gated_mix_fwd_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(v, a, gate_param, out, n, vec);
    BMHRL_POST_LAUNCH("gated_mix_fwd");
    return BMHRL_OK;
}

extern "C" int bmhrl_gated_mix_bwd_f32(const float* g, const float* v, const float* a, const float* gate_param,
                                       float* dv, float* da, float* dgate, float* partial_ws, int ws_floats,
                                       long long n, int accumulate_gate, void* stream) {
    BMHRL_CHECK_ARG(g && v && a && gate_param && dv && da && dgate && partial_ws,
                    "bmhrl_gated_mix_bwd_f32: null pointer");
    BMHRL_CHECK_ARG(ws_floats >= 1, "bmhrl_gated_mix_bwd_f32: workspace too small");
    int blocks = (int)((n + 255) / 256);
    if (blocks > ws_floats) blocks = ws_floats;
    if (blocks > 148 * 4) blocks = 148 * 4;
    if (blocks < 1) blocks = 1;
    cudaStream_t st = (cudaStream_t)stream;
    gated_mix_bwd_kernel<<<blocks, 256, 0, st>>>(g, v, a, gate_param, dv, da, partial_ws, n);
    BMHRL_POST_LAUNCH("gated_mix_bwd");
    gated_mix_bwd_final_kernel<<<1, 256, 0, st>>>(partial_ws, blocks, gate_param, dgate, accumulate_gate);
    BMHRL_POST_LAUNCH("gated_mix_bwd_final");
    return BMHRL_OK;
}
