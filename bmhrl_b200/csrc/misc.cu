// K6 (embedding * sqrt(d) + positional encoding), K7 (goal expansion / segment scans), bias+activation+dropout
// gradient prep for the linear layers, fused multi-tensor Adam and the device-side step/RNG tick.
// Reference: VocabularyEmbedder model/blocks.py:44-48, PositionalEncoder model/blocks.py:105-112,
// Manager.expand_goals model/bm_hrl_agent.py:415-429, biased_kl() manager loop
// epoch_loops/captioning_bmrl_loops.py:299-317, torch.optim.Adam scripts/train_rl_captioning_module.py:81-83.
#include <cuda_bf16.h>
#include "common.cuh"
#include "../../include/bmhrl_b200.h"

namespace bmhrl {

// ------------------------------------------------------------------ K6
// emb[r, :] = W[tok[r], :] * scale ; pos[r, :] = dropout(emb[r, :] + pe[r % L, :])
__global__ void embed_posenc_kernel(const long long* __restrict__ tok, const float* __restrict__ W,
                                    const float* __restrict__ pe, float* __restrict__ emb, float* __restrict__ pos,
                                    long long rows, int L, int D, float scale, float drop_p,
                                    const unsigned long long* __restrict__ rng, unsigned int salt) {
    const long long n = rows * D;
    const float ks = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        long long r = i / D;
        int d = (int)(i % D);
        float e = W[tok[r] * D + d] * scale;
        if (emb) emb[i] = e;
        if (pos) {
            float v = e + pe[(long long)(r % L) * D + d];
            if (drop_p > 0.f) v = dropout_keep(rng, salt, (unsigned long long)i, drop_p) ? v * ks : 0.f;
            pos[i] = v;
        }
    }
}

// out = dropout(x (+ x2) + pe[s])     x: (B, S, D) contiguous
__global__ void add_posenc_kernel(const float* __restrict__ x, const float* __restrict__ x2,
                                  const float* __restrict__ pe, float* __restrict__ out, long long rows, int S, int D,
                                  float drop_p, const unsigned long long* __restrict__ rng, unsigned int salt) {
    const long long n4 = rows * D / 4;  // D % 4 == 0 is checked on the host
    const float ks = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
    const int D4 = D / 4;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        long long r = i / D4;
        int c = (int)(i % D4);
        float4 v = ld_f4(x + 4 * i);
        if (x2) { float4 w = ld_f4(x2 + 4 * i); v.x += w.x; v.y += w.y; v.z += w.z; v.w += w.w; }
        float4 p = ld_f4(pe + ((long long)(r % S) * D4 + c) * 4);
        v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w;
        if (drop_p > 0.f) {
            Philox4 rr = philox_block(rng, salt, (unsigned long long)i);
            v.x = u32_to_unit(rr.x) >= drop_p ? v.x * ks : 0.f;
            v.y = u32_to_unit(rr.y) >= drop_p ? v.y * ks : 0.f;
            v.z = u32_to_unit(rr.z) >= drop_p ? v.z * ks : 0.f;
            v.w = u32_to_unit(rr.w) >= drop_p ? v.w * ks : 0.f;
        }
        st_f4(out + 4 * i, v);
    }
}

// scalar variant for feature sizes that are not a multiple of 4 (same RNG indexing: element i -> counter i)
__global__ void add_posenc_scalar_kernel(const float* __restrict__ x, const float* __restrict__ x2,
                                         const float* __restrict__ pe, float* __restrict__ out, long long rows, int S,
                                         int D, float drop_p, const unsigned long long* __restrict__ rng,
                                         unsigned int salt) {
    const long long n = rows * D;
    const float ks = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        long long r = i / D;
        int d = (int)(i % D);
        float v = x[i] + (x2 ? x2[i] : 0.f) + pe[(long long)(r % S) * D + d];
        if (drop_p > 0.f) v = dropout_keep(rng, salt, (unsigned long long)i, drop_p) ? v * ks : 0.f;
        out[i] = v;
    }
}

// out[r, :Da] = a[r, :], out[r, Da:] = b[r, :]   (torch.cat([x, goal_completion], -1), model/bm_hrl_agent.py:483)
__global__ void concat2_kernel(const float* __restrict__ a, int Da, const float* __restrict__ b, int Db,
                               float* __restrict__ out, long long rows) {
    const int D = Da + Db;
    const long long n = rows * D;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        long long r = i / D;
        int d = (int)(i % D);
        out[i] = d < Da ? a[r * Da + d] : b[r * Db + (d - Da)];
    }
}
__global__ void split2_kernel(const float* __restrict__ g, float* __restrict__ da, int Da, float* __restrict__ db,
                              int Db, long long rows) {
    const int D = Da + Db;
    const long long n = rows * D;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        long long r = i / D;
        int d = (int)(i % D);
        if (d < Da) { if (da) da[r * Da + d] = g[i]; }
        else if (db) db[r * Db + (d - Da)] = g[i];
    }
}

// dx = keep ? dy / (1-p) : 0 with the flat element index as RNG counter
__global__ void dropout_apply_kernel(const float* __restrict__ dy, float* __restrict__ dx, long long n, float drop_p,
                                     const unsigned long long* __restrict__ rng, unsigned int salt) {
    const float ks = 1.f / (1.f - drop_p);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        dx[i] = dropout_keep(rng, salt, (unsigned long long)i, drop_p) ? dy[i] * ks : 0.f;
}

// out = x + dropout(r): generic ResidualConnection tail (model/blocks.py:142-144) for caller-supplied sublayers
__global__ void dropout_add_kernel(const float* __restrict__ x, const float* __restrict__ r, float* __restrict__ out,
                                   long long n, float drop_p, const unsigned long long* __restrict__ rng,
                                   unsigned int salt) {
    const float ks = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float v = r[i];
        if (drop_p > 0.f) v = dropout_keep(rng, salt, (unsigned long long)i, drop_p) ? v * ks : 0.f;
        out[i] = x[i] + v;
    }
}

// dW[tok[r], :] += g[r, :] * scale   (only when the word embedding is unfrozen)
__global__ void embed_bwd_kernel(const long long* __restrict__ tok, const float* __restrict__ g, float* __restrict__ dW,
                                 long long rows, int D, float scale) {
    const long long n = rows * D;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        atomicAdd(&dW[tok[i / D] * D + (i % D)], g[i] * scale);
}

// ------------------------------------------------------------------ K7 goal expansion
// One CTA, one thread per batch row.  map[b,l] = source position whose goal position (b,l) takes, -1 = zero.
__global__ void goal_map_kernel(const int* __restrict__ seg, int* __restrict__ map, int B, int L) {
    extern __shared__ int row_has[];
    for (int b = threadIdx.x; b < B; b += blockDim.x) {
        int any = 0;
        for (int l = 0; l < L; ++l) any |= (seg[b * L + l] != 0);
        row_has[b] = any;
    }
    __syncthreads();
    for (int b = threadIdx.x; b < B; b += blockDim.x) {
        int later = 0;
        for (int bb = b + 1; bb < B; ++bb) later |= row_has[bb];
        int nxt = -1;
        for (int l = L - 1; l >= 0; --l) {
            if (seg[b * L + l] != 0) nxt = l;
            int m;
            if (nxt >= 0) m = nxt;                                   // goal of the next boundary
            else if (row_has[b]) m = later ? -1 : l;                 // tail after the row's last boundary
            else m = (b == 0 && later) ? -1 : l;                     // boundary-free row (row-0 quirk)
            map[b * L + l] = m;
        }
    }
}

__global__ void goal_gather_kernel(const float* __restrict__ x, const float* __restrict__ noise,
                                   const int* __restrict__ map, float* __restrict__ out, long long rows, int L, int D) {
    const long long n = rows * D;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        long long r = i / D;
        int d = (int)(i % D);
        int m = map[r];
        out[i] = m < 0 ? 0.f : x[((r / L) * L + m) * D + d] + (noise ? noise[d] : 0.f);
    }
}

// Manager exploration noise (model/bm_hrl_agent.py:444-452): mean / std over ALL (non-NaN) elements of the goal
// tensor, noise[j] = N(mean/10, std/5) - 0.5 * mean/10, one d_goal vector shared by every position.  eps: optional
// explicit N(0,1) draws (tests); otherwise Box-Muller over Philox.  Single CTA (the tensor is B*L*64 floats).
__global__ void __launch_bounds__(1024) exploration_noise_kernel(const float* __restrict__ x, long long n, int d_goal,
                                                                 const float* __restrict__ eps, float* __restrict__ noise,
                                                                 float* __restrict__ stats,
                                                                 const unsigned long long* __restrict__ rng,
                                                                 unsigned int salt) {
    __shared__ float red[32];
    float s = 0.f, c = 0.f;
    for (long long i = threadIdx.x; i < n; i += blockDim.x) {
        float v = x[i];
        if (v == v) { s += v; c += 1.f; }
    }
    s = block_sum(s, red);
    c = block_sum(c, red);
    const float mean = s / c;
    float q = 0.f;
    for (long long i = threadIdx.x; i < n; i += blockDim.x) {
        float v = x[i];
        if (v == v) { float d = fabsf(v - mean); q += d * d; }
    }
    q = block_sum(q, red);
    const float stdv = sqrtf(q / c);
    const float mu = mean / 10.f, sd = stdv / 5.f;
    if (threadIdx.x == 0 && stats) { stats[0] = mean; stats[1] = stdv; }
    for (int j = threadIdx.x; j < d_goal; j += blockDim.x) {
        float z;
        if (eps) {
            z = eps[j];
        } else {
            Philox4 r = philox_block(rng, salt, (unsigned long long)j);
            float u1 = ((float)(r.x >> 8) + 1.f) * (1.0f / 16777216.0f);  // (0,1]
            float u2 = u32_to_unit(r.y);
            z = sqrtf(-2.f * logf(u1)) * cosf(6.283185307179586f * u2);
        }
        noise[j] = mu + sd * z - 0.5f * mu;
    }
}

// dx[b, l', :] = sum_{l <= l' : map[b,l] == l'} dy[b, l, :]
__global__ void goal_scatter_kernel(const float* __restrict__ dy, const int* __restrict__ map, float* __restrict__ dx,
                                    long long rows, int L, int D) {
    const long long n = rows * D;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        long long r = i / D;
        int d = (int)(i % D);
        long long b = r / L;
        int lp = (int)(r % L);
        float acc = 0.f;
        for (int l = 0; l <= lp; ++l)
            if (map[b * L + l] == lp) acc += dy[(b * L + l) * D + d];
        dx[i] = acc;
    }
}

// ------------------------------------------------------------------ manager segment scan
// seg_prob[b, s..e] = prod p[b, s..e] for each closed segment; 0 elsewhere.
// exp_out: per-segment sums of `expected`; tails follow the same quirks as expand_goals (see goal_map_kernel).
__global__ void segment_scan_fwd_kernel(const float* __restrict__ p, const float* __restrict__ expected,
                                        const int* __restrict__ seg, float* __restrict__ seg_prob,
                                        float* __restrict__ exp_out, int B, int L) {
    extern __shared__ int row_has[];
    for (int b = threadIdx.x; b < B; b += blockDim.x) {
        int any = 0;
        for (int l = 0; l < L; ++l) any |= (seg[b * L + l] != 0);
        row_has[b] = any;
    }
    __syncthreads();
    for (int b = threadIdx.x; b < B; b += blockDim.x) {
        int later = 0;
        for (int bb = b + 1; bb < B; ++bb) later |= row_has[bb];
        int start = 0;
        for (int l = 0; l < L; ++l) {
            seg_prob[b * L + l] = 0.f;
            exp_out[b * L + l] = expected[b * L + l];
        }
        if (row_has[b]) {
            for (int l = 0; l < L; ++l) {
                if (seg[b * L + l] != 0) {
                    float pr = 1.f, su = 0.f;
                    for (int k = start; k <= l; ++k) { pr *= p[b * L + k]; su += expected[b * L + k]; }
                    for (int k = start; k <= l; ++k) { seg_prob[b * L + k] = pr; exp_out[b * L + k] = su; }
                    start = l + 1;
                }
            }
            if (later)
                for (int k = start; k < L; ++k) exp_out[b * L + k] = 0.f;
        } else if (b == 0 && later) {
            for (int k = 0; k < L; ++k) exp_out[k] = 0.f;
        }
    }
}

// dp[b,k] = (sum_{l in seg(k)} g[b,l]) * prod_{j in seg(k), j != k} p[b,j]
__global__ void segment_scan_bwd_kernel(const float* __restrict__ p, const int* __restrict__ seg,
                                        const float* __restrict__ g, float* __restrict__ dp, int B, int L) {
    for (int b = blockIdx.x * blockDim.x + threadIdx.x; b < B; b += gridDim.x * blockDim.x) {
        int start = 0;
        for (int l = 0; l < L; ++l) dp[b * L + l] = 0.f;
        for (int l = 0; l < L; ++l) {
            if (seg[b * L + l] != 0) {
                float G = 0.f;
                for (int k = start; k <= l; ++k) G += g[b * L + k];
                for (int k = start; k <= l; ++k) {
                    float pr = 1.f;
                    for (int j = start; j <= l; ++j)
                        if (j != k) pr *= p[b * L + j];
                    dp[b * L + k] = G * pr;
                }
                start = l + 1;
            }
        }
    }
}

// ------------------------------------------------------------------ linear backward prep
// dZ[m,n] = dY[m,n] * (act==relu ? Y[m,n] > 0 : 1) * (dropout keep ? 1/(1-p) : 0);  db[n] += sum_m dZ[m,n].
// For relu + dropout the forward output Y = drop(relu(z)) is > 0 exactly where both let the gradient through.
// Grid (column blocks, row splits): a block owns CPT*32 columns x rows_per_block rows, its 8 warps stride the rows
// with CPT*4-byte accesses per lane (128-bit when CPT == 4); partial column sums are combined in shared memory and
// added to db with one RED per column per block (db is zeroed by the host wrapper when it does not accumulate).
template <int CPT>
__global__ void __launch_bounds__(256) linear_bwd_prep_kernel(const float* __restrict__ dY, long long lddy,
                                                              const float* __restrict__ Y, long long ldy,
                                                              float* __restrict__ dZ, float* __restrict__ db, int M,
                                                              int N, int act, float drop_p,
                                                              const unsigned long long* __restrict__ rng,
                                                              unsigned int salt, int rows_per_block) {
    __shared__ float sb[8][32 * CPT + 1];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int n0 = (blockIdx.x * 32 + lane) * CPT;
    const int m_begin = blockIdx.y * rows_per_block;
    const int m_end = min(M, m_begin + rows_per_block);
    const float ks = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
    float acc[CPT];
#pragma unroll
    for (int j = 0; j < CPT; ++j) acc[j] = 0.f;
    if (n0 < N) {
        for (int m = m_begin + w; m < m_end; m += 8) {
            float g[CPT], y[CPT];
            if (CPT == 4) {
                float4 t = ld_f4(dY + (long long)m * lddy + n0);
                g[0] = t.x; g[1] = t.y; g[2] = t.z; g[3] = t.w;
                if (act == 1) {
                    float4 u = ld_f4(Y + (long long)m * ldy + n0);
                    y[0] = u.x; y[1] = u.y; y[2] = u.z; y[3] = u.w;
                }
            } else {
                g[0] = dY[(long long)m * lddy + n0];
                if (act == 1) y[0] = Y[(long long)m * ldy + n0];
            }
            if (act != 1 && drop_p > 0.f && CPT == 4) {  // N % 4 == 0 here: the 4 columns share one Philox block
                const Philox4 rb = philox_block(rng, salt, ((unsigned long long)m * N + n0) >> 2);
                g[0] = u32_to_unit(rb.x) >= drop_p ? g[0] * ks : 0.f;
                g[1 % CPT] = u32_to_unit(rb.y) >= drop_p ? g[1 % CPT] * ks : 0.f;
                g[2 % CPT] = u32_to_unit(rb.z) >= drop_p ? g[2 % CPT] * ks : 0.f;
                g[3 % CPT] = u32_to_unit(rb.w) >= drop_p ? g[3 % CPT] * ks : 0.f;
            }
#pragma unroll
            for (int j = 0; j < CPT; ++j) {
                if (act == 1) {
                    g[j] = (y[j] > 0.f) ? g[j] * ks : 0.f;
                } else if (drop_p > 0.f && CPT != 4) {
                    g[j] = dropout_keep(rng, salt, (unsigned long long)m * N + n0 + j, drop_p) ? g[j] * ks : 0.f;
                }
                acc[j] += g[j];
            }
            if (dZ) {
                if (CPT == 4) st_f4(dZ + (long long)m * N + n0, make_float4(g[0], g[1], g[2], g[3]));
                else dZ[(long long)m * N + n0] = g[0];
            }
        }
    }
    if (!db) return;
#pragma unroll
    for (int j = 0; j < CPT; ++j) sb[w][lane * CPT + j] = acc[j];
    __syncthreads();
    for (int c = threadIdx.x; c < 32 * CPT; c += 256) {
        const int n = blockIdx.x * 32 * CPT + c;
        if (n < N) {
            float t = 0.f;
#pragma unroll
            for (int i = 0; i < 8; ++i) t += sb[i][c];
            atomicAdd(db + n, t);
        }
    }
}

// ------------------------------------------------------------------ fused multi-tensor Adam
struct AdamChunk {
    int tensor;
    int start;  // element offset inside the tensor, chunks are ADAM_CHUNK long
};
constexpr int ADAM_CHUNK = 8192;

// Optional weight shadows (bf16 math mode): shadow[t] != NULL makes the kernel also write the UPDATED parameter into a
// second buffer -- bf16 (kind 0; the tensor-core operand copy of a weight matrix, (rows, shadow_cols) with row pitch
// shadow_ld) or fp32 (kind 1; the N-concatenated bias of a fused Q|K|V projection).  Adam touches every updated weight
// anyway, so the shadows cost their own bytes and no extra pass.
__device__ __forceinline__ void adam_chunk(int chunk, float* const* __restrict__ params,
                                           const float* const* __restrict__ grads, float* const* __restrict__ exp_avg,
                                           float* const* __restrict__ exp_avg_sq, const long long* __restrict__ numels,
                                           const int* __restrict__ chunk_tensor, const int* __restrict__ chunk_start,
                                           const float* __restrict__ step_ptr, float lr, float beta1, float beta2, float eps,
                                           float grad_scale, void* const* __restrict__ shadow,
                                           const int* __restrict__ shadow_cols, const int* __restrict__ shadow_ld,
                                           const int* __restrict__ shadow_kind) {
    const int t = chunk_tensor[chunk];
    const long long s0 = chunk_start[chunk];
    const long long n = numels[t];
    float* __restrict__ p = params[t];
    const float* __restrict__ g = grads[t];
    float* __restrict__ m = exp_avg[t];
    float* __restrict__ v = exp_avg_sq[t];
    const float step = step_ptr[t];
    const float bc1 = 1.f - powf(beta1, step);
    const float bc2 = 1.f - powf(beta2, step);
    const float step_size = lr / bc1;
    const float inv_sqrt_bc2 = rsqrtf(bc2);
    long long end = s0 + ADAM_CHUNK;
    if (end > n) end = n;
    void* sh = shadow ? shadow[t] : nullptr;
    const int sk = sh ? shadow_kind[t] : 0;
    const int scols = sh ? shadow_cols[t] : 1, sld = sh ? shadow_ld[t] : 1;
    auto upd = [&](float pi, float gi, float& mi, float& vi) {
        return adam_update(pi, gi, mi, vi, beta1, beta2, eps, grad_scale, step_size, inv_sqrt_bc2);
    };
    const bool vec = ((((uintptr_t)p) | ((uintptr_t)g) | ((uintptr_t)m) | ((uintptr_t)v)) & 15u) == 0 &&
                     (!sh || (sk == 1 ? ((uintptr_t)sh & 15u) == 0 : (((uintptr_t)sh & 7u) == 0 && scols == sld)));
    if (vec) {
        const long long e4 = s0 + ((end - s0) & ~3ll);
        for (long long i = s0 + 4 * threadIdx.x; i < e4; i += 4 * 256) {
            const float4 g4 = ld_f4(g + i), p4 = ld_f4(p + i);
            float4 m4 = ld_f4(m + i), v4 = ld_f4(v + i), o;
            o.x = upd(p4.x, g4.x, m4.x, v4.x);
            o.y = upd(p4.y, g4.y, m4.y, v4.y);
            o.z = upd(p4.z, g4.z, m4.z, v4.z);
            o.w = upd(p4.w, g4.w, m4.w, v4.w);
            st_f4(m + i, m4);
            st_f4(v + i, v4);
            st_f4(p + i, o);
            if (sh) {
                if (sk == 1) {
                    st_f4(reinterpret_cast<float*>(sh) + i, o);
                } else {
                    uint2 w;
                    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w.x) : "f"(o.y), "f"(o.x));
                    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w.y) : "f"(o.w), "f"(o.z));
                    *reinterpret_cast<uint2*>(reinterpret_cast<unsigned short*>(sh) + i) = w;
                }
            }
        }
        if (threadIdx.x < (int)(end - e4)) {  // tail of the last chunk of a tensor whose numel is not a multiple of 4
            const long long i = e4 + threadIdx.x;
            float mi = m[i], vi = v[i];
            const float o = upd(p[i], g[i], mi, vi);
            m[i] = mi; v[i] = vi; p[i] = o;
            if (sh) {
                if (sk == 1) reinterpret_cast<float*>(sh)[i] = o;
                else reinterpret_cast<__nv_bfloat16*>(sh)[i] = __float2bfloat16(o);
            }
        }
        return;
    }
    for (long long i = s0 + threadIdx.x; i < end; i += 256) {
        float mi = m[i], vi = v[i];
        const float o = upd(p[i], g[i], mi, vi);
        m[i] = mi;
        v[i] = vi;
        p[i] = o;
        if (sh) {
            if (sk == 1) reinterpret_cast<float*>(sh)[i] = o;
            else reinterpret_cast<__nv_bfloat16*>(sh)[(i / scols) * sld + (i % scols)] = __float2bfloat16(o);
        }
    }
}

// One CTA per chunk -- or, with a grid limit (bmhrl_adam_set_grid_limit), a grid-stride loop over the chunks.  An update
// that runs BESIDE the rest of backward must not take every thread slot of every SM: the unlimited grid (thousands of
// CTAs, 8 x 256 threads per SM) kept every later launch of the step waiting until it had drained -- 140 us without a
// single kernel start behind the first stage's update (profiles/r02_step_timeline.txt).
__global__ void __launch_bounds__(256) adam_multi_kernel(float* const* __restrict__ params,
                                                         const float* const* __restrict__ grads,
                                                         float* const* __restrict__ exp_avg,
                                                         float* const* __restrict__ exp_avg_sq,
                                                         const long long* __restrict__ numels,
                                                         const int* __restrict__ chunk_tensor,
                                                         const int* __restrict__ chunk_start, int n_chunks,
                                                         const float* __restrict__ step_ptr, float lr, float beta1,
                                                         float beta2, float eps, float grad_scale,
                                                         void* const* __restrict__ shadow,
                                                         const int* __restrict__ shadow_cols,
                                                         const int* __restrict__ shadow_ld,
                                                         const int* __restrict__ shadow_kind) {
    for (int c = blockIdx.x; c < n_chunks; c += gridDim.x)
        adam_chunk(c, params, grads, exp_avg, exp_avg_sq, numels, chunk_tensor, chunk_start, step_ptr, lr, beta1, beta2, eps,
                   grad_scale, shadow, shadow_cols, shadow_ld, shadow_kind);
}

// steps[active[i]] += 1 for the tensors updated this step (torch keeps one step count per parameter and skips
// parameters without a gradient);  rng_state[1] += 1 : dropout / sampling stream
__global__ void tick_kernel(float* steps, const int* active, int n_active, unsigned long long* rng_state) {
    if (steps)
        for (int i = threadIdx.x; i < n_active; i += blockDim.x) steps[active ? active[i] : i] += 1.f;
    if (rng_state && threadIdx.x == 0) rng_state[1] += 1ull;
}

static inline int grid_for(long long n, int threads) {
    long long b = (n + threads - 1) / threads;
    if (b > 148 * 8) b = 148 * 8;
    if (b < 1) b = 1;
    return (int)b;
}

}  // namespace bmhrl
using namespace bmhrl;

extern "C" int bmhrl_embed_posenc_f32(const long long* tokens, const float* W, const float* pe, float* emb, float* pos,
                                      long long rows, int L, int D, float scale, float drop_p,
                                      const unsigned long long* rng_state, unsigned int salt, void* stream) {
    BMHRL_CHECK_ARG(tokens && W && (emb || pos) && (!pos || pe), "bmhrl_embed_posenc_f32: null pointer");
    BMHRL_CHECK_ARG(rows >= 0 && L > 0 && D > 0, "bmhrl_embed_posenc_f32: bad sizes");
    BMHRL_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f && (drop_p == 0.f || rng_state), "bmhrl_embed_posenc_f32: dropout");
    if (rows == 0) return BMHRL_OK;
    embed_posenc_kernel<<<grid_for(rows * D, 256), 256, 0, (cudaStream_t)stream>>>(tokens, W, pe, emb, pos, rows, L, D,
                                                                                   scale, drop_p, rng_state, salt);
    BMHRL_POST_LAUNCH("embed_posenc");
    return BMHRL_OK;
}

extern "C" int bmhrl_add_posenc_f32(const float* x, const float* x2, const float* pe, float* out, long long rows,
                                    int S, int D, float drop_p, const unsigned long long* rng_state,
                                    unsigned int salt, void* stream) {
    BMHRL_CHECK_ARG(x && pe && out, "bmhrl_add_posenc_f32: null pointer");
    BMHRL_CHECK_ARG(rows >= 0 && S > 0 && D > 0, "bmhrl_add_posenc_f32: bad sizes");
    BMHRL_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f && (drop_p == 0.f || rng_state), "bmhrl_add_posenc_f32: dropout");
    if (rows == 0) return BMHRL_OK;
    if (D % 4 != 0 || !(aligned16(x) && aligned16(pe) && aligned16(out) && (!x2 || aligned16(x2)))) {
        add_posenc_scalar_kernel<<<grid_for(rows * D, 256), 256, 0, (cudaStream_t)stream>>>(x, x2, pe, out, rows, S, D,
                                                                                            drop_p, rng_state, salt);
        BMHRL_POST_LAUNCH("add_posenc_scalar");
        return BMHRL_OK;
    }
    add_posenc_kernel<<<grid_for(rows * D / 4, 256), 256, 0, (cudaStream_t)stream>>>(x, x2, pe, out, rows, S, D, drop_p,
                                                                                     rng_state, salt);
    BMHRL_POST_LAUNCH("add_posenc");
    return BMHRL_OK;
}

// Index convention: add_posenc draws one Philox block per float4, i.e. element i uses counter i -- identical to
// dropout_keep(idx = i); dropout_apply therefore reproduces its mask.
extern "C" int bmhrl_dropout_apply_f32(const float* dy, float* dx, long long n, float drop_p,
                                       const unsigned long long* rng_state, unsigned int salt, void* stream) {
    BMHRL_CHECK_ARG(dy && dx && rng_state && drop_p > 0.f && drop_p < 1.f, "bmhrl_dropout_apply_f32: bad arguments");
    if (n <= 0) return BMHRL_OK;
    dropout_apply_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(dy, dx, n, drop_p, rng_state, salt);
    BMHRL_POST_LAUNCH("dropout_apply");
    return BMHRL_OK;
}

extern "C" int bmhrl_dropout_add_f32(const float* x, const float* r, float* out, long long n, float drop_p,
                                     const unsigned long long* rng_state, unsigned int salt, void* stream) {
    BMHRL_CHECK_ARG(x && r && out, "bmhrl_dropout_add_f32: null pointer");
    BMHRL_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f && (drop_p == 0.f || rng_state), "bmhrl_dropout_add_f32: dropout");
    if (n <= 0) return BMHRL_OK;
    dropout_add_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(x, r, out, n, drop_p, rng_state, salt);
    BMHRL_POST_LAUNCH("dropout_add");
    return BMHRL_OK;
}

extern "C" int bmhrl_concat2_f32(const float* a, int Da, const float* b, int Db, float* out, long long rows,
                                 void* stream) {
    BMHRL_CHECK_ARG(a && b && out && Da > 0 && Db > 0 && rows >= 0, "bmhrl_concat2_f32: bad arguments");
    if (rows == 0) return BMHRL_OK;
    concat2_kernel<<<grid_for(rows * (Da + Db), 256), 256, 0, (cudaStream_t)stream>>>(a, Da, b, Db, out, rows);
    BMHRL_POST_LAUNCH("concat2");
    return BMHRL_OK;
}

extern "C" int bmhrl_split2_f32(const float* g, float* da, int Da, float* db, int Db, long long rows, void* stream) {
    BMHRL_CHECK_ARG(g && (da || db) && Da > 0 && Db > 0 && rows >= 0, "bmhrl_split2_f32: bad arguments");
    if (rows == 0) return BMHRL_OK;
    split2_kernel<<<grid_for(rows * (Da + Db), 256), 256, 0, (cudaStream_t)stream>>>(g, da, Da, db, Db, rows);
    BMHRL_POST_LAUNCH("split2");
    return BMHRL_OK;
}

extern "C" int bmhrl_embed_bwd_f32(const long long* tokens, const float* g, float* dW, long long rows, int D,
                                   float scale, void* stream) {
    BMHRL_CHECK_ARG(tokens && g && dW && rows >= 0 && D > 0, "bmhrl_embed_bwd_f32: bad arguments");
    if (rows == 0) return BMHRL_OK;
    embed_bwd_kernel<<<grid_for(rows * D, 256), 256, 0, (cudaStream_t)stream>>>(tokens, g, dW, rows, D, scale);
    BMHRL_POST_LAUNCH("embed_bwd");
    return BMHRL_OK;
}

extern "C" int bmhrl_exploration_noise_f32(const float* x, long long n, int d_goal, const float* eps_normal,
                                           float* noise, float* stats, const unsigned long long* rng_state,
                                           unsigned int salt, void* stream) {
    BMHRL_CHECK_ARG(x && noise && n > 0 && d_goal > 0, "bmhrl_exploration_noise_f32: bad arguments");
    BMHRL_CHECK_ARG(eps_normal || rng_state, "bmhrl_exploration_noise_f32: needs explicit draws or rng_state");
    exploration_noise_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(x, n, d_goal, eps_normal, noise, stats, rng_state,
                                                                   salt);
    BMHRL_POST_LAUNCH("exploration_noise");
    return BMHRL_OK;
}

extern "C" int bmhrl_goal_expand_fwd_f32(const float* x, const float* noise, const int* seg, int* map, float* out,
                                         int B, int L, int D, void* stream) {
    BMHRL_CHECK_ARG(x && seg && map && out && B > 0 && L > 0 && D > 0, "bmhrl_goal_expand_fwd_f32: bad arguments");
    BMHRL_CHECK_ARG(B <= 8192, "bmhrl_goal_expand_fwd_f32: B too large for the single-CTA scan");
    cudaStream_t st = (cudaStream_t)stream;
    goal_map_kernel<<<1, 256, B * sizeof(int), st>>>(seg, map, B, L);
    BMHRL_POST_LAUNCH("goal_map");
    goal_gather_kernel<<<grid_for((long long)B * L * D, 256), 256, 0, st>>>(x, noise, map, out, (long long)B * L, L,
                                                                           D);
    BMHRL_POST_LAUNCH("goal_gather");
    return BMHRL_OK;
}

extern "C" int bmhrl_goal_expand_bwd_f32(const float* dy, const int* map, float* dx, int B, int L, int D,
                                         void* stream) {
    BMHRL_CHECK_ARG(dy && map && dx && B > 0 && L > 0 && D > 0, "bmhrl_goal_expand_bwd_f32: bad arguments");
    goal_scatter_kernel<<<grid_for((long long)B * L * D, 256), 256, 0, (cudaStream_t)stream>>>(dy, map, dx,
                                                                                               (long long)B * L, L, D);
    BMHRL_POST_LAUNCH("goal_scatter");
    return BMHRL_OK;
}

extern "C" int bmhrl_segment_scan_fwd_f32(const float* probs, const float* expected, const int* seg, float* seg_prob,
                                          float* expected_out, int B, int L, void* stream) {
    BMHRL_CHECK_ARG(probs && expected && seg && seg_prob && expected_out && B > 0 && L > 0,
                    "bmhrl_segment_scan_fwd_f32: bad arguments");
    BMHRL_CHECK_ARG(B <= 8192, "bmhrl_segment_scan_fwd_f32: B too large for the single-CTA scan");
    segment_scan_fwd_kernel<<<1, 256, B * sizeof(int), (cudaStream_t)stream>>>(probs, expected, seg, seg_prob,
                                                                               expected_out, B, L);
    BMHRL_POST_LAUNCH("segment_scan_fwd");
    return BMHRL_OK;
}

extern "C" int bmhrl_segment_scan_bwd_f32(const float* probs, const int* seg, const float* g, float* dprobs, int B,
                                          int L, void* stream) {
    BMHRL_CHECK_ARG(probs && seg && g && dprobs && B > 0 && L > 0, "bmhrl_segment_scan_bwd_f32: bad arguments");
    segment_scan_bwd_kernel<<<(B + 127) / 128, 128, 0, (cudaStream_t)stream>>>(probs, seg, g, dprobs, B, L);
    BMHRL_POST_LAUNCH("segment_scan_bwd");
    return BMHRL_OK;
}

extern "C" int bmhrl_linear_bwd_prep_f32(const float* dY, long long lddy, const float* Y, long long ldy, float* dZ,
                                         float* db, int M, int N, int act, float drop_p,
                                         const unsigned long long* rng_state, unsigned int salt, int accumulate_db,
                                         void* stream) {
    BMHRL_CHECK_ARG(dY && (dZ || db) && M >= 0 && N > 0, "bmhrl_linear_bwd_prep_f32: bad arguments");
    BMHRL_CHECK_ARG(act == 0 || (act == 1 && Y), "bmhrl_linear_bwd_prep_f32: relu needs the forward output");
    BMHRL_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f && (drop_p == 0.f || act == 1 || rng_state),
                    "bmhrl_linear_bwd_prep_f32: dropout needs rng_state");
    if (M == 0) return BMHRL_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (db && !accumulate_db && cudaMemsetAsync(db, 0, sizeof(float) * (size_t)N, st) != cudaSuccess) {
        bmhrl_set_error("bmhrl_linear_bwd_prep_f32: memset failed");
        return BMHRL_ERR_CUDA;
    }
    const bool vec = (N % 4 == 0) && (lddy % 4 == 0) && aligned16(dY) && (!dZ || aligned16(dZ)) &&
                     (act != 1 || ((ldy % 4 == 0) && aligned16(Y)));
    const int cols_per_block = vec ? 128 : 32;
    const int col_blocks = (N + cols_per_block - 1) / cols_per_block;
    // One row split whenever a bias gradient is produced: the column sums are then added in a fixed order inside one
    // CTA and the single RED per column lands on a value no other CTA touches -> bit-reproducible bias gradients.
    int splits = db ? 1 : (2 * 148 + col_blocks - 1) / col_blocks;  // without db: ~2 blocks per SM
    const int max_splits = (M + 7) / 8;
    if (splits > max_splits) splits = max_splits;
    const int rows_per_block = ((M + splits - 1) / splits + 7) / 8 * 8;
    splits = (M + rows_per_block - 1) / rows_per_block;
    dim3 grid(col_blocks, splits);
    if (vec)
        linear_bwd_prep_kernel<4><<<grid, 256, 0, st>>>(dY, lddy, Y, ldy, dZ, db, M, N, act, drop_p, rng_state, salt,
                                                        rows_per_block);
    else
        linear_bwd_prep_kernel<1><<<grid, 256, 0, st>>>(dY, lddy, Y, ldy, dZ, db, M, N, act, drop_p, rng_state, salt,
                                                        rows_per_block);
    BMHRL_POST_LAUNCH("linear_bwd_prep");
    return BMHRL_OK;
}

extern "C" int bmhrl_adam_chunk_elems(void) { return ADAM_CHUNK; }

static int g_adam_grid_limit = 0;
extern "C" int bmhrl_adam_set_grid_limit(int max_ctas) {
    g_adam_grid_limit = max_ctas > 0 ? max_ctas : 0;
    return BMHRL_OK;
}
extern "C" int bmhrl_adam_get_grid_limit(void) { return g_adam_grid_limit; }
static inline int adam_grid(int n_chunks) {
    return (g_adam_grid_limit > 0 && n_chunks > g_adam_grid_limit) ? g_adam_grid_limit : n_chunks;
}

extern "C" int bmhrl_adam_multi_f32(float* const* params, const float* const* grads, float* const* exp_avg,
                                    float* const* exp_avg_sq, const long long* numels, const int* chunk_tensor,
                                    const int* chunk_start, int n_chunks, const float* step_counts, float lr,
                                    float beta1, float beta2, float eps, float grad_scale, void* stream) {
    BMHRL_CHECK_ARG(params && grads && exp_avg && exp_avg_sq && numels && chunk_tensor && chunk_start && step_counts,
                    "bmhrl_adam_multi_f32: null pointer");
    if (n_chunks <= 0) return BMHRL_OK;
    adam_multi_kernel<<<adam_grid(n_chunks), 256, 0, (cudaStream_t)stream>>>(params, grads, exp_avg, exp_avg_sq, numels,
                                                                             chunk_tensor, chunk_start, n_chunks, step_counts, lr,
                                                                             beta1, beta2, eps, grad_scale, nullptr, nullptr,
                                                                             nullptr, nullptr);
    BMHRL_POST_LAUNCH("adam_multi");
    return BMHRL_OK;
}

extern "C" int bmhrl_adam_multi_shadow_f32(float* const* params, const float* const* grads, float* const* exp_avg,
                                           float* const* exp_avg_sq, const long long* numels, const int* chunk_tensor,
                                           const int* chunk_start, int n_chunks, const float* step_counts, float lr,
                                           float beta1, float beta2, float eps, float grad_scale, void* const* shadow,
                                           const int* shadow_cols, const int* shadow_ld, const int* shadow_kind,
                                           void* stream) {
    BMHRL_CHECK_ARG(params && grads && exp_avg && exp_avg_sq && numels && chunk_tensor && chunk_start && step_counts,
                    "bmhrl_adam_multi_shadow_f32: null pointer");
    BMHRL_CHECK_ARG(!shadow || (shadow_cols && shadow_ld && shadow_kind), "bmhrl_adam_multi_shadow_f32: shadow tables");
    if (n_chunks <= 0) return BMHRL_OK;
    adam_multi_kernel<<<adam_grid(n_chunks), 256, 0, (cudaStream_t)stream>>>(params, grads, exp_avg, exp_avg_sq, numels,
                                                                             chunk_tensor, chunk_start, n_chunks, step_counts, lr,
                                                                             beta1, beta2, eps, grad_scale, shadow, shadow_cols,
                                                                             shadow_ld, shadow_kind);
    BMHRL_POST_LAUNCH("adam_multi");
    return BMHRL_OK;
}

extern "C" int bmhrl_tick(float* step_counts, const int* active_tensors, int n_active,
                          unsigned long long* rng_state, void* stream) {
    tick_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(step_counts, active_tensors, n_active, rng_state);
    BMHRL_POST_LAUNCH("tick");
    return BMHRL_OK;
}
