// K1 for the TRAINING path in bf16 math mode: fused masked multi-head attention core, forward AND backward, for sequences
// that fit one tile (Sq <= 128, Sk <= 128) -- the caption-side attentions always do (L <= 30), and so do the encoder's at
// the benchmark shapes (Sv = 30, Sa = 80).  Longer sequences take the flash-style kernels (attn_bf16.cu / attn_bf16_bwd.cu).
// Reference: attention() model/multihead_attention.py:7-31 -- softmax(mask_replace(Q K^T / sqrt(d_k), -1e9)) V, dropout on
// the product -- and what autograd derives from it.  Call sites: model/bm_hrl_agent.py:89-94,358-364,481.
//
// One CTA per (batch, head[, 256-wide slice of d_k when d_k = 512]); everything of that (b, h) is resident:
//   TMA     Q, K, V (and dO) bf16 tiles straight out of the N-concatenated projection output (head h = columns
//           [h d_k, (h+1) d_k) of a (B*S, ld) matrix), 128B-swizzled, rows beyond the sequence zero-filled
//   tcgen05 forward : S = Q K^T (SS) -> TMEM;  softmax (thread = query row = TMEM lane) writes the UNNORMALISED
//                     probabilities back into TMEM as packed bf16;  O = P V (A operand from TMEM, V MN-major) -> TMEM;
//                     epilogue: O / l, Philox dropout (element index = row * H d_k + column, the index the epilogue
//                     of the out-projection's dX GEMM regenerates), bf16 store, log-sum-exp store.  No (B,H,Sq,Sk) tensor.
//           backward: phase A (lanes = queries): S = Q K^T, dP = dO V^T;  P = exp2(S - lse), delta = rowsum(P o dP)
//                     (the whole row is resident: no separate rowsum(dO o O) pass and no saved O), dS = P o (dP - delta)
//                     -> TMEM as packed bf16;  dQ = scale * dS K (K MN-major).
//                     phase B (lanes = keys): S^T = K Q^T, dP^T = V dO^T;  P^T and dS^T -> TMEM;  dK = scale * dS^T Q,
//                     dV = P^T dO.  The same four shared-memory tiles serve all seven contractions: a (rows, d_k) tile
//                     with d_k contiguous is the K-major operand of a contraction over d_k and the MN-major operand of a
//                     contraction over its rows.
// Masked scores are REPLACED by -1e9 (a fully masked row is uniform over all Sk keys and sends no gradient to Q / K), as in
// the reference.  dO must already carry the dropout mask of the forward (the dX GEMM of the out-projection applies it).
#include <cuda_bf16.h>
#include "common.cuh"
#include "tc_common.cuh"
#include "attn_bf16_common.cuh"
#include "../../include/bmhrl_b200.h"

namespace bmhrl {

using namespace tc;

constexpr int A16_SLACK = 16 * 1024;  // an M = 128 A operand reads 128 rows of a tile that holds round16(S) of them

struct A16Params {
    int B, H, Sq, Sk, dk, Sqp, Skp, nck, dvt;
    int t_p, t_a, t_cols, two_acc;             // backward TMEM map: dP region, accumulator region, columns allocated
    float scale;
    const unsigned char* mask; int mask_kind, causal;
    float* lse2;                               // (B,H,Sq,2): row maximum m (log2 domain) and 1 / normaliser of the masked
                                               // scaled scores.  Kept apart on purpose: a fully masked row has m = -1e9 * log2(e),
                                               // where fp32 spacing is 128, so m + log2(l) would lose the normaliser
    // forward
    __nv_bfloat16* O; long long ldo;
    float drop_p; const unsigned long long* rng; unsigned int salt;
    // backward
    __nv_bfloat16 *dQ, *dK, *dV; long long lddq, lddk, lddv;
};

// visibility bits of query row q for the 32 keys [k0, k0 + 32)
__device__ __forceinline__ uint32_t a16_vis_row(const A16Params& p, int b, int q, bool row_ok, int k0, int lane) {
    const int rem = p.Sk - k0;
    uint32_t v = rem >= 32 ? 0xffffffffu : (rem <= 0 ? 0u : ((1u << rem) - 1u));
    if (p.mask_kind == 1) {
        const int kk = k0 + lane;
        const bool on = kk < p.Sk && p.mask[(long long)b * p.Sk + kk] != 0;
        v = __ballot_sync(0xffffffffu, on);
    } else if (p.mask_kind == 2) {
        uint32_t m = 0u;
        if (row_ok) {
            const unsigned char* mr = p.mask + ((long long)b * p.Sq + q) * p.Sk + k0;
            const int cnt = min(32, rem);
            for (int i = 0; i < cnt; ++i) m |= (mr[i] != 0 ? 1u : 0u) << i;
        }
        v = m;
    }
    if (p.causal) {
        const int lim = q - k0;  // largest visible key offset inside this piece
        if (lim < 0) v = 0u;
        else if (lim < 31) v &= (2u << lim) - 1u;
    }
    return v;
}

// ------------------------------------------------------------------------------------------------ forward
__global__ void __launch_bounds__(192, 1) attn16_fwd_kernel(const __grid_constant__ CUtensorMap tmQ,
                                                            const __grid_constant__ CUtensorMap tmK,
                                                            const __grid_constant__ CUtensorMap tmV,
                                                            const A16Params p, const int tmem_cols) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t qch = (uint32_t)p.Sqp * 128u, kch = (uint32_t)p.Skp * 128u;   // bytes per 64-column chunk
    const uint32_t sQ = base, sK = sQ + p.nck * qch, sV = sK + p.nck * kch;
    const uint32_t bars = sV + (p.dvt >> 6) * kch + A16_SLACK;
    const uint32_t qk_full = bars, v_full = bars + 8, s_full = bars + 16, p_ready = bars + 24, o_full = bars + 32;
    const uint32_t tmem_slot = bars + 40;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int bh = blockIdx.x, b = bh / p.H, h = bh - b * p.H;
    const int dv0 = blockIdx.y * p.dvt;
    const int col0 = h * p.dk;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmK);
        tma_prefetch_desc(&tmV);
        mbar_init(qk_full, 1);
        mbar_init(v_full, 1);
        mbar_init(s_full, 1);
        mbar_init(p_ready, 128);
        mbar_init(o_full, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    const uint32_t tS = tmem_base, tO = tmem_base + 128;

    if (warp == 0) {
        if (elect_one()) {
            griddep_wait();                 // Q / K / V are the previous kernel's output
            griddep_launch_dependents();
            mbar_arrive_expect_tx(qk_full, p.nck * (qch + kch));
            for (int c = 0; c < p.nck; ++c) {
                tma_load_4d(sQ + c * qch, &tmQ, qk_full, col0 + c * 64, 0, 0, b);
                tma_load_4d(sK + c * kch, &tmK, qk_full, col0 + c * 64, 0, 0, b);
            }
            mbar_arrive_expect_tx(v_full, (p.dvt >> 6) * kch);
            for (int c = 0; c < (p.dvt >> 6); ++c) tma_load_4d(sV + c * kch, &tmV, v_full, col0 + dv0 + c * 64, 0, 0, b);
        }
    } else if (warp == 1) {
        if (elect_one()) {
            const uint32_t idesc_s = make_idesc(1, 128, p.Skp, 0, 0);
            const uint32_t idesc_o = make_idesc(1, 128, p.dvt, 0, 1);
            mbar_wait(qk_full, 0);
            tc_fence_after();
            for (int c = 0; c < p.nck; ++c) {
                const uint64_t a0 = make_smem_desc(sQ + c * qch, 16, 1024), b0 = make_smem_desc(sK + c * kch, 16, 1024);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    mma_bf16_ss(tS, desc_advance(a0, k * 32), desc_advance(b0, k * 32), idesc_s, (c > 0 || k > 0) ? 1u : 0u);
            }
            tc_commit(s_full);
            mbar_wait(p_ready, 0);
            mbar_wait(v_full, 0);
            tc_fence_after();
            const uint64_t v0 = make_smem_desc(sV, kch, 1024);  // MN-major: 64-column chunks kch bytes apart
            for (int kk = 0; kk < (p.Skp >> 4); ++kk)
                mma_bf16_ts(tO, tS + kk * 8, desc_advance(v0, kk * 2048), idesc_o, kk > 0 ? 1u : 0u);
            tc_commit(o_full);
        }
    } else {
        griddep_wait();                     // the mask may be a tensor an earlier kernel of the stream wrote
        const int qt = warp & 3;
        const int q = qt * 32 + lane;
        const bool row_ok = q < p.Sq;
        const uint32_t lane_off = (uint32_t)(qt * 32) << 16;
        constexpr float LOG2E = 1.4426950408889634f;
        const float sl2 = p.scale * LOG2E, neg2 = kNegFill * LOG2E;
        const int nch = (p.Skp + 31) >> 5;
        uint32_t vis[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) vis[c] = c < nch ? a16_vis_row(p, b, q, row_ok, c * 32, lane) : 0u;
        mbar_wait(s_full, 0);
        tc_fence_after();
        float m = -INFINITY;
#pragma unroll 1
        for (int c = 0; c < nch; ++c) {  // pass 1: row maximum
            uint32_t r[32];
            tmem_ld_32x32_b(tS + lane_off + c * 32, r);
            tmem_ld_wait_b();
            const int cnt = min(32, p.Sk - c * 32);
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const float t = ((vis[c] >> i) & 1u) ? __uint_as_float(r[i]) * sl2 : neg2;
                m = fmaxf(m, i < cnt ? t : -INFINITY);
            }
        }
        float l = 0.f;
#pragma unroll 1
        for (int c = 0; c < nch; ++c) {  // pass 2: unnormalised probabilities -> packed bf16 over the scores
            uint32_t r[32], out[16];
            tmem_ld_32x32_b(tS + lane_off + c * 32, r);
            tmem_ld_wait_b();
            const int cnt = min(32, p.Sk - c * 32);
            float e[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const float t = ((vis[c] >> i) & 1u) ? __uint_as_float(r[i]) * sl2 : neg2;
                e[i] = i < cnt ? ex2_approx(t - m) : 0.f;
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                out[i] = pack_bf16x2(e[2 * i], e[2 * i + 1]);
                // the normaliser sums what the tensor core will multiply: the bf16-rounded probabilities
                l += __uint_as_float(out[i] << 16) + __uint_as_float(out[i] & 0xffff0000u);
            }
            tmem_st_32x32_x16(tS + lane_off + c * 16, out);
        }
        tmem_st_wait_b();
        tc_fence_before();
        mbar_arrive(p_ready);
        if (row_ok && blockIdx.y == 0) {
            p.lse2[2 * ((long long)bh * p.Sq + q)] = m;
            p.lse2[2 * ((long long)bh * p.Sq + q) + 1] = 1.f / l;
        }
        // ---- epilogue
        mbar_wait(o_full, 0);
        tc_fence_after();
        const float inv_l = 1.f / l;
        const float ks = p.drop_p > 0.f ? 1.f / (1.f - p.drop_p) : 1.f;
        const long long grow = (long long)b * p.Sq + q;
        __nv_bfloat16* orow = p.O + grow * p.ldo + col0 + dv0;
        const unsigned long long didx = (unsigned long long)grow * ((unsigned long long)p.H * p.dk) + col0 + dv0;
#pragma unroll 1
        for (int c = 0; c < (p.dvt >> 5); ++c) {
            uint32_t a[32];
            tmem_ld_32x32_b(tO + lane_off + c * 32, a);
            tmem_ld_wait_b();
            if (row_ok) {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    float v[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(a[u * 8 + i]) * inv_l;
                    if (p.drop_p > 0.f) {
#pragma unroll
                        for (int g = 0; g < 2; ++g) {
                            const Philox4 rb = philox_block(p.rng, p.salt, (didx + c * 32 + u * 8 + g * 4) >> 2);
                            v[g * 4 + 0] = u32_to_unit(rb.x) >= p.drop_p ? v[g * 4 + 0] * ks : 0.f;
                            v[g * 4 + 1] = u32_to_unit(rb.y) >= p.drop_p ? v[g * 4 + 1] * ks : 0.f;
                            v[g * 4 + 2] = u32_to_unit(rb.z) >= p.drop_p ? v[g * 4 + 2] * ks : 0.f;
                            v[g * 4 + 3] = u32_to_unit(rb.w) >= p.drop_p ? v[g * 4 + 3] * ks : 0.f;
                        }
                    }
                    uint4 v4;
                    v4.x = pack_bf16x2(v[0], v[1]); v4.y = pack_bf16x2(v[2], v[3]);
                    v4.z = pack_bf16x2(v[4], v[5]); v4.w = pack_bf16x2(v[6], v[7]);
                    *reinterpret_cast<uint4*>(orow + c * 32 + u * 8) = v4;
                }
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, tmem_cols);
    }
}

// ------------------------------------------------------------------------------------------------ backward
__global__ void __launch_bounds__(192, 1) attn16_bwd_kernel(const __grid_constant__ CUtensorMap tmQ,
                                                            const __grid_constant__ CUtensorMap tmK,
                                                            const __grid_constant__ CUtensorMap tmV,
                                                            const __grid_constant__ CUtensorMap tmG,
                                                            const A16Params p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t qch = (uint32_t)p.Sqp * 128u, kch = (uint32_t)p.Skp * 128u;
    const uint32_t sQ = base, sG = sQ + p.nck * qch, sK = sG + p.nck * qch, sV = sK + p.nck * kch;
    const uint32_t stats = sV + p.nck * kch + A16_SLACK;       // m[128], 1/l[128], delta[128] of the queries
    const uint32_t bars = stats + 2048;
    const uint32_t ld_full = bars, a_full = bars + 8, ds_ready = bars + 16, dq_full = bars + 24, b_full = bars + 32;
    const uint32_t pt_ready = bars + 40, dk_full = bars + 48, dk_free = bars + 56, dv_full = bars + 64;
    const uint32_t tmem_slot = bars + 72;
    float* st_lse = reinterpret_cast<float*>(smem_raw + (stats - smem_u32(smem_raw)));
    float* st_il = st_lse + 128;
    float* st_del = st_lse + 256;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int bh = blockIdx.x, b = bh / p.H, h = bh - b * p.H;
    const int dv0 = blockIdx.y * p.dvt;
    const int col0 = h * p.dk;
    const bool two_acc = p.two_acc != 0;  // dK and dV accumulators side by side

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmK);
        tma_prefetch_desc(&tmV);
        tma_prefetch_desc(&tmG);
        mbar_init(ld_full, 1);
        mbar_init(a_full, 1);
        mbar_init(ds_ready, 128);
        mbar_init(dq_full, 1);
        mbar_init(b_full, 1);
        mbar_init(pt_ready, 128);
        mbar_init(dk_full, 1);
        mbar_init(dk_free, 128);
        mbar_init(dv_full, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, p.t_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    const uint32_t tS = tmem_base, tP = tmem_base + p.t_p, tA = tmem_base + p.t_a;

    if (warp == 0) {
        if (elect_one()) {
            griddep_wait();                 // dO is the previous kernel's output
            griddep_launch_dependents();
            mbar_arrive_expect_tx(ld_full, 2 * p.nck * (qch + kch));
            for (int c = 0; c < p.nck; ++c) {
                tma_load_4d(sQ + c * qch, &tmQ, ld_full, col0 + c * 64, 0, 0, b);
                tma_load_4d(sK + c * kch, &tmK, ld_full, col0 + c * 64, 0, 0, b);
                tma_load_4d(sG + c * qch, &tmG, ld_full, col0 + c * 64, 0, 0, b);
                tma_load_4d(sV + c * kch, &tmV, ld_full, col0 + c * 64, 0, 0, b);
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            const uint32_t idesc_sk = make_idesc(1, 128, p.Skp, 0, 0);   // lanes = queries, N = keys
            const uint32_t idesc_sq = make_idesc(1, 128, p.Sqp, 0, 0);   // lanes = keys, N = queries
            const uint32_t idesc_f = make_idesc(1, 128, p.dvt, 0, 1);    // accumulating MMAs: B MN-major, N = d_k slice
            const int cz = dv0 >> 6;  // first 64-column chunk of this CTA's d_k slice
            // contraction over d_k: D = X Y^T, X rows on the lanes (chunks xch bytes apart), Y rows = N (ych apart)
            auto dk_mma = [&](uint32_t tD, uint32_t sX, uint32_t xch, uint32_t sY, uint32_t ych, uint32_t idesc) {
                for (int c = 0; c < p.nck; ++c) {
                    const uint64_t a0 = make_smem_desc(sX + c * xch, 16, 1024), b0 = make_smem_desc(sY + c * ych, 16, 1024);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        mma_bf16_ss(tD, desc_advance(a0, k * 32), desc_advance(b0, k * 32), idesc, (c > 0 || k > 0) ? 1u : 0u);
                }
            };
            // accumulate over the rows of an MN-major tile: D = A(tmem, packed bf16) Z, Z = `steps` x 16 rows
            auto row_mma = [&](uint32_t tD, uint32_t tAop, uint32_t sZ, uint32_t zch, int steps) {
                const uint64_t z0 = make_smem_desc(sZ + cz * zch, zch, 1024);
                for (int kk = 0; kk < steps; ++kk)
                    mma_bf16_ts(tD, tAop + kk * 8, desc_advance(z0, kk * 2048), idesc_f, kk > 0 ? 1u : 0u);
            };
            mbar_wait(ld_full, 0);
            tc_fence_after();
            dk_mma(tS, sQ, qch, sK, kch, idesc_sk);   // S  = Q K^T
            dk_mma(tP, sG, qch, sV, kch, idesc_sk);   // dP = dO V^T
            tc_commit(a_full);
            mbar_wait(ds_ready, 0);                   // dS (packed bf16) sits in tS[0 : Skp/2), tP is free
            tc_fence_after();
            row_mma(tA, tS, sK, kch, p.Skp >> 4);     // dQ = dS K
            tc_commit(dq_full);
            dk_mma(tP, sV, kch, sG, qch, idesc_sq);   // dP^T = V dO^T   (tP is not an operand of the dQ MMAs)
            mbar_wait(dq_full, 0);                    // the dQ MMAs have read dS: tS may be overwritten
            tc_fence_after();
            dk_mma(tS, sK, kch, sQ, qch, idesc_sq);   // S^T = K Q^T
            tc_commit(b_full);
            mbar_wait(pt_ready, 0);                   // P^T in tS, dS^T in tP (packed); the dQ accumulator has been drained
            tc_fence_after();
            row_mma(tA, tP, sQ, qch, p.Sqp >> 4);     // dK = dS^T Q
            tc_commit(dk_full);
            if (!two_acc) {
                mbar_wait(dk_free, 0);                // dK drained from the shared accumulator columns
                tc_fence_after();
            }
            row_mma(two_acc ? tA + p.dvt : tA, tS, sG, qch, p.Sqp >> 4);   // dV = P^T dO
            tc_commit(dv_full);
        }
    } else {
        griddep_wait();
        const int qt = warp & 3;
        const int r = qt * 32 + lane;             // phase A: query row; phase B: key row
        const uint32_t lane_off = (uint32_t)(qt * 32) << 16;
        constexpr float LOG2E = 1.4426950408889634f;
        const float sl2 = p.scale * LOG2E, neg2 = kNegFill * LOG2E;
        // ================= phase A: lanes = queries
        {
            const bool row_ok = r < p.Sq;
            const int nch = (p.Skp + 31) >> 5;
            uint32_t vis[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) vis[c] = c < nch ? a16_vis_row(p, b, r, row_ok, c * 32, lane) : 0u;
            const float lse2 = row_ok ? p.lse2[2 * ((long long)bh * p.Sq + r)] : INFINITY;   // row maximum
            const float il = row_ok ? p.lse2[2 * ((long long)bh * p.Sq + r) + 1] : 0.f;       // 1 / normaliser
            mbar_wait(a_full, 0);
            tc_fence_after();
            float delta = 0.f;
#pragma unroll 1
            for (int c = 0; c < nch; ++c) {
                uint32_t s[32], dp[32];
                tmem_ld_32x32_b(tS + lane_off + c * 32, s);
                tmem_ld_32x32_b(tP + lane_off + c * 32, dp);
                tmem_ld_wait_b();
                const int cnt = min(32, p.Sk - c * 32);
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float t = ((vis[c] >> i) & 1u) ? __uint_as_float(s[i]) * sl2 : neg2;
                    const float pr = i < cnt ? ex2_approx(t - lse2) * il : 0.f;
                    delta = fmaf(pr, __uint_as_float(dp[i]), delta);
                }
            }
#pragma unroll 1
            for (int c = 0; c < nch; ++c) {
                uint32_t s[32], dp[32], out[16];
                tmem_ld_32x32_b(tS + lane_off + c * 32, s);
                tmem_ld_32x32_b(tP + lane_off + c * 32, dp);
                tmem_ld_wait_b();
                const int cnt = min(32, p.Sk - c * 32);
                float g[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const bool on = ((vis[c] >> i) & 1u) != 0u && i < cnt;
                    const float pr = ex2_approx(__uint_as_float(s[i]) * sl2 - lse2) * il;
                    g[i] = on ? pr * (__uint_as_float(dp[i]) - delta) : 0.f;   // replaced scores carry no gradient
                }
#pragma unroll
                for (int i = 0; i < 16; ++i) out[i] = pack_bf16x2(g[2 * i], g[2 * i + 1]);
                tmem_st_32x32_x16(tS + lane_off + c * 16, out);
            }
            st_lse[r] = lse2;
            st_il[r] = il;
            st_del[r] = row_ok ? delta : 0.f;
            tmem_st_wait_b();
            tc_fence_before();
            mbar_arrive(ds_ready);
            // dQ epilogue
            mbar_wait(dq_full, 0);
            tc_fence_after();
            __nv_bfloat16* orow = p.dQ + ((long long)b * p.Sq + r) * p.lddq + col0 + dv0;
#pragma unroll 1
            for (int c = 0; c < (p.dvt >> 5); ++c) {
                uint32_t a[32];
                tmem_ld_32x32_b(tA + lane_off + c * 32, a);
                tmem_ld_wait_b();
                if (row_ok) {
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        uint4 v4;
                        v4.x = pack_bf16x2(__uint_as_float(a[u * 8]) * p.scale, __uint_as_float(a[u * 8 + 1]) * p.scale);
                        v4.y = pack_bf16x2(__uint_as_float(a[u * 8 + 2]) * p.scale, __uint_as_float(a[u * 8 + 3]) * p.scale);
                        v4.z = pack_bf16x2(__uint_as_float(a[u * 8 + 4]) * p.scale, __uint_as_float(a[u * 8 + 5]) * p.scale);
                        v4.w = pack_bf16x2(__uint_as_float(a[u * 8 + 6]) * p.scale, __uint_as_float(a[u * 8 + 7]) * p.scale);
                        *reinterpret_cast<uint4*>(orow + c * 32 + u * 8) = v4;
                    }
                }
            }
            tc_fence_before();
        }
        // the four row warps exchange lse2 / delta through shared memory
        asm volatile("bar.sync 1, 128;" ::: "memory");
        // ================= phase B: lanes = keys
        {
            const bool key_ok = r < p.Sk;
            bool key_vis = key_ok;
            if (p.mask_kind == 1) key_vis = key_ok && p.mask[(long long)b * p.Sk + r] != 0;
            const int nch = (p.Sqp + 31) >> 5;
            mbar_wait(b_full, 0);
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < nch; ++c) {
                uint32_t s[32], dp[32], outp[16], outd[16];
                tmem_ld_32x32_b(tS + lane_off + c * 32, s);
                tmem_ld_32x32_b(tP + lane_off + c * 32, dp);
                tmem_ld_wait_b();
                float pt[32], dst[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const int qi = c * 32 + i;
                    bool on = key_vis;
                    if (p.mask_kind == 2)
                        on = key_ok && qi < p.Sq && p.mask[((long long)b * p.Sq + qi) * p.Sk + r] != 0;
                    if (p.causal) on = on && r <= qi;
                    const float l2 = st_lse[qi];      // +inf (and 1/l = 0) for queries out of range: probability 0
                    const float t = on ? __uint_as_float(s[i]) * sl2 : neg2;
                    const float pr = key_ok ? ex2_approx(t - l2) * st_il[qi] : 0.f;
                    pt[i] = pr;
                    dst[i] = on ? pr * (__uint_as_float(dp[i]) - st_del[qi]) : 0.f;
                }
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    outp[i] = pack_bf16x2(pt[2 * i], pt[2 * i + 1]);
                    outd[i] = pack_bf16x2(dst[2 * i], dst[2 * i + 1]);
                }
                tmem_st_32x32_x16(tS + lane_off + c * 16, outp);
                tmem_st_32x32_x16(tP + lane_off + c * 16, outd);
            }
            tmem_st_wait_b();
            tc_fence_before();
            mbar_arrive(pt_ready);
            auto drain = [&](uint32_t tD, __nv_bfloat16* dst, long long ld, float sc) {
                __nv_bfloat16* orow = dst + ((long long)b * p.Sk + r) * ld + col0 + dv0;
#pragma unroll 1
                for (int c = 0; c < (p.dvt >> 5); ++c) {
                    uint32_t a[32];
                    tmem_ld_32x32_b(tD + lane_off + c * 32, a);
                    tmem_ld_wait_b();
                    if (key_ok) {
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            uint4 v4;
                            v4.x = pack_bf16x2(__uint_as_float(a[u * 8]) * sc, __uint_as_float(a[u * 8 + 1]) * sc);
                            v4.y = pack_bf16x2(__uint_as_float(a[u * 8 + 2]) * sc, __uint_as_float(a[u * 8 + 3]) * sc);
                            v4.z = pack_bf16x2(__uint_as_float(a[u * 8 + 4]) * sc, __uint_as_float(a[u * 8 + 5]) * sc);
                            v4.w = pack_bf16x2(__uint_as_float(a[u * 8 + 6]) * sc, __uint_as_float(a[u * 8 + 7]) * sc);
                            *reinterpret_cast<uint4*>(orow + c * 32 + u * 8) = v4;
                        }
                    }
                }
            };
            mbar_wait(dk_full, 0);
            tc_fence_after();
            drain(tA, p.dK, p.lddk, p.scale);
            tc_fence_before();
            mbar_arrive(dk_free);
            mbar_wait(dv_full, 0);
            tc_fence_after();
            drain(two_acc ? tA + p.dvt : tA, p.dV, p.lddv, 1.f);
            tc_fence_before();
        }
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, p.t_cols);
    }
}

// ---------------------------------------------------------------------------------------------- host side
static bool a16_ok16(const void* p, long long ld) { return p && aligned16(p) && ld > 0 && (ld & 7) == 0; }

static int a16_geometry(A16Params& p, int B, int H, int Sq, int Sk, int dk, const char* who) {
    if (!(B > 0 && H > 0 && Sq > 0 && Sk > 0 && Sq <= 128 && Sk <= 128)) {
        bmhrl_set_error("attn16: sequences longer than one 128-row tile take the flash-style kernels (bmhrl_attn_fwd_bf16)");
        return BMHRL_ERR_UNSUPPORTED;
    }
    if (!(dk == 64 || dk == 128 || dk == 256 || dk == 512)) {
        bmhrl_set_error("attn16: d_k must be 64, 128, 256 or 512");
        return BMHRL_ERR_UNSUPPORTED;
    }
    p.B = B; p.H = H; p.Sq = Sq; p.Sk = Sk; p.dk = dk;
    p.Sqp = (Sq + 15) & ~15; p.Skp = (Sk + 15) & ~15;
    p.nck = dk >> 6;
    // Tensor memory is the resource these kernels share with the GEMM CTAs that run beside them (512 columns per SM, an
    // allocation blocks until it fits): slices of the head dimension are chosen so that a CTA needs 256 columns, never
    // the whole SM -- more, smaller CTAs that co-reside with GEMM tiles instead of serialising against them.
    const bool bwd = who[0] == 'b';
    if (!bwd) {
        p.dvt = dk > 128 ? 128 : dk;                       // S (128) + O slice (128)
        p.t_cols = 128 + p.dvt <= 256 ? 256 : 512;
    } else {
        const int cs = ((p.Sqp > p.Skp ? p.Sqp : p.Skp) + 31) & ~31;   // S / S^T and dP / dP^T regions
        p.t_p = cs;
        p.t_a = 2 * cs;
        p.two_acc = 0;
        // cs <= 64: 2 x 64 + 128 = 256 columns.  Longer sequences need all 512 (2 x 128 + 128); 64-wide slices would fit
        // 256 but make four CTAs per (b, h) recompute S and dP -- measured slower (27 vs 21 us at Sq = Sk = 80)
        p.dvt = dk > 128 ? 128 : dk;
        const int need = 2 * cs + p.dvt;
        p.t_cols = need <= 64 ? 64 : need <= 128 ? 128 : need <= 256 ? 256 : 512;
    }
    p.scale = 1.0f / sqrtf((float)dk);
    return BMHRL_OK;
}

}  // namespace bmhrl

using namespace bmhrl;

extern "C" long long bmhrl_attn16_smem_bytes(int backward, int Sq, int Sk, int dk) {
    const long long Sqp = (Sq + 15) & ~15, Skp = (Sk + 15) & ~15, nck = dk >> 6, dvt = dk > 128 ? 128 : dk;
    if (backward) return 2 * nck * (Sqp + Skp) * 128 + A16_SLACK + 2048 + 128 + 1024;
    return nck * (Sqp + Skp) * 128 + (dvt >> 6) * Skp * 128 + A16_SLACK + 128 + 1024;
}

extern "C" int bmhrl_attn16_fwd(const void* Q, long long ldq, const void* K, long long ldk, const void* V, long long ldv,
                                const unsigned char* mask, int mask_kind, int causal, void* O, long long ldo, float* lse2,
                                int B, int H, int Sq, int Sk, int dk, float drop_p, const unsigned long long* rng_state,
                                unsigned int salt, void* stream) {
    BMHRL_CHECK_ARG(a16_ok16(Q, ldq) && a16_ok16(K, ldk) && a16_ok16(V, ldv) && a16_ok16(O, ldo) && lse2,
                    "bmhrl_attn16_fwd: bf16 operands must be 16-byte aligned with row pitches that are multiples of 8");
    BMHRL_CHECK_ARG(mask_kind >= 0 && mask_kind <= 2 && (mask_kind == 0 || mask), "bmhrl_attn16_fwd: mask");
    BMHRL_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f && (drop_p == 0.f || rng_state), "bmhrl_attn16_fwd: dropout");
    A16Params p{};
    int rc = a16_geometry(p, B, H, Sq, Sk, dk, "fwd");
    if (rc != BMHRL_OK) return rc;
    const long long smem = bmhrl_attn16_smem_bytes(0, Sq, Sk, dk);
    if (smem > 220 * 1024) {
        bmhrl_set_error("bmhrl_attn16_fwd: tiles do not fit shared memory");
        return BMHRL_ERR_UNSUPPORTED;
    }
    p.mask = mask; p.mask_kind = mask_kind; p.causal = (causal && mask_kind != 0) ? 1 : 0;
    p.lse2 = lse2; p.O = (__nv_bfloat16*)O; p.ldo = ldo; p.drop_p = drop_p; p.rng = rng_state; p.salt = salt;
    const CUtensorMapDataType BF = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    const unsigned long long D = (unsigned long long)H * dk;
    CUtensorMap tmQ, tmK, tmV;
    bool ok = encode_4d(&tmQ, BF, 2, Q, D, Sq, ldq, 1, 0, B, (unsigned long long)Sq * ldq, 64, p.Sqp) &&
              encode_4d(&tmK, BF, 2, K, D, Sk, ldk, 1, 0, B, (unsigned long long)Sk * ldk, 64, p.Skp) &&
              encode_4d(&tmV, BF, 2, V, D, Sk, ldv, 1, 0, B, (unsigned long long)Sk * ldv, 64, p.Skp);
    if (!ok) {
        bmhrl_set_error("bmhrl_attn16_fwd: cuTensorMapEncodeTiled failed");
        return BMHRL_ERR_CUDA;
    }
    static DeviceOnce attr;
    if (needs_setup(attr)) {
        if (cudaFuncSetAttribute(attn16_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024) != cudaSuccess) {
            bmhrl_set_error("bmhrl_attn16_fwd: cannot raise the shared-memory limit");
            return BMHRL_ERR_CUDA;
        }
        mark_done(attr);
    }
    BMHRL_LAUNCH_PDL("attn16_fwd", attn16_fwd_kernel, dim3(B * H, dk / p.dvt), dim3(192), (size_t)smem, (cudaStream_t)stream,
                     tmQ, tmK, tmV, p, p.t_cols);
    return BMHRL_OK;
}

extern "C" int bmhrl_attn16_bwd(const void* Q, long long ldq, const void* K, long long ldk, const void* V, long long ldv,
                                const void* dO, long long lddo, const float* lse2, const unsigned char* mask, int mask_kind,
                                int causal, void* dQ, long long lddq, void* dK, long long lddk, void* dV, long long lddv,
                                int B, int H, int Sq, int Sk, int dk, void* stream) {
    BMHRL_CHECK_ARG(a16_ok16(Q, ldq) && a16_ok16(K, ldk) && a16_ok16(V, ldv) && a16_ok16(dO, lddo) && lse2,
                    "bmhrl_attn16_bwd: bf16 operands must be 16-byte aligned with row pitches that are multiples of 8");
    BMHRL_CHECK_ARG(a16_ok16(dQ, lddq) && a16_ok16(dK, lddk) && a16_ok16(dV, lddv), "bmhrl_attn16_bwd: gradient outputs");
    BMHRL_CHECK_ARG(mask_kind >= 0 && mask_kind <= 2 && (mask_kind == 0 || mask), "bmhrl_attn16_bwd: mask");
    A16Params p{};
    int rc = a16_geometry(p, B, H, Sq, Sk, dk, "bwd");
    if (rc != BMHRL_OK) return rc;
    const long long smem = bmhrl_attn16_smem_bytes(1, Sq, Sk, dk);
    if (smem > 220 * 1024) {
        bmhrl_set_error("bmhrl_attn16_bwd: tiles do not fit shared memory");
        return BMHRL_ERR_UNSUPPORTED;
    }
    p.mask = mask; p.mask_kind = mask_kind; p.causal = (causal && mask_kind != 0) ? 1 : 0;
    p.lse2 = const_cast<float*>(lse2);
    p.dQ = (__nv_bfloat16*)dQ; p.dK = (__nv_bfloat16*)dK; p.dV = (__nv_bfloat16*)dV;
    p.lddq = lddq; p.lddk = lddk; p.lddv = lddv;
    const CUtensorMapDataType BF = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    const unsigned long long D = (unsigned long long)H * dk;
    CUtensorMap tmQ, tmK, tmV, tmG;
    bool ok = encode_4d(&tmQ, BF, 2, Q, D, Sq, ldq, 1, 0, B, (unsigned long long)Sq * ldq, 64, p.Sqp) &&
              encode_4d(&tmK, BF, 2, K, D, Sk, ldk, 1, 0, B, (unsigned long long)Sk * ldk, 64, p.Skp) &&
              encode_4d(&tmV, BF, 2, V, D, Sk, ldv, 1, 0, B, (unsigned long long)Sk * ldv, 64, p.Skp) &&
              encode_4d(&tmG, BF, 2, dO, D, Sq, lddo, 1, 0, B, (unsigned long long)Sq * lddo, 64, p.Sqp);
    if (!ok) {
        bmhrl_set_error("bmhrl_attn16_bwd: cuTensorMapEncodeTiled failed");
        return BMHRL_ERR_CUDA;
    }
    static DeviceOnce attr;
    if (needs_setup(attr)) {
        if (cudaFuncSetAttribute(attn16_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024) != cudaSuccess) {
            bmhrl_set_error("bmhrl_attn16_bwd: cannot raise the shared-memory limit");
            return BMHRL_ERR_CUDA;
        }
        mark_done(attr);
    }
    BMHRL_LAUNCH_PDL("attn16_bwd", attn16_bwd_kernel, dim3(B * H, dk / p.dvt), dim3(192), (size_t)smem, (cudaStream_t)stream,
                     tmQ, tmK, tmV, tmG, p);
    return BMHRL_OK;
}
