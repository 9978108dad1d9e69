// K1 backward, bf16 operands (config 5, forward + backward): gradients of
//   O = softmax(mask_replace(Q K^T / sqrt(d_k), -1e9)) V          (attention(), model/multihead_attention.py:7-31)
// with respect to Q, K and V on the head-merged (B, S, H*d_k) tensors, flash-style: the (B,H,Sq,Sk) probabilities are
// never stored -- they are recomputed from Q, K and the forward's per-row log-sum-exp, like the scores in the forward.
//   P = exp(S - lse),  dV = P^T dO,  dP = dO V^T,  dS = P o (dP - delta),  delta = rowsum(dO o O),
//   dQ = scale * dS K,  dK = scale * dS^T Q.      Masked positions were REPLACED by a constant: their dS is 0.
//
// One persistent kernel, three instantiations (one launch each), all with the same structure.  An item is a 128-row tile
// of the gradient being produced; the other sequence is streamed through in 64-column steps:
//   DQ : rows = queries.  resident Q, dO tiles;  step j streams K_j and V_j:
//        S = Q K_j^T, dP = dO V_j^T -> TMEM;  dS (bf16) -> TMEM;  dQ += dS K_j        (K_j delivered again, MN-major)
//   DK : rows = keys.     resident K, V tiles;   step i streams Q_i and dO_i:
//        S^T = K Q_i^T, dP^T = V dO_i^T;  dS^T -> TMEM;  dK += dS^T Q_i
//   DV : rows = keys.     resident K tile;       step i streams Q_i and dO_i:
//        S^T = K Q_i^T;  P^T (bf16) -> TMEM;  dV += P^T dO_i
// (dK and dV cannot share a pass at d_k = 256: two 128 x 256 fp32 accumulators are all 512 TMEM columns.)
// 192 threads: warp 0 TMA producer, warp 1 TMEM allocator + tcgen05.mma issuer, warps 2-5 thread = row: read the S / dP
// tiles from TMEM, recompute the probabilities, write dS (or P) back as the bf16 A operand of the accumulating MMA.
// Shared memory (224 KB at d_k = 256): 2 resident 64 KB tiles + two rings of 2 half tiles (16 KB) for the K-major
// operands of S and dP, freed as soon as those MMAs have read them, + ONE 32 KB tile for the MN-major operand of the
// accumulating MMA.  That tile is the same data as the S operand, delivered a second time (from L2): holding the first
// copy until dS exists would leave no room to prefetch the next steps, and the step would wait for a load every time.
// TMEM: S 2 x 64, dP 2 x 64, accumulator 256.
// What bounds it (measured, B=64 S=800: ~2.1 us per 64-column step in every pass, unchanged when the Z delivery or the
// row arithmetic is switched off): shared-memory bandwidth.  An N = 64 MMA re-reads its 4 KB A slice for 2 KB of B every
// 32 tensor-pipe clocks (192 B/clk against 128 B/clk); per step S + dP read 192 KB, the accumulating MMA 32 KB, and TMA
// writes 96 KB.  Wider steps (N = 128 halves the A re-reads) do not fit next to two resident 64 KB tiles and a 256-column
// accumulator; the way out is the CTA-pair form (cta_group::2: M = 256, each SM holds half of every streamed tile).
#include <cuda_bf16.h>
#include "common.cuh"
#include "tc_common.cuh"
#include "attn_bf16_common.cuh"
#include "../../include/bmhrl_b200.h"

namespace bmhrl {

using namespace tc;

constexpr int BW_BM = 128;  // rows per item
constexpr int BW_BN = 64;   // streamed columns per step
constexpr int BW_RES_BYTES = 64 * 1024;
constexpr int BW_P_BYTES = 16 * 1024;  // half of a streamed tile (64 rows x 128 columns), K-major operand of S / dP
constexpr int BW_Z_BYTES = 32 * 1024;  // a streamed tile (64 rows x 256 columns), MN-major operand of the accumulating MMA
constexpr int BW_SMEM = 2 * BW_RES_BYTES + 4 * BW_P_BYTES + BW_Z_BYTES + 1024 + 256;
enum { BW_DQ = 0, BW_DK = 1, BW_DV = 2 };

struct AttnBwdParams {
    __nv_bfloat16* out; long long ldo, bso;  // gradient of the row-side tensor, (B, rows, H*dk)
    const float* lse;                        // (B,H,Sq) from the forward
    const float* delta;                      // (B,H,Sq) rowsum(dO o O)
    int B, H, Sq, Sk, dk;
    int rows, cols;                          // DQ: Sq, Sk;  DK / DV: Sk, Sq
    int nrt, n_items;
    float scale;
    const uint32_t* mbits;                   // (B, mwords) key visibility bits, or NULL (no mask)
    int mwords;
};

// delta[b,h,q] = sum_d dO[b,q,h,d] * O[b,q,h,d]; one warp per (b, q, h)
__global__ void attn_bwd_delta_kernel(const __nv_bfloat16* __restrict__ O, long long ldo, long long bso,
                                      const __nv_bfloat16* __restrict__ dO, long long lddo, long long bsdo,
                                      float* __restrict__ delta, int B, int H, int Sq, int dk) {
    const long long wid = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (wid >= (long long)B * Sq * H) return;
    const int h = (int)(wid % H);
    const long long bq = wid / H;
    const int q = (int)(bq % Sq), b = (int)(bq / Sq);
    const __nv_bfloat16* o = O + (long long)b * bso + (long long)q * ldo + h * dk;
    const __nv_bfloat16* g = dO + (long long)b * bsdo + (long long)q * lddo + h * dk;
    float acc = 0.f;
    for (int d = lane * 2; d < dk; d += 64) {
        const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(o + d));
        const float2 c = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(g + d));
        acc += a.x * c.x + a.y * c.y;
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
    if (lane == 0) delta[((long long)b * H + h) * Sq + q] = acc;
}

template <int MODE>
__global__ void __launch_bounds__(192, 1) attn_bwd_bf16_kernel(const __grid_constant__ CUtensorMap tmRA,
                                                               const __grid_constant__ CUtensorMap tmRB,
                                                               const __grid_constant__ CUtensorMap tmKA,
                                                               const __grid_constant__ CUtensorMap tmKB,
                                                               const __grid_constant__ CUtensorMap tmZ,
                                                               const AttnBwdParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t resA = base, resB = base + BW_RES_BYTES;
    const uint32_t ringA = base + 2 * BW_RES_BYTES;   // parts of the streamed tile that pairs with resident A (S)
    const uint32_t ringB = ringA + 2 * BW_P_BYTES;    // parts of the streamed tile that pairs with resident B (dP)
    const uint32_t tileZ = ringB + 2 * BW_P_BYTES;    // streamed tile as the MN-major operand of the accumulating MMA
    const uint32_t bars = tileZ + BW_Z_BYTES;
    auto full_a = [&](int s) { return bars + 8u * s; };
    auto empty_a = [&](int s) { return bars + 8u * (2 + s); };
    auto full_b = [&](int s) { return bars + 8u * (4 + s); };
    auto empty_b = [&](int s) { return bars + 8u * (6 + s); };
    auto sdp_full = [&](int i) { return bars + 8u * (8 + i); };   // MMA -> row threads: S (and dP) tile complete
    auto ds_ready = [&](int i) { return bars + 8u * (10 + i); };  // row threads -> MMA: dS / P written
    const uint32_t r_full = bars + 8u * 12, r_empty = bars + 8u * 13;
    const uint32_t acc_full = bars + 8u * 14, acc_free = bars + 8u * 15;
    const uint32_t full_z = bars + 8u * 16, empty_z = bars + 8u * 17;
    const uint32_t tmem_slot = bars + 8u * 18;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nvc = p.dk >> 6;                    // 64-column chunks of a tile
    const int nyp = p.dk >= 128 ? 2 : 1;          // parts a K-major streamed tile is delivered in
    const int ypc = nvc / nyp;                    // 64-column chunks per part
    const int nt = (p.cols + BW_BN - 1) / BW_BN;  // steps per item
    constexpr bool TWO_RES = MODE != BW_DV;       // S and dP both computed (two resident operands)

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmRA);
        tma_prefetch_desc(&tmRB);
        tma_prefetch_desc(&tmKA);
        tma_prefetch_desc(&tmKB);
        tma_prefetch_desc(&tmZ);
        for (int s = 0; s < 2; ++s) {
            mbar_init(full_a(s), 1);
            mbar_init(empty_a(s), 1);
            mbar_init(full_b(s), 1);
            mbar_init(empty_b(s), 1);
            mbar_init(sdp_full(s), 1);
            mbar_init(ds_ready(s), 128);
        }
        mbar_init(r_full, 1);
        mbar_init(r_empty, 1);
        mbar_init(acc_full, 1);
        mbar_init(acc_free, 128);
        mbar_init(full_z, 1);
        mbar_init(empty_z, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    const uint32_t tmem_acc = tmem_base + 256;
    auto tmem_s = [&](int buf) { return tmem_base + (uint32_t)(buf * BW_BN); };
    auto tmem_dp = [&](int buf) { return tmem_base + 128u + (uint32_t)(buf * BW_BN); };

    if (warp == 0) {
        if (elect_one()) {  // ===== TMA producer: same order as the MMA issuer consumes -- parts(0) parts(1) Z(0) parts(2) Z(1) ...
            int ga = 0, gb = 0, gz = 0;  // parts / Z tiles loaded so far (all items)
            for (int i = 0, w = blockIdx.x; w < p.n_items; ++i, w += gridDim.x) {
                const int bh = w / p.nrt, rt = w - bh * p.nrt;
                const int b = bh / p.H, h = bh - b * p.H;
                const int col0 = h * p.dk, row0 = rt * BW_BM;
                mbar_wait(r_empty, ((uint32_t)i & 1u) ^ 1u);
                mbar_arrive_expect_tx(r_full, (uint32_t)((TWO_RES ? 2 : 1) * BW_BM * p.dk * 2));
                for (int c = 0; c < nvc; ++c) {
                    tma_load_4d(resA + c * 16384, &tmRA, r_full, col0 + c * 64, row0, 0, b);
                    if (TWO_RES) tma_load_4d(resB + c * 16384, &tmRB, r_full, col0 + c * 64, row0, 0, b);
                }
                auto load_parts = [&](int t) {
                    for (int y = 0; y < nyp; ++y, ++ga) {
                        const int s = ga & 1;
                        mbar_wait(empty_a(s), (((uint32_t)(ga >> 1)) & 1u) ^ 1u);
                        mbar_arrive_expect_tx(full_a(s), (uint32_t)(ypc * 8192));
                        for (int c = 0; c < ypc; ++c)
                            tma_load_4d(ringA + s * BW_P_BYTES + c * 8192, &tmKA, full_a(s), col0 + (y * ypc + c) * 64,
                                        t * BW_BN, 0, b);
                    }
                    if (TWO_RES) {
                        for (int y = 0; y < nyp; ++y, ++gb) {
                            const int s = gb & 1;
                            mbar_wait(empty_b(s), (((uint32_t)(gb >> 1)) & 1u) ^ 1u);
                            mbar_arrive_expect_tx(full_b(s), (uint32_t)(ypc * 8192));
                            for (int c = 0; c < ypc; ++c)
                                tma_load_4d(ringB + s * BW_P_BYTES + c * 8192, &tmKB, full_b(s), col0 + (y * ypc + c) * 64,
                                            t * BW_BN, 0, b);
                        }
                    }
                };
                auto load_z = [&](int t) {
                    mbar_wait(empty_z, ((uint32_t)gz & 1u) ^ 1u);
                    mbar_arrive_expect_tx(full_z, (uint32_t)(nvc * 8192));
                    for (int c = 0; c < nvc; ++c) tma_load_4d(tileZ + c * 8192, &tmZ, full_z, col0 + c * 64, t * BW_BN, 0, b);
                    ++gz;
                };
                load_parts(0);
                if (nt > 1) load_parts(1);
                for (int t = 0; t < nt; ++t) {
                    load_z(t);
                    if (t + 2 < nt) load_parts(t + 2);
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {  // ===== MMA issuer
            const uint32_t idesc_s = make_idesc(1, BW_BM, BW_BN, 0, 0);
            const uint32_t idesc_f = make_idesc(1, BW_BM, p.dk, 0, 1);
            const uint64_t ra_desc = make_smem_desc(resA, 16, 1024);     // K-major, 128-byte rows, 16 KB per 64 columns
            const uint64_t rb_desc = make_smem_desc(resB, 16, 1024);
            const uint64_t ka_desc = make_smem_desc(ringA, 16, 1024);    // parts: K-major B (N = 64 rows), 8 KB per 64 columns
            const uint64_t kb_desc = make_smem_desc(ringB, 16, 1024);
            const uint64_t z_desc = make_smem_desc(tileZ, 8192, 1024);   // MN-major B (K = 64 rows, N = d_k)
            int ga = 0, gb = 0, ns = 0, nf = 0;  // parts consumed; S(+dP) tiles issued; accumulating MMAs issued (all items)
            for (int i = 0, w = blockIdx.x; w < p.n_items; ++i, w += gridDim.x) {
                mbar_wait(r_full, (uint32_t)i & 1u);
                tc_fence_after();
                // contraction over d_k of a resident operand (A) with the parts of a streamed tile (B), into tmem_d
                auto parts_mma = [&](uint64_t a_desc, uint64_t k_desc, bool ring_b, int& g, uint32_t tmem_d) {
                    for (int y = 0; y < nyp; ++y, ++g) {
                        const int s = g & 1;
                        mbar_wait(ring_b ? full_b(s) : full_a(s), ((uint32_t)(g >> 1)) & 1u);
                        tc_fence_after();
                        for (int c = 0; c < ypc; ++c) {
                            const uint64_t a0 = desc_advance(a_desc, (y * ypc + c) * 16384);
                            const uint64_t b0 = desc_advance(k_desc, s * BW_P_BYTES + c * 8192);
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                mma_bf16_ss(tmem_d, desc_advance(a0, k * 32), desc_advance(b0, k * 32), idesc_s,
                                            (y > 0 || c > 0 || k > 0) ? 1u : 0u);
                        }
                        tc_commit(ring_b ? empty_b(s) : empty_a(s));
                    }
                };
                auto sd_op = [&](int t) {
                    const int buf = ns & 1;
                    parts_mma(ra_desc, ka_desc, false, ga, tmem_s(buf));              // S = Q K_j^T   /  S^T = K Q_i^T
                    if (TWO_RES) parts_mma(rb_desc, kb_desc, true, gb, tmem_dp(buf));  // dP = dO V_j^T /  dP^T = V dO_i^T
                    tc_commit(sdp_full(buf));
                    if (t == nt - 1) tc_commit(r_empty);  // the resident tiles have been read for the last time
                    ++ns;
                };
                auto f_op = [&](int t) {  // accumulator += dS (or P), from TMEM, times the Z tile
                    const int buf = nf & 1;
                    mbar_wait(ds_ready(buf), ((uint32_t)(nf >> 1)) & 1u);
                    if (t == 0 && i > 0) mbar_wait(acc_free, ((uint32_t)(i - 1)) & 1u);
                    mbar_wait(full_z, (uint32_t)nf & 1u);
                    tc_fence_after();
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk)
                        mma_bf16_ts(tmem_acc, tmem_s(buf) + kk * 8, desc_advance(z_desc, kk * 2048), idesc_f,
                                    (t > 0 || kk > 0) ? 1u : 0u);
                    tc_commit(empty_z);
                    if (t == nt - 1) tc_commit(acc_full);
                    ++nf;
                };
                sd_op(0);
                if (nt > 1) sd_op(1);
                for (int t = 0; t < nt; ++t) {
                    f_op(t);
                    if (t + 2 < nt) sd_op(t + 2);
                }
            }
        }
    } else {
        // ===== row threads: TMEM lane quarter = warp % 4, thread = row of the item (query for DQ, key for DK / DV)
        const int qtr = warp & 3;
        const uint32_t lane_off = (uint32_t)(qtr * 32) << 16;
        constexpr float LOG2E = 1.4426950408889634f;
        const float sl2 = p.scale * LOG2E;
        const float inv_sk = 1.f / (float)p.Sk;
        int n = 0;  // steps consumed (all items)
        for (int i = 0, w = blockIdx.x; w < p.n_items; ++i, w += gridDim.x) {
            const int bh = w / p.nrt, rt = w - bh * p.nrt;
            const int b = bh / p.H, h = bh - b * p.H;
            const int row = rt * BW_BM + qtr * 32 + lane;
            const bool row_ok = row < p.rows;
            const float* lse_bh = p.lse + (long long)bh * p.Sq;
            const float* del_bh = p.delta + (long long)bh * p.Sq;
            const uint32_t* mrow = p.mbits ? p.mbits + (long long)b * p.mwords : nullptr;
            // a batch row without any visible key: the forward's probabilities are uniform over all Sk keys (every score
            // was replaced by the same constant), and no gradient reaches the scores
            bool uni = false;
            if (mrow) {
                uint32_t any = 0u;
                for (int wd = lane; wd < p.mwords; wd += 32) any |= __ldg(mrow + wd);
                uni = !__any_sync(0xffffffffu, any != 0u);
            }
            float lse2_r = 0.f, del_r = 0.f;  // DQ: statistics of my query row
            bool key_vis = true;              // DK / DV: my key takes part (in range and not masked)
            if (MODE == BW_DQ) {
                if (row_ok) {
                    lse2_r = __ldg(lse_bh + row) * LOG2E;
                    del_r = __ldg(del_bh + row);
                }
            } else {
                key_vis = row_ok && (!mrow || ((__ldg(mrow + (row >> 5)) >> (row & 31)) & 1u) != 0u);
            }
            for (int t = 0; t < nt; ++t, ++n) {
                const int buf = n & 1;
                const int c0 = t * BW_BN;  // first streamed column (key for DQ, query for DK / DV)
                uint32_t vis[2] = {0xffffffffu, 0xffffffffu};   // DQ: visibility of the 64 keys of this step
                float lse2_c[2] = {0.f, 0.f}, del_c[2] = {0.f, 0.f};  // DK / DV: statistics of queries c0+lane, c0+32+lane
                if (MODE == BW_DQ) {
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        const int rem = p.Sk - c0 - hh * 32;
                        uint32_t v = rem >= 32 ? 0xffffffffu : (rem <= 0 ? 0u : ((1u << rem) - 1u));
                        if (mrow) v &= __ldg(mrow + t * 2 + hh);
                        vis[hh] = uni ? 0u : v;
                    }
                } else {
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        const int q = c0 + hh * 32 + lane;
                        const bool q_ok = q < p.Sq;
                        lse2_c[hh] = q_ok ? __ldg(lse_bh + q) * LOG2E : INFINITY;  // out of range: probability 0
                        del_c[hh] = q_ok ? __ldg(del_bh + q) : 0.f;
                    }
                }
                mbar_wait(sdp_full(buf), ((uint32_t)(n >> 1)) & 1u);
                tc_fence_after();
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    uint32_t s[32], dp[32], out[16];
                    tmem_ld_32x32_b(tmem_s(buf) + lane_off + hh * 32, s);
                    if (MODE != BW_DV) tmem_ld_32x32_b(tmem_dp(buf) + lane_off + hh * 32, dp);
                    tmem_ld_wait();
                    float g[32];
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        const float sc = __uint_as_float(s[c]);
                        if (MODE == BW_DQ) {
                            const float pr = ex2_approx(fmaf(sc, sl2, -lse2_r));
                            const float ds = pr * (__uint_as_float(dp[c]) - del_r);
                            g[c] = ((vis[hh] >> c) & 1u) ? ds : 0.f;
                        } else {
                            const float l2 = __shfl_sync(0xffffffffu, lse2_c[hh], c);
                            const float pr = ex2_approx(fmaf(sc, sl2, -l2));
                            if (MODE == BW_DK) {
                                const float dl = __shfl_sync(0xffffffffu, del_c[hh], c);
                                const float ds = pr * (__uint_as_float(dp[c]) - dl);
                                g[c] = (key_vis && !uni) ? ds : 0.f;
                            } else {
                                const float pu = (l2 == INFINITY) ? 0.f : inv_sk;  // uniform row; 0 for queries out of range
                                g[c] = !row_ok ? 0.f : (uni ? pu : (key_vis ? pr : 0.f));
                            }
                        }
                    }
#pragma unroll
                    for (int c = 0; c < 16; ++c) out[c] = pack_bf16x2(g[2 * c], g[2 * c + 1]);
                    tmem_st_32x32_x16(tmem_s(buf) + lane_off + hh * 16, out);  // S columns 32.. are read before 16.. is written
                }
                tmem_st_wait_b();
                tc_fence_before();
                mbar_arrive(ds_ready(buf));
            }
            // ---- epilogue: accumulator (TMEM) -> * scale -> bf16 -> global, head-merged layout
            mbar_wait(acc_full, (uint32_t)i & 1u);
            tc_fence_after();
            const float osc = MODE == BW_DV ? 1.f : p.scale;
            __nv_bfloat16* orow = p.out + (long long)b * p.bso + (long long)row * p.ldo + h * p.dk;
            for (int c = 0; c < (p.dk >> 5); ++c) {
                uint32_t a[32];
                tmem_ld_32x32_b(tmem_acc + lane_off + c * 32, a);
                tmem_ld_wait();
                if (row_ok) {
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        uint4 v4;
                        v4.x = pack_bf16x2(__uint_as_float(a[u * 8]) * osc, __uint_as_float(a[u * 8 + 1]) * osc);
                        v4.y = pack_bf16x2(__uint_as_float(a[u * 8 + 2]) * osc, __uint_as_float(a[u * 8 + 3]) * osc);
                        v4.z = pack_bf16x2(__uint_as_float(a[u * 8 + 4]) * osc, __uint_as_float(a[u * 8 + 5]) * osc);
                        v4.w = pack_bf16x2(__uint_as_float(a[u * 8 + 6]) * osc, __uint_as_float(a[u * 8 + 7]) * osc);
                        *reinterpret_cast<uint4*>(orow + c * 32 + u * 8) = v4;
                    }
                }
            }
            tc_fence_before();
            mbar_arrive(acc_free);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace bmhrl

static int g_attn_bwd_passes = 7;
/* profiling hook: bit mask of the gradient passes bmhrl_attn_bwd_bf16 launches (1 = dQ, 2 = dK, 4 = dV; default 7) */
extern "C" int bmhrl_attn_bwd_bf16_tuning(int passes) {
    BMHRL_CHECK_ARG(passes >= 1 && passes <= 7, "bmhrl_attn_bwd_bf16_tuning: pass mask 1..7");
    g_attn_bwd_passes = passes;
    return BMHRL_OK;
}

extern "C" long long bmhrl_attn_bwd_bf16_ws_bytes(int B, int H, int Sq, int Sk) {
    const long long nd = ((long long)B * H * Sq + 3) & ~3ll;
    return 4ll * (nd + ((B + 3) & ~3) + (long long)B * 4 * ((Sk + 127) / 128));
}

extern "C" int bmhrl_attn_bwd_bf16(const void* Q, long long ldq, long long bsq, const void* K, long long ldk,
                                   long long bsk, const void* V, long long ldv, long long bsv, const void* O,
                                   long long ldo, long long bso, const void* dO, long long lddo, long long bsdo,
                                   const float* lse, const unsigned char* mask, int mask_kind, void* dQ,
                                   long long lddq, long long bsdq, void* dK, long long lddk, long long bsdk, void* dV,
                                   long long lddv, long long bsdv, int B, int H, int Sq, int Sk, int dk, void* ws,
                                   void* stream) {
    using namespace bmhrl;
    using namespace bmhrl::tc;
    BMHRL_CHECK_ARG(Q && K && V && O && dO && lse && dQ && dK && dV && ws, "attn_bwd_bf16: null tensor / workspace");
    BMHRL_CHECK_ARG(B > 0 && H > 0 && Sq > 0 && Sk > 0, "attn_bwd_bf16: empty shape");
    BMHRL_CHECK_ARG(mask_kind >= 0 && mask_kind <= 2 && (mask_kind == 0 || mask), "attn_bwd_bf16: bad mask");
    if (mask_kind == 2) {
        bmhrl_set_error("attn_bwd_bf16: only key-padding (B,1,Sk) masks are supported (the caption attentions train on "
                        "the fp32 / tf32 kernels)");
        return BMHRL_ERR_UNSUPPORTED;
    }
    if (dk != 64 && dk != 128 && dk != 256) {
        bmhrl_set_error("attn_bwd_bf16: d_k must be 64, 128 or 256");
        return BMHRL_ERR_UNSUPPORTED;
    }
    auto ok8 = [](long long v) { return v > 0 && (v & 7) == 0; };
    const long long lds[8] = {ldq, ldk, ldv, ldo, lddo, lddq, lddk, lddv};
    const long long bss[8] = {bsq, bsk, bsv, bso, bsdo, bsdq, bsdk, bsdv};
    const void* ptrs[8] = {Q, K, V, O, dO, dQ, dK, dV};
    for (int i = 0; i < 8; ++i) {
        if (!ok8(lds[i]) || (B > 1 && !ok8(bss[i])) || !aligned16(ptrs[i])) {
            bmhrl_set_error("attn_bwd_bf16: tensors must be 16-byte aligned with pitches that are multiples of 8 elements");
            return BMHRL_ERR_UNSUPPORTED;
        }
    }
    if (get_encode_fn() == nullptr) {
        bmhrl_set_error("attn_bwd_bf16: cuTensorMapEncodeTiled unavailable");
        return BMHRL_ERR_CUDA;
    }
    const CUtensorMapDataType BF = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    const unsigned long long D = (unsigned long long)H * dk;
    // resident (128-row box) and streamed (64-row box) views of the four inputs
    CUtensorMap rQ, rK, rV, rdO, sQ, sK, sV, sdO;
    const bool ok = encode_4d(&rQ, BF, 2, Q, D, Sq, ldq, 1, 0, B, bsq, 64, BW_BM) &&
                    encode_4d(&rdO, BF, 2, dO, D, Sq, lddo, 1, 0, B, bsdo, 64, BW_BM) &&
                    encode_4d(&rK, BF, 2, K, D, Sk, ldk, 1, 0, B, bsk, 64, BW_BM) &&
                    encode_4d(&rV, BF, 2, V, D, Sk, ldv, 1, 0, B, bsv, 64, BW_BM) &&
                    encode_4d(&sQ, BF, 2, Q, D, Sq, ldq, 1, 0, B, bsq, 64, BW_BN) &&
                    encode_4d(&sdO, BF, 2, dO, D, Sq, lddo, 1, 0, B, bsdo, 64, BW_BN) &&
                    encode_4d(&sK, BF, 2, K, D, Sk, ldk, 1, 0, B, bsk, 64, BW_BN) &&
                    encode_4d(&sV, BF, 2, V, D, Sk, ldv, 1, 0, B, bsv, 64, BW_BN);
    if (!ok) {
        bmhrl_set_error("attn_bwd_bf16: cuTensorMapEncodeTiled failed");
        return BMHRL_ERR_CUDA;
    }
    static DeviceOnce attr;
    const int n_sm = device_sm_count();
    if (needs_setup(attr)) {
        const cudaFuncAttribute A = cudaFuncAttributeMaxDynamicSharedMemorySize;
        if (cudaFuncSetAttribute(attn_bwd_bf16_kernel<BW_DQ>, A, BW_SMEM) != cudaSuccess ||
            cudaFuncSetAttribute(attn_bwd_bf16_kernel<BW_DK>, A, BW_SMEM) != cudaSuccess ||
            cudaFuncSetAttribute(attn_bwd_bf16_kernel<BW_DV>, A, BW_SMEM) != cudaSuccess) {
            bmhrl_set_error("attn_bwd_bf16: cannot raise the shared-memory limit");
            return BMHRL_ERR_CUDA;
        }
        mark_done(attr);
    }
    cudaStream_t st = (cudaStream_t)stream;
    float* delta = (float*)ws;
    const long long nd = ((long long)B * H * Sq + 3) & ~3ll;
    int* sk_eff = (int*)ws + nd;
    uint32_t* mbits = (uint32_t*)ws + nd + ((B + 3) & ~3);
    {
        const long long warps = (long long)B * Sq * H;
        const int blocks = (int)((warps * 32 + 255) / 256);
        attn_bwd_delta_kernel<<<blocks, 256, 0, st>>>((const __nv_bfloat16*)O, ldo, bso, (const __nv_bfloat16*)dO, lddo,
                                                      bsdo, delta, B, H, Sq, dk);
        BMHRL_POST_LAUNCH("attn_bwd_delta");
    }
    AttnBwdParams p{};
    p.lse = lse; p.delta = delta; p.B = B; p.H = H; p.Sq = Sq; p.Sk = Sk; p.dk = dk;
    p.scale = 1.0f / sqrtf((float)dk);
    if (mask_kind == 1) {
        p.mwords = 4 * ((Sk + 127) / 128);
        attn_mask_pack(mask, B, Sk, p.mwords, sk_eff, mbits, st);
        BMHRL_POST_LAUNCH("attn_mask_pack");
        p.mbits = mbits;
    }
    auto launch = [&](int mode) {
        AttnBwdParams a = p;
        a.rows = mode == BW_DQ ? Sq : Sk;
        a.cols = mode == BW_DQ ? Sk : Sq;
        a.nrt = (a.rows + BW_BM - 1) / BW_BM;
        a.n_items = B * H * a.nrt;
        const int grid = a.n_items < n_sm ? a.n_items : n_sm;
        if (mode == BW_DQ) {
            a.out = (__nv_bfloat16*)dQ; a.ldo = lddq; a.bso = bsdq;
            attn_bwd_bf16_kernel<BW_DQ><<<grid, 192, BW_SMEM, st>>>(rQ, rdO, sK, sV, sK, a);
        } else if (mode == BW_DK) {
            a.out = (__nv_bfloat16*)dK; a.ldo = lddk; a.bso = bsdk;
            attn_bwd_bf16_kernel<BW_DK><<<grid, 192, BW_SMEM, st>>>(rK, rV, sQ, sdO, sQ, a);
        } else {
            a.out = (__nv_bfloat16*)dV; a.ldo = lddv; a.bso = bsdv;
            attn_bwd_bf16_kernel<BW_DV><<<grid, 192, BW_SMEM, st>>>(rK, rK, sQ, sQ, sdO, a);
        }
    };
    // A gradient whose pass is switched off by the profiling hook comes back as zeros, never as uninitialised memory.
    auto zero_out = [&](void* g, long long ld, long long bs, int rows) -> bool {
        const size_t width = sizeof(__nv_bfloat16) * (size_t)H * dk;
        for (int b = 0; b < B; ++b)
            if (cudaMemset2DAsync((char*)g + sizeof(__nv_bfloat16) * (size_t)b * bs, sizeof(__nv_bfloat16) * (size_t)ld, 0,
                                  width, rows, st) != cudaSuccess)
                return false;
        return true;
    };
    if (g_attn_bwd_passes & 1) {
        launch(BW_DQ);
        BMHRL_POST_LAUNCH("attn_bwd_bf16 dQ");
    } else if (!zero_out(dQ, lddq, bsdq, Sq)) return BMHRL_ERR_CUDA;
    if (g_attn_bwd_passes & 2) {
        launch(BW_DK);
        BMHRL_POST_LAUNCH("attn_bwd_bf16 dK");
    } else if (!zero_out(dK, lddk, bsdk, Sk)) return BMHRL_ERR_CUDA;
    if (g_attn_bwd_passes & 4) {
        launch(BW_DV);
        BMHRL_POST_LAUNCH("attn_bwd_bf16 dV");
    } else if (!zero_out(dV, lddv, bsdv, Sk)) return BMHRL_ERR_CUDA;
    return BMHRL_OK;
}
