// Grouped bf16 tensor-core GEMM for sm_100a: TMA-staged bf16 operands, tcgen05.mma (kind::f16, bf16 x bf16 -> fp32 in
// TMEM), fused linear-layer epilogue, up to GB_MAXP independent problems per launch.
//
// Replaces, in the library's bf16 math mode, the same call sites as gemm_tc.cu: the projections
// model/multihead_attention.py:70-72,89, the FFN model/blocks.py:175-187, Manager.linear model/bm_hrl_agent.py:439-440,
// the vocabulary projection model/bm_hrl_agent.py:463-466 and the dX / dW gradients autograd derives from them.
//
// Why bf16 and why these tiles: the fp32/tf32 kernel is bound by L2 -> SM delivery (B200 delivers ~6300 B/clk chip-wide;
// profiles/r01_ncu_full_gemm_tf32.txt: 83.9 MB moved for 8.1 MB of operands at 128x32 tiles).  Here operands are half
// the bytes and tiles are 128 x 128 (or 128 x 64), so a tile re-reads A N/128 times instead of N/32; when that leaves
// too few tiles to fill the machine, a thread-block cluster of 2 / 4 CTAs splits K of each tile and reduce-scatters the
// partial accumulators through distributed shared memory (no atomics: results are deterministic).
//
//   v(m,n)  = alpha * sum_k A(m,k) B(n,k) + bias[n];  relu;  gate (v = gate(m,n) > 0 ? v * gate_scale : 0);
//             Philox dropout (element index m*N + n);  + resid(m,n);  + beta * C32(m,n)
//   C32(m,n) = v (fp32, optional, optionally split into row segments),  C16(m,n) = bf16(v) (optional)
//
// A is K-major (row m contiguous in k: activations, dZ for dX) or MN-major (row k contiguous in m: dZ^T for dW);
// B is K-major (nn.Linear weight (N,K)) or MN-major (weight in dX = dZ W, X in dW = dZ^T X).  Row pitches are multiples
// of 8 elements (16 bytes); M / N / K tails are zero-filled by TMA.
//
// CTA = 192 threads: warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer (one elected thread), warps 2-5 epilogue
// (tcgen05.ld -> registers -> global, thread = output row).  Tile 128 x BN x 64, runtime-deep mbarrier ring.
#include <cuda_bf16.h>
#include <stdio.h>
#include <stdlib.h>
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/bmhrl_b200.h"

namespace bmhrl {

using namespace tc;

constexpr int GB_BM = 128, GB_BK = 64, GB_MAXP = 4;
constexpr int GB_A_BYTES = GB_BM * 128;  // 16 KB per stage

struct alignas(64) GbProblem {
    CUtensorMap tmA, tmB;
    float* C32; float* seg1; float* seg2;
    __nv_bfloat16* C16;
    const float* bias; const float* resid;
    const __nv_bfloat16* gate;
    const unsigned long long* rng;
    long long ldc32, ldc16, ldr, ldg;
    int M, N, K, num_kb, kb_per_split;
    int a_mn, b_mn, act, seg_rows, b_static;
    int tiles_n, tile_start;   // tiles of this problem are blockIdx.x in [tile_start, tile_start + tiles_m * tiles_n)
    int vec32, vec16, vecr, vecg;  // 128-bit access allowed on C32 rows / 64-bit on C16 / resid / gate
    float alpha, beta, gate_scale, drop_p;
    unsigned int salt;
    // MODE 1 / 2 (vocabulary projection fused with the log-softmax / KL statistics, see the end of this file)
    const long long* v_trg; const long long* v_tok; const float* v_q; float* v_part;
    const float* v_stats; const float* v_dstat;
    int v_pad, v_smode;
};

struct GbParams {
    GbProblem p[GB_MAXP];
    int n, cluster, nstages;
    unsigned long long* trace;   // diagnosis (bmhrl_gemm_bf16_trace): CTA 0 stamps clock64() at its hand-over points; else NULL
};

// slots: 0 entry, 1 set-up done, 2 producer past the grid dependency, 3 first ring requested, 4 first stage landed,
// 5 last MMA committed, 6 accumulator visible to the epilogue, 7 epilogue stores issued, 8 exit, 9 globaltimer at entry,
// 10 globaltimer at exit, 11 first accumulator chunk in registers, 12 first pass staged, 13 first pass stored, 16 + i: stage of k-block i landed (i < 48)
__device__ __forceinline__ void gb_stamp(const GbParams& P, int slot) {
    if (P.trace && blockIdx.x == 0 && blockIdx.z == 0) P.trace[slot] = (unsigned long long)clock64();
}
__device__ __forceinline__ void gb_stamp_ns(const GbParams& P, int slot) {
    if (P.trace && blockIdx.x == 0 && blockIdx.z == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        P.trace[slot] = t;
    }
}

__device__ __forceinline__ uint2 ld_bf16x4(const __nv_bfloat16* p) { return *reinterpret_cast<const uint2*>(p); }
__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack2_bf16(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}

template <int BN, int MODE>
__global__ void __launch_bounds__(192, (BN == 64 ? 3 : 1)) gemm_bf16_tc_kernel(const __grid_constant__ GbParams P) {
    constexpr int B_BYTES = BN * 128;
    constexpr int STAGE_BYTES = GB_A_BYTES + B_BYTES;
    constexpr int TMEM_COLS = BN < 32 ? 32 : BN;
    extern __shared__ uint8_t smem_raw[];
    const int STAGES = P.nstages;
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t recv = base + STAGES * STAGE_BYTES;  // cluster mode: [src rank][col/4][row] float4 partials
    const uint32_t bars = recv + (P.cluster > 1 ? GB_BM * BN * 4 : 0);
    auto full_bar = [&](int s) { return bars + 8u * s; };
    auto empty_bar = [&](int s) { return bars + 8u * (STAGES + s); };
    const uint32_t tmem_full_bar = bars + 8u * (2 * STAGES);
    const uint32_t tmem_slot = bars + 8u * (2 * STAGES + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // which problem / tile
    int pi = 0;
#pragma unroll
    for (int i = 1; i < GB_MAXP; ++i)
        if (i < P.n && (int)blockIdx.x >= P.p[i].tile_start) pi = i;
    const GbProblem& p = P.p[pi];   // a DYNAMICALLY indexed parameter slot: every p.x is an indexed constant load (slow,
                                    // and serial when it feeds a branch); what the hot paths need is copied into
                                    // registers below through a switch, i.e. with static indices
    int c_tile_start = 0, c_tiles_n = 1, c_kb_per_split = 0, c_num_kb = 0, c_a_mn = 0, c_b_mn = 0, c_b_static = 0;
    const CUtensorMap* c_tmA = nullptr;
    const CUtensorMap* c_tmB = nullptr;
#define GB_COMMON(i)                                                                                       \
    case i:                                                                                                \
        c_tile_start = P.p[i].tile_start; c_tiles_n = P.p[i].tiles_n; c_kb_per_split = P.p[i].kb_per_split; \
        c_num_kb = P.p[i].num_kb; c_a_mn = P.p[i].a_mn; c_b_mn = P.p[i].b_mn; c_b_static = P.p[i].b_static; \
        c_tmA = &P.p[i].tmA; c_tmB = &P.p[i].tmB;                                                           \
        break;
    switch (pi) { GB_COMMON(0) GB_COMMON(1) GB_COMMON(2) GB_COMMON(3) }
#undef GB_COMMON
    const int t = blockIdx.x - c_tile_start;
    const int tm = t / c_tiles_n, tn = t - tm * c_tiles_n;
    const int m0 = tm * GB_BM, n0 = tn * BN;
    const int kb0 = blockIdx.z * c_kb_per_split;
    const int kb1 = min(c_num_kb, kb0 + c_kb_per_split);
    const bool has_k = kb1 > kb0;   // a split without k-blocks contributes zeros (cluster mode only)

    if (threadIdx.x == 0) { gb_stamp(P, 0); gb_stamp_ns(P, 9); }
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(c_tmA);
        tma_prefetch_desc(c_tmB);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        mbar_init(tmem_full_bar, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (P.cluster > 1) cluster_sync_all();  // every CTA of the cluster is running before anyone writes into its smem
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    if (threadIdx.x == 0) gb_stamp(P, 1);

    if (warp == 0) {
        if (elect_one()) {  // ===== TMA producer
            const int a_mn = c_a_mn, b_mn = c_b_mn, b_static = c_b_static;
            auto load_a = [&](int s, int kb) {
                const uint32_t sA = base + s * STAGE_BYTES;
                const int k = kb * GB_BK;
                if (!a_mn) {
                    tma_load_2d(sA, c_tmA, full_bar(s), k, m0);            // box {64 k, 128 m}
                } else {
#pragma unroll
                    for (int c = 0; c < GB_BM / 64; ++c)                    // box {64 m, 64 k} per 128-byte MN chunk
                        tma_load_2d(sA + c * 8192, c_tmA, full_bar(s), m0 + c * 64, k);
                }
            };
            auto load_b = [&](int s, int kb) {
                const uint32_t sB = base + s * STAGE_BYTES + GB_A_BYTES;
                const int k = kb * GB_BK;
                if (!b_mn) {
                    tma_load_2d(sB, c_tmB, full_bar(s), k, n0);            // box {64 k, BN n}
                } else {
#pragma unroll
                    for (int c = 0; c < BN / 64; ++c)
                        tma_load_2d(sB + c * 8192, c_tmB, full_bar(s), n0 + c * 64, k);
                }
            };
            // the first ring-full of stages: a static B (weights) is requested BEFORE the grid dependency is resolved --
            // under programmatic dependent launch these loads run while the previous kernel of the stream is finishing
            const int npre = min(STAGES, kb1 - kb0);
            for (int i = 0; i < npre; ++i) {
                mbar_arrive_expect_tx(full_bar(i), STAGE_BYTES);
                if (b_static) load_b(i, kb0 + i);
            }
            griddep_wait();
            griddep_launch_dependents();
            gb_stamp(P, 2);
            for (int i = 0; i < npre; ++i) {
                if (!b_static) load_b(i, kb0 + i);
                load_a(i, kb0 + i);
            }
            gb_stamp(P, 3);
            int s = npre == STAGES ? 0 : npre;
            uint32_t ph = npre == STAGES ? 1u : 0u;
            for (int kb = kb0 + npre; kb < kb1; ++kb) {
                mbar_wait(empty_bar(s), ph ^ 1u);
                mbar_arrive_expect_tx(full_bar(s), STAGE_BYTES);
                load_a(s, kb);
                load_b(s, kb);
                if (++s == STAGES) { s = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {  // ===== MMA issuer
            const uint32_t idesc = make_idesc(1, GB_BM, BN, c_a_mn, c_b_mn);
            const uint64_t ad0 = c_a_mn ? make_smem_desc(base, 8192, 1024) : make_smem_desc(base, 16, 1024);
            const uint64_t bd0 = c_b_mn ? make_smem_desc(base + GB_A_BYTES, 8192, 1024)
                                        : make_smem_desc(base + GB_A_BYTES, 16, 1024);
            const uint32_t a_step = c_a_mn ? 2048u : 32u, b_step = c_b_mn ? 2048u : 32u;
            int s = 0;
            uint32_t ph = 0;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(full_bar(s), ph);
                tc_fence_after();
                if (P.trace) {
                    if (kb == kb0) gb_stamp(P, 4);
                    if (kb - kb0 < 48) gb_stamp(P, 16 + kb - kb0);
                }
                const uint64_t ad = desc_advance(ad0, s * STAGE_BYTES), bd = desc_advance(bd0, s * STAGE_BYTES);
#pragma unroll
                for (int k = 0; k < GB_BK / 16; ++k)  // UMMA_K = 16 for bf16 (32 bytes of K)
                    mma_bf16_ss(tmem_base, desc_advance(ad, k * a_step), desc_advance(bd, k * b_step), idesc,
                                (kb > kb0 || k > 0) ? 1u : 0u);
                tc_commit(empty_bar(s));  // frees the smem slot once these MMAs have read it
                if (++s == STAGES) { s = 0; ph ^= 1u; }
            }
            tc_commit(tmem_full_bar);  // accumulator complete (arrives at once when nothing was issued)
            gb_stamp(P, 5);
        }
    }
    // ===== epilogue warps 2..5: TMEM lane quarter = warp % 4, thread = output row
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int m = m0 + row;
    const uint32_t lane_taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    // every field of the problem that the epilogue loops touch, in registers BEFORE the wait for the accumulator (left as
    // p.x inside the loop bodies they formed a serial chain of ~20 indexed constant loads and branches per 4 output
    // columns: 4-6 us per tile, longer than the main loop -- measured with bmhrl_gemm_bf16_trace)
    int e_M = 0, e_N = 0, e_act = 0, e_seg_rows = 0, e_v_pad = 0, e_vec32 = 0, e_vec16 = 0, e_vecr = 0, e_vecg = 0;
    float e_alpha = 1.f, e_beta = 0.f, e_gate_scale = 1.f, e_drop_p = 0.f;
    unsigned int e_salt = 0;
    const float* e_bias = nullptr;
    const float* e_resid = nullptr;
    const __nv_bfloat16* e_gate = nullptr;
    const unsigned long long* e_rng = nullptr;
    float* e_C32 = nullptr;
    float* e_seg1 = nullptr;
    float* e_seg2 = nullptr;
    __nv_bfloat16* e_C16 = nullptr;
    long long e_ldc32 = 0, e_ldc16 = 0, e_ldr = 0, e_ldg = 0;
#define GB_EPI(i)                                                                                                       \
    case i:                                                                                                             \
        e_M = P.p[i].M; e_N = P.p[i].N; e_act = P.p[i].act; e_seg_rows = P.p[i].seg_rows; e_v_pad = P.p[i].v_pad;       \
        e_vec32 = P.p[i].vec32; e_vec16 = P.p[i].vec16; e_vecr = P.p[i].vecr; e_vecg = P.p[i].vecg;                     \
        e_alpha = P.p[i].alpha; e_beta = P.p[i].beta; e_gate_scale = P.p[i].gate_scale; e_drop_p = P.p[i].drop_p;       \
        e_salt = P.p[i].salt; e_bias = P.p[i].bias; e_resid = P.p[i].resid; e_gate = P.p[i].gate; e_rng = P.p[i].rng;   \
        e_C32 = P.p[i].C32; e_seg1 = P.p[i].seg1; e_seg2 = P.p[i].seg2; e_C16 = P.p[i].C16;                             \
        e_ldc32 = P.p[i].ldc32; e_ldc16 = P.p[i].ldc16; e_ldr = P.p[i].ldr; e_ldg = P.p[i].ldg;                         \
        break;
    if (warp >= 2) {
        switch (pi) { GB_EPI(0) GB_EPI(1) GB_EPI(2) GB_EPI(3) }
    }
#undef GB_EPI
    if (warp >= 2) griddep_wait();   // the epilogue reads bias / residual / gate / C: nothing before the dependency resolves
    const int csz = P.cluster > 1 ? P.cluster : 1;
    const int W = BN / csz;  // columns of the tile this CTA finishes (cluster mode)
    const uint32_t rank = P.cluster > 1 ? cluster_ctarank() : 0u;
    if (warp >= 2) {
        mbar_wait(tmem_full_bar, 0);
        tc_fence_after();
        if (threadIdx.x == 64) gb_stamp(P, 6);
        if (P.cluster > 1) {
            // reduce-scatter, send side: slice r of my partial accumulator goes to CTA r's receive buffer, slot `rank`
            for (uint32_t r = 0; r < (uint32_t)csz; ++r) {
                if (r == rank) continue;
                const uint32_t rbase = mapa_shared(recv, r) + (rank * (uint32_t)(W / 4)) * (GB_BM * 16u);
#pragma unroll 1
                for (int c0 = 0; c0 < W; c0 += 32) {
                    uint32_t v[32];
                    if (has_k) {
                        tmem_ld_32x32(lane_taddr + (uint32_t)(r * W + c0), v);
                        tmem_ld_wait();
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) v[i] = 0u;
                    }
#pragma unroll
                    for (int i4 = 0; i4 < 8; ++i4)
                        st_shared_cluster_v4(rbase + ((uint32_t)(c0 / 4 + i4) * GB_BM + row) * 16u, v[4 * i4],
                                             v[4 * i4 + 1], v[4 * i4 + 2], v[4 * i4 + 3]);
                }
            }
        }
    }
    if (P.cluster > 1) {
        __syncwarp();
        cluster_sync_all();  // release / acquire: all partial slices have landed
    }
    if (MODE == 1 && warp >= 2) {
        const int s_tiles_n = c_tiles_n, s_v_smode = p.v_smode;
        const float* const s_v_q = p.v_q;
        float* const s_v_part = p.v_part;
        const long long* const s_v_trg = p.v_trg;
        // ---- vocabulary statistics of row m over this CTA's BN columns (no output tile is written):
        //   max, sum exp(v - max), sum v, v[target], v[pad], and the sampling arg-max (key, index, logit)
        const int trg = m < e_M ? (int)s_v_trg[m] : -1;
        float mx = -INFINITY, se = 0.f, sv = 0.f, lt = 0.f, lp = 0.f, bk = -INFINITY, bl = 0.f;
        int bi = 0x7fffffff;
        const bool row_ok = m < e_M;
        const bool bias_vec = e_bias && aligned16(e_bias);
#pragma unroll 1
        for (int cc = 0; cc < BN; cc += 32) {   // ONE pass: online max / sum-exp per 32-column chunk
            uint32_t r[32];
            tmem_ld_32x32(lane_taddr + (uint32_t)cc, r);
            tmem_ld_wait();
            const int nb = n0 + cc;
            if (nb >= e_N) break;
            float v[32];
            const bool full = nb + 31 < e_N;
            if (full && bias_vec) {
#pragma unroll
                for (int j4 = 0; j4 < 32; j4 += 4) {
                    const float4 b4 = __ldg(reinterpret_cast<const float4*>(e_bias + nb + j4));
                    v[j4] = __uint_as_float(r[j4]) + b4.x; v[j4 + 1] = __uint_as_float(r[j4 + 1]) + b4.y;
                    v[j4 + 2] = __uint_as_float(r[j4 + 2]) + b4.z; v[j4 + 3] = __uint_as_float(r[j4 + 3]) + b4.w;
                }
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    v[j] = nb + j < e_N ? __uint_as_float(r[j]) + (e_bias ? e_bias[nb + j] : 0.f) : -INFINITY;
            }
            float cm = v[0];
#pragma unroll
            for (int j = 1; j < 32; ++j) cm = fmaxf(cm, v[j]);
            const float nm = fmaxf(mx, cm);
            float cs = 0.f, csum = 0.f;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const bool in = full || nb + j < e_N;
                cs += in ? __expf(v[j] - nm) : 0.f;
                csum += in ? v[j] : 0.f;
            }
            se = se * __expf(mx - nm) + cs;
            mx = nm;
            sv += csum;
            if (trg >= nb && trg < nb + 32) {
#pragma unroll
                for (int j = 0; j < 32; ++j) lt = (nb + j == trg) ? v[j] : lt;
            }
            if (e_v_pad >= nb && e_v_pad < nb + 32) {
#pragma unroll
                for (int j = 0; j < 32; ++j) lp = (nb + j == e_v_pad) ? v[j] : lp;
            }
            if (s_v_smode == 0) {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (v[j] > bk) { bk = v[j]; bi = nb + j; bl = v[j]; }   // ascending n: the first maximum wins
            } else if (row_ok) {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int n = nb + j;
                    if (n < e_N) {
                        float key;
                        if (s_v_smode == 1) {
                            key = v[j] - __logf(s_v_q[(long long)m * e_N + n]);
                        } else {
                            const float u = u32_to_unit(philox_u32(e_rng, e_salt, (unsigned long long)m * e_N + n));
                            key = v[j] - __logf(-__logf(1.f - u));
                        }
                        if (key > bk) { bk = key; bi = n; bl = v[j]; }
                    }
                }
            }
        }
        if (m < e_M) {
            float* o = s_v_part + ((long long)m * s_tiles_n + tn) * 8;
            st_f4(o, make_float4(mx, se, sv, lt));
            st_f4(o + 4, make_float4(lp, bk, __int_as_float(bi), bl));
        }
        tc_fence_before();
    }
    // ---- MODE 0 / 2 epilogue in two phases through shared memory.  Phase A (thread = accumulator row = TMEM lane) moves
    // raw accumulators into a warp-private staging tile; phase B re-reads it with the lanes running ALONG a row, so that
    // every global load (bias, gate, residual, C) and store (C32, C16) of a warp covers whole 128-byte lines.  The
    // row-per-thread form wrote 8..16 bytes into 32 different lines per instruction and cost 4-5 us per tile (measured
    // with bmhrl_gemm_bf16_trace: longer than the 16 k-blocks of the main loop).  The staging tile lives in the operand
    // ring: by the time the accumulator barrier fires every TMA load has landed and every MMA has read its operands.
    constexpr int STG_COLS = 64, STG_LD = STG_COLS + 4;          // +4 floats: conflict-free 128-bit row-strided stores
    constexpr int STG_WARP_FLOATS = 32 * STG_LD + 32 * 8;        // + 8 floats of per-row state (MODE 2)
    float* stg = reinterpret_cast<float*>(smem_raw + (base - smem_u32(smem_raw))) + (size_t)q * STG_WARP_FLOATS;
    if ((MODE == 0 || MODE == 2) && warp >= 2) {
        const float ks = e_drop_p > 0.f ? 1.f / (1.f - e_drop_p) : 1.f;
        if (MODE == 2) {   // per-row statistics / gradients of the row statistics: one row per thread, read in phase B
            float* rs = stg + 32 * STG_LD + lane * 8;
            float lse = 0.f, g0 = 0.f, g1 = 0.f, g2 = 0.f, g3 = 0.f, g4 = 0.f;
            int trg = -1, tok = -1;
            if (m < e_M) {
                lse = p.v_stats[(long long)m * 5];
                const float* g = p.v_dstat + (long long)m * 5;
                g0 = g[0]; g1 = g[1]; g2 = g[2]; g3 = g[3]; g4 = g[4];
                trg = (int)p.v_trg[m];
                tok = p.v_tok ? (int)p.v_tok[m] : -1;
            }
            rs[0] = lse; rs[1] = g0; rs[2] = g1; rs[3] = g2; rs[4] = g3; rs[5] = g4;
            rs[6] = __int_as_float(trg); rs[7] = __int_as_float(tok);
        }
#pragma unroll 1
        for (int p0 = 0; p0 < W; p0 += STG_COLS) {
            const int ncols = min(STG_COLS, W - p0);
            // ---- phase A
#pragma unroll 1
            for (int cc = 0; cc < ncols; cc += 32) {
                const int c0 = (int)rank * W + p0 + cc;
                uint32_t r[32];
                if (has_k) {
                    tmem_ld_32x32(lane_taddr + (uint32_t)c0, r);
                    tmem_ld_wait();
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i) r[i] = 0u;
                }
                if (threadIdx.x == 64 && p0 == 0 && cc == 0) gb_stamp(P, 11);
                if (P.cluster > 1) {
                    for (uint32_t src = 0; src < (uint32_t)csz; ++src) {
                        if (src == rank) continue;
                        const uint32_t sb = recv + (src * (uint32_t)(W / 4)) * (GB_BM * 16u);
#pragma unroll
                        for (int i4 = 0; i4 < 8; ++i4) {
                            float4 tt = ld_shared_v4(sb + ((uint32_t)((p0 + cc) / 4 + i4) * GB_BM + row) * 16u);
                            r[4 * i4] = __float_as_uint(__uint_as_float(r[4 * i4]) + tt.x);
                            r[4 * i4 + 1] = __float_as_uint(__uint_as_float(r[4 * i4 + 1]) + tt.y);
                            r[4 * i4 + 2] = __float_as_uint(__uint_as_float(r[4 * i4 + 2]) + tt.z);
                            r[4 * i4 + 3] = __float_as_uint(__uint_as_float(r[4 * i4 + 3]) + tt.w);
                        }
                    }
                }
                float* srow = stg + lane * STG_LD + cc;
#pragma unroll
                for (int i4 = 0; i4 < 8; ++i4)
                    *reinterpret_cast<float4*>(srow + 4 * i4) =
                        make_float4(__uint_as_float(r[4 * i4]), __uint_as_float(r[4 * i4 + 1]), __uint_as_float(r[4 * i4 + 2]),
                                    __uint_as_float(r[4 * i4 + 3]));
            }
            __syncwarp();
            if (threadIdx.x == 64 && p0 == 0) gb_stamp(P, 12);
            // ---- phase B: lane -> (row rr of this warp's 32, 4 consecutive columns)
            const int sh = ncols == 64 ? 4 : 3;   // float4 per staged row: 16 or 8
            // Fast form (every vector access legal: N % 4 == 0, aligned rows): FOUR items per trip, each optional step of
            // the epilogue applied to all four before the next one -- the tests on the problem's options are uniform and
            // now cost one branch per four items, and the four items are independent instruction streams.  One warp per
            // scheduler running the per-item form below is a serial chain of ~60 dependent instructions per item
            // (2.2 us per 128 x 64 tile, measured).
            const bool fast = MODE == 0 && (e_N & 3) == 0 && (!e_C32 || e_vec32) && (!e_C16 || e_vec16) &&
                              (!e_resid || e_vecr) && (!e_gate || e_vecg) && (!e_bias || aligned16(e_bias));
            if (fast) {
                const int cmask = (1 << sh) - 1;
#pragma unroll 1
                for (int it0 = 0; it0 < (1 << sh); it0 += 4) {
                    float4 v[4];
                    int mm[4], nn[4];
                    bool ok[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int idx = (it0 + i) * 32 + lane;
                        const int rr = idx >> sh, c4 = idx & cmask;
                        mm[i] = m0 + q * 32 + rr;
                        nn[i] = n0 + (int)rank * W + p0 + c4 * 4;
                        ok[i] = mm[i] < e_M && nn[i] < e_N;
                        const float4 a4 = *reinterpret_cast<const float4*>(stg + rr * STG_LD + c4 * 4);
                        v[i] = make_float4(a4.x * e_alpha, a4.y * e_alpha, a4.z * e_alpha, a4.w * e_alpha);
                    }
                    if (e_bias) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            if (ok[i]) {
                                const float4 b4 = __ldg(reinterpret_cast<const float4*>(e_bias + nn[i]));
                                v[i].x += b4.x; v[i].y += b4.y; v[i].z += b4.z; v[i].w += b4.w;
                            }
                    }
                    if (e_act == 1) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            v[i].x = fmaxf(v[i].x, 0.f); v[i].y = fmaxf(v[i].y, 0.f);
                            v[i].z = fmaxf(v[i].z, 0.f); v[i].w = fmaxf(v[i].w, 0.f);
                        }
                    }
                    if (e_gate) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            if (ok[i]) {
                                const uint2 g = ld_bf16x4(e_gate + (long long)mm[i] * e_ldg + nn[i]);
                                v[i].x = bf16_lo(g.x) > 0.f ? v[i].x * e_gate_scale : 0.f;
                                v[i].y = bf16_hi(g.x) > 0.f ? v[i].y * e_gate_scale : 0.f;
                                v[i].z = bf16_lo(g.y) > 0.f ? v[i].z * e_gate_scale : 0.f;
                                v[i].w = bf16_hi(g.y) > 0.f ? v[i].w * e_gate_scale : 0.f;
                            }
                    }
                    if (e_drop_p > 0.f) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            if (ok[i]) {   // N % 4 == 0 and nn % 4 == 0: the 4 elements share one Philox block
                                const unsigned long long e0 = (unsigned long long)mm[i] * e_N + nn[i];
                                const Philox4 rb = philox_block(e_rng, e_salt, e0 >> 2);
                                v[i].x = u32_to_unit(rb.x) >= e_drop_p ? v[i].x * ks : 0.f;
                                v[i].y = u32_to_unit(rb.y) >= e_drop_p ? v[i].y * ks : 0.f;
                                v[i].z = u32_to_unit(rb.z) >= e_drop_p ? v[i].z * ks : 0.f;
                                v[i].w = u32_to_unit(rb.w) >= e_drop_p ? v[i].w * ks : 0.f;
                            }
                    }
                    if (e_resid) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            if (ok[i]) {
                                const float4 tt = ld_f4(e_resid + (long long)mm[i] * e_ldr + nn[i]);
                                v[i].x += tt.x; v[i].y += tt.y; v[i].z += tt.z; v[i].w += tt.w;
                            }
                    }
                    if (e_C32) {
                        float* crow[4];
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            if (e_seg_rows > 0) {
                                const int sgi = mm[i] / e_seg_rows;
                                float* b = sgi == 0 ? e_C32 : (sgi == 1 ? e_seg1 : e_seg2);
                                crow[i] = b + (long long)(mm[i] - sgi * e_seg_rows) * e_ldc32 + nn[i];
                            } else {
                                crow[i] = e_C32 + (long long)mm[i] * e_ldc32 + nn[i];
                            }
                        }
                        if (e_beta != 0.f) {
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                if (ok[i]) {
                                    const float4 tt = ld_f4(crow[i]);
                                    v[i].x += e_beta * tt.x; v[i].y += e_beta * tt.y;
                                    v[i].z += e_beta * tt.z; v[i].w += e_beta * tt.w;
                                }
                        }
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            if (ok[i]) st_f4(crow[i], v[i]);
                    }
                    if (e_C16) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            if (ok[i])
                                *reinterpret_cast<uint2*>(e_C16 + (long long)mm[i] * e_ldc16 + nn[i]) =
                                    make_uint2(pack2_bf16(v[i].x, v[i].y), pack2_bf16(v[i].z, v[i].w));
                    }
                }
            } else
#pragma unroll 1
            for (int idx = lane; idx < (32 << sh); idx += 32) {
                const int rr = idx >> sh, c4 = idx - (rr << sh);
                const int mm = m0 + q * 32 + rr;
                const int n = n0 + (int)rank * W + p0 + c4 * 4;
                if (mm >= e_M || n >= e_N) continue;
                const bool full = n + 3 < e_N;
                const float4 a4 = *reinterpret_cast<const float4*>(stg + rr * STG_LD + c4 * 4);
                float v[4] = {a4.x, a4.y, a4.z, a4.w};
                if (MODE == 2) {
                    // d logits from the gradients of the row statistics {lse, sum v, v[target], v[pad], v[token]}:
                    //   dv(n) = dlse * exp(v(n) - lse) + dsum + [n == target] dlt + [n == pad] dlp + [n == token] dls
                    const float* rs = stg + 32 * STG_LD + rr * 8;
                    const float lse = rs[0], g0 = rs[1], g1 = rs[2], g2 = rs[3], g3 = rs[4], g4 = rs[5];
                    const int trg = __float_as_int(rs[6]), tok = __float_as_int(rs[7]);
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int nn = n + j;
                        const float lg = v[j] + ((e_bias && nn < e_N) ? e_bias[nn] : 0.f);
                        float d = g0 * __expf(lg - lse) + g1;
                        if (nn == trg) d += g2;
                        if (nn == e_v_pad) d += g3;
                        if (nn == tok) d += g4;
                        v[j] = d;
                    }
                    __nv_bfloat16* hrow = e_C16 + (long long)mm * e_ldc16 + n;
                    if (full && e_vec16) {
                        *reinterpret_cast<uint2*>(hrow) = make_uint2(pack2_bf16(v[0], v[1]), pack2_bf16(v[2], v[3]));
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (n + j < e_N) hrow[j] = __float2bfloat16(v[j]);
                    }
                    continue;
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) v[j] *= e_alpha;
                if (e_bias) {
                    if (full && aligned16(e_bias + n)) {
                        const float4 b4 = __ldg(reinterpret_cast<const float4*>(e_bias + n));
                        v[0] += b4.x; v[1] += b4.y; v[2] += b4.z; v[3] += b4.w;
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (n + j < e_N) v[j] += e_bias[n + j];
                    }
                }
                if (e_act == 1) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) v[j] = fmaxf(v[j], 0.f);
                }
                if (e_gate) {
                    const __nv_bfloat16* grow = e_gate + (long long)mm * e_ldg + n;
                    if (full && e_vecg) {
                        const uint2 g = ld_bf16x4(grow);
                        v[0] = bf16_lo(g.x) > 0.f ? v[0] * e_gate_scale : 0.f;
                        v[1] = bf16_hi(g.x) > 0.f ? v[1] * e_gate_scale : 0.f;
                        v[2] = bf16_lo(g.y) > 0.f ? v[2] * e_gate_scale : 0.f;
                        v[3] = bf16_hi(g.y) > 0.f ? v[3] * e_gate_scale : 0.f;
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (n + j < e_N) v[j] = __bfloat162float(grow[j]) > 0.f ? v[j] * e_gate_scale : 0.f;
                    }
                }
                if (e_drop_p > 0.f) {
                    const unsigned long long e0 = (unsigned long long)mm * e_N + n;
                    if ((e0 & 3ull) == 0) {  // the 4 elements share one Philox block
                        const Philox4 rb = philox_block(e_rng, e_salt, e0 >> 2);
                        v[0] = u32_to_unit(rb.x) >= e_drop_p ? v[0] * ks : 0.f;
                        v[1] = u32_to_unit(rb.y) >= e_drop_p ? v[1] * ks : 0.f;
                        v[2] = u32_to_unit(rb.z) >= e_drop_p ? v[2] * ks : 0.f;
                        v[3] = u32_to_unit(rb.w) >= e_drop_p ? v[3] * ks : 0.f;
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            v[j] = u32_to_unit(philox_u32(e_rng, e_salt, e0 + j)) >= e_drop_p ? v[j] * ks : 0.f;
                    }
                }
                if (e_resid) {
                    const float* rrow = e_resid + (long long)mm * e_ldr + n;
                    if (full && e_vecr) {
                        const float4 tt = ld_f4(rrow);
                        v[0] += tt.x; v[1] += tt.y; v[2] += tt.z; v[3] += tt.w;
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (n + j < e_N) v[j] += rrow[j];
                    }
                }
                if (e_C32) {
                    // destination row of the fp32 output (optionally one of up to three row segments)
                    float* crow;
                    if (e_seg_rows > 0) {
                        const int sgi = mm / e_seg_rows;
                        float* b = sgi == 0 ? e_C32 : (sgi == 1 ? e_seg1 : e_seg2);
                        crow = b + (long long)(mm - sgi * e_seg_rows) * e_ldc32 + n;
                    } else {
                        crow = e_C32 + (long long)mm * e_ldc32 + n;
                    }
                    if (full && e_vec32) {
                        if (e_beta != 0.f) {
                            const float4 tt = ld_f4(crow);
                            v[0] += e_beta * tt.x; v[1] += e_beta * tt.y; v[2] += e_beta * tt.z; v[3] += e_beta * tt.w;
                        }
                        st_f4(crow, make_float4(v[0], v[1], v[2], v[3]));
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (n + j < e_N) {
                                if (e_beta != 0.f) v[j] += e_beta * crow[j];
                                crow[j] = v[j];
                            }
                    }
                }
                if (e_C16) {
                    __nv_bfloat16* hrow = e_C16 + (long long)mm * e_ldc16 + n;
                    if (full && e_vec16) {
                        *reinterpret_cast<uint2*>(hrow) = make_uint2(pack2_bf16(v[0], v[1]), pack2_bf16(v[2], v[3]));
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (n + j < e_N) hrow[j] = __float2bfloat16(v[j]);
                    }
                }
            }
            __syncwarp();   // the next pass overwrites the staging tile
            if (threadIdx.x == 64 && p0 == 0) gb_stamp(P, 13);
        }
        tc_fence_before();
        if (threadIdx.x == 64) gb_stamp(P, 7);
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
    if (threadIdx.x == 32) { gb_stamp(P, 8); gb_stamp_ns(P, 10); }
}

// ---------------------------------------------------------------------------------------------- host side
static int g_gb_force_bn = 0, g_gb_force_split = 0, g_gb_force_stages = 0;
static unsigned long long* g_gb_trace = nullptr;

template <int BN, int MODE = 0>
static int launch_gb(const GbParams& P, int total_tiles, int split, cudaStream_t st) {
    const int stage_bytes = GB_A_BYTES + BN * 128;
    const int smem = P.nstages * stage_bytes + 1024 + 8 * (2 * P.nstages + 2) + (P.cluster > 1 ? GB_BM * BN * 4 : 0);
    static DeviceOnce attr;
    if (needs_setup(attr)) {
        cudaError_t e = cudaFuncSetAttribute(gemm_bf16_tc_kernel<BN, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        if (e != cudaSuccess) {
            bmhrl_set_error(cudaGetErrorString(e));
            return BMHRL_ERR_CUDA;
        }
        mark_done(attr);
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(total_tiles, 1, split);
    cfg.blockDim = dim3(192);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[3];
    int na = 0;
    if (P.cluster > 1) {
        at[na].id = cudaLaunchAttributeClusterDimension;
        at[na].val.clusterDim.x = 1;
        at[na].val.clusterDim.y = 1;
        at[na].val.clusterDim.z = P.cluster;
        ++na;
    }
    if (pdl_enabled()) {
        at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    if (launch_priority() != kNoPriority) {
        at[na].id = cudaLaunchAttributePriority;
        at[na].val.priority = launch_priority();
        ++na;
    }
    if (na) {
        cfg.attrs = at;
        cfg.numAttrs = na;
    }
    cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_bf16_tc_kernel<BN, MODE>, P);
    if (e != cudaSuccess) {
        bmhrl_set_error(cudaGetErrorString(e));
        return BMHRL_ERR_CUDA;
    }
    bmhrl_count_launch(1);
    return BMHRL_OK;
}

static bool gb_operand_ok(const void* p, long long ld) { return p && aligned16(p) && ld > 0 && (ld & 7) == 0; }

}  // namespace bmhrl

using namespace bmhrl;

extern "C" int bmhrl_gemm_desc_bytes(void) { return (int)sizeof(bmhrl_gemm_desc); }

extern "C" int bmhrl_gemm_bf16_tuning(int force_bn, int force_split, int force_stages) {
    BMHRL_CHECK_ARG(force_bn == 0 || force_bn == 64 || force_bn == 128, "bmhrl_gemm_bf16_tuning: BN must be 0, 64 or 128");
    BMHRL_CHECK_ARG(force_split == 0 || force_split == 1 || force_split == 2 || force_split == 4,
                    "bmhrl_gemm_bf16_tuning: split must be 0 (cost model), 1, 2 or 4");
    BMHRL_CHECK_ARG(force_stages >= 0 && force_stages <= 6, "bmhrl_gemm_bf16_tuning: stages 0 (auto) .. 6");
    g_gb_force_bn = force_bn;
    g_gb_force_split = force_split;
    g_gb_force_stages = force_stages;
    return BMHRL_OK;
}

extern "C" int bmhrl_gemm_bf16_trace(unsigned long long* device_slots) {
    g_gb_trace = device_slots;
    return BMHRL_OK;
}

extern "C" int bmhrl_gemm_bf16_group(const bmhrl_gemm_desc* d, int n, void* stream) {
    BMHRL_CHECK_ARG(d && n >= 1 && n <= GB_MAXP, "bmhrl_gemm_bf16_group: 1..4 problems per launch");
    if (get_encode_fn() == nullptr) {
        bmhrl_set_error("bmhrl_gemm_bf16_group: cuTensorMapEncodeTiled unavailable");
        return BMHRL_ERR_CUDA;
    }
    // ---- one tile shape / K split for the whole group (cost model over the group's total tile count)
    long long t128 = 0, t64 = 0;
    int max_kb = 0, min_kb = 1 << 30, max_n = 0;
    for (int i = 0; i < n; ++i) {
        const bmhrl_gemm_desc& g = d[i];
        BMHRL_CHECK_ARG(g.M > 0 && g.N > 0 && g.K > 0, "bmhrl_gemm_bf16_group: empty problem");
        BMHRL_CHECK_ARG(gb_operand_ok(g.A, g.lda) && gb_operand_ok(g.B, g.ldb),
                        "bmhrl_gemm_bf16_group: operands must be 16-byte aligned bf16 with row pitches that are multiples of 8");
        BMHRL_CHECK_ARG(g.C32 || g.C16, "bmhrl_gemm_bf16_group: no output");
        BMHRL_CHECK_ARG(g.drop_p >= 0.f && g.drop_p < 1.f && (g.drop_p == 0.f || g.rng_state), "bmhrl_gemm_bf16_group: dropout");
        BMHRL_CHECK_ARG(g.seg_rows == 0 || (g.C32 && g.C32_seg1 && g.seg_rows > 0 && (g.M <= 2 * g.seg_rows || g.C32_seg2) &&
                                            g.M <= 3 * g.seg_rows),
                        "bmhrl_gemm_bf16_group: bad row segmentation");
        const long long tm = (g.M + GB_BM - 1) / GB_BM;
        t128 += tm * ((g.N + 127) / 128);
        t64 += tm * ((g.N + 63) / 64);
        const int kb = (g.K + GB_BK - 1) / GB_BK;
        max_kb = kb > max_kb ? kb : max_kb;
        min_kb = kb < min_kb ? kb : min_kb;
        max_n = g.N > max_n ? g.N : max_n;
    }
    const int n_sm = device_sm_count();
    // Tile width.  Measured inside the training step (profiles/r02_gemm_tile_experiments.txt): these GEMMs are latency
    // chains per CTA (TMA round trip per stage), several of them run side by side on different streams, and what
    // counts is how many CTAs -- bytes in flight -- an SM can hold: 128 x 64 tiles with a 3-deep ring (72 KB, three CTAs
    // per SM next to a co-resident kernel) beat 128 x 128 / 4 stages (132 KB, one CTA per SM) by 12 % of the step.  Only
    // when 128-wide tiles alone fill the machine twice over (large M x N: operand-bandwidth bound) are they worth it.
    int BN = (t128 >= 2ll * n_sm) ? 128 : 64, split = 1;
    {
        const long long tiles = BN == 128 ? t128 : t64;
        // a cluster of `split` CTAs along K when the tiles alone leave most SMs idle and every problem has K to split
        while (split < 4 && tiles * split * 2 <= n_sm + n_sm / 4 && min_kb / (split * 2) >= 2 && BN / (split * 2) >= 32) split *= 2;
    }
    // weight-gradient launches (A and B both MN-major: dW = dZ^T X) are off the step's critical path: what counts for them
    // is SM time, not latency (experiment switch BMHRL_GB_WGRAD_BN = 128: wider tiles, fewer operand bytes per FLOP)
    static const int env_wbn = getenv("BMHRL_GB_WGRAD_BN") ? atoi(getenv("BMHRL_GB_WGRAD_BN")) : 0;
    const bool is_wgrad = d[0].a_mn && d[0].b_mn;
    if (is_wgrad && (env_wbn == 64 || env_wbn == 128) && max_n >= 128) {
        BN = env_wbn;
        split = 1;
    }
    static const int env_bn = getenv("BMHRL_GB_BN") ? atoi(getenv("BMHRL_GB_BN")) : 0;
    static const int env_split = getenv("BMHRL_GB_SPLIT") ? atoi(getenv("BMHRL_GB_SPLIT")) : 0;
    if (env_bn == 64 || env_bn == 128) BN = env_bn;
    if (env_split == 1 || env_split == 2 || env_split == 4) split = env_split;
    if (g_gb_force_bn) BN = g_gb_force_bn;
    if (g_gb_force_split) split = g_gb_force_split;
    while (split > 1 && (BN / split < 32 || min_kb < split)) split >>= 1;
    static const int env_stages = getenv("BMHRL_GB_STAGES") ? atoi(getenv("BMHRL_GB_STAGES")) : 0;  // experiments
    static const int env_wst = getenv("BMHRL_GB_WGRAD_STAGES") ? atoi(getenv("BMHRL_GB_WGRAD_STAGES")) : 0;
    int nstages = g_gb_force_stages ? g_gb_force_stages : (env_stages ? env_stages : (BN == 128 ? 4 : 3));
    if (is_wgrad && env_wst >= 2 && env_wst <= 6 && !g_gb_force_stages) nstages = env_wst;
    {
        // a launch with at most one CTA per SM cannot hide the TMA round trip behind other CTAs of its own: deeper ring
        // (experiment switch BMHRL_GB_DEEP = ring depth for such launches, 0 = off)
        static const int env_deep = getenv("BMHRL_GB_DEEP") ? atoi(getenv("BMHRL_GB_DEEP")) : 0;
        const long long ctas = (BN == 128 ? t128 : t64) * split;
        if (env_deep > nstages && !g_gb_force_stages && !env_stages && ctas <= n_sm) nstages = env_deep > 6 ? 6 : env_deep;
        const int kb_per_cta = (max_kb + split - 1) / split;
        if (nstages > kb_per_cta) nstages = kb_per_cta < 2 ? 2 : kb_per_cta;
        if (split > 1 && nstages > 3) nstages = 3;  // room for the DSMEM receive buffer beside a co-resident kernel
    }

    GbParams P{};
    P.n = n;
    P.cluster = split;
    P.nstages = nstages;
    P.trace = g_gb_trace;
    int tile_start = 0;
    const CUtensorMapDataType BF = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    for (int i = 0; i < n; ++i) {
        const bmhrl_gemm_desc& g = d[i];
        GbProblem& q = P.p[i];
        bool ok;
        if (!g.a_mn) ok = encode_2d(&q.tmA, BF, 2, g.A, g.K, g.M, g.lda, GB_BK, GB_BM);
        else ok = encode_2d(&q.tmA, BF, 2, g.A, g.M, g.K, g.lda, 64, GB_BK);
        if (ok) {
            if (!g.b_mn) ok = encode_2d(&q.tmB, BF, 2, g.B, g.K, g.N, g.ldb, GB_BK, BN);
            else ok = encode_2d(&q.tmB, BF, 2, g.B, g.N, g.K, g.ldb, 64, GB_BK);
        }
        if (!ok) {
            char msg[320];
            snprintf(msg, sizeof(msg), "bmhrl_gemm_bf16_group: cuTensorMapEncodeTiled failed (problem %d: M=%d N=%d K=%d a_mn=%d "
                     "b_mn=%d lda=%lld ldb=%lld BN=%d rc=%d)", i, g.M, g.N, g.K, g.a_mn, g.b_mn, g.lda, g.ldb, BN, g_last_encode_rc);
            bmhrl_set_error(msg);
            return BMHRL_ERR_CUDA;
        }
        q.C32 = g.C32; q.seg1 = g.C32_seg1; q.seg2 = g.C32_seg2; q.C16 = (__nv_bfloat16*)g.C16;
        q.bias = g.bias; q.resid = g.resid; q.gate = (const __nv_bfloat16*)g.gate; q.rng = g.rng_state;
        q.ldc32 = g.ldc32; q.ldc16 = g.ldc16; q.ldr = g.ldr; q.ldg = g.ldg;
        q.M = g.M; q.N = g.N; q.K = g.K;
        q.num_kb = (g.K + GB_BK - 1) / GB_BK;
        q.kb_per_split = (q.num_kb + split - 1) / split;
        q.a_mn = g.a_mn ? 1 : 0; q.b_mn = g.b_mn ? 1 : 0; q.act = g.act; q.seg_rows = g.seg_rows;
        q.b_static = g.b_static ? 1 : 0;
        q.tiles_n = (g.N + BN - 1) / BN;
        q.tile_start = tile_start;
        tile_start += ((g.M + GB_BM - 1) / GB_BM) * q.tiles_n;
        q.vec32 = g.C32 && aligned16(g.C32) && (g.ldc32 & 3) == 0 && (!g.C32_seg1 || aligned16(g.C32_seg1)) &&
                  (!g.C32_seg2 || aligned16(g.C32_seg2));
        q.vec16 = g.C16 && ((uintptr_t)g.C16 & 7u) == 0 && (g.ldc16 & 3) == 0;
        q.vecr = g.resid && aligned16(g.resid) && (g.ldr & 3) == 0;
        q.vecg = g.gate && ((uintptr_t)g.gate & 7u) == 0 && (g.ldg & 3) == 0;
        q.alpha = g.alpha; q.beta = g.beta; q.gate_scale = g.gate_scale; q.drop_p = g.drop_p; q.salt = g.salt;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (BN == 128) return launch_gb<128>(P, tile_start, split, st);
    return launch_gb<64>(P, tile_start, split, st);
}


// ================================================================================================================
// K4: vocabulary projection fused with log-softmax and the KL losses (north_star kernel 3).
// Reference: WorkerCore model/bm_hrl_agent.py:463-466 (Linear 364 -> V + LogSoftmax), LabelSmoothing
// loss/label_smoothing.py:12-32, BiasedKL loss/biased_kl.py:22-53, biased_kl() sampling
// epoch_loops/captioning_bmrl_loops.py:273-285.  The (B*L, V) logits / log-probabilities are never written:
//   forward   the GEMM tiles stay in TMEM; the epilogue reduces each row of a tile to {max, sum exp, sum v, v[target],
//             v[pad], sampling arg-max}; vocab_finalize_kernel merges the V/128 partials of a row into the sufficient
//             statistics {lse, sum v, v[target], v[pad], v[token]} and the sampled / arg-max token.  Both KL losses are
//             closed forms of these five numbers (their target distribution takes at most four distinct values per row).
//   backward  the GEMM is recomputed and the epilogue turns the five statistic gradients into d logits (bf16), the
//             operand of the dX / dW GEMMs.
__global__ void vocab_finalize_kernel(const float* __restrict__ part, int tiles, float* __restrict__ stats,
                                      long long* __restrict__ tok, int M) {
    griddep_wait();
    griddep_launch_dependents();
    const int m = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (m >= M) return;
    const float* pr = part + (long long)m * tiles * 8;
    float mx = -INFINITY;
    for (int t = lane; t < tiles; t += 32) mx = fmaxf(mx, pr[t * 8]);
    mx = warp_max(mx);
    float se = 0.f, sv = 0.f, lt = 0.f, lp = 0.f, bk = -INFINITY, bl = 0.f;
    int bi = 0x7fffffff;
    for (int t = lane; t < tiles; t += 32) {
        const float4 a = ld_f4(pr + t * 8), b = ld_f4(pr + t * 8 + 4);
        se += a.y * __expf(a.x - mx);
        sv += a.z;
        lt += a.w;
        lp += b.x;
        const int i = __float_as_int(b.z);
        if (b.y > bk || (b.y == bk && i < bi)) { bk = b.y; bi = i; bl = b.w; }
    }
    se = warp_sum(se); sv = warp_sum(sv); lt = warp_sum(lt); lp = warp_sum(lp);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ok = __shfl_xor_sync(0xffffffffu, bk, o), ol = __shfl_xor_sync(0xffffffffu, bl, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ok > bk || (ok == bk && oi < bi)) { bk = ok; bi = oi; bl = ol; }
    }
    if (lane == 0) {
        float* o = stats + (long long)m * 5;
        o[0] = mx + logf(se); o[1] = sv; o[2] = lt; o[3] = lp; o[4] = bl;
        if (tok) tok[m] = bi;
    }
}

static int vocab_problem(GbParams& P, const void* x16, long long ldx, const void* w16, long long ldw, const float* bias, int M,
                         int V, int K) {
    if (get_encode_fn() == nullptr) {
        bmhrl_set_error("vocab: cuTensorMapEncodeTiled unavailable");
        return BMHRL_ERR_CUDA;
    }
    P.n = 1; P.cluster = 1; P.trace = nullptr;
    const int num_kb = (K + GB_BK - 1) / GB_BK;
    P.nstages = num_kb < 4 ? (num_kb < 2 ? 2 : num_kb) : 4;
    GbProblem& q = P.p[0];
    const CUtensorMapDataType BF = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    if (!encode_2d(&q.tmA, BF, 2, x16, K, M, ldx, GB_BK, GB_BM) || !encode_2d(&q.tmB, BF, 2, w16, K, V, ldw, GB_BK, 128)) {
        bmhrl_set_error("vocab: cuTensorMapEncodeTiled failed");
        return BMHRL_ERR_CUDA;
    }
    q.bias = bias; q.M = M; q.N = V; q.K = K; q.num_kb = num_kb; q.kb_per_split = num_kb;
    q.b_static = 1;   // the projection weight shadow
    q.tiles_n = (V + 127) / 128; q.tile_start = 0; q.alpha = 1.f;
    return BMHRL_OK;
}

extern "C" long long bmhrl_vocab_stats_ws_floats(int M, int V) { return (long long)M * ((V + 127) / 128) * 8; }

extern "C" int bmhrl_vocab_stats_bf16(const void* x16, long long ldx, const void* w16, long long ldw, const float* bias, int M,
                                      int V, int K, const long long* target, int pad_idx, int sample_mode,
                                      const float* exp_noise, const unsigned long long* rng_state, unsigned int salt,
                                      float* part_ws, float* stats, long long* tokens, void* stream) {
    BMHRL_CHECK_ARG(gb_operand_ok(x16, ldx) && gb_operand_ok(w16, ldw) && target && part_ws && stats && M > 0 && V > 2 && K > 0,
                    "bmhrl_vocab_stats_bf16: bad arguments");
    BMHRL_CHECK_ARG(sample_mode >= 0 && sample_mode <= 2 && (sample_mode != 1 || exp_noise) && (sample_mode != 2 || rng_state),
                    "bmhrl_vocab_stats_bf16: sample_mode 0 arg-max, 1 explicit Exp(1) noise, 2 Philox");
    BMHRL_CHECK_ARG(aligned16(part_ws), "bmhrl_vocab_stats_bf16: workspace alignment");
    GbParams P{};
    int rc = vocab_problem(P, x16, ldx, w16, ldw, bias, M, V, K);
    if (rc != BMHRL_OK) return rc;
    GbProblem& q = P.p[0];
    q.v_trg = target; q.v_q = exp_noise; q.v_part = part_ws; q.v_pad = pad_idx; q.v_smode = sample_mode;
    q.rng = rng_state; q.salt = salt;
    const int tiles = ((M + GB_BM - 1) / GB_BM) * q.tiles_n;
    cudaStream_t st = (cudaStream_t)stream;
    rc = launch_gb<128, 1>(P, tiles, 1, st);
    if (rc != BMHRL_OK) return rc;
    BMHRL_LAUNCH_PDL("vocab_finalize", vocab_finalize_kernel, dim3((M + 7) / 8), dim3(256), 0, st, (const float*)part_ws,
                     q.tiles_n, stats, tokens, M);
    return BMHRL_OK;
}

extern "C" int bmhrl_vocab_dlogits_bf16(const void* x16, long long ldx, const void* w16, long long ldw, const float* bias, int M,
                                        int V, int K, const long long* target, const long long* tokens, int pad_idx,
                                        const float* stats, const float* dstats, void* dlogits16, long long lddl,
                                        void* stream) {
    BMHRL_CHECK_ARG(gb_operand_ok(x16, ldx) && gb_operand_ok(w16, ldw) && target && stats && dstats && dlogits16 && M > 0 &&
                        V > 2 && K > 0 && lddl >= V && (lddl & 7) == 0,
                    "bmhrl_vocab_dlogits_bf16: bad arguments");
    GbParams P{};
    int rc = vocab_problem(P, x16, ldx, w16, ldw, bias, M, V, K);
    if (rc != BMHRL_OK) return rc;
    GbProblem& q = P.p[0];
    q.v_trg = target; q.v_tok = tokens; q.v_pad = pad_idx; q.v_stats = stats; q.v_dstat = dstats;
    q.C16 = (__nv_bfloat16*)dlogits16; q.ldc16 = lddl;
    q.vec16 = ((uintptr_t)dlogits16 & 7u) == 0 && (lddl & 3) == 0;
    const int tiles = ((M + GB_BM - 1) / GB_BM) * q.tiles_n;
    return launch_gb<128, 2>(P, tiles, 1, (cudaStream_t)stream);
}
