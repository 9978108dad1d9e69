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
    int a_mn, b_mn, act, seg_rows;
    int tiles_n, tile_start;   // tiles of this problem are blockIdx.x in [tile_start, tile_start + tiles_m * tiles_n)
    int vec32, vec16, vecr, vecg;  // 128-bit access allowed on C32 rows / 64-bit on C16 / resid / gate
    float alpha, beta, gate_scale, drop_p;
    unsigned int salt;
};

struct GbParams {
    GbProblem p[GB_MAXP];
    int n, cluster, nstages;
};

__device__ __forceinline__ uint2 ld_bf16x4(const __nv_bfloat16* p) { return *reinterpret_cast<const uint2*>(p); }
__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack2_bf16(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}

template <int BN>
__global__ void __launch_bounds__(192) gemm_bf16_tc_kernel(const __grid_constant__ GbParams P) {
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
    const GbProblem& p = P.p[pi];
    const int t = blockIdx.x - p.tile_start;
    const int tm = t / p.tiles_n, tn = t - tm * p.tiles_n;
    const int m0 = tm * GB_BM, n0 = tn * BN;
    const int kb0 = blockIdx.z * p.kb_per_split;
    const int kb1 = min(p.num_kb, kb0 + p.kb_per_split);
    const bool has_k = kb1 > kb0;   // a split without k-blocks contributes zeros (cluster mode only)

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmA);
        tma_prefetch_desc(&p.tmB);
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

    if (warp == 0) {
        if (elect_one()) {  // ===== TMA producer
            int s = 0;
            uint32_t ph = 0;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(empty_bar(s), ph ^ 1u);
                mbar_arrive_expect_tx(full_bar(s), STAGE_BYTES);
                const uint32_t sA = base + s * STAGE_BYTES, sB = sA + GB_A_BYTES;
                const int k = kb * GB_BK;
                if (!p.a_mn) {
                    tma_load_2d(sA, &p.tmA, full_bar(s), k, m0);            // box {64 k, 128 m}
                } else {
#pragma unroll
                    for (int c = 0; c < GB_BM / 64; ++c)                    // box {64 m, 64 k} per 128-byte MN chunk
                        tma_load_2d(sA + c * 8192, &p.tmA, full_bar(s), m0 + c * 64, k);
                }
                if (!p.b_mn) {
                    tma_load_2d(sB, &p.tmB, full_bar(s), k, n0);            // box {64 k, BN n}
                } else {
#pragma unroll
                    for (int c = 0; c < BN / 64; ++c)
                        tma_load_2d(sB + c * 8192, &p.tmB, full_bar(s), n0 + c * 64, k);
                }
                if (++s == STAGES) { s = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {  // ===== MMA issuer
            const uint32_t idesc = make_idesc(1, GB_BM, BN, p.a_mn, p.b_mn);
            const uint64_t ad0 = p.a_mn ? make_smem_desc(base, 8192, 1024) : make_smem_desc(base, 16, 1024);
            const uint64_t bd0 = p.b_mn ? make_smem_desc(base + GB_A_BYTES, 8192, 1024)
                                        : make_smem_desc(base + GB_A_BYTES, 16, 1024);
            const uint32_t a_step = p.a_mn ? 2048u : 32u, b_step = p.b_mn ? 2048u : 32u;
            int s = 0;
            uint32_t ph = 0;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(full_bar(s), ph);
                tc_fence_after();
                const uint64_t ad = desc_advance(ad0, s * STAGE_BYTES), bd = desc_advance(bd0, s * STAGE_BYTES);
#pragma unroll
                for (int k = 0; k < GB_BK / 16; ++k)  // UMMA_K = 16 for bf16 (32 bytes of K)
                    mma_bf16_ss(tmem_base, desc_advance(ad, k * a_step), desc_advance(bd, k * b_step), idesc,
                                (kb > kb0 || k > 0) ? 1u : 0u);
                tc_commit(empty_bar(s));  // frees the smem slot once these MMAs have read it
                if (++s == STAGES) { s = 0; ph ^= 1u; }
            }
            tc_commit(tmem_full_bar);  // accumulator complete (arrives at once when nothing was issued)
        }
    }
    // ===== epilogue warps 2..5: TMEM lane quarter = warp % 4, thread = output row
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int m = m0 + row;
    const uint32_t lane_taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    const int csz = P.cluster > 1 ? P.cluster : 1;
    const int W = BN / csz;  // columns of the tile this CTA finishes (cluster mode)
    const uint32_t rank = P.cluster > 1 ? cluster_ctarank() : 0u;
    if (warp >= 2) {
        mbar_wait(tmem_full_bar, 0);
        tc_fence_after();
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
    if (warp >= 2) {
        const float ks = p.drop_p > 0.f ? 1.f / (1.f - p.drop_p) : 1.f;
        // destination row of the fp32 output (optionally one of up to three row segments)
        float* c32row = nullptr;
        if (p.C32 && m < p.M) {
            if (p.seg_rows > 0) {
                const int sgi = m / p.seg_rows;
                float* b = sgi == 0 ? p.C32 : (sgi == 1 ? p.seg1 : p.seg2);
                c32row = b + (long long)(m - sgi * p.seg_rows) * p.ldc32;
            } else {
                c32row = p.C32 + (long long)m * p.ldc32;
            }
        }
#pragma unroll 1
        for (int cc = 0; cc < W; cc += 32) {
            const int c0 = (int)rank * W + cc;
            uint32_t r[32];
            if (has_k) {
                tmem_ld_32x32(lane_taddr + (uint32_t)c0, r);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) r[i] = 0u;
            }
            if (P.cluster > 1) {
                for (uint32_t src = 0; src < (uint32_t)csz; ++src) {
                    if (src == rank) continue;
                    const uint32_t sb = recv + (src * (uint32_t)(W / 4)) * (GB_BM * 16u);
#pragma unroll
                    for (int i4 = 0; i4 < 8; ++i4) {
                        float4 tt = ld_shared_v4(sb + ((uint32_t)(cc / 4 + i4) * GB_BM + row) * 16u);
                        r[4 * i4] = __float_as_uint(__uint_as_float(r[4 * i4]) + tt.x);
                        r[4 * i4 + 1] = __float_as_uint(__uint_as_float(r[4 * i4 + 1]) + tt.y);
                        r[4 * i4 + 2] = __float_as_uint(__uint_as_float(r[4 * i4 + 2]) + tt.z);
                        r[4 * i4 + 3] = __float_as_uint(__uint_as_float(r[4 * i4 + 3]) + tt.w);
                    }
                }
            }
            const int nb = n0 + c0;
            if (m < p.M && nb < p.N) {
                const float* rrow = p.resid ? p.resid + (long long)m * p.ldr + nb : nullptr;
                const __nv_bfloat16* grow = p.gate ? p.gate + (long long)m * p.ldg + nb : nullptr;
                float* crow = c32row ? c32row + nb : nullptr;
                __nv_bfloat16* hrow = p.C16 ? p.C16 + (long long)m * p.ldc16 + nb : nullptr;
#pragma unroll
                for (int j4 = 0; j4 < 32; j4 += 4) {
                    const int n = nb + j4;
                    if (n >= p.N) break;
                    const bool full = n + 3 < p.N;
                    float v[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) v[j] = p.alpha * __uint_as_float(r[j4 + j]);
                    if (p.bias) {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (n + j < p.N) v[j] += p.bias[n + j];
                    }
                    if (p.act == 1) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) v[j] = fmaxf(v[j], 0.f);
                    }
                    if (grow) {
                        if (full && p.vecg) {
                            const uint2 g = ld_bf16x4(grow + j4);
                            v[0] = bf16_lo(g.x) > 0.f ? v[0] * p.gate_scale : 0.f;
                            v[1] = bf16_hi(g.x) > 0.f ? v[1] * p.gate_scale : 0.f;
                            v[2] = bf16_lo(g.y) > 0.f ? v[2] * p.gate_scale : 0.f;
                            v[3] = bf16_hi(g.y) > 0.f ? v[3] * p.gate_scale : 0.f;
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (n + j < p.N)
                                    v[j] = __bfloat162float(grow[j4 + j]) > 0.f ? v[j] * p.gate_scale : 0.f;
                        }
                    }
                    if (p.drop_p > 0.f) {
                        const unsigned long long e0 = (unsigned long long)m * p.N + n;
                        if ((e0 & 3ull) == 0) {  // the 4 elements share one Philox block
                            const Philox4 rb = philox_block(p.rng, p.salt, e0 >> 2);
                            v[0] = u32_to_unit(rb.x) >= p.drop_p ? v[0] * ks : 0.f;
                            v[1] = u32_to_unit(rb.y) >= p.drop_p ? v[1] * ks : 0.f;
                            v[2] = u32_to_unit(rb.z) >= p.drop_p ? v[2] * ks : 0.f;
                            v[3] = u32_to_unit(rb.w) >= p.drop_p ? v[3] * ks : 0.f;
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                v[j] = u32_to_unit(philox_u32(p.rng, p.salt, e0 + j)) >= p.drop_p ? v[j] * ks : 0.f;
                        }
                    }
                    if (rrow) {
                        if (full && p.vecr) {
                            const float4 tt = ld_f4(rrow + j4);
                            v[0] += tt.x; v[1] += tt.y; v[2] += tt.z; v[3] += tt.w;
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (n + j < p.N) v[j] += rrow[j4 + j];
                        }
                    }
                    if (crow) {
                        if (full && p.vec32) {
                            if (p.beta != 0.f) {
                                const float4 tt = ld_f4(crow + j4);
                                v[0] += p.beta * tt.x; v[1] += p.beta * tt.y; v[2] += p.beta * tt.z; v[3] += p.beta * tt.w;
                            }
                            st_f4(crow + j4, make_float4(v[0], v[1], v[2], v[3]));
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (n + j < p.N) {
                                    if (p.beta != 0.f) v[j] += p.beta * crow[j4 + j];
                                    crow[j4 + j] = v[j];
                                }
                        }
                    }
                    if (hrow) {
                        if (full && p.vec16) {
                            *reinterpret_cast<uint2*>(hrow + j4) = make_uint2(pack2_bf16(v[0], v[1]), pack2_bf16(v[2], v[3]));
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (n + j < p.N) hrow[j4 + j] = __float2bfloat16(v[j]);
                        }
                    }
                }
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ---------------------------------------------------------------------------------------------- host side
static int g_gb_force_bn = 0, g_gb_force_split = 0, g_gb_force_stages = 0;

template <int BN>
static int launch_gb(const GbParams& P, int total_tiles, int split, cudaStream_t st) {
    const int stage_bytes = GB_A_BYTES + BN * 128;
    const int smem = P.nstages * stage_bytes + 1024 + 8 * (2 * P.nstages + 2) + (P.cluster > 1 ? GB_BM * BN * 4 : 0);
    static DeviceOnce attr;
    if (needs_setup(attr)) {
        cudaError_t e = cudaFuncSetAttribute(gemm_bf16_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
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
    cudaLaunchAttribute at[1];
    if (P.cluster > 1) {
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 1;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = P.cluster;
        cfg.attrs = at;
        cfg.numAttrs = 1;
    }
    cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_bf16_tc_kernel<BN>, P);
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
    int BN = 128, split = 1;
    if (max_n <= 64) BN = 64;
    else if (t128 < n_sm / 2 && t64 <= 2 * n_sm && min_kb <= 4) BN = 64;  // short K: nothing to split, use narrow tiles
    {
        const long long tiles = BN == 128 ? t128 : t64;
        // a cluster of `split` CTAs along K when the tiles alone leave most SMs idle and every problem has K to split
        while (split < 4 && tiles * split * 2 <= n_sm + n_sm / 4 && min_kb / (split * 2) >= 2 && BN / (split * 2) >= 32) split *= 2;
    }
    static const int env_bn = getenv("BMHRL_GB_BN") ? atoi(getenv("BMHRL_GB_BN")) : 0;
    static const int env_split = getenv("BMHRL_GB_SPLIT") ? atoi(getenv("BMHRL_GB_SPLIT")) : 0;
    if (env_bn == 64 || env_bn == 128) BN = env_bn;
    if (env_split == 1 || env_split == 2 || env_split == 4) split = env_split;
    if (g_gb_force_bn) BN = g_gb_force_bn;
    if (g_gb_force_split) split = g_gb_force_split;
    while (split > 1 && (BN / split < 32 || min_kb < split)) split >>= 1;
    static const int env_stages = getenv("BMHRL_GB_STAGES") ? atoi(getenv("BMHRL_GB_STAGES")) : 0;  // experiments
    int nstages = g_gb_force_stages ? g_gb_force_stages : (env_stages ? env_stages : (BN == 128 ? 4 : 6));
    {
        const int kb_per_cta = (max_kb + split - 1) / split;
        if (nstages > kb_per_cta) nstages = kb_per_cta < 2 ? 2 : kb_per_cta;
        if (split > 1 && nstages > 3) nstages = 3;  // room for the DSMEM receive buffer beside a co-resident kernel
    }

    GbParams P{};
    P.n = n;
    P.cluster = split;
    P.nstages = nstages;
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
