// Tensor-core GEMM for sm_100a: TMA-staged operands, tcgen05.mma (kind::tf32) accumulating in TMEM, fused
// linear-layer epilogue.  Replaces the same call sites as gemm_simt.cu (projections model/multihead_attention.py:
// 70-72,89; FFN model/blocks.py:175-187; vocab projection model/bm_hrl_agent.py:463-466 and their dX / dW
// gradients) when the library's math mode is TF32.
//
// The operands stay the fp32 tensors PyTorch owns -- no bf16 shadow copies, no conversion kernels: TMA reads fp32
// tiles straight into 128B-swizzled shared memory (converted to tf32 on the fly by the tensor map's TFLOAT32 type),
// so awkward row lengths (300, 364, 10172: all 16-byte multiples in fp32) need no padded storage and every
// K / M / N tail is zero-filled by TMA.
//
//   C(m,n) = [beta*C +] resid + dropout(act(alpha * sum_k A(m,k) B(k,n) + bias[n]))      or, split-K:  C += alpha * partial
//
// A is K-major (row m contiguous in k: activations, dZ for dX) or MN-major (row k contiguous in m: dZ^T for dW);
// B is K-major (nn.Linear weight (N,K) in forward) or MN-major (weight in dX = dZ W, X in dW = dZ^T X).
//
// CTA = 192 threads: warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer (one elected thread), warps 2-5
// epilogue (tcgen05.ld -> registers -> global).  Tile 128 x BN x 32 (32 fp32 = one 128-byte swizzle row),
// STAGES-deep mbarrier ring.  One output tile (x one K split) per CTA.
#include <stdio.h>
#include <stdlib.h>
#include "common.cuh"
#include "tc_common.cuh"
#include "gemm_simt.cuh"
#include "../../include/bmhrl_b200.h"

namespace bmhrl {

using namespace tc;

constexpr int TC_BM = 128, TC_BK = 32;
constexpr int TC_A_BYTES = TC_BM * 128;  // 16 KB per stage

struct TcEpilogue {
    float* C;
    long long ldc;
    int M, N;
    int num_kb, kb_per_split, split_k;
    int a_mn, b_mn;
    float alpha, beta;
    const float* bias;
    int act;
    const float* resid;
    long long ldr;
    float drop_p;
    const unsigned long long* rng;
    unsigned int salt;
    int vec_ok;  // C (and resid) rows allow 128-bit access
    int cluster;             // > 1: the split_k CTAs of a tile form a thread-block cluster along z and reduce their
                             // partial accumulators through distributed shared memory (no atomics, any epilogue)
    int batched, Hdiv;       // blockIdx.z = z * split_k + split; z -> (outer = z / Hdiv, inner = z % Hdiv)
    long long bCo, bCi;      // C offset of batch entry z, elements
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(192) gemm_tf32_tc_kernel(const __grid_constant__ CUtensorMap tmA,
                                                           const __grid_constant__ CUtensorMap tmB,
                                                           const TcEpilogue p) {
    constexpr int B_BYTES = BN * 128;
    constexpr int STAGE_BYTES = TC_A_BYTES + B_BYTES;
    constexpr int TMEM_COLS = BN < 32 ? 32 : BN;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B tiles need 1024-byte alignment
    const uint32_t recv = base + STAGES * STAGE_BYTES;            // cluster mode: [src rank][col/4][row] float4 partials
    const uint32_t bars = recv + (p.cluster > 1 ? TC_BM * BN * 4 : 0);  // full[STAGES], empty[STAGES], tmem_full, tmem ptr
    auto full_bar = [&](int s) { return bars + 8u * s; };
    auto empty_bar = [&](int s) { return bars + 8u * (STAGES + s); };
    const uint32_t tmem_full_bar = bars + 8u * (2 * STAGES);
    const uint32_t tmem_slot = bars + 8u * (2 * STAGES + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.y * TC_BM, n0 = blockIdx.x * BN;
    const int zb = blockIdx.z / p.split_k;                 // batch entry
    const int kb0 = (blockIdx.z - zb * p.split_k) * p.kb_per_split;
    const int kb1 = min(p.num_kb, kb0 + p.kb_per_split);
    const int z_out = zb / p.Hdiv, z_in = zb - z_out * p.Hdiv;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
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
    if (p.cluster > 1) cluster_sync_all();  // every CTA of the cluster is running before anyone writes into its smem
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

    if (warp == 0) {
        if (elect_one()) {  // ===== TMA producer
            for (int kb = kb0, it = 0; kb < kb1; ++kb, ++it) {
                const int s = it % STAGES;
                const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
                mbar_wait(empty_bar(s), ph ^ 1u);
                mbar_arrive_expect_tx(full_bar(s), STAGE_BYTES);
                const uint32_t sA = base + s * STAGE_BYTES, sB = sA + TC_A_BYTES;
                const int k = kb * TC_BK;
                if (p.batched) {
                    if (!p.a_mn) {
                        tma_load_4d(sA, &tmA, full_bar(s), k, m0, z_in, z_out);
                    } else {
#pragma unroll
                        for (int c = 0; c < TC_BM / 32; ++c)
                            tma_load_4d(sA + c * (TC_BK * 128), &tmA, full_bar(s), m0 + c * 32, k, z_in, z_out);
                    }
                    if (!p.b_mn) {
                        tma_load_4d(sB, &tmB, full_bar(s), k, n0, z_in, z_out);
                    } else {
#pragma unroll
                        for (int c = 0; c < BN / 32; ++c)
                            tma_load_4d(sB + c * (TC_BK * 128), &tmB, full_bar(s), n0 + c * 32, k, z_in, z_out);
                    }
                    continue;
                }
                if (!p.a_mn) {
                    tma_load_2d(sA, &tmA, full_bar(s), k, m0);  // box {32 k, 128 m}
                } else {
#pragma unroll
                    for (int c = 0; c < TC_BM / 32; ++c)        // box {32 m, 32 k} per 128-byte MN chunk
                        tma_load_2d(sA + c * (TC_BK * 128), &tmA, full_bar(s), m0 + c * 32, k);
                }
                if (!p.b_mn) {
                    tma_load_2d(sB, &tmB, full_bar(s), k, n0);  // box {32 k, BN n}
                } else {
#pragma unroll
                    for (int c = 0; c < BN / 32; ++c)
                        tma_load_2d(sB + c * (TC_BK * 128), &tmB, full_bar(s), n0 + c * 32, k);
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {  // ===== MMA issuer
            const uint32_t idesc = make_idesc(2, TC_BM, BN, p.a_mn, p.b_mn);
            // stage-0 descriptors; a stage / K step further on is one add (see desc_advance)
            const uint64_t ad0 = p.a_mn ? make_smem_desc(base, TC_BK * 128, 512, 1) : make_smem_desc(base, 16, 1024);
            const uint64_t bd0 = p.b_mn ? make_smem_desc(base + TC_A_BYTES, TC_BK * 128, 512, 1)
                                        : make_smem_desc(base + TC_A_BYTES, 16, 1024);
            const uint32_t a_step = p.a_mn ? 1024u : 32u, b_step = p.b_mn ? 1024u : 32u;
            for (int kb = kb0, it = 0; kb < kb1; ++kb, ++it) {
                const int s = it % STAGES;
                const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
                mbar_wait(full_bar(s), ph);
                tc_fence_after();
                const uint64_t ad = desc_advance(ad0, s * STAGE_BYTES), bd = desc_advance(bd0, s * STAGE_BYTES);
#pragma unroll
                for (int k = 0; k < TC_BK / 8; ++k)  // UMMA_K = 8 for tf32 (32 bytes of K)
                    mma_tf32_ss(tmem_base, desc_advance(ad, k * a_step), desc_advance(bd, k * b_step), idesc,
                                (it > 0 || k > 0) ? 1u : 0u);
                tc_commit(empty_bar(s));  // frees the smem slot once these MMAs have read it
            }
            tc_commit(tmem_full_bar);  // accumulator complete
        }
    }
    // ===== epilogue warps 2..5: TMEM lane quarter = warp % 4, thread = output row
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int m = m0 + row;
    const uint32_t lane_taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    const int csz = p.cluster > 1 ? p.cluster : 1;
    const int W = BN / csz;                         // columns of the tile this CTA finishes (cluster mode)
    const uint32_t rank = p.cluster > 1 ? cluster_ctarank() : 0u;
    if (warp >= 2) {
        mbar_wait(tmem_full_bar, 0);
        tc_fence_after();
        if (p.cluster > 1) {
            // reduce-scatter, send side: slice r of my partial accumulator goes to CTA r's receive buffer, slot `rank`
            for (uint32_t r = 0; r < (uint32_t)csz; ++r) {
                if (r == rank) continue;
                const uint32_t rbase = mapa_shared(recv, r) + (rank * (uint32_t)(W / 4)) * (TC_BM * 16u);
#pragma unroll 1
                for (int c0 = 0; c0 < W; c0 += 32) {
                    uint32_t v[32];
                    tmem_ld_32x32(lane_taddr + (uint32_t)(r * W + c0), v);
                    tmem_ld_wait();
#pragma unroll
                    for (int i4 = 0; i4 < 8; ++i4)
                        st_shared_cluster_v4(rbase + ((uint32_t)(c0 / 4 + i4) * TC_BM + row) * 16u, v[4 * i4],
                                             v[4 * i4 + 1], v[4 * i4 + 2], v[4 * i4 + 3]);
                }
            }
        }
    }
    if (p.cluster > 1) {
        __syncwarp();
        cluster_sync_all();  // release / acquire: all partial slices have landed
    }
    if (warp >= 2) {
        const float ks = p.drop_p > 0.f ? 1.f / (1.f - p.drop_p) : 1.f;
#pragma unroll 1
        for (int cc = 0; cc < W; cc += 32) {
            const int c0 = (int)rank * W + cc;
            uint32_t r[32];
            tmem_ld_32x32(lane_taddr + (uint32_t)c0, r);
            tmem_ld_wait();
            if (p.cluster > 1) {
                for (uint32_t src = 0; src < (uint32_t)csz; ++src) {
                    if (src == rank) continue;
                    const uint32_t sb = recv + (src * (uint32_t)(W / 4)) * (TC_BM * 16u);
#pragma unroll
                    for (int i4 = 0; i4 < 8; ++i4) {
                        float4 t = ld_shared_v4(sb + ((uint32_t)(cc / 4 + i4) * TC_BM + row) * 16u);
                        r[4 * i4] = __float_as_uint(__uint_as_float(r[4 * i4]) + t.x);
                        r[4 * i4 + 1] = __float_as_uint(__uint_as_float(r[4 * i4 + 1]) + t.y);
                        r[4 * i4 + 2] = __float_as_uint(__uint_as_float(r[4 * i4 + 2]) + t.z);
                        r[4 * i4 + 3] = __float_as_uint(__uint_as_float(r[4 * i4 + 3]) + t.w);
                    }
                }
            }
            const int nb = n0 + c0;
            if (m < p.M && nb < p.N) {
                float* crow = p.C + (long long)z_out * p.bCo + (long long)z_in * p.bCi + (long long)m * p.ldc + nb;
                const float* rrow = p.resid ? p.resid + (long long)m * p.ldr + nb : nullptr;
                if (p.split_k > 1 && p.cluster <= 1) {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (nb + j < p.N) atomicAdd(crow + j, p.alpha * __uint_as_float(r[j]));
                } else {
#pragma unroll
                    for (int j4 = 0; j4 < 32; j4 += 4) {
                        float v[4];
                        const unsigned long long e0 = ((unsigned long long)zb * p.M + m) * p.N + nb + j4;
                        const bool one_block = p.drop_p > 0.f && (e0 & 3ull) == 0;  // 4 elements, one Philox block
                        Philox4 rb{};
                        if (one_block) rb = philox_block(p.rng, p.salt, e0 >> 2);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const int n = nb + j4 + j;
                            float x = p.alpha * __uint_as_float(r[j4 + j]);
                            if (p.bias && n < p.N) x += p.bias[n];
                            if (p.act == 1) x = fmaxf(x, 0.f);
                            if (p.drop_p > 0.f) {
                                const uint32_t u = one_block ? (j == 0 ? rb.x : j == 1 ? rb.y : j == 2 ? rb.z : rb.w)
                                                             : philox_u32(p.rng, p.salt, e0 + j);
                                x = u32_to_unit(u) >= p.drop_p ? x * ks : 0.f;
                            }
                            v[j] = x;
                        }
                        const int n = nb + j4;
                        if (p.vec_ok && n + 3 < p.N) {
                            if (rrow) { float4 t = ld_f4(rrow + j4); v[0] += t.x; v[1] += t.y; v[2] += t.z; v[3] += t.w; }
                            if (p.beta != 0.f) {
                                float4 t = ld_f4(crow + j4);
                                v[0] += p.beta * t.x; v[1] += p.beta * t.y; v[2] += p.beta * t.z; v[3] += p.beta * t.w;
                            }
                            st_f4(crow + j4, make_float4(v[0], v[1], v[2], v[3]));
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                if (n + j < p.N) {
                                    float x = v[j];
                                    if (rrow) x += rrow[j4 + j];
                                    if (p.beta != 0.f) x += p.beta * crow[j4 + j];
                                    crow[j4 + j] = x;
                                }
                            }
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
static int g_math_mode = 0;  // 0: exact fp32 SIMT; 1: TF32 tensor cores (tcgen05)
static int g_force_bn = 0, g_force_split = 0;
static long long g_tc_batched_min_macs = 4ll << 20;

static bool tc_operand_ok(const float* p, long long ld) { return aligned16(p) && ld > 0 && (ld & 3) == 0; }

bool tc_gemm_supported(const TcGemmArgs& a) {
    if (a.M <= 0 || a.N <= 0 || a.K <= 0) return false;
    if (a.batch > 1) {
        if (a.resid || a.bias) return false;
        if (((a.bAo | a.bAi | a.bBo | a.bBi) & 3) != 0 || a.bAo < 0 || a.bAi < 0 || a.bBo < 0 || a.bBi < 0) return false;
        if (a.Hdiv < 1 || a.batch % a.Hdiv != 0) return false;
        if ((a.Hdiv > 1 && (a.bAi == 0 || a.bBi == 0)) || (a.batch > a.Hdiv && (a.bAo == 0 || a.bBo == 0))) return false;
    }
    return tc_operand_ok(a.A, a.lda) && tc_operand_ok(a.B, a.ldb) && get_encode_fn() != nullptr;
}

template <int BN, int STAGES>
static int launch_tc(const CUtensorMap& tmA, const CUtensorMap& tmB, const TcEpilogue& ep, dim3 grid, cudaStream_t st) {
    constexpr int SMEM_BASE = STAGES * (TC_A_BYTES + BN * 128) + 1024 + 8 * (2 * STAGES + 2);
    constexpr int SMEM_MAX = SMEM_BASE + TC_BM * BN * 4;  // + the cluster receive buffer
    static DeviceOnce attr;
    if (needs_setup(attr)) {
        cudaError_t e = cudaFuncSetAttribute(gemm_tf32_tc_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             SMEM_MAX);
        if (e != cudaSuccess) {
            bmhrl_set_error(cudaGetErrorString(e));
            return BMHRL_ERR_CUDA;
        }
        mark_done(attr);
    }
    if (ep.cluster > 1) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = grid;
        cfg.blockDim = dim3(192);
        cfg.dynamicSmemBytes = SMEM_MAX;
        cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 1;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = ep.cluster;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_tf32_tc_kernel<BN, STAGES>, tmA, tmB, ep);
        if (e != cudaSuccess) {
            bmhrl_set_error(cudaGetErrorString(e));
            return BMHRL_ERR_CUDA;
        }
        bmhrl_count_launch(1);
        return BMHRL_OK;
    }
    gemm_tf32_tc_kernel<BN, STAGES><<<grid, 192, SMEM_BASE, st>>>(tmA, tmB, ep);
    BMHRL_POST_LAUNCH("gemm_tf32_tc");
    return BMHRL_OK;
}

int launch_gemm_tc(const TcGemmArgs& a, cudaStream_t st) {
    const int tiles_m = (a.M + TC_BM - 1) / TC_BM;
    const int num_kb = (a.K + TC_BK - 1) / TC_BK;
    // tile width: the widest BN that still gives >= ~1 CTA per SM; narrow tiles re-read A but fill the machine
    const int nbatch = a.batch > 1 ? a.batch : 1;
    // Configuration: every SM ingests ~70 GB/s from L2 (measured), so the time of these small GEMMs is
    //   slots_per_SM * bytes_per_CTA / 70 GB/s + fixed costs,  bytes_per_CTA = (128 + BN) * 128 B * k-blocks per CTA.
    // Wide tiles cut the bytes, a cluster of `cs` CTAs along K (DSMEM reduce-scatter in the epilogue) keeps all SMs busy.
    const bool plain = !a.bias && a.act == 0 && !a.resid && a.drop_p == 0.f;
    int BN = 128, split = 1, cluster = 1;
    if (g_force_bn || g_force_split) {
        if (g_force_bn) BN = g_force_bn;
        else {
            while (BN > 32 && (long long)nbatch * tiles_m * ((a.N + BN - 1) / BN) < 120) BN >>= 1;
            if (a.N <= 32) BN = 32; else if (a.N <= 64 && BN > 64) BN = 64;
        }
        if (g_force_split < 0 && nbatch == 1) {  // forced cluster size
            cluster = -g_force_split;
            while (cluster > 1 && (BN / cluster < 32 || num_kb < cluster)) cluster >>= 1;
            split = cluster;
        } else if (g_force_split > 0 && plain && nbatch == 1) {
            split = g_force_split;  // atomic split-K
        }
    } else {
        // Measured on B200 (tools/microbench.py gemm): a lone CTA per SM ingests ~70 GB/s from L2, so when 128-wide
        // tiles alone would leave most SMs idle and K is long, a cluster of 2 / 4 CTAs splits K of each 128x128 tile
        // (2.5x fewer bytes per SM than 128x32 tiles); otherwise narrow tiles with 2 co-resident CTAs per SM win.
        const long long t128 = (long long)tiles_m * ((a.N + 127) / 128);
        static const int allow_cluster = getenv("BMHRL_TC_CLUSTER") ? atoi(getenv("BMHRL_TC_CLUSTER")) : 0;
        static const int min_tiles = getenv("BMHRL_TC_MIN_TILES") ? atoi(getenv("BMHRL_TC_MIN_TILES")) : 120;
        if (allow_cluster && nbatch == 1 && a.N > 64 && num_kb >= 16 && t128 * 2 <= 148) {
            BN = 128;
            cluster = t128 * 4 <= 160 ? 4 : 2;
            split = cluster;
        } else {
            while (BN > 32 && (long long)nbatch * tiles_m * ((a.N + BN - 1) / BN) < min_tiles) BN >>= 1;
            if (a.N <= 32) BN = 32; else if (a.N <= 64 && BN > 64) BN = 64;
            if (plain && nbatch == 1) {
                const long long tiles = (long long)tiles_m * ((a.N + BN - 1) / BN);
                while (tiles * split < 100 && num_kb / (split * 2) >= 4) split *= 2;
            }
        }
    }
    if (split > num_kb) split = num_kb;
    const int tiles_n = (a.N + BN - 1) / BN;
    int kb_per = (num_kb + split - 1) / split;
    if (cluster > 1) {
        if ((long long)kb_per * (cluster - 1) >= num_kb) { cluster = 1; split = 1; kb_per = num_kb; }  // no empty CTA
    } else {
        split = (num_kb + kb_per - 1) / kb_per;  // no empty splits
    }

    CUtensorMap tmA, tmB;
    bool ok;
    const CUtensorMapDataType TF = CU_TENSOR_MAP_DATA_TYPE_TFLOAT32;
    if (nbatch > 1) {
        const unsigned long long hd = a.Hdiv, ob = nbatch / a.Hdiv;
        if (!a.a_mn) ok = encode_4d(&tmA, TF, 4, a.A, a.K, a.M, a.lda, hd, a.bAi, ob, a.bAo, TC_BK, TC_BM);
        else ok = encode_4d(&tmA, TF, 4, a.A, a.M, a.K, a.lda, hd, a.bAi, ob, a.bAo, 32, TC_BK,
                            CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
        if (ok) {
            if (!a.b_mn) ok = encode_4d(&tmB, TF, 4, a.B, a.K, a.N, a.ldb, hd, a.bBi, ob, a.bBo, TC_BK, BN);
            else ok = encode_4d(&tmB, TF, 4, a.B, a.N, a.K, a.ldb, hd, a.bBi, ob, a.bBo, 32, TC_BK,
                                CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
        }
    } else {
    if (!a.a_mn) ok = encode_2d(&tmA, CU_TENSOR_MAP_DATA_TYPE_TFLOAT32, 4, a.A, a.K, a.M, a.lda, TC_BK, TC_BM);
    else ok = encode_2d(&tmA, CU_TENSOR_MAP_DATA_TYPE_TFLOAT32, 4, a.A, a.M, a.K, a.lda, 32, TC_BK,
                        CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
    if (ok) {
        if (!a.b_mn) ok = encode_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_TFLOAT32, 4, a.B, a.K, a.N, a.ldb, TC_BK, BN);
        else ok = encode_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_TFLOAT32, 4, a.B, a.N, a.K, a.ldb, 32, TC_BK,
                            CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
    }
    }
    if (!ok) {
        char msg[384];
        snprintf(msg, sizeof(msg), "cuTensorMapEncodeTiled failed (M=%d N=%d K=%d batch=%d Hdiv=%d a_mn=%d b_mn=%d lda=%lld "
                 "ldb=%lld bAo=%lld bAi=%lld bBo=%lld bBi=%lld BN=%d rc=%d ptr=%p A=%p B=%p)", a.M, a.N, a.K, nbatch, a.Hdiv, a.a_mn, a.b_mn, a.lda,
                 a.ldb, a.bAo, a.bAi, a.bBo, a.bBi, BN, g_last_encode_rc, g_last_encode_ptr, (const void*)a.A, (const void*)a.B);
        bmhrl_set_error(msg);
        return BMHRL_ERR_CUDA;
    }
    TcEpilogue ep{};
    ep.C = a.C; ep.ldc = a.ldc; ep.M = a.M; ep.N = a.N;
    ep.num_kb = num_kb; ep.kb_per_split = kb_per; ep.split_k = split;
    ep.a_mn = a.a_mn; ep.b_mn = a.b_mn; ep.alpha = a.alpha; ep.beta = a.beta;
    ep.bias = a.bias; ep.act = a.act; ep.resid = a.resid; ep.ldr = a.ldr;
    ep.drop_p = a.drop_p; ep.rng = a.rng; ep.salt = a.salt;
    ep.vec_ok = aligned16(a.C) && (a.ldc & 3) == 0 && (!a.resid || (aligned16(a.resid) && (a.ldr & 3) == 0)) &&
                (nbatch == 1 || (((a.bCo | a.bCi) & 3) == 0));
    ep.cluster = cluster;
    ep.batched = nbatch > 1; ep.Hdiv = nbatch > 1 ? a.Hdiv : 1; ep.bCo = a.bCo; ep.bCi = a.bCi;
    if (split > 1 && cluster <= 1 && a.beta == 0.f) {  // partial sums are accumulated atomically into a zeroed C
        if (a.ldc == a.N) {
            if (cudaMemsetAsync(a.C, 0, sizeof(float) * (size_t)a.M * a.N, st) != cudaSuccess) return BMHRL_ERR_CUDA;
        } else {
            if (cudaMemset2DAsync(a.C, a.ldc * sizeof(float), 0, a.N * sizeof(float), a.M, st) != cudaSuccess)
                return BMHRL_ERR_CUDA;
        }
    }
    dim3 grid(tiles_n, tiles_m, split * nbatch);
    if (BN == 128) return launch_tc<128, 4>(tmA, tmB, ep, grid, st);
    if (BN == 64) return launch_tc<64, 4>(tmA, tmB, ep, grid, st);
    return launch_tc<32, 5>(tmA, tmB, ep, grid, st);
}

int math_mode() { return g_math_mode; }

int launch_gemm_auto(GemmParams& p, int batch, cudaStream_t st) {
    if (g_math_mode == 1 && !p.bias && !p.resid && p.act == 0 && (p.sAk == 1 || p.sAm == 1) &&
        (p.sBk == 1 || p.sBn == 1) && (p.beta == 0.f || (p.beta == 1.f && p.drop_p == 0.f))) {
        TcGemmArgs t{};
        t.A = p.A; t.a_mn = (p.sAk == 1) ? 0 : 1; t.lda = t.a_mn ? p.sAk : p.sAm;
        t.B = p.B; t.b_mn = (p.sBk == 1) ? 0 : 1; t.ldb = t.b_mn ? p.sBk : p.sBn;
        t.C = p.C; t.ldc = p.ldc; t.M = p.M; t.N = p.N; t.K = p.K; t.alpha = p.alpha; t.beta = p.beta;
        t.drop_p = p.drop_p; t.rng = p.rng; t.salt = p.salt;
        t.batch = batch; t.Hdiv = p.Hdiv < 1 ? 1 : p.Hdiv;
        t.bAo = p.bAo; t.bAi = p.bAi; t.bBo = p.bBo; t.bBi = p.bBi; t.bCo = p.bCo; t.bCi = p.bCi;
        // batched attention-sized contractions: below ~4 M MACs per entry a tensor-core CTA is all prologue /
        // epilogue latency and the SIMT kernel (48 KB of shared memory, many CTAs per SM) is faster
        const bool tiny = batch > 1 && (long long)p.M * p.N * p.K < g_tc_batched_min_macs;
        if (!tiny && tc_gemm_supported(t)) return launch_gemm_tc(t, st);
    }
    gemm_set_vec_flags(p);
    return launch_gemm_simt(p, batch, st);
}

}  // namespace bmhrl

using namespace bmhrl;

extern "C" int bmhrl_set_math_mode(int mode) {
    BMHRL_CHECK_ARG(mode == 0 || mode == 1, "bmhrl_set_math_mode: 0 = fp32 (SIMT), 1 = tf32 (tcgen05)");
    g_math_mode = mode;
    return BMHRL_OK;
}
extern "C" int bmhrl_get_math_mode(void) { return g_math_mode; }
extern "C" int bmhrl_tc_tuning(int force_bn, int force_split) {
    BMHRL_CHECK_ARG(force_bn == 0 || force_bn == 32 || force_bn == 64 || force_bn == 128, "bmhrl_tc_tuning: BN");
    BMHRL_CHECK_ARG(force_split >= -4 && force_split != -3, "bmhrl_tc_tuning: split (>0 atomic split-K, -2/-4 cluster)");
    g_force_bn = force_bn;
    g_force_split = force_split;
    g_tc_batched_min_macs = (force_bn || force_split) ? 0 : (4ll << 20);  // forced tiling: always the tensor-core kernel
    return BMHRL_OK;
}
