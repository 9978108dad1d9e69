// K1 fused masked multi-head attention core, forward, on the 5th-gen tensor cores (sm_100a).
// Reference: attention() model/multihead_attention.py:7-31 -- softmax(mask_replace(Q K^T / sqrt(d_k), -1e9)) V, dropout
// on the product -- for the head-merged (B, S, H*d_k) tensors the projections produce.
//
// One CTA = one (batch, head, 128-query tile[, 256-wide slice of d_k for d_k = 512]).  192 threads:
//   warp 0     TMA producer: fp32 Q / K / V tiles straight from the projection outputs (no bf16 shadow copies; the
//              tensor map's TFLOAT32 type converts on the fly), 128B-swizzled, 5-deep mbarrier ring of 32 KB stages
//   warp 1     TMEM allocator + tcgen05.mma issuer (one thread): S = Q K^T (kind::tf32, SS) into a double-buffered
//              128-column TMEM tile, then O += P V with P read FROM TMEM (A operand in tensor memory) and V MN-major
//   warps 2-5  softmax: thread = query row = TMEM lane.  tcgen05.ld the scores, scale, replace masked scores by -1e9,
//              exponentiate, tcgen05.st the probabilities back over the scores (they become the A operand of P V) and
//              write them to global memory for the backward pass; finally scale / drop / store O.
// Sequences longer than one 128-key tile run two passes over the key tiles (pass 1: row max and normaliser with the
// online rescaling, pass 2: probabilities and P V) -- Q K^T is recomputed on the otherwise idle tensor cores instead
// of rescaling the 128 x 256 fp32 accumulator through registers, and the normalised probabilities can be emitted in
// one sweep.  Masked scores are REPLACED (not added to), so a fully masked row is uniform over all Sk keys, exactly
// like the reference.
#include "common.cuh"
#include "tc_common.cuh"
#include "gemm_simt.cuh"
#include "../../include/bmhrl_b200.h"

namespace bmhrl {

using namespace tc;

constexpr int AT_BM = 128;        // queries per CTA
constexpr int AT_BN = 128;        // keys per S tile
constexpr int AT_STAGES = 5;
constexpr int AT_STAGE_BYTES = 32 * 1024;
constexpr int AT_SMEM = AT_STAGES * AT_STAGE_BYTES + 1024 + 256;

struct AttnFwdParams {
    float* O; long long ldo, bso;
    float* P; int ldp;                 // (B,H,Sq,ldp) normalised probabilities (saved for backward)
    const unsigned char* mask; int mask_kind, causal;
    int B, H, Sq, Sk, dk, dvt;         // dvt: width of this CTA's slice of the value / output head dimension
    float scale, drop_p;
    const unsigned long long* rng; unsigned int salt;
};

__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        :
        : "r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
          "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]),
          "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]),
          "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[tmem] * B[smem], kind::tf32 (A: one fp32/tf32 element per 32-bit TMEM cell, row = lane, k = column)
__device__ __forceinline__ void mma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        :
        : "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

__global__ void __launch_bounds__(192) attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ,
                                                          const __grid_constant__ CUtensorMap tmK,
                                                          const __grid_constant__ CUtensorMap tmV,
                                                          const AttnFwdParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars = base + AT_STAGES * AT_STAGE_BYTES;
    auto full_bar = [&](int s) { return bars + 8u * s; };
    auto empty_bar = [&](int s) { return bars + 8u * (AT_STAGES + s); };
    auto s_full = [&](int i) { return bars + 8u * (2 * AT_STAGES + i); };      // MMA -> softmax: S tile i complete
    auto s_rel = [&](int i) { return bars + 8u * (2 * AT_STAGES + 2 + i); };   // softmax -> MMA: tile i read / P written
    const uint32_t o_full = bars + 8u * (2 * AT_STAGES + 4);
    const uint32_t tmem_slot = bars + 8u * (2 * AT_STAGES + 5);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * AT_BM;
    const int bh = blockIdx.y, b = bh / p.H, h = bh - b * p.H;
    const int dv0 = blockIdx.z * p.dvt;
    const int nkt = (p.Sk + AT_BN - 1) / AT_BN;
    const bool two_pass = nkt > 1;
    const int ncq = p.dk >> 5;          // 32-wide d_k chunks of Q K^T
    const int nvb = p.dvt >> 5;         // 32-wide column boxes of a V stage
    const int col_q = h * p.dk;         // first column of this head inside the merged (.., H*dk) tensors

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmK);
        tma_prefetch_desc(&tmV);
        for (int s = 0; s < AT_STAGES; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        mbar_init(s_full(0), 1);
        mbar_init(s_full(1), 1);
        mbar_init(s_rel(0), 128);
        mbar_init(s_rel(1), 128);
        mbar_init(o_full, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    const uint32_t tmem_o = tmem_base + 256;

    if (warp == 0) {
        if (elect_one()) {  // ===== TMA producer: same op order as the MMA issuer below
            int it = 0;
            auto load_qk = [&](int j) {
                for (int c = 0; c < ncq; ++c, ++it) {
                    const int s = it % AT_STAGES;
                    mbar_wait(empty_bar(s), (((uint32_t)(it / AT_STAGES)) & 1u) ^ 1u);
                    mbar_arrive_expect_tx(full_bar(s), 2 * AT_BM * 128);
                    const uint32_t sA = base + s * AT_STAGE_BYTES;
                    tma_load_4d(sA, &tmQ, full_bar(s), col_q + c * 32, q0, 0, b);
                    tma_load_4d(sA + AT_BM * 128, &tmK, full_bar(s), col_q + c * 32, j * AT_BN, 0, b);
                }
            };
            auto load_v = [&](int j) {
                const int nkc = (min(AT_BN, p.Sk - j * AT_BN) + 31) >> 5;
                for (int kc = 0; kc < nkc; ++kc, ++it) {
                    const int s = it % AT_STAGES;
                    mbar_wait(empty_bar(s), (((uint32_t)(it / AT_STAGES)) & 1u) ^ 1u);
                    mbar_arrive_expect_tx(full_bar(s), (uint32_t)nvb * 4096u);
                    const uint32_t sV = base + s * AT_STAGE_BYTES;
                    for (int n = 0; n < nvb; ++n)
                        tma_load_4d(sV + n * 4096, &tmV, full_bar(s), col_q + dv0 + n * 32, j * AT_BN + kc * 32, 0, b);
                }
            };
            if (two_pass)
                for (int j = 0; j < nkt; ++j) load_qk(j);
            load_qk(0);
            if (nkt > 1) load_qk(1);
            for (int j = 0; j < nkt; ++j) {
                load_v(j);
                if (j + 2 < nkt) load_qk(j + 2);
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {  // ===== MMA issuer
            const uint32_t idesc_s = make_idesc(2, AT_BM, AT_BN, 0, 0);
            const uint32_t idesc_o = make_idesc(2, AT_BM, p.dvt, 0, 1);
            const uint64_t qk_desc = make_smem_desc(base, 16, 1024);      // stage 0, K-major; later stages: desc_advance
            const uint64_t v_desc = make_smem_desc(base, 4096, 512, 1);   // stage 0, MN-major fp32 (BASE32B layout)
            int it = 0, n = 0;
            int waited[2] = {0, 0};
            auto wait_rel = [&](int buf, int count) {
                while (waited[buf] < count) {
                    mbar_wait(s_rel(buf), (uint32_t)waited[buf] & 1u);
                    ++waited[buf];
                }
                tc_fence_after();
            };
            auto s_op = [&]() {  // S[n & 1] = Q K_j^T
                const int buf = n & 1;
                wait_rel(buf, n >> 1);  // the previous tile in this buffer has been consumed
                for (int c = 0; c < ncq; ++c, ++it) {
                    const int s = it % AT_STAGES;
                    mbar_wait(full_bar(s), ((uint32_t)(it / AT_STAGES)) & 1u);
                    tc_fence_after();
                    const uint64_t da = desc_advance(qk_desc, s * AT_STAGE_BYTES), db = desc_advance(da, AT_BM * 128);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        mma_tf32_ss(tmem_base + buf * AT_BN, desc_advance(da, k * 32), desc_advance(db, k * 32), idesc_s,
                                    (c > 0 || k > 0) ? 1u : 0u);
                    tc_commit(empty_bar(s));
                }
                tc_commit(s_full(buf));
                ++n;
            };
            auto pv_op = [&](int j, int sn) {  // O += P_j V_j, P_j sits in the S buffer of S-op number sn
                const int buf = sn & 1;
                wait_rel(buf, (sn >> 1) + 1);
                const int nkc = (min(AT_BN, p.Sk - j * AT_BN) + 31) >> 5;
                for (int kc = 0; kc < nkc; ++kc, ++it) {
                    const int s = it % AT_STAGES;
                    mbar_wait(full_bar(s), ((uint32_t)(it / AT_STAGES)) & 1u);
                    tc_fence_after();
                    const uint64_t dv = desc_advance(v_desc, s * AT_STAGE_BYTES);
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk)
                        mma_tf32_ts(tmem_o, tmem_base + buf * AT_BN + kc * 32 + kk * 8, desc_advance(dv, kk * 1024),
                                    idesc_o, (j > 0 || kc > 0 || kk > 0) ? 1u : 0u);
                    tc_commit(empty_bar(s));
                }
            };
            if (two_pass)
                for (int j = 0; j < nkt; ++j) s_op();
            const int n2 = n;
            s_op();
            if (nkt > 1) s_op();
            for (int j = 0; j < nkt; ++j) {
                pv_op(j, n2 + j);
                if (j + 2 < nkt) s_op();
            }
            tc_commit(o_full);
        }
    } else {  // ===== softmax / epilogue warps: TMEM lane quarter = warp % 4, thread = query row
        const int qt = warp & 3;
        const int q = q0 + qt * 32 + lane;
        const bool row_ok = q < p.Sq;
        const uint32_t lane_base = tmem_base + ((uint32_t)(qt * 32) << 16);
        constexpr float LOG2E = 1.4426950408889634f;
        const float sl2 = p.scale * LOG2E;
        const float neg2 = kNegFill * LOG2E;  // the replaced score, in the log2 domain
        float m = -INFINITY, l = 0.f;
        const int n_ops = two_pass ? 2 * nkt : 1;
        for (int n = 0; n < n_ops; ++n) {
            const bool pass1 = two_pass && n < nkt;
            const int j = two_pass ? (pass1 ? n : n - nkt) : 0;
            const int buf = n & 1;
            const int keys = min(AT_BN, p.Sk - j * AT_BN);
            // visibility bits of this row for the 128 keys of tile j
            uint32_t vis[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int kk0 = c * 32;
                uint32_t valid = keys - kk0 >= 32 ? 0xffffffffu : (keys - kk0 <= 0 ? 0u : ((1u << (keys - kk0)) - 1u));
                uint32_t v = valid;
                if (p.mask_kind == 1) {
                    const int kk = j * AT_BN + kk0 + lane;
                    const bool on = kk < p.Sk && p.mask[(long long)b * p.Sk + kk] != 0;
                    v = __ballot_sync(0xffffffffu, on);
                } else if (p.mask_kind == 2) {
                    v = 0u;
                    if (row_ok) {
                        const unsigned char* mr = p.mask + ((long long)b * p.Sq + q) * p.Sk + j * AT_BN + kk0;
                        const int cnt = min(32, keys - kk0);
                        for (int i = 0; i < cnt; ++i) v |= (mr[i] != 0 ? 1u : 0u) << i;
                    }
                }
                if (p.causal) {
                    const int lim = q - (j * AT_BN + kk0);  // largest visible key offset inside this piece
                    if (lim < 0) v = 0u;
                    else if (lim < 31) v &= (2u << lim) - 1u;
                }
                vis[c] = v;
            }
            mbar_wait(s_full(buf), ((uint32_t)(n >> 1)) & 1u);
            tc_fence_after();
            const uint32_t taddr = lane_base + buf * AT_BN;
            if (pass1 || !two_pass) {  // running max / normaliser
#pragma unroll 1
                for (int c = 0; c < 4; ++c) {
                    const int cnt = min(32, keys - c * 32);
                    if (cnt <= 0) break;
                    uint32_t r[32];
                    tmem_ld_32x32(taddr + c * 32, r);
                    tmem_ld_wait();
                    float pm = -INFINITY;
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        float t = ((vis[c] >> i) & 1u) ? __uint_as_float(r[i]) * sl2 : neg2;
                        t = i < cnt ? t : -INFINITY;
                        r[i] = __float_as_uint(t);
                        pm = fmaxf(pm, t);
                    }
                    const float mn = fmaxf(m, pm);
                    float acc = 0.f;
#pragma unroll
                    for (int i = 0; i < 32; ++i) acc += exp2f(__uint_as_float(r[i]) - mn);
                    l = l * exp2f(m - mn) + acc;
                    m = mn;
                }
            }
            if (!pass1) {  // probabilities: back into TMEM (A operand of P V) and out to global memory
                const float inv_l = 1.f / l;
                float* prow = p.P + ((long long)bh * p.Sq + q) * p.ldp + j * AT_BN;
#pragma unroll 1
                for (int c = 0; c < 4; ++c) {
                    const int cnt = min(32, keys - c * 32);
                    if (cnt <= 0) break;
                    uint32_t r[32];
                    tmem_ld_32x32(taddr + c * 32, r);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const float t = ((vis[c] >> i) & 1u) ? __uint_as_float(r[i]) * sl2 : neg2;
                        const float pr = i < cnt ? exp2f(t - m) * inv_l : 0.f;
                        r[i] = __float_as_uint(pr);
                    }
                    tmem_st_32x32(taddr + c * 32, r);
                    if (row_ok) {
#pragma unroll
                        for (int i4 = 0; i4 < 32; i4 += 4)
                            if (j * AT_BN + c * 32 + i4 < p.ldp)
                                st_f4(prow + c * 32 + i4,
                                      make_float4(__uint_as_float(r[i4]), __uint_as_float(r[i4 + 1]),
                                                  __uint_as_float(r[i4 + 2]), __uint_as_float(r[i4 + 3])));
                    }
                }
                tmem_st_wait();
            }
            tc_fence_before();
            mbar_arrive(s_rel(buf));
        }
        // ---- epilogue: O tile (TMEM) -> dropout -> global, head-merged layout
        mbar_wait(o_full, 0);
        tc_fence_after();
        const float ks = p.drop_p > 0.f ? 1.f / (1.f - p.drop_p) : 1.f;
        float* orow = p.O + (long long)b * p.bso + (long long)q * p.ldo + col_q + dv0;
        const unsigned long long didx = ((unsigned long long)bh * p.Sq + q) * p.dk + dv0;
#pragma unroll 1
        for (int c = 0; c < nvb; ++c) {
            uint32_t r[32];
            tmem_ld_32x32(lane_base + 256 + c * 32, r);
            tmem_ld_wait();
            if (row_ok) {
#pragma unroll
                for (int i4 = 0; i4 < 32; i4 += 4) {
                    float v[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) v[i] = __uint_as_float(r[i4 + i]);
                    if (p.drop_p > 0.f) {  // didx is a multiple of 4: the four elements share one Philox block
                        const Philox4 rb = philox_block(p.rng, p.salt, (didx + c * 32 + i4) >> 2);
                        v[0] = u32_to_unit(rb.x) >= p.drop_p ? v[0] * ks : 0.f;
                        v[1] = u32_to_unit(rb.y) >= p.drop_p ? v[1] * ks : 0.f;
                        v[2] = u32_to_unit(rb.z) >= p.drop_p ? v[2] * ks : 0.f;
                        v[3] = u32_to_unit(rb.w) >= p.drop_p ? v[3] * ks : 0.f;
                    }
                    st_f4(orow + c * 32 + i4, make_float4(v[0], v[1], v[2], v[3]));
                }
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------------------------- host side
static int g_attn_fused = 1;

// Returns BMHRL_OK / error, or 1 when the shape / layout is outside what the fused kernel handles.
int launch_attn_fwd_tc(const float* Q, long long ldq, long long bsq, const float* K, long long ldk, long long bsk,
                       const float* V, long long ldv, long long bsv, const unsigned char* mask, int mask_kind, int causal,
                       float* O, long long ldo, long long bso, float* P, int ldp, int B, int H, int Sq, int Sk, int dk,
                       float drop_p, const unsigned long long* rng, unsigned int salt, cudaStream_t st) {
    if (!g_attn_fused || get_encode_fn() == nullptr) return 1;
    if (dk % 32 != 0 || (dk > 256 && dk % 256 != 0) || dk < 32) return 1;
    auto ok4 = [](long long v) { return v > 0 && (v & 3) == 0; };
    if (!ok4(ldq) || !ok4(ldk) || !ok4(ldv) || !ok4(ldo) || (B > 1 && (!ok4(bsq) || !ok4(bsk) || !ok4(bsv) || !ok4(bso))))
        return 1;
    if (!aligned16(Q) || !aligned16(K) || !aligned16(V) || !aligned16(O) || !aligned16(P) || (ldp & 3)) return 1;
    const int dvt = dk > 256 ? 256 : dk;
    const CUtensorMapDataType TF = CU_TENSOR_MAP_DATA_TYPE_TFLOAT32;
    CUtensorMap tmQ, tmK, tmV;
    const unsigned long long D = (unsigned long long)H * dk;
    bool ok = encode_4d(&tmQ, TF, 4, Q, D, Sq, ldq, 1, 0, B, bsq, 32, AT_BM) &&
              encode_4d(&tmK, TF, 4, K, D, Sk, ldk, 1, 0, B, bsk, 32, AT_BN) &&
              encode_4d(&tmV, TF, 4, V, D, Sk, ldv, 1, 0, B, bsv, 32, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
    if (!ok) {
        bmhrl_set_error("attention: cuTensorMapEncodeTiled failed");
        return BMHRL_ERR_CUDA;
    }
    static DeviceOnce attr;
    if (needs_setup(attr)) {
        if (cudaFuncSetAttribute(attn_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM) != cudaSuccess) {
            bmhrl_set_error("attention: cannot raise the shared-memory limit");
            return BMHRL_ERR_CUDA;
        }
        mark_done(attr);
    }
    AttnFwdParams p{};
    p.O = O; p.ldo = ldo; p.bso = bso; p.P = P; p.ldp = ldp; p.mask = mask; p.mask_kind = mask_kind; p.causal = causal;
    p.B = B; p.H = H; p.Sq = Sq; p.Sk = Sk; p.dk = dk; p.dvt = dvt; p.scale = 1.0f / sqrtf((float)dk);
    p.drop_p = drop_p; p.rng = rng; p.salt = salt;
    dim3 grid((Sq + AT_BM - 1) / AT_BM, B * H, dk / dvt);
    attn_fwd_tc_kernel<<<grid, 192, AT_SMEM, st>>>(tmQ, tmK, tmV, p);
    BMHRL_POST_LAUNCH("attn_fwd_tc");
    return BMHRL_OK;
}

}  // namespace bmhrl

extern "C" int bmhrl_attn_tuning(int fused) {
    bmhrl::g_attn_fused = fused ? 1 : 0;
    return BMHRL_OK;
}
