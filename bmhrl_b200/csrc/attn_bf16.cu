// K1, bf16 operands (config 5: long-clip encoder attention sweep, B=64, S=300..800, d_k=256): fused masked multi-head
// attention forward as ONE persistent kernel on the 5th-gen tensor cores (sm_100a).
// Reference: attention() model/multihead_attention.py:7-31 -- softmax(mask_replace(Q K^T / sqrt(d_k), -1e9)) V on the
// head-merged (B, S, H*d_k) tensors the projections produce.  Flash-style: the (B,H,Sq,Sk) score tensor never exists
// in HBM (the reference materialises it and sweeps it three times); the per-row log-sum-exp is kept instead.
//
// Grid = one CTA per SM; CTA c walks the work items c, c+grid, ... ; item = (batch, head, 128-query tile).  192 threads:
//   warp 0     TMA producer.  Q tile (128 x d_k bf16, resident for the whole item) + a 5-deep ring of 32 KB stages that
//              carries K tiles (128 keys x 128 d_k columns) and V tiles (64 keys x d_k) in the order the MMA warp
//              consumes them.  It runs ahead across items, so the next item's Q/K arrive under this item's P V.
//   warp 1     TMEM allocator + tcgen05.mma issuer (one thread).  S = Q K^T (kind::f16, bf16 x bf16 -> fp32, both
//              operands K-major in shared memory) into a double-buffered 128-column TMEM tile; O += P V with P read
//              FROM TMEM (bf16, two keys per 32-bit cell) and V MN-major in shared memory, accumulator 128 x d_k fp32.
//              Issue order S0 S1 PV0 S2 PV1 ... keeps Q K^T of tile j+1 under the softmax of tile j.
//   warps 2-5  softmax + epilogue, thread = query row = TMEM lane: no cross-thread traffic for row max / row sum.
//              ONE pass over the keys: probabilities are taken against a running maximum that is only raised when the
//              new tile maximum exceeds it by more than 2^8 (exact arithmetic -- the normaliser carries the same
//              offset); only then is the O accumulator rescaled through registers.
// Masked scores are REPLACED by -1e9 (not added to), so a fully masked row is uniform over all Sk keys, exactly like
// the reference.  TMEM: 2 x 128 columns S/P + 256 columns O = 512.
#include <cuda_bf16.h>
#include "common.cuh"
#include "tc_common.cuh"
#include "attn_bf16_common.cuh"
#include "../../include/bmhrl_b200.h"

namespace bmhrl {

using namespace tc;

constexpr int AB_BM = 128;   // queries per item
constexpr int AB_BN = 128;   // keys per S tile
constexpr int AB_STAGES = 4;
constexpr int AB_STAGE_BYTES = 32 * 1024;
constexpr int AB_Q_BYTES = 64 * 1024;  // 128 x 256 bf16
constexpr int AB_OST_BYTES = 4 * 4096;  // per softmax warp: 32 rows x 128 B of output, transposed for full-line global stores
constexpr int AB_SMEM = AB_Q_BYTES + AB_STAGES * AB_STAGE_BYTES + 1024 + 256 + AB_OST_BYTES;
// automatic choice: never the CTA-pair form yet -- measured 270 us vs 183 us at B=64, S=800 (the softmax warps, not the
// K / V delivery, bound the kernel, and the pair adds cross-CTA hand-offs to that chain); bmhrl_attn_bf16_tuning(2) selects it
constexpr int AB_PAIR_MIN_QTILES = 1 << 30;
constexpr float AB_LAZY = 8.f;  // log2 of the growth the running maximum may lag behind by

struct AttnBf16Params {
    __nv_bfloat16* O; long long ldo, bso;
    float* lse;                        // (B,H,Sq) natural-log sum of exp of the (masked, scaled) scores
    const unsigned char* mask; int mask_kind, causal;
    int B, H, Sq, Sk, dk;
    int nqt, n_items;
    float scale;
    // key-padding masks only (mask_kind 1, not causal), from attn_mask_pack_kernel: keys beyond the last visible one of a
    // batch row get probability exactly 0 (unless the whole row is masked: sk_eff = Sk), so their tiles are skipped
    const int* sk_eff;          // (B) or NULL
    const uint32_t* mbits;      // (B, mwords) visibility bits, 4 words per 128-key tile, or NULL
    int mwords;
};

// One CTA per batch row: bit-packs the (B,Sk) key-padding mask and finds the effective key count.
__global__ void attn_mask_pack_kernel(const unsigned char* __restrict__ mask, int Sk, int mwords, int* __restrict__ sk_eff,
                                      uint32_t* __restrict__ mbits) {
    const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    __shared__ int last;
    if (threadIdx.x == 0) last = -1;
    __syncthreads();
    int mine = -1;
    for (int wd = warp; wd < mwords; wd += nw) {
        const int k = wd * 32 + lane;
        const bool on = k < Sk && mask[(long long)b * Sk + k] != 0;
        const uint32_t bits = __ballot_sync(0xffffffffu, on);
        if (lane == 0) mbits[(long long)b * mwords + wd] = bits;
        if (bits) mine = max(mine, wd * 32 + 31 - __clz(bits));
    }
    if (lane == 0 && mine >= 0) atomicMax(&last, mine);
    __syncthreads();
    if (threadIdx.x == 0) sk_eff[b] = last >= 0 ? last + 1 : Sk;
}

void attn_mask_pack(const unsigned char* mask, int B, int Sk, int mwords, int* sk_eff, uint32_t* mbits, cudaStream_t stream) {
    attn_mask_pack_kernel<<<B, 128, 0, stream>>>(mask, Sk, mwords, sk_eff, mbits);
}

// rare path of the lazy softmax: multiply this warp's 32 accumulator rows (nchunks x 32 fp32 columns) by alpha[row]
__device__ __noinline__ void rescale_o(uint32_t taddr, int nchunks, float alpha) {
    for (int c = 0; c < nchunks; ++c) {
        uint32_t o[32];
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(o[0]), "=r"(o[1]), "=r"(o[2]), "=r"(o[3]), "=r"(o[4]), "=r"(o[5]), "=r"(o[6]), "=r"(o[7]), "=r"(o[8]),
              "=r"(o[9]), "=r"(o[10]), "=r"(o[11]), "=r"(o[12]), "=r"(o[13]), "=r"(o[14]), "=r"(o[15]), "=r"(o[16]),
              "=r"(o[17]), "=r"(o[18]), "=r"(o[19]), "=r"(o[20]), "=r"(o[21]), "=r"(o[22]), "=r"(o[23]), "=r"(o[24]),
              "=r"(o[25]), "=r"(o[26]), "=r"(o[27]), "=r"(o[28]), "=r"(o[29]), "=r"(o[30]), "=r"(o[31])
            : "r"(taddr + c * 32)
            : "memory");
        tmem_ld_wait_b();
#pragma unroll
        for (int t = 0; t < 32; ++t) o[t] = __float_as_uint(__uint_as_float(o[t]) * alpha);
        tmem_st_32x32_b(taddr + c * 32, o);
    }
    tmem_st_wait_b();
}

// ---- CTA pair (cta_group::2): one MMA spans the two SMs of a cluster; only the leader CTA (cluster rank 0) issues
__device__ __forceinline__ void mma2_bf16_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        :
        : "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void mma2_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        :
        : "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrives once on the barrier at the same shared-memory offset in BOTH CTAs when the MMAs issued so far have completed
__device__ __forceinline__ void tc_commit2(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((unsigned short)3) : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t dst_smem, uint32_t cols) {  // one warp in EACH CTA of the pair
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// TMA load into MY shared memory whose completion bytes are counted on a barrier of the LEADER CTA (cluster address)
__device__ __forceinline__ void tma2_load_4d(uint32_t dst, const CUtensorMap* m, uint32_t cluster_bar, int c0, int c1,
                                             int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%3, %4, %5, %6}], [%2];"
        :
        : "r"(dst), "l"(reinterpret_cast<uint64_t>(m)), "r"(cluster_bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar) {  // barrier possibly in the peer CTA
    // default semantics (as CUTLASS' ClusterBarrier::arrive): what crosses the CTAs is tensor memory, ordered by the
    // tcgen05 fences around the barrier; an explicit .release.cluster costs a MEMBAR / ERRBAR per arrival
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}
// wait on a LOCAL barrier whose arrivals may come from the peer CTA (acquire at cluster scope); bounded like mbar_wait
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
    uint32_t ok = 0;
#pragma unroll 1
    for (uint32_t spin = 0; spin < (1u << 26); ++spin) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"  // .acquire.cluster would invalidate L1 (CCTL.IVALL) per poll
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
        if (ok) return;
    }
    __trap();
}

// PAIR: two CTAs of a cluster (one SM pair) work on two neighbouring query tiles of the same (batch, head) with
// cta_group::2 MMAs (M = 256): every K tile is split by keys and every V tile by value columns between the two shared
// memories, so each SM takes in HALF of the K / V bytes -- the delivery of K / V from L2 is what bounds the one-CTA form
// at long sequences.  Only the leader (cluster rank 0) issues MMAs; completion is multicast to the barriers of both
// CTAs; TMA loads of both CTAs are counted on the leader's barriers; softmax threads of the peer arrive remotely.
// MASKED = false: no mask and not causal -- the visibility code (and its instruction-cache footprint) is compiled out.
template <bool PAIR, bool MASKED>
__global__ void __launch_bounds__(192, 1) attn_fwd_bf16_kernel(const __grid_constant__ CUtensorMap tmQ,
                                                               const __grid_constant__ CUtensorMap tmK,
                                                               const __grid_constant__ CUtensorMap tmV,
                                                               const AttnBf16Params p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t sQ = base;
    const uint32_t ring = base + AB_Q_BYTES;
    const uint32_t bars = ring + AB_STAGES * AB_STAGE_BYTES;
    auto full_bar = [&](int s) { return bars + 8u * s; };
    auto empty_bar = [&](int s) { return bars + 8u * (AB_STAGES + s); };
    auto s_full = [&](int i) { return bars + 8u * (2 * AB_STAGES + i); };     // MMA -> softmax: S tile complete
    auto s_rel = [&](int i) { return bars + 8u * (2 * AB_STAGES + 2 + i); };  // softmax -> MMA: P written (O rescaled)
    const uint32_t q_full = bars + 8u * (2 * AB_STAGES + 4);
    const uint32_t q_empty = bars + 8u * (2 * AB_STAGES + 5);
    const uint32_t pv_bar = bars + 8u * (2 * AB_STAGES + 6);   // MMA -> softmax: P V of one tile complete
    const uint32_t o_free = bars + 8u * (2 * AB_STAGES + 7);   // epilogue -> MMA: O accumulator drained
    const uint32_t tmem_slot = bars + 8u * (2 * AB_STAGES + 8);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    auto item_sk = [&](int b) { return p.sk_eff ? __ldg(p.sk_eff + b) : p.Sk; };  // keys that can get weight
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;            // 0 = leader of the pair
    const int w_first = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;   // first item / item stride of this CTA (pair)
    const int w_step = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    const int tiles_bh = PAIR ? (p.nqt + 1) >> 1 : p.nqt;           // items per (batch, head)
    auto item_qt = [&](int tq) { return PAIR ? 2 * tq + (int)rank : tq; };  // query tile of THIS CTA inside the item
    // the leader's copy of a barrier of mine (same offset), as a shared::cluster address
    const uint32_t lead_bars = PAIR ? mapa_shared(bars, 0) : bars;
    auto lead = [&](uint32_t bar) { return lead_bars + (bar - bars); };
    auto wait_x = [&](uint32_t bar, uint32_t parity) {  // barriers that the leader's multicast commits arrive on
        if (PAIR) mbar_wait_cluster(bar, parity);
        else mbar_wait(bar, parity);
    };
    const int kcols = p.dk < 128 ? p.dk : 128;   // d_k columns of one K stage
    const int nkh = p.dk / kcols;                // K stages per key tile
    const int ncs = kcols >> 6;                  // 64-column (128-byte) chunks per K stage
    const int nvc = p.dk >> 6;                   // 64-column chunks of a V stage / of the Q tile

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmK);
        tma_prefetch_desc(&tmV);
        for (int s = 0; s < AB_STAGES; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        mbar_init(s_full(0), 1);
        mbar_init(s_full(1), 1);
        mbar_init(s_rel(0), PAIR ? 256 : 128);   // softmax threads of both CTAs arrive on the leader's copy
        mbar_init(s_rel(1), PAIR ? 256 : 128);
        mbar_init(q_full, 1);
        mbar_init(q_empty, 1);
        mbar_init(pv_bar, 1);
        mbar_init(o_free, PAIR ? 256 : 128);
        fence_mbar_init();
    }
    if (warp == 1) {
        if (PAIR) tmem_alloc2(tmem_slot, 512);
        else tmem_alloc(tmem_slot, 512);
    }
    tc_fence_before();
    if (PAIR) cluster_sync_all();  // the peer's barriers are initialised before anything is signalled on them
    else __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    const uint32_t tmem_o = tmem_base + 256;

    if (warp == 0) {
        if (elect_one()) {  // ===== TMA producer (same op order as the MMA issuer)
            int it = 0;
            for (int i = 0, w = w_first; w < p.n_items; ++i, w += w_step) {
                const int bh = w / tiles_bh, qt = item_qt(w - bh * tiles_bh);
                const int b = bh / p.H, h = bh - b * p.H;
                const int q0 = qt * AB_BM, col_q = h * p.dk;
                const int Sk = item_sk(b), nkt = (Sk + AB_BN - 1) / AB_BN;
                wait_x(q_empty, ((uint32_t)i & 1u) ^ 1u);
                if (PAIR) {
                    // K stage: MY 64 keys of the tile x d_k; V stage: the 128 keys of the tile x MY half of the columns
                    if (rank == 0) mbar_arrive_expect_tx(q_full, (uint32_t)(2 * AB_BM * p.dk * 2));
                    for (int c = 0; c < nvc; ++c)
                        tma2_load_4d(sQ + c * 16384, &tmQ, lead(q_full), col_q + c * 64, q0, 0, b);
                    auto load2 = [&](bool is_k, int j) {
                        const int s = it % AB_STAGES;
                        mbar_wait_cluster(empty_bar(s), (((uint32_t)(it / AB_STAGES)) & 1u) ^ 1u);
                        if (rank == 0) mbar_arrive_expect_tx(full_bar(s), (uint32_t)(2 * 64 * p.dk * 2));
                        const uint32_t st = ring + s * AB_STAGE_BYTES;
                        if (is_k) {
                            for (int c = 0; c < nvc; ++c)
                                tma2_load_4d(st + c * 8192, &tmK, lead(full_bar(s)), col_q + c * 64,
                                             j * AB_BN + (int)rank * 64, 0, b);
                        } else {
                            for (int c = 0; c < (nvc >> 1); ++c)
                                tma2_load_4d(st + c * 16384, &tmV, lead(full_bar(s)),
                                             col_q + (int)rank * (p.dk >> 1) + c * 64, j * AB_BN, 0, b);
                        }
                        ++it;
                    };
                    load2(true, 0);
                    if (nkt > 1) load2(true, 1);
                    for (int j = 0; j < nkt; ++j) {
                        load2(false, j);
                        if (j + 2 < nkt) load2(true, j + 2);
                    }
                    continue;
                }
                mbar_arrive_expect_tx(q_full, (uint32_t)(AB_BM * p.dk * 2));
                for (int c = 0; c < nvc; ++c) tma_load_4d(sQ + c * 16384, &tmQ, q_full, col_q + c * 64, q0, 0, b);
                auto load_k = [&](int j) {
                    for (int hf = 0; hf < nkh; ++hf, ++it) {
                        const int s = it % AB_STAGES;
                        mbar_wait(empty_bar(s), (((uint32_t)(it / AB_STAGES)) & 1u) ^ 1u);
                        mbar_arrive_expect_tx(full_bar(s), (uint32_t)(ncs * 16384));
                        const uint32_t sK = ring + s * AB_STAGE_BYTES;
                        for (int cc = 0; cc < ncs; ++cc)
                            tma_load_4d(sK + cc * 16384, &tmK, full_bar(s), col_q + hf * kcols + cc * 64, j * AB_BN, 0, b);
                    }
                };
                auto load_v = [&](int j) {
                    const int nkc = (min(AB_BN, Sk - j * AB_BN) + 63) >> 6;
                    for (int kc = 0; kc < nkc; ++kc, ++it) {
                        const int s = it % AB_STAGES;
                        mbar_wait(empty_bar(s), (((uint32_t)(it / AB_STAGES)) & 1u) ^ 1u);
                        mbar_arrive_expect_tx(full_bar(s), (uint32_t)(nvc * 8192));
                        const uint32_t sV = ring + s * AB_STAGE_BYTES;
                        for (int n = 0; n < nvc; ++n)
                            tma_load_4d(sV + n * 8192, &tmV, full_bar(s), col_q + n * 64, j * AB_BN + kc * 64, 0, b);
                    }
                };
                load_k(0);
                if (nkt > 1) load_k(1);
                for (int j = 0; j < nkt; ++j) {
                    load_v(j);
                    if (j + 2 < nkt) load_k(j + 2);
                }
            }
        }
    } else if (warp == 1) {
        if (PAIR && rank == 0 && elect_one()) {  // ===== MMA issuer of the pair (leader only): M = 256 over both CTAs
            const uint32_t idesc_s = make_idesc(1, 2 * AB_BM, AB_BN, 0, 0);
            const uint32_t idesc_o = make_idesc(1, 2 * AB_BM, p.dk, 0, 1);
            const uint64_t q_desc = make_smem_desc(sQ, 16, 1024);         // per CTA: its 128 query rows, K-major
            const uint64_t k_desc = make_smem_desc(ring, 16, 1024);       // per CTA: its 64 keys, K-major, 8 KB per 64 columns
            const uint64_t v_desc = make_smem_desc(ring, 16384, 1024);    // per CTA: its d_k/2 columns, MN-major, 16 KB chunks
            int it = 0, ns = 0, np = 0;
            for (int i = 0, w = w_first; w < p.n_items; ++i, w += w_step) {
                const int Sk = item_sk((w / tiles_bh) / p.H), nkt = (Sk + AB_BN - 1) / AB_BN;
                mbar_wait_cluster(q_full, (uint32_t)i & 1u);
                tc_fence_after();
                auto s_op = [&](bool last) {
                    const int buf = ns & 1;
                    const int s = it % AB_STAGES;
                    mbar_wait_cluster(full_bar(s), ((uint32_t)(it / AB_STAGES)) & 1u);
                    tc_fence_after();
                    for (int c = 0; c < nvc; ++c) {
                        const uint64_t a0 = desc_advance(q_desc, c * 16384);
                        const uint64_t b0 = desc_advance(k_desc, s * AB_STAGE_BYTES + c * 8192);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            mma2_bf16_ss(tmem_base + buf * AB_BN, desc_advance(a0, k * 32), desc_advance(b0, k * 32),
                                         idesc_s, (c > 0 || k > 0) ? 1u : 0u);
                    }
                    tc_commit2(empty_bar(s));
                    ++it;
                    tc_commit2(s_full(buf));
                    if (last) tc_commit2(q_empty);
                    ++ns;
                };
                auto pv_op = [&](int j) {
                    const int buf = np & 1;
                    mbar_wait_cluster(s_rel(buf), ((uint32_t)(np >> 1)) & 1u);
                    if (j == 0 && i > 0) mbar_wait_cluster(o_free, ((uint32_t)(i - 1)) & 1u);
                    tc_fence_after();
                    const int nk16 = (min(AB_BN, Sk - j * AB_BN) + 15) >> 4;  // 16-key MMA steps
                    const int s = it % AB_STAGES;
                    mbar_wait_cluster(full_bar(s), ((uint32_t)(it / AB_STAGES)) & 1u);
                    tc_fence_after();
                    const uint64_t v0 = desc_advance(v_desc, s * AB_STAGE_BYTES);
                    for (int kk = 0; kk < nk16; ++kk)
                        mma2_bf16_ts(tmem_o, tmem_base + buf * AB_BN + kk * 8, desc_advance(v0, kk * 2048), idesc_o,
                                     (j > 0 || kk > 0) ? 1u : 0u);
                    tc_commit2(empty_bar(s));
                    ++it;
                    tc_commit2(pv_bar);
                    ++np;
                };
                s_op(nkt == 1);
                if (nkt > 1) s_op(nkt == 2);
                for (int j = 0; j < nkt; ++j) {
                    pv_op(j);
                    if (j + 2 < nkt) s_op(j + 3 == nkt);
                }
            }
        } else if (!PAIR && elect_one()) {  // ===== MMA issuer
            const uint32_t idesc_full = make_idesc(1, AB_BM, AB_BN, 0, 0);
            const uint32_t idesc_o = make_idesc(1, AB_BM, p.dk, 0, 1);
            const uint64_t q_desc = make_smem_desc(sQ, 16, 1024);        // K-major, 128-byte rows, 8-row groups
            const uint64_t ring_desc = make_smem_desc(ring, 16, 1024);   // K tiles, same layout
            const uint64_t v_desc = make_smem_desc(ring, 8192, 1024);    // V tiles: MN-major, 64-column chunks 8 KB apart
            int it = 0, ns = 0, np = 0;  // ring position, S tiles issued, P V tiles issued (all items)
            for (int i = 0, w = w_first; w < p.n_items; ++i, w += w_step) {
                const int Sk = item_sk((w / tiles_bh) / p.H), nkt = (Sk + AB_BN - 1) / AB_BN;
                mbar_wait(q_full, (uint32_t)i & 1u);
                tc_fence_after();
                auto s_op = [&](int j) {  // S[ns & 1] = Q K_j^T; in-order execution behind the P V that read this buffer
                    const int buf = ns & 1;
                    // a partial last tile only computes the score columns of its keys (N rounded up to 16)
                    const int keys = min(AB_BN, Sk - j * AB_BN);
                    const uint32_t idesc_s = keys == AB_BN ? idesc_full : make_idesc(1, AB_BM, (keys + 15) & ~15, 0, 0);
                    for (int hf = 0; hf < nkh; ++hf, ++it) {
                        const int s = it % AB_STAGES;
                        mbar_wait(full_bar(s), ((uint32_t)(it / AB_STAGES)) & 1u);
                        tc_fence_after();
                        for (int cc = 0; cc < ncs; ++cc) {
                            const uint64_t a0 = desc_advance(q_desc, (hf * ncs + cc) * 16384);
                            const uint64_t b0 = desc_advance(ring_desc, s * AB_STAGE_BYTES + cc * 16384);
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                mma_bf16_ss(tmem_base + buf * AB_BN, desc_advance(a0, k * 32), desc_advance(b0, k * 32),
                                            idesc_s, (hf > 0 || cc > 0 || k > 0) ? 1u : 0u);
                        }
                        tc_commit(empty_bar(s));
                    }
                    tc_commit(s_full(buf));
                    if (j == nkt - 1) tc_commit(q_empty);  // every Q K^T of this item has completed: Q may be overwritten
                    ++ns;
                };
                auto pv_op = [&](int j) {  // O += P_j V_j
                    const int buf = np & 1;
                    mbar_wait(s_rel(buf), ((uint32_t)(np >> 1)) & 1u);
                    if (j == 0 && i > 0) mbar_wait(o_free, ((uint32_t)(i - 1)) & 1u);
                    tc_fence_after();
                    const int keys = min(AB_BN, Sk - j * AB_BN);
                    const int nkc = (keys + 63) >> 6, nk16 = (keys + 15) >> 4;  // V stages; 16-key MMA steps
                    for (int kc = 0; kc < nkc; ++kc, ++it) {
                        const int s = it % AB_STAGES;
                        mbar_wait(full_bar(s), ((uint32_t)(it / AB_STAGES)) & 1u);
                        tc_fence_after();
                        const uint64_t v0 = desc_advance(v_desc, s * AB_STAGE_BYTES);
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk)
                            if (kc * 4 + kk < nk16)
                                mma_bf16_ts(tmem_o, tmem_base + buf * AB_BN + kc * 32 + kk * 8, desc_advance(v0, kk * 2048),
                                            idesc_o, (j > 0 || kc > 0 || kk > 0) ? 1u : 0u);
                        tc_commit(empty_bar(s));
                    }
                    tc_commit(pv_bar);
                    ++np;
                };
                s_op(0);
                if (nkt > 1) s_op(1);
                for (int j = 0; j < nkt; ++j) {
                    pv_op(j);
                    if (j + 2 < nkt) s_op(j + 2);
                }
            }
        }
    } else {  // ===== softmax / epilogue warps: TMEM lane quarter = warp % 4, thread = query row
        const int qtr = warp & 3;
        const uint32_t lane_base = tmem_base + ((uint32_t)(qtr * 32) << 16);
        constexpr float LOG2E = 1.4426950408889634f;
        const float sl2 = p.scale * LOG2E;
        const float neg2 = kNegFill * LOG2E;  // the replaced score, in the log2 domain
        int n = 0;                            // tiles consumed (all items) == index of the next P V
        for (int i = 0, w = w_first; w < p.n_items; ++i, w += w_step) {
            const int bh = w / tiles_bh, qt = item_qt(w - bh * tiles_bh);
            const int b = bh / p.H, h = bh - b * p.H;
            const int q = qt * AB_BM + qtr * 32 + lane;
            const bool row_ok = q < p.Sq;
            float m = -INFINITY, l = 0.f;
            const int Sk = item_sk(b), nkt = (Sk + AB_BN - 1) / AB_BN;
            const uint32_t* mrow = (MASKED && p.mbits) ? p.mbits + (long long)b * p.mwords : nullptr;
            uint32_t vw[4] = {0u, 0u, 0u, 0u};  // packed key-padding bits of the NEXT tile (loaded one tile ahead)
            if (mrow) {
#pragma unroll
                for (int c = 0; c < 4; ++c) vw[c] = __ldg(mrow + c);
            }
            for (int j = 0; j < nkt; ++j, ++n) {
                const int buf = n & 1;
                const int keys = min(AB_BN, Sk - j * AB_BN);
                uint32_t vis[4];  // visibility bits of this row for the 128 keys of tile j
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int kk0 = c * 32;
                    uint32_t v = keys - kk0 >= 32 ? 0xffffffffu : (keys - kk0 <= 0 ? 0u : ((1u << (keys - kk0)) - 1u));
                    if (!MASKED) {
                        vis[c] = v;
                        continue;
                    }
                    if (mrow) {
                        v &= vw[c];
                    } else if (p.mask_kind == 1) {
                        const int kk = j * AB_BN + kk0 + lane;
                        const bool on = kk < p.Sk && p.mask[(long long)b * p.Sk + kk] != 0;
                        v = __ballot_sync(0xffffffffu, on);
                    } else if (p.mask_kind == 2) {
                        v = 0u;
                        if (row_ok) {
                            const unsigned char* mr = p.mask + ((long long)b * p.Sq + q) * p.Sk + j * AB_BN + kk0;
                            const int cnt = min(32, keys - kk0);
                            for (int t = 0; t < cnt; ++t) v |= (mr[t] != 0 ? 1u : 0u) << t;
                        }
                    }
                    if (p.causal) {
                        const int lim = q - (j * AB_BN + kk0);
                        if (lim < 0) v = 0u;
                        else if (lim < 31) v &= (2u << lim) - 1u;
                    }
                    vis[c] = v;
                }
                if (mrow && j + 1 < nkt) {
#pragma unroll
                    for (int c = 0; c < 4; ++c) vw[c] = __ldg(mrow + (j + 1) * 4 + c);
                }
                wait_x(s_full(buf), ((uint32_t)(n >> 1)) & 1u);
                tc_fence_after();
                const uint32_t taddr = lane_base + buf * AB_BN;
                // Scores stay in the raw domain: max over raw scores, one FFMA (score * scale*log2e - m) feeds ex2.
                const bool fast = MASKED ? __all_sync(0xffffffffu, keys == AB_BN && (vis[0] & vis[1] & vis[2] & vis[3]) == 0xffffffffu)
                                         : keys == AB_BN;
                const int nch = ((keys + 63) >> 6) << 1;  // 32-key chunks covered by the V stages of this tile
                bool raise;
                float alpha = 1.f;
                if (fast) {  // every key of the tile visible to every row of the warp: straight-line, one TMEM read
                    uint32_t r[128];
#pragma unroll
                    for (int c = 0; c < 4; ++c) tmem_ld_32x32_b(taddr + c * 32, r + c * 32);
                    tmem_ld_wait();
                    float mx4[4] = {__uint_as_float(r[0]), __uint_as_float(r[1]), __uint_as_float(r[2]), __uint_as_float(r[3])};
#pragma unroll
                    for (int t = 4; t < 128; t += 4) {  // four independent chains instead of one 127-deep one
#pragma unroll
                        for (int u = 0; u < 4; ++u) mx4[u] = fmaxf(mx4[u], __uint_as_float(r[t + u]));
                    }
                    const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
                    const float mt = mx * sl2;
                    raise = mt > m + AB_LAZY;  // first tile: m = -inf
                    if (raise) {
                        alpha = ex2_approx(m - mt);
                        m = mt;
                        l *= alpha;
                    }
                    const float nm = -m;
                    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int t = 0; t < 128; t += 4) {
                        const float p0 = ex2_approx(fmaf(__uint_as_float(r[t]), sl2, nm));
                        const float p1 = ex2_approx(fmaf(__uint_as_float(r[t + 1]), sl2, nm));
                        const float p2 = ex2_approx(fmaf(__uint_as_float(r[t + 2]), sl2, nm));
                        const float p3 = ex2_approx(fmaf(__uint_as_float(r[t + 3]), sl2, nm));
                        acc[0] += p0;
                        acc[1] += p1;
                        acc[2] += p2;
                        acc[3] += p3;
                        r[t >> 1] = pack_bf16x2(p0, p1);
                        r[(t >> 1) + 1] = pack_bf16x2(p2, p3);
                    }
                    l += (acc[0] + acc[1]) + (acc[2] + acc[3]);
                    if (j > 0) {  // P V of the previous tile: every phase of pv_bar is waited for, in order
                        wait_x(pv_bar, ((uint32_t)(n - 1)) & 1u);
                        tc_fence_after();
                    }
                    if (j > 0 && __any_sync(0xffffffffu, raise)) rescale_o(lane_base + 256, p.dk >> 5, alpha);
                    tmem_st_32x32_b(taddr, r);
                    tmem_st_32x32_b(taddr + 32, r + 32);
                } else {  // masked / partial tile: two rolled sweeps over 32-key chunks (small code, selects per element)
                    const float neg_raw = kNegFill / p.scale;  // * scale*log2e gives the replaced score in the log2 domain
                    float mx = -INFINITY;
#pragma unroll 1
                    for (int c = 0; c < nch; ++c) {
                        const uint32_t v = c == 0 ? vis[0] : c == 1 ? vis[1] : c == 2 ? vis[2] : vis[3];
                        const int rem = keys - c * 32;
                        const uint32_t valid = rem >= 32 ? 0xffffffffu : (rem <= 0 ? 0u : ((1u << rem) - 1u));
                        uint32_t r[32];
                        tmem_ld_32x32_b(taddr + c * 32, r);
                        tmem_ld_wait();
#pragma unroll
                        for (int t = 0; t < 32; ++t) {
                            float x = ((v >> t) & 1u) ? __uint_as_float(r[t]) : neg_raw;
                            x = ((valid >> t) & 1u) ? x : -INFINITY;
                            mx = fmaxf(mx, x);
                        }
                    }
                    const float mt = mx * sl2;
                    raise = mt > m + AB_LAZY;
                    if (raise) {
                        alpha = ex2_approx(m - mt);
                        m = mt;
                        l *= alpha;
                    }
                    const float nm = -m;
                    if (j > 0) {
                        wait_x(pv_bar, ((uint32_t)(n - 1)) & 1u);
                        tc_fence_after();
                    }
                    if (j > 0 && __any_sync(0xffffffffu, raise)) rescale_o(lane_base + 256, p.dk >> 5, alpha);
                    float acc = 0.f;
#pragma unroll 1
                    for (int c = 0; c < nch; ++c) {
                        const uint32_t v = c == 0 ? vis[0] : c == 1 ? vis[1] : c == 2 ? vis[2] : vis[3];
                        const int rem = keys - c * 32;
                        const uint32_t valid = rem >= 32 ? 0xffffffffu : (rem <= 0 ? 0u : ((1u << rem) - 1u));
                        uint32_t r[32];
                        tmem_ld_32x32_b(taddr + c * 32, r);
                        tmem_ld_wait();
#pragma unroll
                        for (int t = 0; t < 32; t += 2) {
                            float x0 = ((v >> t) & 1u) ? __uint_as_float(r[t]) : neg_raw;
                            float x1 = ((v >> (t + 1)) & 1u) ? __uint_as_float(r[t + 1]) : neg_raw;
                            x0 = ((valid >> t) & 1u) ? x0 : -INFINITY;
                            x1 = ((valid >> (t + 1)) & 1u) ? x1 : -INFINITY;
                            const float p0 = ex2_approx(fmaf(x0, sl2, nm));
                            const float p1 = ex2_approx(fmaf(x1, sl2, nm));
                            acc += p0 + p1;
                            r[t >> 1] = pack_bf16x2(p0, p1);
                        }
                        tmem_st_32x32_x16(taddr + c * 16, r);  // packed P never reaches the unread S chunks
                    }
                    l += acc;
                }
                tmem_st_wait_b();
                tc_fence_before();
                if (PAIR) mbar_arrive_cluster(lead(s_rel(buf)));
                else mbar_arrive(s_rel(buf));
            }
            // ---- epilogue: O tile (TMEM) / l -> bf16 -> global, head-merged layout; log-sum-exp
            wait_x(pv_bar, ((uint32_t)(n - 1)) & 1u);
            tc_fence_after();
            const float inv_l = 1.f / l;
            // thread = row out of TMEM, but a row-per-thread 16-byte store touches 32 different lines per instruction
            // (the LSU then costs more than the MMAs of a short item): transpose 64-column chunks through shared
            // memory (16-byte units XOR-swizzled by row) and store 4 rows x 128 B = 4 full lines per instruction.
            uint8_t* ost = smem_raw + (bars - smem_u32(smem_raw)) + 256 + qtr * 4096;
            const int q_w0 = qt * AB_BM + qtr * 32;  // first query row of this warp
            __nv_bfloat16* obase = p.O + (long long)b * p.bso + h * p.dk;
            {
                const int nc64 = p.dk >> 6;  // 64-column chunks; the TMEM read of chunk c+1 is in flight under the stores of c
                uint32_t o[64];
                tmem_ld_32x32_b(lane_base + 256, o);
                tmem_ld_32x32_b(lane_base + 256 + 32, o + 32);
#pragma unroll 1
                for (int c = 0; c < nc64; ++c) {
                    uint4 u[8];
                    tmem_ld_wait();
#pragma unroll
                    for (int u8 = 0; u8 < 8; ++u8) {
                        const int t = u8 * 8;
                        u[u8].x = pack_bf16x2(__uint_as_float(o[t]) * inv_l, __uint_as_float(o[t + 1]) * inv_l);
                        u[u8].y = pack_bf16x2(__uint_as_float(o[t + 2]) * inv_l, __uint_as_float(o[t + 3]) * inv_l);
                        u[u8].z = pack_bf16x2(__uint_as_float(o[t + 4]) * inv_l, __uint_as_float(o[t + 5]) * inv_l);
                        u[u8].w = pack_bf16x2(__uint_as_float(o[t + 6]) * inv_l, __uint_as_float(o[t + 7]) * inv_l);
                    }
                    if (c + 1 < nc64) {
                        tmem_ld_32x32_b(lane_base + 256 + (c + 1) * 64, o);
                        tmem_ld_32x32_b(lane_base + 256 + (c + 1) * 64 + 32, o + 32);
                    }
                    __syncwarp();  // the previous chunk has been read out of the staging tile
#pragma unroll
                    for (int u8 = 0; u8 < 8; ++u8)
                        *reinterpret_cast<uint4*>(ost + lane * 128 + ((u8 ^ (lane & 7)) << 4)) = u[u8];
                    __syncwarp();
#pragma unroll
                    for (int it = 0; it < 8; ++it) {
                        const int r = it * 4 + (lane >> 3), u8 = lane & 7;
                        const uint4 v4 = *reinterpret_cast<const uint4*>(ost + r * 128 + ((u8 ^ (r & 7)) << 4));
                        if (q_w0 + r < p.Sq)
                            *reinterpret_cast<uint4*>(obase + (long long)(q_w0 + r) * p.ldo + c * 64 + u8 * 8) = v4;
                    }
                }
            }
            if (row_ok && p.lse) p.lse[(long long)bh * p.Sq + q] = (m + log2f(l)) * 0.6931471805599453f;
            tc_fence_before();
            if (PAIR) mbar_arrive_cluster(lead(o_free));
            else mbar_arrive(o_free);
        }
    }
    tc_fence_before();
    if (PAIR) cluster_sync_all();  // the leader's MMAs read the peer's shared / tensor memory: leave together
    else __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        if (PAIR) tmem_dealloc2(tmem_base, 512);
        else tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace bmhrl

static int g_attn_bf16_pair = 0;
/* test / tuning hook: 0 = automatic, 1 = always the one-CTA kernel, 2 = the CTA-pair kernel whenever d_k allows */
extern "C" int bmhrl_attn_bf16_tuning(int pair_mode) {
    BMHRL_CHECK_ARG(pair_mode >= 0 && pair_mode <= 2, "bmhrl_attn_bf16_tuning: mode 0..2");
    g_attn_bf16_pair = pair_mode;
    return BMHRL_OK;
}

extern "C" long long bmhrl_attn_fwd_bf16_ws_bytes(int B, int Sk) {
    return 4ll * (((B + 3) & ~3) + (long long)B * 4 * ((Sk + 127) / 128));
}

extern "C" int bmhrl_attn_fwd_bf16(const void* Q, long long ldq, long long bsq, const void* K, long long ldk,
                                   long long bsk, const void* V, long long ldv, long long bsv,
                                   const unsigned char* mask, int mask_kind, int causal, void* O, long long ldo,
                                   long long bso, float* lse, int B, int H, int Sq, int Sk, int dk, void* ws,
                                   void* stream) {
    using namespace bmhrl;
    using namespace bmhrl::tc;
    BMHRL_CHECK_ARG(Q && K && V && O, "attn_fwd_bf16: null tensor");
    BMHRL_CHECK_ARG(B > 0 && H > 0 && Sq > 0 && Sk > 0, "attn_fwd_bf16: empty shape");
    BMHRL_CHECK_ARG(mask_kind >= 0 && mask_kind <= 2 && (mask_kind == 0 || mask), "attn_fwd_bf16: bad mask");
    if (dk != 64 && dk != 128 && dk != 256) {
        bmhrl_set_error("attn_fwd_bf16: d_k must be 64, 128 or 256");
        return BMHRL_ERR_UNSUPPORTED;
    }
    auto ok8 = [](long long v) { return v > 0 && (v & 7) == 0; };
    if (!ok8(ldq) || !ok8(ldk) || !ok8(ldv) || !ok8(ldo) || (B > 1 && (!ok8(bsq) || !ok8(bsk) || !ok8(bsv) || !ok8(bso))) ||
        !aligned16(Q) || !aligned16(K) || !aligned16(V) || !aligned16(O)) {
        bmhrl_set_error("attn_fwd_bf16: tensors must be 16-byte aligned with pitches that are multiples of 8 elements");
        return BMHRL_ERR_UNSUPPORTED;
    }
    if (get_encode_fn() == nullptr) {
        bmhrl_set_error("attn_fwd_bf16: cuTensorMapEncodeTiled unavailable");
        return BMHRL_ERR_CUDA;
    }
    const CUtensorMapDataType BF = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    CUtensorMap tmQ, tmK, tmV;
    const unsigned long long D = (unsigned long long)H * dk;
    bool ok = encode_4d(&tmQ, BF, 2, Q, D, Sq, ldq, 1, 0, B, bsq, 64, AB_BM) &&
              encode_4d(&tmK, BF, 2, K, D, Sk, ldk, 1, 0, B, bsk, 64, AB_BN) &&
              encode_4d(&tmV, BF, 2, V, D, Sk, ldv, 1, 0, B, bsv, 64, 64);
    if (!ok) {
        bmhrl_set_error("attn_fwd_bf16: cuTensorMapEncodeTiled failed");
        return BMHRL_ERR_CUDA;
    }
    static DeviceOnce attr;
    const int n_sm = device_sm_count();
    if (needs_setup(attr)) {
        const cudaFuncAttribute A = cudaFuncAttributeMaxDynamicSharedMemorySize;
        if (cudaFuncSetAttribute(attn_fwd_bf16_kernel<false, false>, A, AB_SMEM) != cudaSuccess ||
            cudaFuncSetAttribute(attn_fwd_bf16_kernel<false, true>, A, AB_SMEM) != cudaSuccess ||
            cudaFuncSetAttribute(attn_fwd_bf16_kernel<true, false>, A, AB_SMEM) != cudaSuccess ||
            cudaFuncSetAttribute(attn_fwd_bf16_kernel<true, true>, A, AB_SMEM) != cudaSuccess) {
            bmhrl_set_error("attn_fwd_bf16: cannot raise the shared-memory limit");
            return BMHRL_ERR_CUDA;
        }
        mark_done(attr);
    }
    AttnBf16Params p{};
    p.O = (__nv_bfloat16*)O; p.ldo = ldo; p.bso = bso; p.lse = lse; p.mask = mask; p.mask_kind = mask_kind;
    p.causal = causal; p.B = B; p.H = H; p.Sq = Sq; p.Sk = Sk; p.dk = dk;
    p.nqt = (Sq + AB_BM - 1) / AB_BM;
    p.n_items = B * H * p.nqt;
    p.scale = 1.0f / sqrtf((float)dk);
    if (ws && mask_kind == 1 && !causal) {  // key-padding mask: bit-pack it and trim fully masked tail tiles per batch row
        p.mwords = 4 * ((Sk + AB_BN - 1) / AB_BN);
        int* sk_eff = (int*)ws;
        uint32_t* mbits = (uint32_t*)ws + ((B + 3) & ~3);
        attn_mask_pack(mask, B, Sk, p.mwords, sk_eff, mbits, (cudaStream_t)stream);
        BMHRL_POST_LAUNCH("attn_mask_pack");
        p.sk_eff = sk_eff;
        p.mbits = mbits;
    }
    const bool masked = mask_kind != 0 || causal != 0;
    const bool use_pair = (dk == 128 || dk == 256) && n_sm >= 2 &&
                          (g_attn_bf16_pair == 2 || (g_attn_bf16_pair == 0 && p.nqt >= AB_PAIR_MIN_QTILES));
    if (use_pair) {  // CTA pairs: K split by keys (box 64 x 64), V by columns (box 64 x 128), two query tiles per item
        ok = encode_4d(&tmK, BF, 2, K, D, Sk, ldk, 1, 0, B, bsk, 64, 64) &&
             encode_4d(&tmV, BF, 2, V, D, Sk, ldv, 1, 0, B, bsv, 64, AB_BN);
        if (!ok) {
            bmhrl_set_error("attn_fwd_bf16: cuTensorMapEncodeTiled failed (pair)");
            return BMHRL_ERR_CUDA;
        }
        p.n_items = B * H * ((p.nqt + 1) >> 1);
        const int pairs = p.n_items < (n_sm >> 1) ? p.n_items : (n_sm >> 1);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2 * pairs);
        cfg.blockDim = dim3(192);
        cfg.dynamicSmemBytes = AB_SMEM;
        cfg.stream = (cudaStream_t)stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        const cudaError_t e = masked ? cudaLaunchKernelEx(&cfg, attn_fwd_bf16_kernel<true, true>, tmQ, tmK, tmV, p)
                                     : cudaLaunchKernelEx(&cfg, attn_fwd_bf16_kernel<true, false>, tmQ, tmK, tmV, p);
        if (e != cudaSuccess) {
            bmhrl_set_error(cudaGetErrorString(cudaGetLastError()));
            return BMHRL_ERR_CUDA;
        }
        bmhrl_count_launch(1);
        return BMHRL_OK;
    }
    const int grid = p.n_items < n_sm ? p.n_items : n_sm;
    if (masked) attn_fwd_bf16_kernel<false, true><<<grid, 192, AB_SMEM, (cudaStream_t)stream>>>(tmQ, tmK, tmV, p);
    else attn_fwd_bf16_kernel<false, false><<<grid, 192, AB_SMEM, (cudaStream_t)stream>>>(tmQ, tmK, tmV, p);
    BMHRL_POST_LAUNCH("attn_fwd_bf16");
    return BMHRL_OK;
}
