// fp32 attention core (K1, exact-fp32 path): softmax(mask_replace(Q K^T / sqrt(d_k), -1e9)) V with dropout on the
// product, forward and backward.  Reference: attention() model/multihead_attention.py:7-31.
//
// This path materialises the (B,H,Sq,Sk) probability tensor like the reference does and runs the four
// contractions through the batched fp32 SIMT GEMM; it is the parity anchor (1e-3 fp32 tolerance) for the fused
// tcgen05 kernel in attn_tc.cu and serves the short sequences of the B=16 training step.
//
// Mask semantics (bit-for-bit the reference's): a masked score is REPLACED by -1e9, so a fully masked row is a
// uniform distribution over all Sk keys (padding included); the gradient through a replaced score is zero.
#include "common.cuh"
#include "../../include/bmhrl_b200.h"
#include "gemm_simt.cuh"

namespace bmhrl {

__device__ __forceinline__ bool key_visible(const unsigned char* mask, int mask_kind, int causal, int b, int q, int k,
                                            int Sq, int Sk) {
    bool vis = true;
    if (mask_kind == 1) vis = mask[(long long)b * Sk + k] != 0;
    else if (mask_kind == 2) vis = mask[((long long)b * Sq + q) * Sk + k] != 0;
    if (causal && k > q) vis = false;
    return vis;
}

// One warp per (b,h,q) row of S (leading dimension ldp >= Sk); in place.
__global__ void __launch_bounds__(128) masked_softmax_fwd_kernel(float* __restrict__ S,
                                                                 const unsigned char* __restrict__ mask,
                                                                 int mask_kind, int causal, int B, int H, int Sq,
                                                                 int Sk, int ldp) {
    const long long row = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= (long long)B * H * Sq) return;
    const int q = (int)(row % Sq);
    const int b = (int)(row / ((long long)H * Sq));
    float* s = S + row * ldp;
    float mx = -INFINITY;
    for (int k = lane; k < Sk; k += 32) {
        float v = s[k];
        if (mask_kind != 0 && !key_visible(mask, mask_kind, causal, b, q, k, Sq, Sk)) v = kNegFill;
        s[k] = v;
        mx = fmaxf(mx, v);
    }
    mx = warp_max(mx);
    float sum = 0.f;
    for (int k = lane; k < Sk; k += 32) {
        float e = expf(s[k] - mx);
        s[k] = e;
        sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    for (int k = lane; k < Sk; k += 32) s[k] *= inv;
    for (int k = Sk + lane; k < ldp; k += 32) s[k] = 0.f;
}

// dS = scale * P * (dP - sum_k P dP), zero where the score was replaced; in place on dP.
__global__ void __launch_bounds__(128) masked_softmax_bwd_kernel(float* __restrict__ dP, const float* __restrict__ P,
                                                                 const unsigned char* __restrict__ mask,
                                                                 int mask_kind, int causal, int B, int H, int Sq,
                                                                 int Sk, int ldp, float scale) {
    const long long row = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= (long long)B * H * Sq) return;
    const int q = (int)(row % Sq);
    const int b = (int)(row / ((long long)H * Sq));
    float* d = dP + row * ldp;
    const float* p = P + row * ldp;
    float dot = 0.f;
    for (int k = lane; k < Sk; k += 32) dot += p[k] * d[k];
    dot = warp_sum(dot);
    for (int k = lane; k < Sk; k += 32) {
        float v = scale * p[k] * (d[k] - dot);
        if (mask_kind != 0 && !key_visible(mask, mask_kind, causal, b, q, k, Sq, Sk)) v = 0.f;
        d[k] = v;
    }
    for (int k = Sk + lane; k < ldp; k += 32) d[k] = 0.f;
}

// Gradient through the dropout applied to the per-head attention output; the keep decision is indexed exactly like
// the PV GEMM epilogue: idx = ((b*H + h)*Sq + s)*dk + d.
__global__ void attn_dropout_bwd_kernel(const float* __restrict__ dO, long long ldo, long long bso,
                                        float* __restrict__ out, int B, int H, int Sq, int dk, float p,
                                        const unsigned long long* __restrict__ rng, unsigned int salt, int vec) {
    const long long n = (long long)B * H * Sq * dk;
    const float ks = 1.f / (1.f - p);
    const long long stride = (long long)gridDim.x * blockDim.x;
    if (vec) {  // dk % 4 == 0 and 16-byte aligned rows: 4 elements = one Philox block, one 128-bit access
        for (long long i4 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i4 < (n >> 2); i4 += stride) {
            const long long i = i4 << 2;
            int d = (int)(i % dk);
            long long t = i / dk;
            int s = (int)(t % Sq);
            t /= Sq;
            int h = (int)(t % H);
            int b = (int)(t / H);
            float4 g = ld_f4(dO + (long long)b * bso + (long long)s * ldo + h * dk + d);
            const Philox4 rb = philox_block(rng, salt, (unsigned long long)i4);
            g.x = u32_to_unit(rb.x) >= p ? g.x * ks : 0.f;
            g.y = u32_to_unit(rb.y) >= p ? g.y * ks : 0.f;
            g.z = u32_to_unit(rb.z) >= p ? g.z * ks : 0.f;
            g.w = u32_to_unit(rb.w) >= p ? g.w * ks : 0.f;
            st_f4(out + ((long long)b * Sq + s) * ((long long)H * dk) + h * dk + d, g);
        }
        return;
    }
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        int d = (int)(i % dk);
        long long t = i / dk;
        int s = (int)(t % Sq);
        t /= Sq;
        int h = (int)(t % H);
        int b = (int)(t / H);
        float g = dO[(long long)b * bso + (long long)s * ldo + h * dk + d];
        out[((long long)b * Sq + s) * ((long long)H * dk) + h * dk + d] = dropout_keep(rng, salt, i, p) ? g * ks : 0.f;
    }
}

}  // namespace bmhrl
using namespace bmhrl;

static inline int round4(int v) { return (v + 3) & ~3; }

extern "C" long long bmhrl_mha_core_probs_floats(int B, int H, int Sq, int Sk) {
    return (long long)B * H * Sq * round4(Sk);
}

extern "C" int bmhrl_mha_core_fwd_f32(const float* Q, long long ldq, long long bsq, const float* K, long long ldk,
                                      long long bsk, const float* V, long long ldv, long long bsv,
                                      const unsigned char* mask, int mask_kind, int causal, float* O, long long ldo,
                                      long long bso, float* P, int B, int H, int Sq, int Sk, int dk, float drop_p,
                                      const unsigned long long* rng_state, unsigned int salt, void* stream) {
    BMHRL_CHECK_ARG(Q && K && V && O && P, "bmhrl_mha_core_fwd_f32: null pointer");
    BMHRL_CHECK_ARG(mask_kind >= 0 && mask_kind <= 2, "bmhrl_mha_core_fwd_f32: mask_kind must be 0,1,2");
    BMHRL_CHECK_ARG(mask_kind == 0 || mask, "bmhrl_mha_core_fwd_f32: mask pointer missing");
    BMHRL_CHECK_ARG(!causal || Sq == Sk, "bmhrl_mha_core_fwd_f32: causal needs Sq == Sk");
    BMHRL_CHECK_ARG(B > 0 && H > 0 && Sq > 0 && Sk > 0 && dk > 0, "bmhrl_mha_core_fwd_f32: bad sizes");
    BMHRL_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f && (drop_p == 0.f || rng_state), "bmhrl_mha_core_fwd_f32: dropout");
    cudaStream_t st = (cudaStream_t)stream;
    const int ldp = round4(Sk);
    const float scale = 1.0f / sqrtf((float)dk);
    // a causal mask without an explicit mask is still a mask: the reference only applies tril when mask is given,
    // callers pass mask_kind 0 + causal 0 for "mask is None".
    int rc;
    if (math_mode() == 1) {  // tensor-core mode: one fused tcgen05 kernel (attn_tc.cu)
        rc = launch_attn_fwd_tc(Q, ldq, bsq, K, ldk, bsk, V, ldv, bsv, mask, mask_kind, causal, O, ldo, bso, P, ldp, B, H,
                                Sq, Sk, dk, drop_p, rng_state, salt, st);
        if (rc != 1) return rc;
    }
    {   // S = scale * Q K^T
        GemmParams p{};
        p.A = Q; p.sAm = ldq; p.sAk = 1; p.bAo = bsq; p.bAi = dk;
        p.B = K; p.sBk = 1; p.sBn = ldk; p.bBo = bsk; p.bBi = dk;
        p.C = P; p.ldc = ldp; p.bCo = (long long)H * Sq * ldp; p.bCi = (long long)Sq * ldp;
        p.M = Sq; p.N = Sk; p.K = dk; p.Hdiv = H; p.alpha = scale; p.beta = 0.f;
        if ((rc = launch_gemm_auto(p, B * H, st)) != BMHRL_OK) return rc;
    }
    {
        long long rows = (long long)B * H * Sq;
        masked_softmax_fwd_kernel<<<(unsigned)((rows + 3) / 4), 128, 0, st>>>(P, mask, mask_kind, causal, B, H, Sq, Sk,
                                                                             ldp);
        BMHRL_POST_LAUNCH("masked_softmax_fwd");
    }
    {   // O = dropout(P V), written head-merged (B, Sq, H*dk)
        GemmParams p{};
        p.A = P; p.sAm = ldp; p.sAk = 1; p.bAo = (long long)H * Sq * ldp; p.bAi = (long long)Sq * ldp;
        p.B = V; p.sBk = ldv; p.sBn = 1; p.bBo = bsv; p.bBi = dk;
        p.C = O; p.ldc = ldo; p.bCo = bso; p.bCi = dk;
        p.M = Sq; p.N = dk; p.K = Sk; p.Hdiv = H; p.alpha = 1.f; p.beta = 0.f;
        p.drop_p = drop_p; p.rng = rng_state; p.salt = salt;
        if ((rc = launch_gemm_auto(p, B * H, st)) != BMHRL_OK) return rc;
    }
    return BMHRL_OK;
}

extern "C" int bmhrl_mha_core_bwd_f32(const float* dO, long long ldo, long long bso, const float* Q, long long ldq,
                                      long long bsq, const float* K, long long ldk, long long bsk, const float* V,
                                      long long ldv, long long bsv, const float* P, const unsigned char* mask,
                                      int mask_kind, int causal, float* dQ, long long lddq, long long bsdq, float* dK,
                                      long long lddk, long long bsdk, float* dV, long long lddv, long long bsdv,
                                      float* ws_dP, float* ws_dO, int B, int H, int Sq, int Sk, int dk, float drop_p,
                                      const unsigned long long* rng_state, unsigned int salt, void* stream) {
    BMHRL_CHECK_ARG(dO && Q && K && V && P && dQ && dK && dV && ws_dP, "bmhrl_mha_core_bwd_f32: null pointer");
    BMHRL_CHECK_ARG(mask_kind >= 0 && mask_kind <= 2 && (mask_kind == 0 || mask), "bmhrl_mha_core_bwd_f32: mask");
    BMHRL_CHECK_ARG(drop_p == 0.f || (ws_dO && rng_state), "bmhrl_mha_core_bwd_f32: dropout needs ws_dO and rng_state");
    cudaStream_t st = (cudaStream_t)stream;
    const int ldp = round4(Sk);
    const float scale = 1.0f / sqrtf((float)dk);
    const long long pbo = (long long)H * Sq * ldp, pbi = (long long)Sq * ldp;
    int rc;
    const float* g = dO;
    long long ldg = ldo, bsg = bso;
    if (drop_p > 0.f) {
        long long n = (long long)B * H * Sq * dk;
        const int vec = (dk % 4 == 0) && (ldo % 4 == 0) && (bso % 4 == 0) && aligned16(dO) && aligned16(ws_dO);
        const long long items = vec ? n / 4 : n;
        int blocks = (int)((items + 255) / 256);
        if (blocks > 148 * 8) blocks = 148 * 8;
        attn_dropout_bwd_kernel<<<blocks, 256, 0, st>>>(dO, ldo, bso, ws_dO, B, H, Sq, dk, drop_p, rng_state, salt, vec);
        BMHRL_POST_LAUNCH("attn_dropout_bwd");
        g = ws_dO; ldg = (long long)H * dk; bsg = (long long)Sq * H * dk;
    }
    {   // dV = P^T g     (Sk x dk) = (Sk x Sq)(Sq x dk)
        GemmParams p{};
        p.A = P; p.sAm = 1; p.sAk = ldp; p.bAo = pbo; p.bAi = pbi;
        p.B = g; p.sBk = ldg; p.sBn = 1; p.bBo = bsg; p.bBi = dk;
        p.C = dV; p.ldc = lddv; p.bCo = bsdv; p.bCi = dk;
        p.M = Sk; p.N = dk; p.K = Sq; p.Hdiv = H; p.alpha = 1.f;
        if ((rc = launch_gemm_auto(p, B * H, st)) != BMHRL_OK) return rc;
    }
    {   // dP = g V^T     (Sq x Sk) = (Sq x dk)(dk x Sk)
        GemmParams p{};
        p.A = g; p.sAm = ldg; p.sAk = 1; p.bAo = bsg; p.bAi = dk;
        p.B = V; p.sBk = 1; p.sBn = ldv; p.bBo = bsv; p.bBi = dk;
        p.C = ws_dP; p.ldc = ldp; p.bCo = pbo; p.bCi = pbi;
        p.M = Sq; p.N = Sk; p.K = dk; p.Hdiv = H; p.alpha = 1.f;
        if ((rc = launch_gemm_auto(p, B * H, st)) != BMHRL_OK) return rc;
    }
    {
        long long rows = (long long)B * H * Sq;
        masked_softmax_bwd_kernel<<<(unsigned)((rows + 3) / 4), 128, 0, st>>>(ws_dP, P, mask, mask_kind, causal, B, H, Sq,
                                                                             Sk, ldp, scale);
        BMHRL_POST_LAUNCH("masked_softmax_bwd");
    }
    {   // dQ = dS K      (Sq x dk) = (Sq x Sk)(Sk x dk)
        GemmParams p{};
        p.A = ws_dP; p.sAm = ldp; p.sAk = 1; p.bAo = pbo; p.bAi = pbi;
        p.B = K; p.sBk = ldk; p.sBn = 1; p.bBo = bsk; p.bBi = dk;
        p.C = dQ; p.ldc = lddq; p.bCo = bsdq; p.bCi = dk;
        p.M = Sq; p.N = dk; p.K = Sk; p.Hdiv = H; p.alpha = 1.f;
        if ((rc = launch_gemm_auto(p, B * H, st)) != BMHRL_OK) return rc;
    }
    {   // dK = dS^T Q    (Sk x dk) = (Sk x Sq)(Sq x dk)
        GemmParams p{};
        p.A = ws_dP; p.sAm = 1; p.sAk = ldp; p.bAo = pbo; p.bAi = pbi;
        p.B = Q; p.sBk = ldq; p.sBn = 1; p.bBo = bsq; p.bBi = dk;
        p.C = dK; p.ldc = lddk; p.bCo = bsdk; p.bCi = dk;
        p.M = Sk; p.N = dk; p.K = Sq; p.Hdiv = H; p.alpha = 1.f;
        if ((rc = launch_gemm_auto(p, B * H, st)) != BMHRL_OK) return rc;
    }
    return BMHRL_OK;
}
