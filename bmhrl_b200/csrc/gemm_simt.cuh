// Parameter block of the fp32 SIMT GEMM (gemm_simt.cu); shared with the fp32 attention path.
#pragma once
#include "common.cuh"

namespace bmhrl {

struct GemmParams {
    const float* A;
    const float* B;
    float* C;
    int M, N, K;
    long long sAm, sAk, sBk, sBn, ldc;
    int Hdiv;
    long long bAo, bAi, bBo, bBi, bCo, bCi;
    float alpha, beta;
    const float* bias;
    int act;
    const float* resid;
    long long ldr, bRo, bRi;
    float drop_p;
    const unsigned long long* rng;
    unsigned int salt;
    int vecA, vecB, vecC;  // 128-bit access allowed (alignment checked on host)
};

static inline bool mult4(long long v) { return (v & 3) == 0; }

// Fills the 128-bit-access flags from pointer alignment and strides.
static inline void gemm_set_vec_flags(GemmParams& p) {
    p.vecA = aligned16(p.A) && (p.sAm == 1 || p.sAk == 1) && mult4(p.sAm == 1 ? p.sAk : p.sAm) && mult4(p.bAo) &&
             mult4(p.bAi);
    p.vecB = aligned16(p.B) && (p.sBk == 1 || p.sBn == 1) && mult4(p.sBk == 1 ? p.sBn : p.sBk) && mult4(p.bBo) &&
             mult4(p.bBi);
    p.vecC = aligned16(p.C) && mult4(p.ldc) && mult4(p.bCo) && mult4(p.bCi);
}

int launch_gemm_simt(const GemmParams& p, int batch, cudaStream_t st);

// ---- tensor-core path (gemm_tc.cu)
struct TcGemmArgs {
    const float* A; long long lda; int a_mn;  // a_mn = 0: A(m,k) at A[m*lda + k]; 1: A(m,k) at A[k*lda + m]
    const float* B; long long ldb; int b_mn;  // b_mn = 0: B(k,n) at B[n*ldb + k]; 1: B(k,n) at B[k*ldb + n]
    float* C; long long ldc;
    int M, N, K;
    float alpha, beta;
    const float* bias; int act; const float* resid; long long ldr;
    float drop_p; const unsigned long long* rng; unsigned int salt;
    // batched problems: z in [0, batch): outer = z / Hdiv, inner = z % Hdiv; operand offsets outer*b?o + inner*b?i
    // elements (batch <= 1: a single problem, everything below ignored)
    int batch, Hdiv;
    long long bAo, bAi, bBo, bBi, bCo, bCi;
};
// Routes a GemmParams problem to the tensor-core kernel when the math mode asks for it and TMA can address the
// operands, else to the fp32 SIMT kernel.
int launch_gemm_auto(GemmParams& p, int batch, cudaStream_t st);
// fused tcgen05 attention forward (attn_tc.cu); returns 1 when the shape is outside what the fused kernel handles
int launch_attn_fwd_tc(const float* Q, long long ldq, long long bsq, const float* K, long long ldk, long long bsk,
                       const float* V, long long ldv, long long bsv, const unsigned char* mask, int mask_kind, int causal,
                       float* O, long long ldo, long long bso, float* P, int ldp, int B, int H, int Sq, int Sk, int dk,
                       float drop_p, const unsigned long long* rng, unsigned int salt, cudaStream_t st);
int math_mode();                                  // 0 exact fp32 (SIMT), 1 tf32 (tcgen05)
bool tc_gemm_supported(const TcGemmArgs& a);      // alignment / pitch constraints of TMA
int launch_gemm_tc(const TcGemmArgs& a, cudaStream_t st);

}  // namespace bmhrl
