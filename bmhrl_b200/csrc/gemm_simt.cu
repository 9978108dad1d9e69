// fp32 SIMT GEMM with fused linear-layer epilogue (bias, ReLU, dropout, residual) and (outer,inner)
// batch strides.  This is the exact-fp32 path of the library: it backs the fp32 parity mode of every
// contraction on the hot path (projections model/multihead_attention.py:70-72,89; FFN model/blocks.py:
// 175-187; vocab projection model/bm_hrl_agent.py:463-466; and, batched over (b,h), the QK^T / PV
// contractions of model/multihead_attention.py:13,25 and their gradients).  The tensor-core (tcgen05)
// kernels in gemm_tc.cu / attn_tc.cu are validated against it.
//
// C[z](m,n) = beta * C[z](m,n) + resid[z](m,n) + dropout(act(alpha * sum_k A[z](m,k) B[z](k,n) + bias[n]))
//
// A(m,k) and B(k,n) are addressed through element strides, so every transpose combination is one kernel;
// loads are 128-bit whenever the unit-stride dimension allows it.  Tile 64x64x16, 256 threads, 4x4 per
// thread, register-prefetch double buffering (one __syncthreads per k-tile).
#include "common.cuh"
#include "../../include/bmhrl_b200.h"
#include "gemm_simt.cuh"

namespace bmhrl {

constexpr int BM = 64, BN = 64, BK = 16, PAD = 4;


// Loads a 4-element strip of a tile: element i lives at p + i*s_in for i in [0,4), valid iff pos+i < lim.
__device__ __forceinline__ float4 load_strip(const float* p, long long s_in, int pos, int lim, bool row_ok, int vec) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!row_ok) return v;
    if (vec && s_in == 1 && pos + 3 < lim) return ld_f4(p);
    if (pos + 0 < lim) v.x = p[0];
    if (pos + 1 < lim) v.y = p[s_in];
    if (pos + 2 < lim) v.z = p[2 * s_in];
    if (pos + 3 < lim) v.w = p[3 * s_in];
    return v;
}

template <bool A_KMAJOR, bool B_KMAJOR>
__global__ void __launch_bounds__(256) gemm_simt_kernel(const GemmParams p) {
    __shared__ __align__(16) float As[2][BK][BM + PAD];
    __shared__ __align__(16) float Bs[2][BK][BN + PAD];

    const int tid = threadIdx.x;
    const int z = blockIdx.z;
    const int zo = z / p.Hdiv, zi = z % p.Hdiv;
    const float* A = p.A + zo * p.bAo + zi * p.bAi;
    const float* B = p.B + zo * p.bBo + zi * p.bBi;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;

    // ---- load mapping
    // K-major operand: thread -> (row = tid/4, k-strip = (tid%4)*4); MN-major: (k = tid/16, strip = (tid%16)*4)
    int a_r, a_c, b_r, b_c;
    if (A_KMAJOR) { a_r = tid >> 2; a_c = (tid & 3) << 2; } else { a_r = tid >> 4; a_c = (tid & 15) << 2; }
    if (B_KMAJOR) { b_r = tid >> 2; b_c = (tid & 3) << 2; } else { b_r = tid >> 4; b_c = (tid & 15) << 2; }

    auto load_a = [&](int k0) -> float4 {
        if (A_KMAJOR) {  // a_r = m within tile, a_c = k offset; strip runs along k
            int m = m0 + a_r, k = k0 + a_c;
            return load_strip(A + (long long)m * p.sAm + (long long)k * p.sAk, p.sAk, k, p.K, m < p.M, p.vecA);
        } else {  // a_r = k within tile, a_c = m offset; strip runs along m
            int k = k0 + a_r, m = m0 + a_c;
            return load_strip(A + (long long)m * p.sAm + (long long)k * p.sAk, p.sAm, m, p.M, k < p.K, p.vecA);
        }
    };
    auto load_b = [&](int k0) -> float4 {
        if (B_KMAJOR) {  // b_r = n within tile, b_c = k offset
            int n = n0 + b_r, k = k0 + b_c;
            return load_strip(B + (long long)k * p.sBk + (long long)n * p.sBn, p.sBk, k, p.K, n < p.N, p.vecB);
        } else {  // b_r = k within tile, b_c = n offset
            int k = k0 + b_r, n = n0 + b_c;
            return load_strip(B + (long long)k * p.sBk + (long long)n * p.sBn, p.sBn, n, p.N, k < p.K, p.vecB);
        }
    };
    auto store_a = [&](int buf, float4 v) {
        if (A_KMAJOR) {
            As[buf][a_c + 0][a_r] = v.x; As[buf][a_c + 1][a_r] = v.y;
            As[buf][a_c + 2][a_r] = v.z; As[buf][a_c + 3][a_r] = v.w;
        } else {
            st_f4(&As[buf][a_r][a_c], v);
        }
    };
    auto store_b = [&](int buf, float4 v) {
        if (B_KMAJOR) {
            Bs[buf][b_c + 0][b_r] = v.x; Bs[buf][b_c + 1][b_r] = v.y;
            Bs[buf][b_c + 2][b_r] = v.z; Bs[buf][b_c + 3][b_r] = v.w;
        } else {
            st_f4(&Bs[buf][b_r][b_c], v);
        }
    };

    const int ty = tid >> 4, tx = tid & 15;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    const int nk = (p.K + BK - 1) / BK;
    float4 ra = load_a(0), rb = load_b(0);
    store_a(0, ra);
    store_b(0, rb);
    __syncthreads();
    for (int t = 0; t < nk; ++t) {
        const int buf = t & 1;
        if (t + 1 < nk) { ra = load_a((t + 1) * BK); rb = load_b((t + 1) * BK); }
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            float4 a = ld_f4(&As[buf][kk][ty * 4]);
            float4 b = ld_f4(&Bs[buf][kk][tx * 4]);
            float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        if (t + 1 < nk) {
            store_a(buf ^ 1, ra);
            store_b(buf ^ 1, rb);
        }
        __syncthreads();
    }

    // ---- epilogue
    float* C = p.C + zo * p.bCo + zi * p.bCi;
    const float* R = p.resid ? p.resid + zo * p.bRo + zi * p.bRi : nullptr;
    const float keep_scale = p.drop_p > 0.f ? 1.f / (1.f - p.drop_p) : 1.f;
    const int nb = n0 + tx * 4;
    float bia[4] = {0.f, 0.f, 0.f, 0.f};
    if (p.bias) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (nb + j < p.N) bia[j] = p.bias[nb + j];
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty * 4 + i;
        if (m >= p.M) continue;
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float x = p.alpha * acc[i][j] + bia[j];
            if (p.act == 1) x = fmaxf(x, 0.f);
            v[j] = x;
        }
        if (p.drop_p > 0.f) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                unsigned long long idx = ((unsigned long long)z * p.M + m) * (unsigned long long)p.N + (nb + j);
                v[j] = dropout_keep(p.rng, p.salt, idx, p.drop_p) ? v[j] * keep_scale : 0.f;
            }
        }
        float* crow = C + (long long)m * p.ldc + nb;
        if (R) {
            const float* rrow = R + (long long)m * p.ldr + nb;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (nb + j < p.N) v[j] += rrow[j];
        }
        if (p.beta != 0.f) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (nb + j < p.N) v[j] += p.beta * crow[j];
        }
        if (p.vecC && nb + 3 < p.N) {
            st_f4(crow, make_float4(v[0], v[1], v[2], v[3]));
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (nb + j < p.N) crow[j] = v[j];
        }
    }
}

int launch_gemm_simt(const GemmParams& p, int batch, cudaStream_t st) {
    if (p.M <= 0 || p.N <= 0 || batch <= 0) return BMHRL_OK;
    dim3 grid((p.N + BN - 1) / BN, (p.M + BM - 1) / BM, batch);
    const bool ak = (p.sAk == 1), bk = (p.sBk == 1);
    if (ak && bk) gemm_simt_kernel<true, true><<<grid, 256, 0, st>>>(p);
    else if (ak && !bk) gemm_simt_kernel<true, false><<<grid, 256, 0, st>>>(p);
    else if (!ak && bk) gemm_simt_kernel<false, true><<<grid, 256, 0, st>>>(p);
    else gemm_simt_kernel<false, false><<<grid, 256, 0, st>>>(p);
    BMHRL_POST_LAUNCH("gemm_simt");
    return BMHRL_OK;
}


// ------------------------------------------------------------------------------------------------------------------
// Skinny linear layer: y = act(x W^T + b) + resid for M <= 32 rows (token-by-token decoding: M = B * prefix length or
// M = B).  The tcgen05 GEMM pays its fixed costs (tensor-map fetch, TMEM allocation, barrier set-up, a 128-row MMA
// tile that is > 75 % padding) on every one of the ~45 projections per generated token; here one warp owns one output
// column: it streams its row of W once (128-bit loads), the <= 32 rows of x sit in shared memory in K chunks, and the
// lanes' partial sums are reduced with shuffles.  Exact fp32 (FMA), so it is at least as accurate as the tf32 path it
// stands in for.  Tensor-core math mode only: in fp32 mode every linear layer stays on ONE kernel so that row results
// do not depend on how many rows a call has (the incremental critic relies on that, rnn.cu).
constexpr int SK_WARPS = 8;  // output columns per CTA
constexpr int SK_KC = 256;   // K chunk staged in shared memory

template <int MT>
__global__ void __launch_bounds__(SK_WARPS * 32) linear_skinny_kernel(const float* __restrict__ X, long long ldx,
                                                                     const float* __restrict__ W,
                                                                     const float* __restrict__ bias,
                                                                     const float* __restrict__ resid, long long ldr,
                                                                     float* __restrict__ Y, long long ldy, int M, int N,
                                                                     int K, int act) {
    __shared__ float4 xs[MT][SK_KC / 4];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = blockIdx.x * SK_WARPS + warp;
    const bool valid = n < N;
    const float* wrow = W + (long long)(valid ? n : 0) * K;
    float acc[MT];
#pragma unroll
    for (int m = 0; m < MT; ++m) acc[m] = 0.f;
    for (int k0 = 0; k0 < K; k0 += SK_KC) {
        const int q = min(SK_KC, K - k0) >> 2;  // float4 per row in this chunk (K % 4 == 0)
        __syncthreads();
        for (int i = threadIdx.x; i < MT * (SK_KC / 4); i += SK_WARPS * 32) {
            const int m = i / (SK_KC / 4), c = i % (SK_KC / 4);
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (m < M && c < q) v = ld_f4(X + (long long)m * ldx + k0 + 4 * c);
            xs[m][c] = v;
        }
        __syncthreads();
        for (int c = lane; c < q; c += 32) {
            const float4 w = ld_f4(wrow + k0 + 4 * c);
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                const float4 x = xs[m][c];
                acc[m] = fmaf(w.x, x.x, acc[m]);
                acc[m] = fmaf(w.y, x.y, acc[m]);
                acc[m] = fmaf(w.z, x.z, acc[m]);
                acc[m] = fmaf(w.w, x.w, acc[m]);
            }
        }
    }
    float mine = 0.f;
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        const float s = warp_sum(acc[m]);
        if (lane == m) mine = s;
    }
    if (!valid || lane >= M) return;
    float v = mine + (bias ? bias[n] : 0.f);
    if (act == 1) v = fmaxf(v, 0.f);
    if (resid) v += resid[(long long)lane * ldr + n];
    Y[(long long)lane * ldy + n] = v;
}

static int g_skinny_linear = -1;  // -1: not read yet (BMHRL_SKINNY_LINEAR=0 disables)

static bool skinny_linear_enabled() {
    if (g_skinny_linear < 0) {
        const char* e = getenv("BMHRL_SKINNY_LINEAR");
        g_skinny_linear = (e && e[0] == '0') ? 0 : 1;
    }
    return g_skinny_linear == 1;
}

static int launch_linear_skinny(const float* X, long long ldx, const float* W, const float* bias, const float* resid,
                                long long ldr, float* Y, long long ldy, int M, int N, int K, int act, cudaStream_t st) {
    if (M <= 0 || N <= 0) return BMHRL_OK;
    const dim3 grid((N + SK_WARPS - 1) / SK_WARPS);
    if (M <= 8) linear_skinny_kernel<8><<<grid, SK_WARPS * 32, 0, st>>>(X, ldx, W, bias, resid, ldr, Y, ldy, M, N, K, act);
    else if (M <= 16)
        linear_skinny_kernel<16><<<grid, SK_WARPS * 32, 0, st>>>(X, ldx, W, bias, resid, ldr, Y, ldy, M, N, K, act);
    else linear_skinny_kernel<32><<<grid, SK_WARPS * 32, 0, st>>>(X, ldx, W, bias, resid, ldr, Y, ldy, M, N, K, act);
    BMHRL_POST_LAUNCH("linear_skinny");
    return BMHRL_OK;
}


}  // namespace bmhrl

using namespace bmhrl;

extern "C" int bmhrl_gemm_f32(const float* A, long long sAm, long long sAk, const float* B, long long sBk,
                              long long sBn, float* C, long long ldc, int M, int N, int K, int batch, int Hdiv,
                              long long bAo, long long bAi, long long bBo, long long bBi, long long bCo,
                              long long bCi, float alpha, float beta, void* stream) {
    BMHRL_CHECK_ARG(A && B && C, "bmhrl_gemm_f32: null pointer");
    BMHRL_CHECK_ARG(M >= 0 && N >= 0 && K >= 0 && batch >= 0 && Hdiv >= 1, "bmhrl_gemm_f32: bad sizes");
    GemmParams p{};
    p.A = A; p.B = B; p.C = C; p.M = M; p.N = N; p.K = K;
    p.sAm = sAm; p.sAk = sAk; p.sBk = sBk; p.sBn = sBn; p.ldc = ldc;
    p.Hdiv = Hdiv; p.bAo = bAo; p.bAi = bAi; p.bBo = bBo; p.bBi = bBi; p.bCo = bCo; p.bCi = bCi;
    p.alpha = alpha; p.beta = beta;
    if (math_mode() == 1 && batch == 1 && (sAk == 1 || sAm == 1) && (sBk == 1 || sBn == 1) && (beta == 0.f || beta == 1.f)) {
        TcGemmArgs t{};
        t.A = A; t.a_mn = (sAk == 1) ? 0 : 1; t.lda = t.a_mn ? sAk : sAm;
        t.B = B; t.b_mn = (sBk == 1) ? 0 : 1; t.ldb = t.b_mn ? sBk : sBn;
        t.C = C; t.ldc = ldc; t.M = M; t.N = N; t.K = K; t.alpha = alpha; t.beta = beta;
        if (tc_gemm_supported(t)) return launch_gemm_tc(t, (cudaStream_t)stream);
    }
    gemm_set_vec_flags(p);
    return launch_gemm_simt(p, batch, (cudaStream_t)stream);
}

extern "C" int bmhrl_set_skinny_linear(int on) {
    g_skinny_linear = on ? 1 : 0;
    return BMHRL_OK;
}

extern "C" int bmhrl_linear_fwd_f32(const float* X, long long ldx, const float* W, const float* bias,
                                    const float* resid, long long ldr, float* Y, long long ldy, int M, int N,
                                    int K, int act, float drop_p, const unsigned long long* rng_state,
                                    unsigned int salt, void* stream) {
    BMHRL_CHECK_ARG(X && W && Y, "bmhrl_linear_fwd_f32: null pointer");
    BMHRL_CHECK_ARG(M >= 0 && N >= 0 && K >= 0, "bmhrl_linear_fwd_f32: bad sizes");
    BMHRL_CHECK_ARG(act == 0 || act == 1, "bmhrl_linear_fwd_f32: act must be 0 (none) or 1 (relu)");
    BMHRL_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "bmhrl_linear_fwd_f32: drop_p out of range");
    BMHRL_CHECK_ARG(drop_p == 0.f || rng_state, "bmhrl_linear_fwd_f32: dropout needs rng_state");
    GemmParams p{};
    p.A = X; p.B = W; p.C = Y; p.M = M; p.N = N; p.K = K;
    p.sAm = ldx; p.sAk = 1; p.sBk = 1; p.sBn = K; p.ldc = ldy;  // W is (N,K) row-major: B(k,n) = W[n*K + k]
    p.Hdiv = 1;
    p.alpha = 1.f; p.beta = 0.f; p.bias = bias; p.act = act; p.resid = resid; p.ldr = ldr;
    p.drop_p = drop_p; p.rng = rng_state; p.salt = salt;
    if (math_mode() == 1 && M <= 32 && K > 0 && drop_p == 0.f && (K & 3) == 0 && (ldx & 3) == 0 &&
        ((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(W)) & 15) == 0 && skinny_linear_enabled())
        return launch_linear_skinny(X, ldx, W, bias, resid, ldr, Y, ldy, M, N, K, act, (cudaStream_t)stream);
    if (math_mode() == 1) {
        TcGemmArgs t{};
        t.A = X; t.lda = ldx; t.a_mn = 0; t.B = W; t.ldb = K; t.b_mn = 0; t.C = Y; t.ldc = ldy;
        t.M = M; t.N = N; t.K = K; t.alpha = 1.f; t.beta = 0.f; t.bias = bias; t.act = act; t.resid = resid; t.ldr = ldr;
        t.drop_p = drop_p; t.rng = rng_state; t.salt = salt;
        if (tc_gemm_supported(t)) return launch_gemm_tc(t, (cudaStream_t)stream);
    }
    gemm_set_vec_flags(p);
    return launch_gemm_simt(p, 1, (cudaStream_t)stream);
}
