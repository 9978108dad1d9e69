/* bmhrl_b200.h -- C ABI of libbmhrl_b200.so (sm_100a).
 *
 * The reference (Berghojo/bmhrl) is pure Python/PyTorch and has no FFI of its own: its "plugin boundary" for
 * the hot path is the nn.Module surface (model/multihead_attention.py, model/blocks.py, model/bm_hrl_agent.py,
 * loss/label_smoothing.py, loss/biased_kl.py).  The Python modules in bmhrl_b200/ keep that surface and call the
 * entry points below through ctypes; each entry point cites the reference code it replaces (paths relative to the
 * reference root).  INTEGRATION.md shows the binding a maintainer of the reference would add.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless stated; tensors are fp32, row-major, contiguous unless a leading
 *     dimension / stride argument says otherwise; token / index tensors are int64 (torch.long), segment labels and
 *     maps int32, masks uint8 (torch.bool storage);
 *   - `stream` is a cudaStream_t; all work is enqueued on it, nothing synchronises, nothing allocates: workspaces are
 *     passed in by the caller (so every call is CUDA-graph capturable);
 *   - return value 0 = ok, negative = error (BMHRL_ERR_*), message via bmhrl_last_error(); never throws / exits;
 *   - dropout / sampling randomness is counter based (Philox4x32-10): rng_state points to two device uint64
 *     {seed, step}; `salt` names the call site; forward and backward of an op regenerate the same mask;
 *   - there is no CPU implementation behind any of these symbols.
 */
#ifndef BMHRL_B200_H
#define BMHRL_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define BMHRL_ABI_VERSION 1

/* ---- library state ---------------------------------------------------------------------------------------- */
int bmhrl_abi_version(void);
int bmhrl_check_device(void);             /* 0 iff the current device is sm_100 */
const char* bmhrl_last_error(void);       /* thread-local message of the last failing call */
long long bmhrl_launch_count(void);       /* kernels launched by this library since the last reset */
void bmhrl_reset_launch_count(void);
/* A new non-blocking CUDA stream on the current device (cudaStreamCreateWithFlags).  The host layer forks its
 * concurrent branches onto streams made here instead of framework-pooled ones: a pool hands the same handle out again
 * after a few dozen requests, and two "different" side streams that are one CUDA stream defeat the per-stream ownership
 * the caching allocator relies on while a CUDA graph is being captured (a block freed by one branch was re-used by an
 * unordered branch).  Replaces nothing in the reference (its nn.DataParallel threads share the default stream). */
int bmhrl_stream_create(void** stream);
/* Programmatic dependent launch of the chain kernels of the bf16 math mode (GEMM, one-tile attention, LayerNorm, cast,
 * column sums): with it, the next kernel of a stream becomes resident and runs its prologue -- and the TMA loads of
 * operands no earlier launch produces (weights: bmhrl_gemm_desc.b_static) -- while the current one is still running;
 * every such kernel executes griddepcontrol.wait before it touches anything a predecessor wrote.  1 (default; the
 * environment variable BMHRL_PDL=0 sets the initial value) / 0. */
int bmhrl_set_pdl(int on);
int bmhrl_get_pdl(void);
/* use != 0: every following launch of the chain kernels (GEMM, attention, LayerNorm, cast) carries
 * cudaLaunchAttributePriority = priority (CUDA stream-priority scale: lower number = placed first); use == 0: no attribute.
 * Takes effect inside CUDA graphs instantiated with cudaGraphInstantiateFlagUseNodePriority (train.StepRunner does that
 * when asked to: the kernels of the backward chain then overtake queued weight-gradient CTAs). */
int bmhrl_set_launch_priority(int use, int priority);

/* ---- math mode ------------------------------------------------------------------------------------------------ */
/* 0 (default): every contraction in exact fp32 on the SIMT pipes (parity bound 1e-3, in practice ~1e-6).
 * 1: projection / FFN / vocabulary / gradient GEMMs on the 5th-gen tensor cores (tcgen05.mma kind::tf32, TMA-staged
 *    fp32 operands, fp32 accumulation in TMEM; parity bound 2e-2 like the north_star's reduced-precision bound). */
int bmhrl_set_math_mode(int mode);
int bmhrl_get_math_mode(void);
/* tuning / test hook of the tensor-core GEMM: force the tile width (0 = cost model, 32, 64, 128) and the K split
 * (0 = cost model; > 0: split-K with RED accumulation, plain GEMMs only; -2 / -4: thread-block cluster of that many
 * CTAs along K, partial accumulators reduce-scattered through distributed shared memory, any epilogue). */
int bmhrl_tc_tuning(int force_bn, int force_split);
/* test hook: 0 = run the attention core as separate (batched tcgen05) contractions + softmax kernels even in math
 * mode 1; 1 (default) = the fused tcgen05 attention kernel. */
int bmhrl_attn_tuning(int fused);

/* ---- contractions (fp32 SIMT path, gemm_simt.cu) ------------------------------------------------------------ */
/* C[z](m,n) = beta*C + alpha * sum_k A[z](m,k) B[z](k,n);  A(m,k) at A + m*sAm + k*sAk, B(k,n) at B + k*sBk + n*sBn.
 * z in [0,batch): outer = z / Hdiv, inner = z % Hdiv, offsets outer*b?o + inner*b?i.
 * Replaces the aten mm/bmm calls behind nn.Linear backward (dX = dY W, dW = dY^T X). */
int bmhrl_gemm_f32(const float* A, long long sAm, long long sAk, const float* B, long long sBk, long long sBn,
                   float* C, long long ldc, int M, int N, int K, int batch, int Hdiv, long long bAo, long long bAi,
                   long long bBo, long long bBi, long long bCo, long long bCi, float alpha, float beta, void* stream);

/* Tensor-core math mode only: linear layers with M <= 32 rows and no dropout (token-by-token decoding) run on a
 * one-warp-per-output-column exact-fp32 kernel instead of a mostly empty 128-row tcgen05 tile.  1 (default) / 0; the
 * environment variable BMHRL_SKINNY_LINEAR=0 sets the initial value. */
int bmhrl_set_skinny_linear(int on);
/* Y = resid + dropout(act(X W^T + bias)); W is (N,K) like nn.Linear.weight.  act: 0 none, 1 ReLU.
 * Replaces linear_Q2d/K2d/V2d/d2Q (model/multihead_attention.py:70-72,89), fc1/fc2 (model/blocks.py:175-187,
 * with the ResidualConnection add of model/blocks.py:144 fused), Manager.linear (model/bm_hrl_agent.py:439-440),
 * WorkerCore.projection (model/bm_hrl_agent.py:464). */
int bmhrl_linear_fwd_f32(const float* X, long long ldx, const float* W, const float* bias, const float* resid,
                         long long ldr, float* Y, long long ldy, int M, int N, int K, int act, float drop_p,
                         const unsigned long long* rng_state, unsigned int salt, void* stream);

/* dZ = dY * relu'(Y) * dropout-mask/(1-p) (dZ may be NULL when act == 0 and drop_p == 0); db[n] (+)= sum_m dZ[m,n]
 * (accumulate_db != 0: added to db, which then is the parameter's slice of the flat gradient buffer). */
int bmhrl_linear_bwd_prep_f32(const float* dY, long long lddy, const float* Y, long long ldy, float* dZ, float* db,
                              int M, int N, int act, float drop_p, const unsigned long long* rng_state,
                              unsigned int salt, int accumulate_db, void* stream);

/* ---- contractions, bf16 operands (gemm_bf16.cu): the library's bf16 math mode ------------------------------------ */
/* One problem of a grouped launch.  Every pointer is a device pointer; bf16 operands are 16-byte aligned with row pitches
 * (lda / ldb, in ELEMENTS) that are multiples of 8; tails of M / N / K need no padding (TMA zero-fills them).
 *   v(m,n) = alpha * sum_k A(m,k) B(n,k) + bias[n];  act (1 = ReLU);  gate: v = gate(m,n) > 0 ? v * gate_scale : 0
 *            (the ReLU / dropout gradient gate: the saved forward output drop(relu(z)) is > 0 exactly where both let
 *            the gradient through);  Philox dropout with element index m*N + n;  + resid(m,n);  + beta * C32(m,n)
 *   C32 (fp32, optional) and / or C16 (bf16, optional) receive v.
 * a_mn = 0: A(m,k) at A + m*lda + k (activations; dZ in dX = dZ W);  a_mn = 1: A(m,k) at A + k*lda + m (dZ^T in
 * dW = dZ^T X).  b_mn = 0: B(n,k) at B + n*ldb + k (nn.Linear.weight (N,K));  b_mn = 1: B(n,k) at B + k*ldb + n (the
 * weight in dX, X in dW).  seg_rows > 0: row m of the fp32 output goes to C32 / C32_seg1 / C32_seg2 for m / seg_rows =
 * 0 / 1 / 2, at row m % seg_rows (dW of N-concatenated Q|K|V weights lands in three separate gradient slices).
 * Replaces: linear_Q2d/K2d/V2d/d2Q (model/multihead_attention.py:70-72,89) -- Q|K|V as ONE N-concatenated problem --,
 * fc1/fc2 with the ResidualConnection add (model/blocks.py:175-187,144), Manager.linear (model/bm_hrl_agent.py:439-440),
 * WorkerCore.projection (model/bm_hrl_agent.py:464) and the aten mm calls of their backward. */
typedef struct bmhrl_gemm_desc {
    const void* A; long long lda; int a_mn;
    const void* B; long long ldb; int b_mn;
    int M, N, K;
    float alpha;
    const float* bias;
    int act;
    const void* gate; long long ldg; float gate_scale;
    float drop_p; const unsigned long long* rng_state; unsigned int salt;
    const float* resid; long long ldr;
    float beta;
    float* C32; long long ldc32;
    void* C16; long long ldc16;
    int seg_rows; float* C32_seg1; float* C32_seg2;
    int b_static;   /* != 0: B is not written by any earlier launch of the stream since the last full dependency (a weight
                       shadow inside one training step): its tiles may be fetched before griddepcontrol.wait */
} bmhrl_gemm_desc;
int bmhrl_gemm_desc_bytes(void);   /* sizeof(bmhrl_gemm_desc): bindings check their mirror of the struct against it */
/* 1..4 independent problems in ONE launch (tcgen05.mma kind::f16 into TMEM, TMA-staged operands, 128 x 128 / 128 x 64
 * tiles; when the tiles alone leave most SMs idle a thread-block cluster of 2 / 4 CTAs splits K of every tile and reduces
 * the partial accumulators through distributed shared memory: no atomics, bit-reproducible). */
int bmhrl_gemm_bf16_group(const bmhrl_gemm_desc* problems, int n_problems, void* stream);
/* tuning / test hook: force the tile width (0 = cost model, 64, 128), the K split (0 = cost model, 1, 2, 4) and the
 * pipeline depth (0 = auto, 2..6). */
int bmhrl_gemm_bf16_tuning(int force_bn, int force_split, int force_stages);
/* diagnosis hook: with a device buffer of 64 slots set, CTA 0 of every following bmhrl_gemm_bf16_group launch stamps
 * clock64() at its hand-over points (slot map: csrc/gemm_bf16.cu, gb_stamp); NULL switches it off (the default). */
int bmhrl_gemm_bf16_trace(unsigned long long* device_slots);

/* ---- K2: residual + LayerNorm, gated mix (norm.cu) ---------------------------------------------------------- */
/* z = x (+ resid); y = LayerNorm(z) (eps, affine).  z_out may be NULL.  ResidualConnection model/blocks.py:135-144,
 * normCA/normCV model/bm_hrl_agent.py:107-108, post-LN model/encoder.py:64-68, model/decoder.py:77-99. */
int bmhrl_layernorm_fwd_f32(const float* x, const float* resid, const float* gamma, const float* beta, float* z_out,
                            float* y, float* mean, float* rstd, int rows, int D, float eps, void* stream);
/* dx = LN'(dy) (+ dres); dgamma, dbeta (both or neither; dx may be NULL; accumulate_affine != 0: added in place). */
int bmhrl_layernorm_bwd_f32(const float* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                            const float* dres, float* dx, float* dgamma, float* dbeta, int rows, int D,
                            int accumulate_affine, void* stream);
/* The dx rows of the same backward, plus dx16 = bf16(dropout(dx)) (Philox, element index row*D + col, like
 * bmhrl_cast_bf16): in bf16 math mode the backward of the block that produced x starts from exactly that matrix, so the
 * separate mask + cast launch at the head of every block's backward disappears.  D % 4 == 0. */
int bmhrl_layernorm_bwd_twin_f32(const float* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                                 const float* dres, float* dx, void* dx16, long long lddx16, float drop_p,
                                 const unsigned long long* rng_state, unsigned int salt, int rows, int D, void* stream);
/* out = s*v + (1-s)*a, s = sigmoid(clamp(*gate_param,-2,2)); model/bm_hrl_agent.py:111-114. */
int bmhrl_gated_mix_fwd_f32(const float* v, const float* a, const float* gate_param, float* out, long long n,
                            void* stream);
int bmhrl_gated_mix_bwd_f32(const float* g, const float* v, const float* a, const float* gate_param, float* dv,
                            float* da, float* dgate, float* partial_ws, int ws_floats, long long n, int accumulate_gate,
                            void* stream);

/* ---- K1: attention core, exact-fp32 path (attn_simt.cu) ----------------------------------------------------- */
/* attention() model/multihead_attention.py:7-31.  Q/K/V/O are (B, S, H*dk) views: row pitch ld?, batch pitch bs?,
 * head h at column h*dk.  mask_kind 0 none, 1 key padding uint8 (B,Sk), 2 full uint8 (B,Sq,Sk); causal adds k<=q
 * (only together with a mask, like the reference).  P: workspace AND saved tensor, bmhrl_mha_core_probs_floats()
 * floats.  Dropout is applied to the per-head output P V. */
long long bmhrl_mha_core_probs_floats(int B, int H, int Sq, int Sk);
int bmhrl_mha_core_fwd_f32(const float* Q, long long ldq, long long bsq, const float* K, long long ldk, long long bsk,
                           const float* V, long long ldv, long long bsv, const unsigned char* mask, int mask_kind,
                           int causal, float* O, long long ldo, long long bso, float* P, int B, int H, int Sq, int Sk,
                           int dk, float drop_p, const unsigned long long* rng_state, unsigned int salt, void* stream);
/* ws_dP: probs_floats() floats; ws_dO: B*Sq*H*dk floats, only needed when drop_p > 0. */
int bmhrl_mha_core_bwd_f32(const float* dO, long long ldo, long long bso, const float* Q, long long ldq, long long bsq,
                           const float* K, long long ldk, long long bsk, const float* V, long long ldv, long long bsv,
                           const float* P, const unsigned char* mask, int mask_kind, int causal, float* dQ,
                           long long lddq, long long bsdq, float* dK, long long lddk, long long bsdk, float* dV,
                           long long lddv, long long bsdv, float* ws_dP, float* ws_dO, int B, int H, int Sq, int Sk,
                           int dk, float drop_p, const unsigned long long* rng_state, unsigned int salt, void* stream);

/* ---- K1, bf16 operands: fused flash-style forward (attn_bf16.cu) --------------------------------------------- */
/* attention() model/multihead_attention.py:7-31 on bf16 head-merged (B, S, H*dk) tensors (BASELINE.json config 5: the
 * long-clip encoder attention sweep).  One persistent tcgen05 kernel: S = Q K^T and O += P V on the tensor cores
 * (bf16 x bf16 -> fp32 in TMEM), one-pass softmax with the reference's REPLACE-by--1e9 masking; the (B,H,Sq,Sk)
 * score tensor is never written.  Pitches ld?/bs? in ELEMENTS (multiples of 8).  mask_kind / causal as above.
 * O bf16 (B,Sq,H*dk); lse (may be NULL): fp32 (B,H,Sq) log-sum-exp of the masked scaled scores.  dk in {64,128,256};
 * no dropout (inference / sweep path). */
int bmhrl_attn_fwd_bf16(const void* Q, long long ldq, long long bsq, const void* K, long long ldk, long long bsk,
                        const void* V, long long ldv, long long bsv, const unsigned char* mask, int mask_kind,
                        int causal, void* O, long long ldo, long long bso, float* lse, int B, int H, int Sq, int Sk,
                        int dk, void* ws, void* stream);
/* ws (may be NULL): bmhrl_attn_fwd_bf16_ws_bytes() bytes.  With a key-padding mask the kernel then reads the mask
 * bit-packed and skips, per batch row, the key tiles behind the last visible key (their probabilities are exactly 0
 * unless the whole row is masked, which keeps all Sk keys like the reference). */
long long bmhrl_attn_fwd_bf16_ws_bytes(int B, int Sk);
/* test / tuning hook: 0 = automatic (currently always one CTA per query tile: it is the faster form), 1 = one CTA per
 * query tile, 2 = CTA pairs (cta_group::2 MMAs over two SMs, K / V tiles split between their shared memories) whenever
 * d_k is 128 or 256. */
int bmhrl_attn_bf16_tuning(int pair_mode);

/* ---- K1, bf16 operands: flash-style backward (attn_bf16_bwd.cu) ---------------------------------------------- */
/* Gradients of attention() model/multihead_attention.py:7-31 (what autograd derives from :22-30 in the reference) with
 * respect to the bf16 head-merged Q (B,Sq,H*dk), K, V (B,Sk,H*dk), given O and lse of bmhrl_attn_fwd_bf16 and dO:
 *   P = exp(S - lse), dV = P^T dO, dP = dO V^T, dS = P o (dP - rowsum(dO o O)), dQ = dS K / sqrt(dk), dK = dS^T Q / sqrt(dk).
 * The (B,H,Sq,Sk) probabilities are recomputed on the tensor cores, never stored.  Masked scores were REPLACED by -1e9
 * in the forward: no gradient reaches them (a batch row with no visible key still sends its uniform weights to dV).
 * mask_kind 0 (none) or 1 ((B,Sk) key padding); 2 (full (B,Sq,Sk) masks) returns BMHRL_ERR_UNSUPPORTED.  Pitches in
 * ELEMENTS (multiples of 8); dQ / dK / dV bf16, fully overwritten; ws: bmhrl_attn_bwd_bf16_ws_bytes() bytes, required.
 * Four launches: rowsum(dO o O), then one persistent tcgen05 kernel per gradient (dQ, dK, dV). */
int bmhrl_attn_bwd_bf16(const void* Q, long long ldq, long long bsq, const void* K, long long ldk, long long bsk,
                        const void* V, long long ldv, long long bsv, const void* O, long long ldo, long long bso,
                        const void* dO, long long lddo, long long bsdo, const float* lse, const unsigned char* mask,
                        int mask_kind, void* dQ, long long lddq, long long bsdq, void* dK, long long lddk,
                        long long bsdk, void* dV, long long lddv, long long bsdv, int B, int H, int Sq, int Sk, int dk,
                        void* ws, void* stream);
long long bmhrl_attn_bwd_bf16_ws_bytes(int B, int H, int Sq, int Sk);
/* profiling hook: bit mask of the gradient passes to launch (1 = dQ, 2 = dK, 4 = dV); default 7 = all.  The gradient of
 * a pass that is switched off is returned as ZEROS (never uninitialised memory). */
int bmhrl_attn_bwd_bf16_tuning(int passes);

/* ---- K4 fused: vocabulary projection + log-softmax + KL-loss statistics, bf16 operands (gemm_bf16.cu) ---------------- */
/* WorkerCore (model/bm_hrl_agent.py:463-466: Linear 364 -> V, LogSoftmax) together with what LabelSmoothing
 * (loss/label_smoothing.py:12-32), BiasedKL (loss/biased_kl.py:22-53) and the token sampling of biased_kl()
 * (epoch_loops/captioning_bmrl_loops.py:273-285) read from its output -- WITHOUT materialising the (M, V) logits or
 * log-probabilities: the logits tile lives in TMEM, the GEMM epilogue reduces it per row, a tiny second kernel merges
 * the V/128 partials.  stats (M,5) fp32 = {lse, sum_v logit, logit[target], logit[pad_idx], logit[token]}; tokens (M)
 * int64 (may be NULL) = arg-max of logit (sample_mode 0), of logit - log(exp_noise) (1: explicit Exp(1) draws (M,V),
 * the arithmetic of Categorical.sample) or of logit - log(q), q Philox Exp(1) (2).  Both KL losses are closed forms of
 * the five statistics (bmhrl_b200/ops16.py: kl_rowsum_from_stats).  part_ws: bmhrl_vocab_stats_ws_floats() floats. */
long long bmhrl_vocab_stats_ws_floats(int M, int V);
int bmhrl_vocab_stats_bf16(const void* x16, long long ldx, const void* w16, long long ldw, const float* bias, int M, int V,
                           int K, const long long* target, int pad_idx, int sample_mode, const float* exp_noise,
                           const unsigned long long* rng_state, unsigned int salt, float* part_ws, float* stats,
                           long long* tokens, void* stream);
/* Backward: recomputes the logits tile by tile and writes d logits (M, V) bf16 (row pitch lddl) from the gradients of the
 * five statistics, dstats (M,5):  dl(n) = dstats[0] exp(logit(n) - lse) + dstats[1] + [n == target] dstats[2] +
 * [n == pad_idx] dstats[3] + [n == token] dstats[4].  dX / dW follow as ordinary bmhrl_gemm_bf16_group problems. */
int bmhrl_vocab_dlogits_bf16(const void* x16, long long ldx, const void* w16, long long ldw, const float* bias, int M, int V,
                             int K, const long long* target, const long long* tokens, int pad_idx, const float* stats,
                             const float* dstats, void* dlogits16, long long lddl, void* stream);

/* ---- K1, bf16 math mode, TRAINING path: one-tile fused attention, forward and backward (attn16.cu) ------------------- */
/* attention() model/multihead_attention.py:7-31 and its gradients for sequences that fit one tile (Sq, Sk <= 128; the
 * caption-side attentions always do, the encoder's do at the benchmark shapes), one CTA per (batch, head): S = Q K^T and
 * O = P V on tcgen05 with the probabilities handed over in TMEM; nothing of size (B,H,Sq,Sk) is ever stored.
 * Q (B*Sq, ldq), K / V (B*Sk, ld?), O (B*Sq, ldo): bf16 matrices, head h = columns [h dk, (h+1) dk) -- typically column
 * slices of the N-concatenated Q|K|V projection output; pitches in ELEMENTS (multiples of 8).  dk in {64,128,256,512}.
 * mask_kind / causal as bmhrl_mha_core_fwd_f32.  lse2: fp32 (B,H,Sq,2), per query row the maximum (base-2 domain) and the reciprocal
 * normaliser of the masked scaled scores (saved for backward; kept apart because a fully masked row has maximum -1e9).  Dropout (Philox, element index row*H*dk + column) is applied to the output like the reference.
 * Returns BMHRL_ERR_UNSUPPORTED when the tiles do not fit (callers then take bmhrl_attn_fwd_bf16 / bmhrl_attn_bwd_bf16). */
int bmhrl_attn16_fwd(const void* Q, long long ldq, const void* K, long long ldk, const void* V, long long ldv,
                     const unsigned char* mask, int mask_kind, int causal, void* O, long long ldo, float* lse2, int B,
                     int H, int Sq, int Sk, int dk, float drop_p, const unsigned long long* rng_state, unsigned int salt,
                     void* stream);
/* dQ (B*Sq, lddq), dK / dV (B*Sk, ldd?) bf16 from Q, K, V, lse2 and dO.  dO must ALREADY carry the forward's dropout mask
 * and 1/(1-p) scale (the dX GEMM of the out-projection applies it in its epilogue: bmhrl_gemm_desc.drop_p with the same
 * salt).  delta = rowsum(P o dP) is formed in-kernel (the whole row is resident), so neither O nor P is needed. */
int bmhrl_attn16_bwd(const void* Q, long long ldq, const void* K, long long ldk, const void* V, long long ldv,
                     const void* dO, long long lddo, const float* lse2, const unsigned char* mask, int mask_kind,
                     int causal, void* dQ, long long lddq, void* dK, long long lddk, void* dV, long long lddv, int B, int H,
                     int Sq, int Sk, int dk, void* stream);
/* dynamic shared memory the one-tile kernels need (forward: backward = 0); above 220 KB the shape is unsupported */
long long bmhrl_attn16_smem_bytes(int backward, int Sq, int Sk, int dk);

/* ---- bf16 math mode: HBM-bound helpers (aux16.cu); all deterministic (no floating-point atomics) ------------------- */
/* y = bf16(LayerNorm(x)) with row pitch ldy (elements): the A operand of the projection that follows the norm
 * (ResidualConnection model/blocks.py:135-144, normCA/normCV model/bm_hrl_agent.py:107-108); mean / rstd saved for
 * bmhrl_layernorm_bwd_f32. */
int bmhrl_layernorm_fwd_bf16(const float* x, const float* gamma, const float* beta, void* y, long long ldy, float* mean,
                             float* rstd, int rows, int D, float eps, void* stream);
/* y = bf16(dropout(x)) (Philox, element index m*cols + n; drop_p = 0: plain cast) and, when db != NULL, the bias gradient
 * db[n] (+)= sum_m dropout(x)[m,n] in a fixed summation order: the head of every linear layer's backward (dZ from the
 * fp32 residual-stream gradient).  y may be NULL (only db). */
int bmhrl_cast_bf16(const float* x, long long ldx, void* y, long long ldy, int rows, int cols, float drop_p,
                    const unsigned long long* rng_state, unsigned int salt, float* db, int accumulate_db, void* stream);
/* Column sums of a bf16 matrix (bias gradient of a dZ a GEMM / attention epilogue emitted): column n is written to
 * out{n / seg_cols}[n % seg_cols] (seg_cols = 0: everything to out0) -- the three bias slices of a fused Q|K|V. */
int bmhrl_colsum_bf16(const void* z, long long ldz, int rows, int cols, float* out0, float* out1, float* out2,
                      int seg_cols, int accumulate, void* stream);

/* ---- K4 / K5: vocabulary log-softmax, KL losses, sampling (loss.cu) ----------------------------------------- */
int bmhrl_log_softmax_fwd_f32(const float* x, float* y, long long rows, int V, void* stream); /* bm_hrl_agent.py:465 */
int bmhrl_log_softmax_bwd_f32(const float* dy, const float* y, float* dx, long long rows, int V, void* stream);
/* mode 0: LabelSmoothing loss/label_smoothing.py:12-32; mode 1: BiasedKL loss/biased_kl.py:22-53.
 * out (rows,V) element-wise divergence and/or rowsum (rows) -- either may be NULL. */
int bmhrl_kl_loss_fwd_f32(int mode, const float* pred, const long long* trg, const long long* biased_trg,
                          const float* biased_offset, float* out, float* rowsum, long long rows, int V, float smoothing,
                          int pad_idx, void* stream);
/* upstream gradient either element-wise (dout) or per row (drow); doffset only for mode 1 (may be NULL). */
int bmhrl_kl_loss_bwd_f32(int mode, const float* pred, const long long* trg, const long long* biased_trg,
                          const float* biased_offset, const float* dout, const float* drow, float* dpred,
                          float* doffset, long long rows, int V, float smoothing, int pad_idx, void* stream);
/* The same two losses as ROW SUMS, in closed form from the five per-row statistics of bmhrl_vocab_stats_bf16 (the target
 * distribution is constant outside the pad, ground-truth and sampled-token columns): rowsum (rows); backward gives dstats
 * (rows,5) and, for mode 1, doffset (rows; may be NULL). */
int bmhrl_kl_stats_fwd_f32(int mode, const float* stats, const long long* trg, const long long* biased_trg,
                           const float* biased_offset, float* rowsum, long long rows, int V, float smoothing, int pad_idx,
                           void* stream);
int bmhrl_kl_stats_bwd_f32(int mode, const float* stats, const long long* trg, const long long* biased_trg,
                           const float* biased_offset, const float* drow, float* dstats, float* doffset, long long rows,
                           int V, float smoothing, int pad_idx, void* stream);
/* biased_kl() epoch_loops/captioning_bmrl_loops.py:273-285.  mode 0 argmax, 1 Categorical sample with explicit
 * Exp(1) noise (rows,V), 2 Categorical sample with in-kernel Philox noise.  probs[row] = exp(pred[row, token]). */
int bmhrl_sample_rows_f32(int mode, const float* pred, const float* exp_noise, long long* tokens, float* probs,
                          long long rows, int V, const unsigned long long* rng_state, unsigned int salt, void* stream);
int bmhrl_scatter_prob_grad_f32(float* dpred, const long long* tokens, const float* probs, const float* g,
                                long long rows, int V, void* stream);
/* Reinforce loss/biased_kl.py:69-75: out[row] = log(clamp(probs[row, action[row]], eps, 1-eps)). */
int bmhrl_reinforce_gather_fwd_f32(const float* probs, const long long* action, float* out, long long rows, int V,
                                   float eps, void* stream);
int bmhrl_reinforce_gather_bwd_f32(const float* probs, const long long* action, const float* g, float* dprobs,
                                   long long rows, int V, float eps, void* stream);

/* ---- frozen SegmentCritic recurrences, forward only (rnn.cu) ------------------------------------------------- */
/* One nn.LSTM (cell 0, gates i,f,g,o) or nn.GRU (cell 1, gates r,z,n) layer, batch_first, zero initial state
 * (model/bm_hrl_agent.py:192-193,206,209).  gx (B,L,NG*H) = x W_ih^T + b_ih for all steps (bmhrl_linear_fwd_f32);
 * out (B,L,H); ws: bmhrl_rnn_ws_floats(B,H) floats, 16-byte aligned (cell state, the grid-barrier counter and the
 * tagged h exchange buffer).  One cooperative launch per layer: W_hh stays resident in shared memory across all L
 * steps, one grid barrier per step (or, after bmhrl_rnn_set_exchange(1) and for even H <= 608: h_t travels between
 * the CTAs as 64-bit {value, step tag} words that the consumers poll, no barrier).  Falls back to one launch per step
 * when the layer does not fit one co-resident grid, or after bmhrl_rnn_set_persistent(0).  All variants give identical
 * bits.  bmhrl_rnn_set_units_per_cta(n): at least n hidden units per CTA (0 = as few as fill the GPU): a layer that
 * runs BESIDE other work (the frozen critic of the next batch under the current step) is a latency chain, and
 * n = 16 confines it to H / 16 SMs at nearly the same duration. */
long long bmhrl_rnn_ws_floats(int B, int H);
int bmhrl_rnn_set_persistent(int on);
int bmhrl_rnn_set_units_per_cta(int units);
int bmhrl_rnn_set_exchange(int mode);   /* 1 / 0: force the exchange / the barrier kernel; -1: BMHRL_RNN_EXCHANGE (default 0) */
/* diagnosis hook: CTA 0 of the exchange kernel stamps clock64() per time step into 8 slots per step (step start, hint
 * seen, h tile loaded, barrier passed, dot products done, gates stored; 60 steps -> 480 slots); NULL = off (default). */
int bmhrl_rnn_trace(unsigned long long* device_slots);
int bmhrl_rnn_layer_f32(int cell, const float* gx, const float* W_hh, const float* b_hh, float* out, float* ws, int B,
                        int L, int H, void* stream);
/* The same layer without any allocation by the caller or the library: the workspace is a file-scope device array of
 * bmhrl_rnn_noalloc_max_floats() = 262 144 floats (1 MB, held by every GPU that loads the library until exit).  No
 * counterpart in the reference: added for callers that cannot call cudaMalloc
 * (CUDA-graph capture, serving loops).  Shapes with bmhrl_rnn_ws_floats(B, H) above the maximum return BMHRL_ERR_ARG.
 * Calls no cudaMalloc / cudaFree.  All calls share the array (it keeps the last call's contents; what must start at zero
 * is cleared per call, stream-ordered): calls that overlap on the GPU race -- serialise them.  Same bits as
 * bmhrl_rnn_layer_f32. */
long long bmhrl_rnn_noalloc_max_floats(void);
int bmhrl_rnn_layer_noalloc_f32(int cell, const float* gx, const float* W_hh, const float* b_hh, float* out, int B, int L,
                                int H, void* stream);
/* ONE time step t of the same layer on caller-kept buffers laid out for L steps: reads h_{t-1} from out[:, t-1, :] and
 * the LSTM cell state from c_prev (B,H) (both taken as zero at t == 0), gx[:, t, :]; writes out[:, t, :] and the new
 * cell state to c_next (B,H; may be c_prev; GRU: both null).  Distinct c_prev / c_next make the step repeatable.
 * Greedy decoding appends one token per step (captioning_bmrl_loops.py:127-152), and the critic is causal: calling this
 * for t = 0, 1, ... gives the bits bmhrl_rnn_layer_f32 gives for the whole prefix, at the cost of one step per token. */
int bmhrl_rnn_step_f32(int cell, const float* gx, const float* W_hh, const float* b_hh, float* out, const float* c_prev,
                       float* c_next, int B, int L, int H, int t, void* stream);
/* The whole recurrent part of the critic -- n_lstm nn.LSTM layers, AReLU (may be NULL: none), n_gru nn.GRU layers
 * (model/bm_hrl_agent.py:192-193,204-211) -- in ONE cooperative launch: cells (layer, step) with layer + step = w are
 * independent and are computed side by side, L + n_layers - 1 grid barriers instead of n_layers * L.  Input AND
 * recurrent halves of the gate pre-activations in exact fp32 FMAs, weights streamed from L2.  w_ih / w_hh / b_ih / b_hh:
 * HOST arrays of n_lstm + n_gru device pointers (nn.LSTM / nn.GRU parameter layout, layer 0 has E input features).
 * ws: bmhrl_critic_wavefront_ws_floats() floats; the hidden states of layer l are ws[l*B*L*H ...] as (B, L, H): the last
 * layer's block is the output.  Returns 1 (nothing launched) when the grid cannot be co-resident. */
long long bmhrl_critic_wavefront_ws_floats(int n_layers, int n_lstm, int B, int L, int H);
int bmhrl_critic_wavefront_f32(const float* emb, int E, const float* const* w_ih, const float* const* w_hh,
                               const float* const* b_ih, const float* const* b_hh, int n_lstm, int n_gru,
                               const float* arelu_alpha, const float* arelu_beta, float* ws, int B, int L, int H,
                               void* stream);
/* AReLU model/bm_hrl_agent.py:13-23 (alpha, beta: device scalars). */
int bmhrl_arelu_f32(const float* x, const float* alpha, const float* beta, float* y, long long n, void* stream);
/* scores[row] = x[row,:] . w + b; labels[row] = sigmoid(score) > threshold (model/bm_hrl_agent.py:213,638-640). */
int bmhrl_segment_labels_f32(const float* x, const float* w, const float* b, float threshold, float* scores,
                             int* labels, long long rows, int D, void* stream);

/* ---- K6 / K7 and step plumbing (misc.cu) --------------------------------------------------------------------- */
/* emb = W[tok]*scale (model/blocks.py:44-48); pos = dropout(emb + pe[l]) (model/blocks.py:105-112). */
int bmhrl_embed_posenc_f32(const long long* tokens, const float* W, const float* pe, float* emb, float* pos,
                           long long rows, int L, int D, float scale, float drop_p,
                           const unsigned long long* rng_state, unsigned int salt, void* stream);
int bmhrl_embed_bwd_f32(const long long* tokens, const float* g, float* dW, long long rows, int D, float scale,
                        void* stream);
/* out = dropout(x (+ x2) + pe[s]); x2 fuses V = rgb + flow (epoch_loops/captioning_bmrl_loops.py:498). */
int bmhrl_add_posenc_f32(const float* x, const float* x2, const float* pe, float* out, long long rows, int S, int D,
                         float drop_p, const unsigned long long* rng_state, unsigned int salt, void* stream);
/* out = cat([a, b], -1) (model/bm_hrl_agent.py:483) and its gradient split (da / db may be NULL). */
int bmhrl_concat2_f32(const float* a, int Da, const float* b, int Db, float* out, long long rows, void* stream);
int bmhrl_split2_f32(const float* g, float* da, int Da, float* db, int Db, long long rows, void* stream);
/* out = x + dropout(r): tail of ResidualConnection model/blocks.py:142-144 for arbitrary sublayers. */
int bmhrl_dropout_add_f32(const float* x, const float* r, float* out, long long n, float drop_p,
                          const unsigned long long* rng_state, unsigned int salt, void* stream);
int bmhrl_dropout_apply_f32(const float* dy, float* dx, long long n, float drop_p, const unsigned long long* rng_state,
                            unsigned int salt, void* stream);
/* Manager exploration noise model/bm_hrl_agent.py:444-452: noise[j] = N(mean/10, std/5) - 0.5*mean/10 with mean/std
 * over all non-NaN elements of x (n floats); eps_normal: optional explicit N(0,1) draws (d_goal), else Philox;
 * stats (optional): {mean, std}. */
int bmhrl_exploration_noise_f32(const float* x, long long n, int d_goal, const float* eps_normal, float* noise,
                                float* stats, const unsigned long long* rng_state, unsigned int salt, void* stream);
/* Manager.expand_goals model/bm_hrl_agent.py:415-429 without the host loop, applied to x + noise (noise (D) may be
 * NULL); map (B,L) int32 is saved for bwd. */
int bmhrl_goal_expand_fwd_f32(const float* x, const float* noise, const int* seg, int* map, float* out, int B, int L,
                              int D, void* stream);
int bmhrl_goal_expand_bwd_f32(const float* dy, const int* map, float* dx, int B, int L, int D, void* stream);
/* manager branch of biased_kl() epoch_loops/captioning_bmrl_loops.py:299-317. */
int bmhrl_segment_scan_fwd_f32(const float* probs, const float* expected, const int* seg, float* seg_prob,
                               float* expected_out, int B, int L, void* stream);
int bmhrl_segment_scan_bwd_f32(const float* probs, const int* seg, const float* g, float* dprobs, int B, int L,
                               void* stream);
/* torch.optim.Adam (scripts/train_rl_captioning_module.py:81-83), one launch for all tensors.  The pointer tables,
 * numels and chunk tables live in device memory; chunk c covers elements [chunk_start[c], +bmhrl_adam_chunk_elems())
 * of tensor chunk_tensor[c].  step_counts: one device float per tensor holding its (already incremented) step
 * number -- torch keeps a step per parameter and skips parameters that have no gradient in a phase. */
int bmhrl_adam_chunk_elems(void);
int bmhrl_adam_multi_f32(float* const* params, const float* const* grads, float* const* exp_avg,
                         float* const* exp_avg_sq, const long long* numels, const int* chunk_tensor,
                         const int* chunk_start, int n_chunks, const float* step_counts, float lr, float beta1,
                         float beta2, float eps, float grad_scale, void* stream);
/* Same update, and additionally every tensor t with shadow[t] != NULL gets its UPDATED values written to shadow[t]:
 * shadow_kind[t] = 0: bf16 matrix (numel / shadow_cols[t] rows of shadow_cols[t] elements, row pitch shadow_ld[t]) -- the
 * tensor-core operand copy of a weight in bf16 math mode; 1: fp32 flat copy (N-concatenated Q|K|V bias).  Tables are
 * device arrays indexed like params. */
int bmhrl_adam_multi_shadow_f32(float* const* params, const float* const* grads, float* const* exp_avg,
                                float* const* exp_avg_sq, const long long* numels, const int* chunk_tensor,
                                const int* chunk_start, int n_chunks, const float* step_counts, float lr, float beta1,
                                float beta2, float eps, float grad_scale, void* const* shadow, const int* shadow_cols,
                                const int* shadow_ld, const int* shadow_kind, void* stream);
/* step_counts[active_tensors[i]] += 1 for i < n_active (active_tensors NULL = identity) and rng_state[1] += 1, on
 * the device; step_counts / rng_state may be NULL. */
/* Grid limit of the Adam kernels (bmhrl_adam_multi*_f32, bmhrl_adam_peer_f32): 0 = one CTA per 8192-element chunk; n > 0 = at
 * most n CTAs in a grid-stride loop.  For an update that runs BESIDE other work (a finished stage's update under the rest
 * of backward): the unlimited grid takes every thread slot of every SM until it has drained. */
int bmhrl_adam_set_grid_limit(int max_ctas);
int bmhrl_adam_get_grid_limit(void);
int bmhrl_tick(float* step_counts, const int* active_tensors, int n_active, unsigned long long* rng_state,
               void* stream);

/* ---- data parallel over NVLink peer memory (peer.cu): gradient sum + Adam + parameter broadcast in ONE kernel ----
 * Replaces nn.DataParallel's gradient reduction + torch.optim.Adam on every replica
 * (scripts/train_rl_captioning_module.py:81-83,97-99) when there is one process per GPU on one NVLink / NVSwitch node.
 * Set-up (host, once): every rank allocates its arenas with bmhrl_peer_alloc (cudaMalloc, zeroed), publishes their IPC
 * handles (bmhrl_peer_get_handle, 64 bytes each, exchanged by the caller, e.g. torch.distributed.all_gather_object) and
 * maps the other ranks' arenas with bmhrl_peer_open_handle.  All ranks lay their arenas out identically, so a peer's
 * copy of a local address is that address + (peer arena base - local arena base): the *_off_bytes arrays below hold
 * that difference per rank (0 for the caller's own rank), at most bmhrl_peer_max_ranks() entries. */
int bmhrl_peer_max_ranks(void);
int bmhrl_peer_slots(void);
long long bmhrl_peer_flag_bytes(void);     /* size of the flag table bmhrl_peer_signal / _wait work on (zero-initialised) */
int bmhrl_peer_alloc(long long bytes, void** out_ptr);
int bmhrl_peer_free(void* ptr);
int bmhrl_peer_handle_bytes(void);
int bmhrl_peer_get_handle(const void* ptr, unsigned char* handle);
int bmhrl_peer_open_handle(const unsigned char* handle, void** out_ptr);
int bmhrl_peer_close_handle(void* ptr);
/* Cross-GPU ordering, stream-ordered and CUDA-graph capturable.  signal: "everything enqueued before this on `stream` is
 * done on my GPU" -> the slot's (only growing) count is stored into every rank's flag table.  wait: blocks `stream` (one
 * warp spinning, never the consumer grid) until every rank's count of the slot has reached the count this rank's own
 * last signal of that slot sent: call signal before wait, once each per step and slot, on every rank. */
int bmhrl_peer_signal(unsigned long long* flags, const long long* flag_off_bytes, int n_ranks, int rank, int slot,
                      void* stream);
int bmhrl_peer_wait(const unsigned long long* flags, int n_ranks, int slot, void* stream);
/* bmhrl_adam_multi_shadow_f32 on the chunks THIS rank owns (the caller lists them; chunk = bmhrl_adam_chunk_elems()
 * elements): gradient = sum over ranks 0 .. n_ranks-1 of that rank's flat gradient (peer loads, fixed order), Adam on the
 * local moments, new parameter (+ its bf16 / fp32 shadow) stored into every rank's parameter (shadow) arena. */
int bmhrl_adam_peer_f32(float* const* params, const float* const* grads, float* const* exp_avg, float* const* exp_avg_sq,
                        const long long* numels, const int* chunk_tensor, const int* chunk_start, int n_chunks,
                        const float* step_counts, float lr, float beta1, float beta2, float eps, float grad_scale,
                        void* const* shadow, const int* shadow_cols, const int* shadow_ld, const int* shadow_kind,
                        const long long* grad_off_bytes, const long long* param_off_bytes, const long long* shadow_off_bytes,
                        int n_ranks, int owner_only_master, void* stream);
/* owner_only_master != 0: the fp32 value of a weight that has a bf16 shadow is stored on the owner only (the replicas
 * compute with the shadow; 4 of 6 bytes per element and peer saved); the replicas' fp32 copies are stale until the owners
 * push theirs with this call (same chunk lists as the owner's bmhrl_adam_peer_f32 calls; checkpoints, mode switches). */
int bmhrl_peer_bcast_params_f32(float* const* params, const long long* numels, const int* chunk_tensor,
                                const int* chunk_start, int n_chunks, const long long* param_off_bytes, int n_ranks,
                                void* stream);

/* ---- input side of the step (pipeline.cu) --------------------------------------------------------------------------- */
/* pad_sequence + rgb + flow + padding mask in one pass over a ragged batch that arrived in ONE host->device copy:
 * a / b (b may be NULL): the clips' frames back to back, (total_frames, D) fp32; offsets (B+1) int32 frame offsets;
 * out (B, S, D) = a + b on the frames of clip b, pad_a + pad_b on its padding rows; mask (B, S) uint8 (may be NULL) =
 * first component of the PADDED a != 0.  Replaces pad_sequence (captioning_datasets/captioning_dataset.py:296-299), the
 * V = rgb + flow add (epoch_loops/captioning_bmrl_loops.py:497-499) and the feature half of make_masks
 * (model/masking.py:18-47).  The reference pads rgb / audio with the vocabulary's pad index (1), so with pad_a = 1 the
 * padding rows stay VISIBLE exactly as they do there; pad_a = 0 gives the length mask. */
int bmhrl_pad_features_f32(const float* a, const float* b, const int* offsets, float* out, unsigned char* mask, int B, int S,
                           int D, float pad_a, float pad_b, void* stream);
/* make_masks (model/masking.py:18-47, modality audio_video) + the token count of the loss normaliser
 * (captioning_bmrl_loops.py:1156) in one launch: V_mask (B,Sv) = rgb[b,s,0] != 0, A_mask (B,Sa) = audio[b,s,0] != 0,
 * C_mask (B,L,L) = (captions[b,j] != pad) && j <= i (NULL: not wanted), *n_tokens = #{targets != pad} (NULL: not wanted).
 * Masks are bytes 0 / 1 (torch.bool storage); strides in elements. */
int bmhrl_make_masks(const float* rgb, long long rgb_batch_stride, long long rgb_frame_stride, int Sv, const float* audio,
                     long long audio_batch_stride, long long audio_frame_stride, int Sa, const long long* captions, int L,
                     long long pad_idx, int B, unsigned char* V_mask, unsigned char* A_mask, unsigned char* C_mask,
                     const long long* targets, long long n_targets, long long* n_tokens, void* stream);
/* generate_synonyms (epoch_loops/captioning_bmrl_loops.py:510-528) on the device: per row, up to and including the first
 * end token (which becomes pad_idx and ends the row's scan): with probability p a token is replaced -- by pad_idx when
 * u2 < 0.8, by a random word of [2, V-1] when u2 >= 0.9, by itself otherwise.  u1 / u2 (fp32) and rand_words (int64),
 * all (B, L): explicit draws per position (reproduces the reference for a given draw table); all NULL: Philox. */
int bmhrl_generate_synonyms(const long long* captions, long long* out, int B, int L, float p, int V, int end_idx, int pad_idx,
                            const float* u1, const float* u2, const long long* rand_words,
                            const unsigned long long* rng_state, unsigned int salt, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BMHRL_B200_H */
