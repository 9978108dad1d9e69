// K4 / K5: log-softmax over the vocabulary, label-smoothed KL and biased KL (element-wise and row-sum forms),
// Categorical sampling / argmax with probability gather, and the REINFORCE log-prob gather.
// Reference: WorkerCore model/bm_hrl_agent.py:463-466; LabelSmoothing loss/label_smoothing.py:12-32;
// BiasedKL loss/biased_kl.py:22-53; Reinforce loss/biased_kl.py:69-81; biased_kl() sampling
// epoch_loops/captioning_bmrl_loops.py:273-285.
// All kernels are HBM-bound streaming passes over the (rows, V) log-prob tensor: one CTA per row, coalesced
// 128-bit loads where the row pitch allows (V = 10172 floats = 40688 B is a 16-byte multiple).
#include "common.cuh"
#include "../../include/bmhrl_b200.h"

namespace bmhrl {

constexpr int ROW_THREADS = 256;

// ------------------------------------------------------------------ log-softmax
__global__ void __launch_bounds__(ROW_THREADS) log_softmax_fwd_kernel(const float* __restrict__ x,
                                                                      float* __restrict__ y, int V) {
    __shared__ float red[32];
    const float* xr = x + (long long)blockIdx.x * V;
    float* yr = y + (long long)blockIdx.x * V;
    float mx = -INFINITY;
    for (int i = threadIdx.x; i < V; i += ROW_THREADS) mx = fmaxf(mx, xr[i]);
    mx = block_max(mx, red);
    float s = 0.f;
    for (int i = threadIdx.x; i < V; i += ROW_THREADS) s += expf(xr[i] - mx);
    s = block_sum(s, red);
    const float lse = mx + logf(s);
    for (int i = threadIdx.x; i < V; i += ROW_THREADS) yr[i] = xr[i] - lse;
}

// dx = dy - exp(y) * sum(dy)
__global__ void __launch_bounds__(ROW_THREADS) log_softmax_bwd_kernel(const float* __restrict__ dy,
                                                                      const float* __restrict__ y,
                                                                      float* __restrict__ dx, int V) {
    __shared__ float red[32];
    const float* dyr = dy + (long long)blockIdx.x * V;
    const float* yr = y + (long long)blockIdx.x * V;
    float* dxr = dx + (long long)blockIdx.x * V;
    float s = 0.f;
    for (int i = threadIdx.x; i < V; i += ROW_THREADS) s += dyr[i];
    s = block_sum(s, red);
    for (int i = threadIdx.x; i < V; i += ROW_THREADS) dxr[i] = dyr[i] - expf(yr[i]) * s;
}

// ------------------------------------------------------------------ target distributions
struct KlRow {
    int mode;       // 0 label smoothing, 1 biased KL
    bool pad_row;   // target == pad_idx
    int trg, samp, pad_idx;
    float u;        // smoothing / (V - 2)
    float t_trg;    // mass at the ground-truth token
    float t_samp;   // mass ADDED at the sampled token (biased KL)
    float add;      // +1e-8 (biased KL) or 0
};

__device__ __forceinline__ float kl_target(const KlRow& r, int v) {
    if (r.pad_row) return r.add;
    float t = (v == r.pad_idx) ? 0.f : (v == r.trg ? r.t_trg : r.u);
    if (r.mode == 1 && v == r.samp) t += r.t_samp;
    return t + r.add;
}

__device__ __forceinline__ float xlogy(float t) { return t == 0.f ? 0.f : t * logf(t); }

// Rows whose target is pad are zeroed by the reference only `if mask.sum() > 0` where mask holds the flat row INDICES
// of the pad targets (loss/label_smoothing.py:26-30, loss/biased_kl.py:46-50): when flat row 0 is the only pad row,
// nothing is zeroed.  Only the CTA of row 0 ever has to look at the other rows.
__device__ __forceinline__ KlRow make_row(int mode, const long long* trg, const long long* samp, const float* offset,
                                          long long row, long long rows, int V, float smoothing, int pad_idx) {
    KlRow r;
    r.mode = mode;
    r.trg = (int)trg[row];
    r.pad_idx = pad_idx;
    r.pad_row = (r.trg == pad_idx);
    if (r.pad_row && row == 0) {
        bool later = false;
        for (long long i = 1; i < rows && !later; ++i) later = ((int)trg[i] == pad_idx);
        r.pad_row = later;
    }
    r.u = smoothing / (float)(V - 2);
    if (mode == 0) {
        r.samp = -1;
        r.t_trg = 1.f - smoothing;
        r.t_samp = 0.f;
        r.add = 0.f;
    } else {
        const float a = offset[row];
        r.samp = (int)samp[row];
        r.t_trg = (1.f - smoothing) * (1.f - a);
        r.t_samp = a * (1.f - smoothing);
        r.add = 1e-8f;
    }
    return r;
}

// Element-wise divergence out[row, v] = xlogy(t, t) - t * pred[row, v]; when rowsum != nullptr the per-row sum
// is produced instead (out may then be nullptr).
__global__ void __launch_bounds__(ROW_THREADS) kl_fwd_kernel(int mode, const float* __restrict__ pred,
                                                             const long long* __restrict__ trg,
                                                             const long long* __restrict__ samp,
                                                             const float* __restrict__ offset,
                                                             float* __restrict__ out, float* __restrict__ rowsum,
                                                             long long rows, int V, float smoothing, int pad_idx) {
    __shared__ float red[32];
    const long long row = blockIdx.x;
    const KlRow r = make_row(mode, trg, samp, offset, row, rows, V, smoothing, pad_idx);
    const float* pr = pred + row * V;
    float acc = 0.f;
    for (int v = threadIdx.x; v < V; v += ROW_THREADS) {
        const float t = kl_target(r, v);
        const float e = xlogy(t) - t * pr[v];
        if (out) out[row * V + v] = e;
        acc += e;
    }
    if (rowsum) {
        acc = block_sum(acc, red);
        if (threadIdx.x == 0) rowsum[row] = acc;
    }
}

// dpred[row, v] = -t * g,  g = dout[row, v] (element-wise form) or drow[row] (row-sum form).
// Biased KL also returns d/d offset (the target distribution is differentiable in the reference):
//   doffset = (1-eps) * ( g_s (ln t_s + 1 - pred_s) - g_t (ln t_t + 1 - pred_t) ),  0 for pad rows.
__global__ void __launch_bounds__(ROW_THREADS) kl_bwd_kernel(int mode, const float* __restrict__ pred,
                                                             const long long* __restrict__ trg,
                                                             const long long* __restrict__ samp,
                                                             const float* __restrict__ offset,
                                                             const float* __restrict__ dout,
                                                             const float* __restrict__ drow,
                                                             float* __restrict__ dpred, float* __restrict__ doffset,
                                                             long long rows, int V, float smoothing, int pad_idx) {
    const long long row = blockIdx.x;
    const KlRow r = make_row(mode, trg, samp, offset, row, rows, V, smoothing, pad_idx);
    const float grow = drow ? drow[row] : 0.f;
    if (dpred) {
        for (int v = threadIdx.x; v < V; v += ROW_THREADS) {
            const float g = dout ? dout[row * V + v] : grow;
            dpred[row * V + v] = -kl_target(r, v) * g;
        }
    }
    if (mode == 1 && doffset && threadIdx.x == 0) {
        float d = 0.f;
        if (!r.pad_row) {
            const float tt = kl_target(r, r.trg), ts = kl_target(r, r.samp);
            const float gt = dout ? dout[row * V + r.trg] : grow;
            const float gs = dout ? dout[row * V + r.samp] : grow;
            const float pt = pred[row * V + r.trg], ps = pred[row * V + r.samp];
            d = (1.f - smoothing) * (gs * (logf(ts) + 1.f - ps) - gt * (logf(tt) + 1.f - pt));
        }
        doffset[row] = d;
    }
}

// ------------------------------------------------------------------ KL row sums from the vocabulary statistics
// The fused vocabulary kernel (gemm_bf16.cu, MODE 1) leaves five numbers per row: lse, S = sum_v logit, and the logits
// of the target, pad and sampled-token columns.  The target distribution of both KL losses is constant (b = u + add)
// outside at most three "special" columns, so with p_c = logit_c - lse and sum_v p_v = S - V lse
//   row = sum_{c special} [xlogy(d_c) - d_c p_c] + (V - n) xlogy(b) - b (sum_v p_v - sum_{c special} p_c).
// One thread per row; the same case analysis as kl_target() above (pad column first, then target, then the ADDED sampled
// mass; zeroed pad rows; the reference's flat-row-0 quirk).
struct KlSpec {
    bool pad_row;
    int n;             // number of special columns (1..3)
    float d[3], pc[3]; // their target mass (incl. add) and log-probability
    int which[3];      // statistic index of the column's logit: 2 target, 3 pad, 4 token
    float b, add;
    int sigma;         // special-column slot that received the sampled mass (-1: none / merged with the target)
    int tslot;         // slot of the ground-truth column (-1 when it coincides with the pad column)
};

__device__ __forceinline__ KlSpec kl_spec(const KlRow& r, const float* st) {
    KlSpec k;
    const float lse = st[0];
    k.pad_row = r.pad_row;
    k.add = r.add;
    k.b = r.u + r.add;
    k.n = 0;
    k.sigma = -1;
    k.tslot = -1;
    // pad column
    k.d[0] = r.add; k.pc[0] = st[3] - lse; k.which[0] = 3; k.n = 1;
    if (r.trg != r.pad_idx) {  // ground-truth column (when the target IS the pad column the scatter was overwritten by 0)
        k.d[1] = r.t_trg + r.add; k.pc[1] = st[2] - lse; k.which[1] = 2; k.tslot = 1; k.n = 2;
    }
    if (r.mode == 1) {
        if (r.samp == r.pad_idx) { k.d[0] += r.t_samp; k.sigma = 0; }
        else if (r.samp == r.trg) { k.d[k.tslot] += r.t_samp; k.sigma = -1; }   // (1-eps)(1-a) + a(1-eps): constant in a
        else { k.d[k.n] = r.u + r.t_samp + r.add; k.pc[k.n] = st[4] - lse; k.which[k.n] = 4; k.sigma = k.n; ++k.n; }
    }
    return k;
}

__global__ void kl_stats_fwd_kernel(int mode, const float* __restrict__ stats, const long long* __restrict__ trg,
                                    const long long* __restrict__ samp, const float* __restrict__ offset,
                                    float* __restrict__ rowsum, long long rows, int V, float smoothing, int pad_idx) {
    const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= rows) return;
    const KlRow r = make_row(mode, trg, samp, offset, row, rows, V, smoothing, pad_idx);
    const float* st = stats + row * 5;
    const float sum_p = st[1] - (float)V * st[0];
    if (r.pad_row) {
        rowsum[row] = (float)V * xlogy(r.add) - r.add * sum_p;
        return;
    }
    const KlSpec k = kl_spec(r, st);
    float acc = 0.f, ps = 0.f;
    for (int i = 0; i < k.n; ++i) { acc += xlogy(k.d[i]) - k.d[i] * k.pc[i]; ps += k.pc[i]; }
    rowsum[row] = acc + (float)(V - k.n) * xlogy(k.b) - k.b * (sum_p - ps);
}

// dstats (rows,5) = drow * d row / d {lse, S, logit_t, logit_pad, logit_tok};  doffset (mode 1) = drow * d row / d a
__global__ void kl_stats_bwd_kernel(int mode, const float* __restrict__ stats, const long long* __restrict__ trg,
                                    const long long* __restrict__ samp, const float* __restrict__ offset,
                                    const float* __restrict__ drow, float* __restrict__ dstats,
                                    float* __restrict__ doffset, long long rows, int V, float smoothing, int pad_idx) {
    const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= rows) return;
    const KlRow r = make_row(mode, trg, samp, offset, row, rows, V, smoothing, pad_idx);
    const float* st = stats + row * 5;
    const float g = drow[row];
    float ds[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
    float da = 0.f;
    if (r.pad_row) {
        ds[0] = r.add * (float)V;
        ds[1] = -r.add;
    } else {
        const KlSpec k = kl_spec(r, st);
        float mass = k.b * (float)(V - k.n);
        ds[1] = -k.b;
        for (int i = 0; i < k.n; ++i) {
            mass += k.d[i];
            ds[k.which[i]] += k.b - k.d[i];
        }
        ds[0] = mass;  // d row / d lse = total target mass
        if (mode == 1 && k.sigma >= 0) {
            const float dsig = logf(k.d[k.sigma]) + 1.f - k.pc[k.sigma];
            const float dt = k.tslot >= 0 ? logf(k.d[k.tslot]) + 1.f - k.pc[k.tslot] : 0.f;
            da = (1.f - smoothing) * (dsig - dt);
        }
    }
#pragma unroll
    for (int i = 0; i < 5; ++i) dstats[row * 5 + i] = g * ds[i];
    if (doffset) doffset[row] = g * da;
}

// ------------------------------------------------------------------ sampling (K5)
// mode 0: argmax over exp(pred)                      (manager, captioning_bmrl_loops.py:283)
// mode 1: argmax((exp(pred)/sum) / q) with explicit q (worker; q ~ Exp(1) supplied by the caller -- the exact
//         arithmetic of torch.multinomial's one-sample path, so tokens are bit-reproducible for a given q)
// mode 2: like 1 with q = -log(u) drawn in-kernel from Philox (throughput path)
// First maximum wins on ties (ascending index), like torch.argmax on CPU.
__global__ void __launch_bounds__(ROW_THREADS) sample_rows_kernel(int mode, const float* __restrict__ pred,
                                                                  const float* __restrict__ qnoise,
                                                                  long long* __restrict__ tok,
                                                                  float* __restrict__ prob, int V,
                                                                  const unsigned long long* __restrict__ rng,
                                                                  unsigned int salt) {
    __shared__ float red[32];
    __shared__ float sval[ROW_THREADS / 32];
    __shared__ int sidx[ROW_THREADS / 32];
    const long long row = blockIdx.x;
    const float* pr = pred + row * V;
    float inv = 1.f;
    if (mode != 0) {
        float s = 0.f;
        for (int v = threadIdx.x; v < V; v += ROW_THREADS) s += expf(pr[v]);
        s = block_sum(s, red);
        inv = s;  // divide (not multiply by a reciprocal) to mirror probs / probs.sum()
    }
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int v = threadIdx.x; v < V; v += ROW_THREADS) {
        float val = expf(pr[v]);
        if (mode == 1) {
            val = (val / inv) / qnoise[row * V + v];
        } else if (mode == 2) {
            float u = u32_to_unit(philox_u32(rng, salt, (unsigned long long)row * V + v));
            val = (val / inv) / (-logf(1.f - u));  // 1-u in (0,1] -> q in [0, inf); q == 0 only if u == 0
        }
        if (val > best) { best = val; bi = v; }
    }
    // warp then block arg-max, ties -> smaller index
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float ov = __shfl_xor_sync(0xffffffffu, best, o);
        int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { sval[w] = best; sidx[w] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < ROW_THREADS / 32; ++i)
            if (sval[i] > best || (sval[i] == best && sidx[i] < bi)) { best = sval[i]; bi = sidx[i]; }
        tok[row] = bi;
        prob[row] = expf(pr[bi]);
    }
}

// dpred[row, tok] += g * prob  (d exp(pred_tok) / d pred_tok); dpred must be zero-initialised or hold a gradient.
__global__ void scatter_prob_grad_kernel(float* __restrict__ dpred, const long long* __restrict__ tok,
                                         const float* __restrict__ prob, const float* __restrict__ g, long long rows,
                                         int V) {
    long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < rows) dpred[r * V + tok[r]] += g[r] * prob[r];
}

// ------------------------------------------------------------------ REINFORCE gather
// out[row] = log(clamp(probs[row, action], eps, 1-eps))
__global__ void reinforce_gather_fwd_kernel(const float* __restrict__ probs, const long long* __restrict__ action,
                                            float* __restrict__ out, long long rows, int V, float eps) {
    long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < rows) out[r] = logf(fminf(fmaxf(probs[r * V + action[r]], eps), 1.f - eps));
}
// dprobs[row, action] = g / p inside the clamp range (boundaries included, like torch.clamp), else 0
__global__ void reinforce_gather_bwd_kernel(const float* __restrict__ probs, const long long* __restrict__ action,
                                            const float* __restrict__ g, float* __restrict__ dprobs, long long rows,
                                            int V, float eps) {
    long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < rows) {
        float p = probs[r * V + action[r]];
        dprobs[r * V + action[r]] = (p >= eps && p <= 1.f - eps) ? g[r] / p : 0.f;
    }
}

}  // namespace bmhrl
using namespace bmhrl;

extern "C" int bmhrl_log_softmax_fwd_f32(const float* x, float* y, long long rows, int V, void* stream) {
    BMHRL_CHECK_ARG(x && y && V > 0 && rows >= 0, "bmhrl_log_softmax_fwd_f32: bad arguments");
    if (rows == 0) return BMHRL_OK;
    log_softmax_fwd_kernel<<<(unsigned)rows, ROW_THREADS, 0, (cudaStream_t)stream>>>(x, y, V);
    BMHRL_POST_LAUNCH("log_softmax_fwd");
    return BMHRL_OK;
}

extern "C" int bmhrl_log_softmax_bwd_f32(const float* dy, const float* y, float* dx, long long rows, int V,
                                         void* stream) {
    BMHRL_CHECK_ARG(dy && y && dx && V > 0 && rows >= 0, "bmhrl_log_softmax_bwd_f32: bad arguments");
    if (rows == 0) return BMHRL_OK;
    log_softmax_bwd_kernel<<<(unsigned)rows, ROW_THREADS, 0, (cudaStream_t)stream>>>(dy, y, dx, V);
    BMHRL_POST_LAUNCH("log_softmax_bwd");
    return BMHRL_OK;
}

extern "C" int bmhrl_kl_loss_fwd_f32(int mode, const float* pred, const long long* trg, const long long* biased_trg,
                                     const float* biased_offset, float* out, float* rowsum, long long rows, int V,
                                     float smoothing, int pad_idx, void* stream) {
    BMHRL_CHECK_ARG(mode == 0 || mode == 1, "bmhrl_kl_loss_fwd_f32: mode must be 0 (label smoothing) or 1 (biased KL)");
    BMHRL_CHECK_ARG(pred && trg && (out || rowsum) && V > 2 && rows >= 0, "bmhrl_kl_loss_fwd_f32: bad arguments");
    BMHRL_CHECK_ARG(mode == 0 || (biased_trg && biased_offset), "bmhrl_kl_loss_fwd_f32: biased KL needs trg/offset");
    if (rows == 0) return BMHRL_OK;
    kl_fwd_kernel<<<(unsigned)rows, ROW_THREADS, 0, (cudaStream_t)stream>>>(mode, pred, trg, biased_trg, biased_offset,
                                                                           out, rowsum, rows, V, smoothing, pad_idx);
    BMHRL_POST_LAUNCH("kl_fwd");
    return BMHRL_OK;
}

extern "C" int bmhrl_kl_loss_bwd_f32(int mode, const float* pred, const long long* trg, const long long* biased_trg,
                                     const float* biased_offset, const float* dout, const float* drow, float* dpred,
                                     float* doffset, long long rows, int V, float smoothing, int pad_idx,
                                     void* stream) {
    BMHRL_CHECK_ARG(mode == 0 || mode == 1, "bmhrl_kl_loss_bwd_f32: mode");
    BMHRL_CHECK_ARG(pred && trg && (dout || drow) && V > 2 && rows >= 0, "bmhrl_kl_loss_bwd_f32: bad arguments");
    BMHRL_CHECK_ARG(mode == 0 || (biased_trg && biased_offset), "bmhrl_kl_loss_bwd_f32: biased KL needs trg/offset");
    if (rows == 0) return BMHRL_OK;
    kl_bwd_kernel<<<(unsigned)rows, ROW_THREADS, 0, (cudaStream_t)stream>>>(
        mode, pred, trg, biased_trg, biased_offset, dout, drow, dpred, doffset, rows, V, smoothing, pad_idx);
    BMHRL_POST_LAUNCH("kl_bwd");
    return BMHRL_OK;
}

extern "C" int bmhrl_sample_rows_f32(int mode, const float* pred, const float* exp_noise, long long* tokens,
                                     float* probs, long long rows, int V, const unsigned long long* rng_state,
                                     unsigned int salt, void* stream) {
    BMHRL_CHECK_ARG(mode >= 0 && mode <= 2, "bmhrl_sample_rows_f32: mode must be 0,1,2");
    BMHRL_CHECK_ARG(pred && tokens && probs && V > 0 && rows >= 0, "bmhrl_sample_rows_f32: bad arguments");
    BMHRL_CHECK_ARG(mode != 1 || exp_noise, "bmhrl_sample_rows_f32: mode 1 needs the Exp(1) noise tensor");
    BMHRL_CHECK_ARG(mode != 2 || rng_state, "bmhrl_sample_rows_f32: mode 2 needs rng_state");
    if (rows == 0) return BMHRL_OK;
    sample_rows_kernel<<<(unsigned)rows, ROW_THREADS, 0, (cudaStream_t)stream>>>(mode, pred, exp_noise, tokens, probs,
                                                                                V, rng_state, salt);
    BMHRL_POST_LAUNCH("sample_rows");
    return BMHRL_OK;
}

extern "C" int bmhrl_scatter_prob_grad_f32(float* dpred, const long long* tokens, const float* probs, const float* g,
                                           long long rows, int V, void* stream) {
    BMHRL_CHECK_ARG(dpred && tokens && probs && g && V > 0 && rows >= 0, "bmhrl_scatter_prob_grad_f32: bad arguments");
    if (rows == 0) return BMHRL_OK;
    scatter_prob_grad_kernel<<<(unsigned)((rows + 127) / 128), 128, 0, (cudaStream_t)stream>>>(dpred, tokens, probs, g,
                                                                                               rows, V);
    BMHRL_POST_LAUNCH("scatter_prob_grad");
    return BMHRL_OK;
}

extern "C" int bmhrl_reinforce_gather_fwd_f32(const float* probs, const long long* action, float* out, long long rows,
                                              int V, float eps, void* stream) {
    BMHRL_CHECK_ARG(probs && action && out && V > 0 && rows >= 0, "bmhrl_reinforce_gather_fwd_f32: bad arguments");
    if (rows == 0) return BMHRL_OK;
    reinforce_gather_fwd_kernel<<<(unsigned)((rows + 127) / 128), 128, 0, (cudaStream_t)stream>>>(probs, action, out,
                                                                                                  rows, V, eps);
    BMHRL_POST_LAUNCH("reinforce_gather_fwd");
    return BMHRL_OK;
}

extern "C" int bmhrl_reinforce_gather_bwd_f32(const float* probs, const long long* action, const float* g,
                                              float* dprobs, long long rows, int V, float eps, void* stream) {
    BMHRL_CHECK_ARG(probs && action && g && dprobs && V > 0 && rows >= 0,
                    "bmhrl_reinforce_gather_bwd_f32: bad arguments");
    if (rows == 0) return BMHRL_OK;
    reinforce_gather_bwd_kernel<<<(unsigned)((rows + 127) / 128), 128, 0, (cudaStream_t)stream>>>(probs, action, g,
                                                                                                  dprobs, rows, V, eps);
    BMHRL_POST_LAUNCH("reinforce_gather_bwd");
    return BMHRL_OK;
}

extern "C" int bmhrl_kl_stats_fwd_f32(int mode, const float* stats, const long long* trg, const long long* biased_trg,
                                      const float* biased_offset, float* rowsum, long long rows, int V, float smoothing,
                                      int pad_idx, void* stream) {
    BMHRL_CHECK_ARG(mode == 0 || mode == 1, "bmhrl_kl_stats_fwd_f32: mode must be 0 (label smoothing) or 1 (biased KL)");
    BMHRL_CHECK_ARG(stats && trg && rowsum && V > 2 && rows >= 0, "bmhrl_kl_stats_fwd_f32: bad arguments");
    BMHRL_CHECK_ARG(mode == 0 || (biased_trg && biased_offset), "bmhrl_kl_stats_fwd_f32: biased KL needs trg/offset");
    if (rows == 0) return BMHRL_OK;
    kl_stats_fwd_kernel<<<(unsigned)((rows + 127) / 128), 128, 0, (cudaStream_t)stream>>>(mode, stats, trg, biased_trg,
                                                                                          biased_offset, rowsum, rows, V,
                                                                                          smoothing, pad_idx);
    BMHRL_POST_LAUNCH("kl_stats_fwd");
    return BMHRL_OK;
}

extern "C" int bmhrl_kl_stats_bwd_f32(int mode, const float* stats, const long long* trg, const long long* biased_trg,
                                      const float* biased_offset, const float* drow, float* dstats, float* doffset,
                                      long long rows, int V, float smoothing, int pad_idx, void* stream) {
    BMHRL_CHECK_ARG(mode == 0 || mode == 1, "bmhrl_kl_stats_bwd_f32: mode");
    BMHRL_CHECK_ARG(stats && trg && drow && dstats && V > 2 && rows >= 0, "bmhrl_kl_stats_bwd_f32: bad arguments");
    BMHRL_CHECK_ARG(mode == 0 || (biased_trg && biased_offset), "bmhrl_kl_stats_bwd_f32: biased KL needs trg/offset");
    if (rows == 0) return BMHRL_OK;
    kl_stats_bwd_kernel<<<(unsigned)((rows + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        mode, stats, trg, biased_trg, biased_offset, drow, dstats, doffset, rows, V, smoothing, pad_idx);
    BMHRL_POST_LAUNCH("kl_stats_bwd");
    return BMHRL_OK;
}
