// Input side of the training step: what the reference does per sample in Python on the
// host -- pad_sequence of the cropped feature stacks, rgb + flow, the padding masks
// (captioning_datasets/captioning_dataset.py:296-299, epoch_loops/captioning_bmrl_loops.py:497-499, model/masking.py:18-47)
// and the per-token caption corruption `generate_synonyms` (epoch_loops/captioning_bmrl_loops.py:510-528) -- as two kernels
// over ONE staged host->device copy of the ragged batch.
#include "common.cuh"
#include "../../include/bmhrl_b200.h"

namespace bmhrl {

// packed: the B clips' frames back to back, (total_frames, D) fp32; off[b] .. off[b+1] = the frames of clip b.
// out (B, S, D): frames then `pad_value` rows;  out = a (+ b) on valid frames, pad_a (+ pad_b) on padding rows.
// mask (B, S) uint8 = (a-part of out[b, s, 0] != 0): evaluated on the PADDED first operand like the reference does
// (rgb is padded with the vocabulary's pad index, 1, so its padding rows are NOT masked -- reference behaviour, kept).
__global__ void pad_features_kernel(const float* __restrict__ a, const float* __restrict__ b, const int* __restrict__ off,
                                    float* __restrict__ out, unsigned char* __restrict__ mask, int B, int S, int D4,
                                    float pad_a, float pad_b) {
    const long long n = (long long)B * S * D4;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % D4);
        const long long bs = i / D4;
        const int s = (int)(bs % S), bb = (int)(bs / S);
        const int o0 = off[bb], len = off[bb + 1] - o0;
        float4 v;
        float first;
        if (s < len) {
            v = ld_f4(a + ((long long)(o0 + s) * D4 + c) * 4);
            first = v.x;
            if (b) {
                const float4 w = ld_f4(b + ((long long)(o0 + s) * D4 + c) * 4);
                v.x += w.x; v.y += w.y; v.z += w.z; v.w += w.w;
            }
        } else {
            const float pv = pad_a + (b ? pad_b : 0.f);
            v = make_float4(pv, pv, pv, pv);
            first = pad_a;
        }
        st_f4(out + i * 4, v);
        if (c == 0 && mask) mask[bs] = first != 0.f ? 1 : 0;
    }
}

// One thread per caption row (the scan stops at the first end token, so a row is sequential):
//   token == end_idx           -> pad_idx, and the rest of the row is left untouched
//   u1 < p:  u2 < 0.8          -> pad_idx
//            u2 >= 0.9         -> a random word in [2, V-1]
//            otherwise         -> unchanged
// u1 / u2 / r: explicit draws per position (bit-reproducible against a given draw table), or NULL: Philox.
__global__ void generate_synonyms_kernel(const long long* __restrict__ cap, long long* __restrict__ out, int B, int L,
                                         float p, int V, int end_idx, int pad_idx, const float* __restrict__ u1,
                                         const float* __restrict__ u2, const long long* __restrict__ r,
                                         const unsigned long long* __restrict__ rng, unsigned int salt) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B) return;
    bool stopped = false;
    for (int j = 0; j < L; ++j) {
        const long long w = cap[(long long)i * L + j];
        long long o = w;
        if (!stopped) {
            if (w == end_idx) {
                o = pad_idx;
                stopped = true;
            } else {
                float a, c;
                long long rw;
                if (u1) {
                    a = u1[(long long)i * L + j];
                    c = u2[(long long)i * L + j];
                    rw = r[(long long)i * L + j];
                } else {
                    const Philox4 q = philox_block(rng, salt, (unsigned long long)i * L + j);
                    a = u32_to_unit(q.x);
                    c = u32_to_unit(q.y);
                    rw = 2 + (long long)(q.z % (unsigned)(V - 2));
                }
                if (a < p) {
                    if (c < 0.8f) o = pad_idx;
                    else if (c >= 0.9f) o = rw;
                }
            }
        }
        out[(long long)i * L + j] = o;
    }
}

// The three masks of make_masks (model/masking.py:18-47, modality audio_video) and the token count of the loss
// normaliser in ONE launch (the reference builds them with ~12 small tensor ops in front of every forward pass):
//   V_mask[b, s]    = rgb[b, s, 0]   != 0        (rgb only, before flow is added)
//   A_mask[b, s]    = audio[b, s, 0] != 0
//   C_mask[b, i, j] = (cap[b, j] != pad) && j <= i
//   *n_tokens       = #{ y != pad }                (captioning_bmrl_loops.py:1156: the normaliser of the summed loss)
__global__ void make_masks_kernel(const float* __restrict__ rgb, long long rgb_bs, long long rgb_ss, int Sv,
                                  const float* __restrict__ aud, long long aud_bs, long long aud_ss, int Sa,
                                  const long long* __restrict__ cap, int L, long long pad, int B,
                                  unsigned char* __restrict__ vm, unsigned char* __restrict__ am,
                                  unsigned char* __restrict__ cm, const long long* __restrict__ y, long long n_y,
                                  long long* __restrict__ n_tokens, int mask_blocks) {
    if ((int)blockIdx.x < mask_blocks) {
        const long long nv = (long long)B * Sv, na = (long long)B * Sa, nc = cm ? (long long)B * L * L : 0;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv + na + nc;
             i += (long long)mask_blocks * blockDim.x) {
            if (i < nv) {
                vm[i] = rgb[(i / Sv) * rgb_bs + (i % Sv) * rgb_ss] != 0.f;
            } else if (i < nv + na) {
                const long long k = i - nv;
                am[k] = aud[(k / Sa) * aud_bs + (k % Sa) * aud_ss] != 0.f;
            } else {
                const long long k = i - nv - na;
                const int j = (int)(k % L), q = (int)((k / L) % L), b = (int)(k / ((long long)L * L));
                cm[k] = (j <= q && cap[(long long)b * L + j] != pad) ? 1 : 0;
            }
        }
        return;
    }
    // the last block: token count (integers: any order gives the same sum)
    __shared__ long long red[256];
    long long c = 0;
    for (long long i = threadIdx.x; i < n_y; i += blockDim.x) c += y[i] != pad;
    red[threadIdx.x] = c;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) *n_tokens = red[0];
}

}  // namespace bmhrl
using namespace bmhrl;

extern "C" int bmhrl_make_masks(const float* rgb, long long rgb_batch_stride, long long rgb_frame_stride, int Sv,
                                const float* audio, long long audio_batch_stride, long long audio_frame_stride, int Sa,
                                const long long* captions, int L, long long pad_idx, int B, unsigned char* V_mask,
                                unsigned char* A_mask, unsigned char* C_mask, const long long* targets, long long n_targets,
                                long long* n_tokens, void* stream) {
    BMHRL_CHECK_ARG(rgb && audio && V_mask && A_mask && B > 0 && Sv > 0 && Sa > 0, "bmhrl_make_masks: bad arguments");
    BMHRL_CHECK_ARG(!C_mask || (captions && L > 0), "bmhrl_make_masks: C_mask needs the captions");
    BMHRL_CHECK_ARG(!n_tokens || (targets && n_targets >= 0), "bmhrl_make_masks: n_tokens needs the targets");
    const long long n = (long long)B * (Sv + Sa) + (C_mask ? (long long)B * L * L : 0);
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 4) blocks = 148 * 4;
    make_masks_kernel<<<(int)blocks + (n_tokens ? 1 : 0), 256, 0, (cudaStream_t)stream>>>(
        rgb, rgb_batch_stride, rgb_frame_stride, Sv, audio, audio_batch_stride, audio_frame_stride, Sa, captions, L, pad_idx, B,
        V_mask, A_mask, C_mask, targets, n_targets, n_tokens, (int)blocks);
    BMHRL_POST_LAUNCH("make_masks");
    return BMHRL_OK;
}

extern "C" int bmhrl_pad_features_f32(const float* a, const float* b, const int* offsets, float* out, unsigned char* mask,
                                      int B, int S, int D, float pad_a, float pad_b, void* stream) {
    BMHRL_CHECK_ARG(a && offsets && out && B > 0 && S > 0 && D > 0 && D % 4 == 0, "bmhrl_pad_features_f32: bad arguments");
    BMHRL_CHECK_ARG(aligned16(a) && aligned16(out) && (!b || aligned16(b)), "bmhrl_pad_features_f32: 16-byte alignment");
    const long long n = (long long)B * S * (D / 4);
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    pad_features_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(a, b, offsets, out, mask, B, S, D / 4, pad_a, pad_b);
    BMHRL_POST_LAUNCH("pad_features");
    return BMHRL_OK;
}

extern "C" int bmhrl_generate_synonyms(const long long* captions, long long* out, int B, int L, float p, int V, int end_idx,
                                       int pad_idx, const float* u1, const float* u2, const long long* rand_words,
                                       const unsigned long long* rng_state, unsigned int salt, void* stream) {
    BMHRL_CHECK_ARG(captions && out && B > 0 && L > 0 && V > 2 && p >= 0.f && p <= 1.f, "bmhrl_generate_synonyms: bad arguments");
    BMHRL_CHECK_ARG((u1 && u2 && rand_words) || (!u1 && !u2 && !rand_words && rng_state),
                    "bmhrl_generate_synonyms: either all three draw tables or rng_state");
    generate_synonyms_kernel<<<(B + 63) / 64, 64, 0, (cudaStream_t)stream>>>(captions, out, B, L, p, V, end_idx, pad_idx, u1,
                                                                             u2, rand_words, rng_state, salt);
    BMHRL_POST_LAUNCH("generate_synonyms");
    return BMHRL_OK;
}
