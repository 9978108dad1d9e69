// Frozen SegmentCritic recurrences (model/bm_hrl_agent.py:186-215): nn.LSTM(300->600, 4 layers) -> AReLU ->
// nn.GRU(600->600, 2 layers) -> AReLU -> Linear(600->1), then sigmoid > threshold -> int32 segment labels
// (model/bm_hrl_agent.py:638-640).  Forward only (the critic runs under no_grad) and exact fp32: the labels are
// integers that must match the reference bit for bit, so nothing here uses reduced precision or fast-math.
//
// The input contribution gx = x W_ih^T + b_ih of ALL time steps is one GEMM (bmhrl_linear_fwd_f32); the recurrent
// part is one fused kernel per time step: h_{t-1} W_hh^T, the gate non-linearities and the state update.  One
// warp owns one hidden unit (its 4 / 3 gate rows of W_hh, read once, coalesced) for up to 16 batch rows whose
// h_{t-1} sits in shared memory.
#include <stdlib.h>
#include <atomic>
#include "common.cuh"
#include "../../include/bmhrl_b200.h"

namespace bmhrl {

constexpr int RNN_BT = 16;     // batch rows per block
constexpr int RNN_WARPS = 8;   // hidden units per block

__device__ __forceinline__ float sigmoid_exact(float x) { return 1.f / (1.f + expf(-x)); }

__device__ __forceinline__ void load_h_tile(float* hs, const float* out, int b0, int nb, int L, int H, int t);

// NG = 4: LSTM (gate order i,f,g,o; b_hh added to the pre-activations); NG = 3: GRU (r,z,n; b_hh inside r*(...)).
template <int NG>
__global__ void __launch_bounds__(RNN_WARPS * 32) rnn_step_kernel(const float* __restrict__ gx,    // (B,L,NG*H)
                                                                 const float* __restrict__ W_hh,  // (NG*H,H)
                                                                 const float* __restrict__ b_hh,  // (NG*H)
                                                                 float* __restrict__ out,         // (B,L,H)
                                                                 const float* c_prev,             // (B,H) LSTM only
                                                                 float* c_next,                   // may be c_prev
                                                                 int B, int L, int H, int t) {
    extern __shared__ float hs[];  // RNN_BT x H
    const int b0 = blockIdx.y * RNN_BT;
    const int nb = min(RNN_BT, B - b0);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int j = blockIdx.x * RNN_WARPS + warp;
    if (t > 0) load_h_tile(hs, out, b0, nb, L, H, t - 1);
    __syncthreads();
    if (j >= H) return;
    float acc[NG][RNN_BT];
#pragma unroll
    for (int g = 0; g < NG; ++g)
#pragma unroll
        for (int bb = 0; bb < RNN_BT; ++bb) acc[g][bb] = 0.f;
    if (t > 0) {
        for (int k = lane; k < H; k += 32) {
            float w[NG];
#pragma unroll
            for (int g = 0; g < NG; ++g) w[g] = W_hh[((long long)g * H + j) * H + k];
#pragma unroll
            for (int bb = 0; bb < RNN_BT; ++bb) {
                float hv = hs[bb * H + k];
#pragma unroll
                for (int g = 0; g < NG; ++g) acc[g][bb] = fmaf(w[g], hv, acc[g][bb]);
            }
        }
    }
    float mine[NG];
#pragma unroll
    for (int g = 0; g < NG; ++g) mine[g] = 0.f;
#pragma unroll
    for (int g = 0; g < NG; ++g)
#pragma unroll
        for (int bb = 0; bb < RNN_BT; ++bb) {
            float s = warp_sum(acc[g][bb]);
            if (lane == bb) mine[g] = s;
        }
    if (lane >= nb) return;
    const int b = b0 + lane;
    const float* gxr = gx + ((long long)b * L + t) * (NG * H);
    float hprev = t > 0 ? hs[lane * H + j] : 0.f;
    float hnew;
    if (NG == 4) {
        float pi = gxr[j] + b_hh[j] + mine[0];
        float pf = gxr[H + j] + b_hh[H + j] + mine[1];
        float pg = gxr[2 * H + j] + b_hh[2 * H + j] + mine[2];
        float po = gxr[3 * H + j] + b_hh[3 * H + j] + mine[NG - 1];
        float cprev = t > 0 ? c_prev[(long long)b * H + j] : 0.f;
        float cn = sigmoid_exact(pf) * cprev + sigmoid_exact(pi) * tanhf(pg);
        c_next[(long long)b * H + j] = cn;
        hnew = sigmoid_exact(po) * tanhf(cn);
    } else {
        float r = sigmoid_exact(gxr[j] + mine[0] + b_hh[j]);
        float z = sigmoid_exact(gxr[H + j] + mine[1] + b_hh[H + j]);
        float n = tanhf(gxr[2 * H + j] + r * (mine[2] + b_hh[2 * H + j]));
        hnew = (1.f - z) * n + z * hprev;
    }
    out[((long long)b * L + t) * H + j] = hnew;
}

// ---------------------------------------------------------------------------------------------------------------
// Persistent variant: ONE cooperative launch per layer.  CTA c owns `upb` hidden units (one warp each); their NG
// rows of W_hh are loaded into shared memory once and stay there for all L time steps (5.76 MB of LSTM weights
// spread over ~120 SMs = 48 KB each), so a time step costs one grid-wide barrier + the broadcast of h_{t-1} (B x H
// floats through L2) instead of a kernel launch that re-streams W_hh.  Batch rows are processed in groups of 16 that
// share the resident weights.  Arithmetic order per (unit, batch row) is identical to rnn_step_kernel, so both
// variants give bit-identical states.
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// All CTAs of a cooperative launch are co-resident; `counter` only grows (zeroed by the host before the launch).
__device__ __forceinline__ void grid_barrier(unsigned* counter, unsigned target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1u);
        unsigned spins = 0;
        while (ld_acquire_u32(counter) < target) {
            if (++spins > (1u << 28)) __trap();  // a scheduling bug must fault, never hang the GPU
        }
        __threadfence();
    }
    __syncthreads();
}

// hs[bb][k] = out[b0 + bb][t][k] for bb < nb (0 beyond): L2 loads (.cg: the rows were written by other SMs in this
// same launch), 4 independent 128-bit requests in flight per thread.
__device__ __forceinline__ void load_h_tile(float* hs, const float* out, int b0, int nb, int L, int H, int t) {
    if ((H & 3) == 0 && aligned16(out)) {
        const int H4 = H >> 2, total = RNN_BT * H4;
        for (int i0 = threadIdx.x; i0 < total; i0 += 4 * blockDim.x) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + u * blockDim.x;
                v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (i < total) {
                    const int bb = i / H4, k4 = i - bb * H4;
                    if (bb < nb)
                        v[u] = __ldcg(reinterpret_cast<const float4*>(out + ((long long)(b0 + bb) * L + t) * H) + k4);
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + u * blockDim.x;
                if (i < total) reinterpret_cast<float4*>(hs)[i] = v[u];
            }
        }
    } else {
        for (int i = threadIdx.x; i < RNN_BT * H; i += blockDim.x) {
            int bb = i / H, k = i - bb * H;
            hs[i] = bb < nb ? __ldcg(out + ((long long)(b0 + bb) * L + t) * H + k) : 0.f;
        }
    }
}

template <int NG>
__global__ void __launch_bounds__(512) rnn_layer_persistent_kernel(const float* __restrict__ gx,    // (B,L,NG*H)
                                                                   const float* __restrict__ W_hh,  // (NG*H,H)
                                                                   const float* __restrict__ b_hh,  // (NG*H)
                                                                   float* out,                      // (B,L,H)
                                                                   float* c,                        // (B,H) LSTM
                                                                   unsigned* bar, int B, int L, int H, int upb) {
    extern __shared__ float smem[];
    float* Ws = smem;                                 // [upb][NG][H]
    float* hs = smem + (size_t)upb * NG * H;          // [RNN_BT][H]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int j = blockIdx.x * upb + warp;
    const bool unit_ok = warp < upb && j < H;
    if (unit_ok) {
        if ((H & 3) == 0 && aligned16(W_hh)) {  // 128-bit loads, NG*4 independent requests in flight per lane
            const int H4 = H >> 2;
            for (int k4 = lane; k4 < H4; k4 += 128) {
                float4 v[NG][4];
#pragma unroll
                for (int g = 0; g < NG; ++g)
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        if (k4 + 32 * u < H4)
                            v[g][u] = __ldg(reinterpret_cast<const float4*>(W_hh + ((long long)g * H + j) * H) + k4 + 32 * u);
#pragma unroll
                for (int g = 0; g < NG; ++g)
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        if (k4 + 32 * u < H4)
                            reinterpret_cast<float4*>(Ws + ((size_t)warp * NG + g) * H)[k4 + 32 * u] = v[g][u];
            }
        } else {
#pragma unroll
            for (int g = 0; g < NG; ++g)
                for (int k = lane; k < H; k += 32)
                    Ws[((size_t)warp * NG + g) * H + k] = W_hh[((long long)g * H + j) * H + k];
        }
    }
    float bh[NG];
#pragma unroll
    for (int g = 0; g < NG; ++g) bh[g] = unit_ok ? b_hh[g * H + j] : 0.f;
    const float* Wr = Ws + (size_t)warp * NG * H;
    __syncthreads();
    for (int t = 0; t < L; ++t) {
        for (int b0 = 0; b0 < B; b0 += RNN_BT) {
            const int nb = min(RNN_BT, B - b0);
            // input contributions of this step do not depend on h: request them before waiting on the h broadcast
            float gxv[NG];
#pragma unroll
            for (int g = 0; g < NG; ++g)
                gxv[g] = (unit_ok && lane < nb) ? __ldg(gx + ((long long)(b0 + lane) * L + t) * (NG * H) + g * H + j) : 0.f;
            if (t > 0) load_h_tile(hs, out, b0, nb, L, H, t - 1);
            __syncthreads();
            if (unit_ok) {
                float acc[NG][RNN_BT];
#pragma unroll
                for (int g = 0; g < NG; ++g)
#pragma unroll
                    for (int bb = 0; bb < RNN_BT; ++bb) acc[g][bb] = 0.f;
                if (t > 0) {
                    for (int k = lane; k < H; k += 32) {
                        float w[NG];
#pragma unroll
                        for (int g = 0; g < NG; ++g) w[g] = Wr[g * H + k];
#pragma unroll
                        for (int bb = 0; bb < RNN_BT; ++bb) {
                            float hv = hs[bb * H + k];
#pragma unroll
                            for (int g = 0; g < NG; ++g) acc[g][bb] = fmaf(w[g], hv, acc[g][bb]);
                        }
                    }
                }
                float mine[NG];
#pragma unroll
                for (int g = 0; g < NG; ++g) mine[g] = 0.f;
#pragma unroll
                for (int g = 0; g < NG; ++g)
#pragma unroll
                    for (int bb = 0; bb < RNN_BT; ++bb) {
                        float s = warp_sum(acc[g][bb]);
                        if (lane == bb) mine[g] = s;
                    }
                if (lane < nb) {
                    const int b = b0 + lane;
                    float hprev = t > 0 ? hs[lane * H + j] : 0.f;
                    float hnew;
                    if (NG == 4) {
                        float pi = gxv[0] + bh[0] + mine[0];
                        float pf = gxv[1] + bh[1] + mine[1];
                        float pg = gxv[2] + bh[2] + mine[2];
                        float po = gxv[NG - 1] + bh[NG - 1] + mine[NG - 1];
                        float cprev = t > 0 ? c[(long long)b * H + j] : 0.f;
                        float cn = sigmoid_exact(pf) * cprev + sigmoid_exact(pi) * tanhf(pg);
                        c[(long long)b * H + j] = cn;
                        hnew = sigmoid_exact(po) * tanhf(cn);
                    } else {
                        float r = sigmoid_exact(gxv[0] + mine[0] + bh[0]);
                        float z = sigmoid_exact(gxv[1] + mine[1] + bh[1]);
                        float n = tanhf(gxv[2] + r * (mine[2] + bh[2]));
                        hnew = (1.f - z) * n + z * hprev;
                    }
                    out[((long long)b * L + t) * H + j] = hnew;
                }
            }
            __syncthreads();  // hs is overwritten by the next batch group / time step
        }
        if (t + 1 < L) grid_barrier(bar, (unsigned)(t + 1) * gridDim.x);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Exchange variant of the persistent layer: no grid barrier.  A time step of the barrier kernel above is a chain of
// L2 round trips (atomic arrive, poll, then the h broadcast) -- ~9 us at B = 16, H = 600.  Here every h value travels
// as ONE 64-bit word {fp32 bits, tag = t + 1} (single-copy atomic, so the tag vouches for the value next to it): the
// producer stores it, the consumers poll the words themselves; data and "ready" arrive in the same round trip and a slow
// CTA only delays the words it owes.  X is double-buffered by step parity (a CTA can only produce step t + 1 after it
// has consumed ALL of step t - 1 ... so nobody still reads the buffer that step t + 1 overwrites) and zeroed by the
// host side before the launch (tags start at 1).  The grid is still launched cooperatively: consumers spin on words
// owed by other CTAs, so all of them must be resident.
// A warp owns one hidden unit x RNN2_RB batch rows (2 warps per unit instead of 1: half the serial FMA chain); the
// per-lane partial sums run over k = lane, lane + 32, ... exactly like rnn_step_kernel, and the 32 (gate, row) totals of
// a warp are reduced by a TRANSPOSED butterfly -- same pairs, same order (xor 16, 8, 4, 2, 1) as warp_sum, so the
// results are bit-identical to the other two variants, with 31 shuffles instead of 160.
constexpr int RNN2_RB = 8;
constexpr int RNN2_NGRP = RNN_BT / RNN2_RB;

__device__ __forceinline__ void st_tagged(unsigned long long* p, float v, unsigned tag) {
    const unsigned long long w = ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v);
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
__device__ __forceinline__ void ld_tagged2(const unsigned long long* p, unsigned long long& a, unsigned long long& b) {
    // .cg: served by L2 (the point of coherence), never by this SM's L1 -- and, measured, several times the throughput of
    // ld.relaxed.gpu for a 77 KB tile (each 64-bit word is still read whole: value and tag cannot be torn apart)
    asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}

// hs[bb][k] = value part of X[b0 + bb][k] once its tag equals `tag`, rows bb < nb (H even: two words per 128-bit load;
// one load per row in flight per thread, all rows at once)
__device__ __forceinline__ void load_h_tagged(float* hs, const unsigned long long* X, int b0, int nb, int H, unsigned tag) {
    const int H2 = H >> 1;
    for (int k2 = threadIdx.x; k2 < H2; k2 += blockDim.x) {
        unsigned long long qa[RNN_BT], qb[RNN_BT];
        unsigned pending = nb >= RNN_BT ? 0xffffu : ((1u << nb) - 1u);
        unsigned spins = 0;
        while (pending) {
#pragma unroll
            for (int bb = 0; bb < RNN_BT; ++bb)
                if (pending >> bb & 1u) ld_tagged2(X + (long long)(b0 + bb) * H + 2 * k2, qa[bb], qb[bb]);
#pragma unroll
            for (int bb = 0; bb < RNN_BT; ++bb)
                if ((pending >> bb & 1u) && (unsigned)(qa[bb] >> 32) == tag && (unsigned)(qb[bb] >> 32) == tag) {
                    *reinterpret_cast<float2*>(hs + bb * H + 2 * k2) =
                        make_float2(__uint_as_float((unsigned)qa[bb]), __uint_as_float((unsigned)qb[bb]));
                    pending &= ~(1u << bb);
                }
            if (pending && ++spins > (1u << 24)) __trap();  // a scheduling bug must fault, never hang the GPU
        }
    }
}

constexpr int RNN2_KI = 19;        // k-slots per lane: the W_hh row of a (unit, gate) lives in registers, H <= 32 * 19 = 608
constexpr int RNN2_MAXT = 320;     // 5 units x 2 row groups x 32 lanes

template <int NG>
__global__ void __launch_bounds__(RNN2_MAXT) rnn_layer_exchange_kernel(const float* __restrict__ gx,    // (B,L,NG*H)
                                                                 const float* __restrict__ W_hh,  // (NG*H,H)
                                                                 const float* __restrict__ b_hh,  // (NG*H)
                                                                 float* out,                      // (B,L,H)
                                                                 float* c,                        // (B,H) LSTM
                                                                 unsigned long long* X,           // (2,B,H) tagged h
                                                                 int B, int L, int H, int upb,
                                                                 unsigned long long* trace) {
    extern __shared__ float smem[];
    float* hs = smem;                                 // [2][RNN_BT][H]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int u = warp / RNN2_NGRP, grp = warp - u * RNN2_NGRP;
    const int j = blockIdx.x * upb + u;
    const bool unit_ok = u < upb && j < H;
    // the NG rows of W_hh of this warp's unit stay in REGISTERS for all L steps, lane l holding k = l, l + 32, ...: the
    // dot products then read only h from shared memory (with W there too the loop was bound by LDS bandwidth: 12 loads
    // per 32 FMAs)
    float wreg[NG][RNN2_KI];
#pragma unroll
    for (int g = 0; g < NG; ++g)
#pragma unroll
        for (int i = 0; i < RNN2_KI; ++i) {
            const int k = lane + 32 * i;
            wreg[g][i] = (unit_ok && k < H) ? __ldg(W_hh + ((long long)g * H + j) * H + k) : 0.f;
        }
    for (int i = threadIdx.x; i < 2 * RNN_BT * H; i += blockDim.x) hs[i] = 0.f;
    float bh[NG];
#pragma unroll
    for (int g = 0; g < NG; ++g) bh[g] = unit_ok ? b_hh[g * H + j] : 0.f;
    const int r_mine = grp * RNN2_RB + (lane & (RNN2_RB - 1));   // the row whose gates this lane finishes (lanes < RNN2_RB)
    __syncthreads();
    for (int t = 0; t < L; ++t) {
        float* hcur = hs + (size_t)(t & 1) * RNN_BT * H;   // h_{t-1} of the current row group
        const unsigned long long* Xprev = X + (size_t)((t + 1) & 1) * B * H;
        unsigned long long* Xnext = X + (size_t)(t & 1) * B * H;
        for (int b0 = 0; b0 < B; b0 += RNN_BT) {
            const int nb = min(RNN_BT, B - b0);
            const bool mine_ok = unit_ok && lane < RNN2_RB && r_mine < nb;
            const int b = b0 + r_mine;
            // this step's input contributions and cell state do not depend on h: request them before polling for h
            float gxv[NG], cprev = 0.f;
#pragma unroll
            for (int g = 0; g < NG; ++g)
                gxv[g] = mine_ok ? __ldg(gx + ((long long)b * L + t) * (NG * H) + g * H + j) : 0.f;
            if (NG == 4 && mine_ok && t > 0) cprev = c[(long long)b * H + j];
            const bool stamp = trace && blockIdx.x == 0 && threadIdx.x == 0 && t < 60 && b0 == 0;
            if (stamp) trace[t * 8 + 0] = clock64();
            if (t > 0) {
                // a hint, not a guard: thread i watches ONE word of producer CTA i (its last unit, the last row of the
                // group) so that the CTA polls ~1 KB per round trip instead of hammering L2 with the whole h tile; the
                // tags checked by load_h_tagged are what makes the values safe to use
                for (int pc = threadIdx.x; pc < (int)gridDim.x; pc += blockDim.x) {
                    const unsigned long long* w = Xprev + (long long)(b0 + nb - 1) * H + min(pc * upb + upb - 1, H - 1);
                    unsigned long long q;
                    unsigned spins = 0;
                    do {
                        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(q) : "l"(w) : "memory");
                        if (++spins > (1u << 24)) __trap();
                    } while ((unsigned)(q >> 32) != (unsigned)t);
                }
                __syncthreads();
                if (stamp) trace[t * 8 + 1] = clock64();
                load_h_tagged(hcur, Xprev, b0, nb, H, (unsigned)t);
            }
            if (stamp) trace[t * 8 + 2] = clock64();
            __syncthreads();
            if (stamp) trace[t * 8 + 3] = clock64();
            float v[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = 0.f;
            if (t > 0 && unit_ok) {
                const float* hg = hcur + (size_t)grp * RNN2_RB * H;
#pragma unroll
                for (int i = 0; i < RNN2_KI; ++i) {
                    const int k = lane + 32 * i;
                    if (k < H) {
#pragma unroll
                        for (int r = 0; r < RNN2_RB; ++r) {
                            const float hv = hg[r * H + k];
#pragma unroll
                            for (int g = 0; g < NG; ++g) v[g * RNN2_RB + r] = fmaf(wreg[g][i], hv, v[g * RNN2_RB + r]);
                        }
                    }
                }
            }
            if (stamp) trace[t * 8 + 4] = clock64();
            // transposed butterfly: afterwards lane l holds the warp-wide total of value l = gate * RNN2_RB + row
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const bool up = (lane & o) != 0;
#pragma unroll
                for (int i = 0; i < o; ++i) {
                    const float keep = up ? v[i + o] : v[i];
                    const float send = up ? v[i] : v[i + o];
                    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
                }
            }
            float mine[NG];
#pragma unroll
            for (int g = 0; g < NG; ++g) mine[g] = __shfl_sync(0xffffffffu, v[0], g * RNN2_RB + (lane & (RNN2_RB - 1)));
            if (mine_ok) {
                const float hprev = t > 0 ? hcur[r_mine * H + j] : 0.f;
                float hnew;
                if (NG == 4) {
                    float pi = gxv[0] + bh[0] + mine[0];
                    float pf = gxv[1] + bh[1] + mine[1];
                    float pg = gxv[2] + bh[2] + mine[2];
                    float po = gxv[NG - 1] + bh[NG - 1] + mine[NG - 1];
                    float cn = sigmoid_exact(pf) * cprev + sigmoid_exact(pi) * tanhf(pg);
                    c[(long long)b * H + j] = cn;
                    hnew = sigmoid_exact(po) * tanhf(cn);
                } else {
                    float r = sigmoid_exact(gxv[0] + mine[0] + bh[0]);
                    float z = sigmoid_exact(gxv[1] + mine[1] + bh[1]);
                    float n = tanhf(gxv[2] + r * (mine[2] + bh[2]));
                    hnew = (1.f - z) * n + z * hprev;
                }
                if (t + 1 < L) st_tagged(Xnext + (long long)b * H + j, hnew, (unsigned)(t + 1));
                out[((long long)b * L + t) * H + j] = hnew;
            }
            if (stamp) trace[t * 8 + 5] = clock64();
            if (B > RNN_BT) __syncthreads();   // several row groups share the h tile of this step
        }
    }
}

// AReLU (model/bm_hrl_agent.py:13-23): relu(x) * (1 + sigmoid(beta)) - relu(-x) * clamp(alpha, 0.01, 0.99)
__global__ void arelu_kernel(const float* __restrict__ x, const float* __restrict__ alpha_p,
                             const float* __restrict__ beta_p, float* __restrict__ y, long long n) {
    const float alpha = fminf(fmaxf(alpha_p[0], 0.01f), 0.99f);
    const float beta = 1.f + sigmoid_exact(beta_p[0]);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float v = x[i];
        y[i] = fmaxf(v, 0.f) * beta - fmaxf(-v, 0.f) * alpha;
    }
}

// score = x . w + b (one warp per row); label = sigmoid(score) > thr.
__global__ void __launch_bounds__(128) segment_label_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                            const float* __restrict__ b, float thr,
                                                            float* __restrict__ scores, int* __restrict__ labels,
                                                            long long rows, int D) {
    const long long row = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float* xr = x + row * D;
    float s = 0.f;
    for (int k = lane; k < D; k += 32) s = fmaf(xr[k], w[k], s);
    s = warp_sum(s);
    if (lane == 0) {
        s += b[0];
        if (scores) scores[row] = s;
        labels[row] = sigmoid_exact(s) > thr ? 1 : 0;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Wavefront variant for the whole critic: ALL recurrent layers in ONE cooperative launch.  Cell (layer l, step t)
// depends on (l, t-1) and (l-1, t), so the cells with l + t = w are independent: wavefront w computes them side by
// side and the launch needs L + n_layers - 1 grid barriers instead of n_layers * L (25 instead of 120 for the critic at
// L = 20) -- the per-layer kernels above are latency chains (barrier + broadcast of h per step), not arithmetic.
// A CTA is bound to one layer and a block of its hidden units; a warp computes, for one unit and 16 batch rows, the NG
// gate pre-activations W_ih x + W_hh h (both halves here: the input half of layer l at step t exists only one
// wavefront earlier) in exact fp32 FMAs, with the weights streamed from L2 (all six layers' 60 MB stay L2-resident).
struct CriticWave {
    const float* emb; int E;                       // (B, L, E) input of layer 0
    const float* w_ih[8]; const float* w_hh[8]; const float* b_ih[8]; const float* b_hh[8];
    int n_lstm, n_layers;
    const float* alpha; const float* beta;         // AReLU between the LSTM and the GRU stack (device scalars)
    float* h;                                      // (n_layers, B, L, H) hidden states of every layer
    float* c;                                      // (n_lstm, B, H) LSTM cell states
    unsigned* bar;
    int B, L, H, cpl, upc;                         // CTAs per layer, hidden units per CTA
};

template <int NG>
__device__ __forceinline__ void wave_unit(const CriticWave& p, int l, int t, int j, int b0, int nb, int Kin,
                                          const float* xs, const float* hs, int lane) {
    const int H = p.H;
    const float* Wih = p.w_ih[l];
    const float* Whh = p.w_hh[l];
    float acc[NG][RNN_BT];
    float gi[NG], gh[NG];
#pragma unroll
    for (int g = 0; g < NG; ++g) { gi[g] = 0.f; gh[g] = 0.f; }
    // ---- input half: W_ih[g*H + j, :] . x[bb, :]
#pragma unroll
    for (int g = 0; g < NG; ++g)
#pragma unroll
        for (int bb = 0; bb < RNN_BT; ++bb) acc[g][bb] = 0.f;
    for (int k4 = lane; k4 < (Kin >> 2); k4 += 32) {
        float4 w[NG];
#pragma unroll
        for (int g = 0; g < NG; ++g) w[g] = __ldg(reinterpret_cast<const float4*>(Wih + ((long long)g * H + j) * Kin) + k4);
#pragma unroll
        for (int bb = 0; bb < RNN_BT; ++bb) {
            const float4 xv = reinterpret_cast<const float4*>(xs + bb * Kin)[k4];
#pragma unroll
            for (int g = 0; g < NG; ++g) {
                acc[g][bb] = fmaf(w[g].x, xv.x, acc[g][bb]);
                acc[g][bb] = fmaf(w[g].y, xv.y, acc[g][bb]);
                acc[g][bb] = fmaf(w[g].z, xv.z, acc[g][bb]);
                acc[g][bb] = fmaf(w[g].w, xv.w, acc[g][bb]);
            }
        }
    }
#pragma unroll
    for (int g = 0; g < NG; ++g)
#pragma unroll
        for (int bb = 0; bb < RNN_BT; ++bb) {
            const float s = warp_sum(acc[g][bb]);
            if (lane == bb) gi[g] = s;
        }
    // ---- recurrent half: W_hh[g*H + j, :] . h_{t-1}[bb, :]
    if (t > 0) {
#pragma unroll
        for (int g = 0; g < NG; ++g)
#pragma unroll
            for (int bb = 0; bb < RNN_BT; ++bb) acc[g][bb] = 0.f;
        for (int k4 = lane; k4 < (H >> 2); k4 += 32) {
            float4 w[NG];
#pragma unroll
            for (int g = 0; g < NG; ++g) w[g] = __ldg(reinterpret_cast<const float4*>(Whh + ((long long)g * H + j) * H) + k4);
#pragma unroll
            for (int bb = 0; bb < RNN_BT; ++bb) {
                const float4 hv = reinterpret_cast<const float4*>(hs + bb * H)[k4];
#pragma unroll
                for (int g = 0; g < NG; ++g) {
                    acc[g][bb] = fmaf(w[g].x, hv.x, acc[g][bb]);
                    acc[g][bb] = fmaf(w[g].y, hv.y, acc[g][bb]);
                    acc[g][bb] = fmaf(w[g].z, hv.z, acc[g][bb]);
                    acc[g][bb] = fmaf(w[g].w, hv.w, acc[g][bb]);
                }
            }
        }
#pragma unroll
        for (int g = 0; g < NG; ++g)
#pragma unroll
            for (int bb = 0; bb < RNN_BT; ++bb) {
                const float s = warp_sum(acc[g][bb]);
                if (lane == bb) gh[g] = s;
            }
    }
    if (lane >= nb) return;
    const int b = b0 + lane;
    const float* bi = p.b_ih[l];
    const float* bh = p.b_hh[l];
    float* hout = p.h + (((long long)l * p.B + b) * p.L + t) * H + j;
    const float hprev = t > 0 ? hs[lane * H + j] : 0.f;
    if (NG == 4) {
        const float pi = (gi[0] + bi[j]) + bh[j] + gh[0];
        const float pf = (gi[1] + bi[H + j]) + bh[H + j] + gh[1];
        const float pg = (gi[2] + bi[2 * H + j]) + bh[2 * H + j] + gh[2];
        const float po = (gi[NG - 1] + bi[3 * H + j]) + bh[3 * H + j] + gh[NG - 1];
        float* cp = p.c + ((long long)l * p.B + b) * H + j;
        const float cprev = t > 0 ? *cp : 0.f;
        const float cn = sigmoid_exact(pf) * cprev + sigmoid_exact(pi) * tanhf(pg);
        *cp = cn;
        *hout = sigmoid_exact(po) * tanhf(cn);
    } else {
        const float r = sigmoid_exact((gi[0] + bi[j]) + gh[0] + bh[j]);
        const float z = sigmoid_exact((gi[1] + bi[H + j]) + gh[1] + bh[H + j]);
        const float n = tanhf((gi[2] + bi[2 * H + j]) + r * (gh[2] + bh[2 * H + j]));
        *hout = (1.f - z) * n + z * hprev;
    }
}

__global__ void __launch_bounds__(256) critic_wavefront_kernel(const CriticWave p) {
    extern __shared__ float smem[];
    const int H = p.H;
    const int l = blockIdx.x / p.cpl, ci = blockIdx.x - l * p.cpl;
    const int Kin = l == 0 ? p.E : H;
    float* xs = smem;                       // [RNN_BT][Kin]
    float* hs = smem + RNN_BT * (p.E > H ? p.E : H);   // [RNN_BT][H]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool lstm = l < p.n_lstm;
    const bool act_in = (l == p.n_lstm) && p.alpha != nullptr;   // AReLU on the input of the first GRU layer
    float aa = 0.f, ab = 0.f;
    if (act_in) {
        aa = fminf(fmaxf(p.alpha[0], 0.01f), 0.99f);
        ab = 1.f + sigmoid_exact(p.beta[0]);
    }
    const int n_waves = p.L + p.n_layers - 1;
    for (int w = 0; w < n_waves; ++w) {
        const int t = w - l;
        if (t >= 0 && t < p.L) {
            for (int b0 = 0; b0 < p.B; b0 += RNN_BT) {
                const int nb = min(RNN_BT, p.B - b0);
                // x tile: the embedding (layer 0) or the previous layer's h at the SAME step (written one wavefront ago)
                const int K4 = Kin >> 2;
                for (int i = threadIdx.x; i < RNN_BT * K4; i += blockDim.x) {
                    const int bb = i / K4, k4 = i - bb * K4;
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (bb < nb) {
                        if (l == 0) {
                            v = __ldg(reinterpret_cast<const float4*>(p.emb + ((long long)(b0 + bb) * p.L + t) * p.E) + k4);
                        } else {
                            v = __ldcg(reinterpret_cast<const float4*>(p.h + ((((long long)(l - 1) * p.B + b0 + bb) * p.L + t) * H)) + k4);
                            if (act_in) {
                                v.x = fmaxf(v.x, 0.f) * ab - fmaxf(-v.x, 0.f) * aa;
                                v.y = fmaxf(v.y, 0.f) * ab - fmaxf(-v.y, 0.f) * aa;
                                v.z = fmaxf(v.z, 0.f) * ab - fmaxf(-v.z, 0.f) * aa;
                                v.w = fmaxf(v.w, 0.f) * ab - fmaxf(-v.w, 0.f) * aa;
                            }
                        }
                    }
                    reinterpret_cast<float4*>(xs)[i] = v;
                }
                if (t > 0) {
                    const int H4 = H >> 2;
                    for (int i = threadIdx.x; i < RNN_BT * H4; i += blockDim.x) {
                        const int bb = i / H4, k4 = i - bb * H4;
                        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (bb < nb)
                            v = __ldcg(reinterpret_cast<const float4*>(p.h + ((((long long)l * p.B + b0 + bb) * p.L + t - 1) * H)) + k4);
                        reinterpret_cast<float4*>(hs)[i] = v;
                    }
                }
                __syncthreads();
                for (int u = warp; u < p.upc; u += 8) {
                    const int j = ci * p.upc + u;
                    if (j < H) {
                        if (lstm) wave_unit<4>(p, l, t, j, b0, nb, Kin, xs, hs, lane);
                        else wave_unit<3>(p, l, t, j, b0, nb, Kin, xs, hs, lane);
                    }
                }
                __syncthreads();
            }
        }
        if (w + 1 < n_waves) grid_barrier(p.bar, (unsigned)(w + 1) * gridDim.x);
    }
}

}  // namespace bmhrl

using namespace bmhrl;

extern "C" long long bmhrl_critic_wavefront_ws_floats(int n_layers, int n_lstm, int B, int L, int H) {
    return (long long)n_layers * B * L * H + (long long)n_lstm * B * H + 4;
}

extern "C" int bmhrl_critic_wavefront_f32(const float* emb, int E, const float* const* w_ih, const float* const* w_hh,
                                          const float* const* b_ih, const float* const* b_hh, int n_lstm, int n_gru,
                                          const float* arelu_alpha, const float* arelu_beta, float* ws, int B, int L, int H,
                                          void* stream) {
    const int n_layers = n_lstm + n_gru;
    BMHRL_CHECK_ARG(emb && w_ih && w_hh && b_ih && b_hh && ws, "bmhrl_critic_wavefront_f32: null pointer");
    BMHRL_CHECK_ARG(n_layers >= 1 && n_layers <= 8 && n_lstm >= 0 && n_gru >= 0, "bmhrl_critic_wavefront_f32: 1..8 layers");
    BMHRL_CHECK_ARG(B > 0 && L > 0 && H > 0 && E > 0 && (H & 3) == 0 && (E & 3) == 0 && aligned16(emb) && aligned16(ws),
                    "bmhrl_critic_wavefront_f32: sizes must be multiples of 4, pointers 16-byte aligned");
    CriticWave p{};
    p.emb = emb; p.E = E;
    for (int l = 0; l < n_layers; ++l) {
        BMHRL_CHECK_ARG(w_ih[l] && w_hh[l] && b_ih[l] && b_hh[l] && aligned16(w_ih[l]) && aligned16(w_hh[l]),
                        "bmhrl_critic_wavefront_f32: layer weights");
        p.w_ih[l] = w_ih[l]; p.w_hh[l] = w_hh[l]; p.b_ih[l] = b_ih[l]; p.b_hh[l] = b_hh[l];
    }
    p.n_lstm = n_lstm; p.n_layers = n_layers; p.alpha = arelu_alpha; p.beta = arelu_beta;
    p.h = ws;
    p.c = ws + (long long)n_layers * B * L * H;
    p.bar = reinterpret_cast<unsigned*>(ws + (long long)n_layers * B * L * H + (long long)n_lstm * B * H);
    p.B = B; p.L = L; p.H = H;
    const int n_sm = device_sm_count();
    p.cpl = n_sm / n_layers;
    if (p.cpl < 1) return 1;
    p.upc = (H + p.cpl - 1) / p.cpl;
    const size_t smem = sizeof(float) * RNN_BT * ((size_t)(E > H ? E : H) + H);
    static DeviceOnce attr;
    if (needs_setup(attr)) {
        if (cudaFuncSetAttribute(critic_wavefront_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) != cudaSuccess) {
            bmhrl_set_error("critic wavefront: cannot raise the shared-memory limit");
            return BMHRL_ERR_CUDA;
        }
        mark_done(attr);
    }
    if (smem > 200 * 1024) return 1;
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, critic_wavefront_kernel, 256, smem) != cudaSuccess || per_sm < 1)
        return 1;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(p.bar, 0, sizeof(unsigned), st) != cudaSuccess) {
        bmhrl_set_error("critic wavefront: memset failed");
        return BMHRL_ERR_CUDA;
    }
    void* args[] = {(void*)&p};
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)critic_wavefront_kernel, dim3(p.cpl * n_layers), dim3(256), args,
                                                smem, st);
    if (e != cudaSuccess) {
        bmhrl_set_error(cudaGetErrorString(e));
        return BMHRL_ERR_CUDA;
    }
    bmhrl_count_launch(1);
    return BMHRL_OK;
}

static int g_rnn_persistent = 1;
static int g_rnn_min_upb = 0;   // bmhrl_rnn_set_units_per_cta

// Returns BMHRL_OK / error, or 1 when the shape cannot run as one co-resident grid.
template <int NG>
static int launch_rnn_persistent_t(const float* gx, const float* W_hh, const float* b_hh, float* out, float* ws, int B,
                                   int L, int H, cudaStream_t st) {
    const int n_sm = device_sm_count();
    static DeviceOnce attr;
    if (needs_setup(attr)) {
        if (cudaFuncSetAttribute(rnn_layer_persistent_kernel<NG>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 220 * 1024) != cudaSuccess) {
            bmhrl_set_error("rnn persistent: cannot raise the shared-memory limit");
            return BMHRL_ERR_CUDA;
        }
        mark_done(attr);
    }
    // smallest units-per-block whose grid is co-resident (1 CTA per SM is assumed; smaller smem only helps)
    int upb = (H + n_sm - 1) / n_sm;
    // BMHRL_RNN_UPB: at least that many hidden units per CTA -> fewer, fatter CTAs (a step is a latency chain, not
    // arithmetic: the layer then leaves the other SMs to the kernels that run beside the frozen critic)
    static const int env_upb = getenv("BMHRL_RNN_UPB") ? atoi(getenv("BMHRL_RNN_UPB")) : 0;
    const int want_upb = env_upb > 0 ? env_upb : g_rnn_min_upb;
    if (want_upb > upb) upb = want_upb > 16 ? 16 : want_upb;
    for (; upb <= 16; ++upb) {
        const size_t smem = sizeof(float) * ((size_t)upb * NG * H + (size_t)RNN_BT * H);
        if (smem > 220 * 1024) return 1;
        int per_sm = 0;
        const int threads = upb * 32 < 256 ? 256 : upb * 32;  // warps beyond `upb` only help broadcasting h
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rnn_layer_persistent_kernel<NG>, threads, smem) !=
            cudaSuccess) {
            bmhrl_set_error("rnn persistent: occupancy query failed");
            return BMHRL_ERR_CUDA;
        }
        const int grid = (H + upb - 1) / upb;
        if (per_sm * n_sm < grid) continue;
        float* c = ws;
        unsigned* bar = reinterpret_cast<unsigned*>(ws + (size_t)B * H);
        if (cudaMemsetAsync(bar, 0, sizeof(unsigned), st) != cudaSuccess) {
            bmhrl_set_error("rnn persistent: memset failed");
            return BMHRL_ERR_CUDA;
        }
        void* args[] = {(void*)&gx, (void*)&W_hh, (void*)&b_hh, (void*)&out, (void*)&c, (void*)&bar,
                        (void*)&B,  (void*)&L,    (void*)&H,    (void*)&upb};
        cudaError_t e = cudaLaunchCooperativeKernel((const void*)rnn_layer_persistent_kernel<NG>, dim3(grid),
                                                    dim3(threads), args, smem, st);
        if (e != cudaSuccess) {
            bmhrl_set_error(cudaGetErrorString(e));
            return BMHRL_ERR_CUDA;
        }
        bmhrl_count_launch(1);
        return BMHRL_OK;
    }
    return 1;
}

// offset (in floats) of the tagged exchange buffer inside the workspace: behind the cell state and the barrier words
static inline long long rnn_ws_x_offset(int B, int H) { return (((long long)B * H + 32) + 3) & ~3LL; }
static unsigned long long* g_rnn_trace = nullptr;   // diagnosis: bmhrl_rnn_trace
static int g_rnn_exchange = -1;   // -1: environment BMHRL_RNN_EXCHANGE (default on), 0 / 1: forced by bmhrl_rnn_set_exchange

// Returns BMHRL_OK / error, or 1 when the shape cannot run as one co-resident grid of the exchange kernel.
template <int NG>
static int launch_rnn_exchange_t(const float* gx, const float* W_hh, const float* b_hh, float* out, float* ws, int B, int L,
                                 int H, cudaStream_t st) {
    if ((H & 1) || H > 32 * RNN2_KI || !aligned16(ws)) return 1;
    const int n_sm = device_sm_count();
    static DeviceOnce attr;
    if (needs_setup(attr)) {
        if (cudaFuncSetAttribute(rnn_layer_exchange_kernel<NG>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024) !=
            cudaSuccess) {
            bmhrl_set_error("rnn exchange: cannot raise the shared-memory limit");
            return BMHRL_ERR_CUDA;
        }
        mark_done(attr);
    }
    int upb = (H + n_sm - 1) / n_sm;
    static const int env_upb = getenv("BMHRL_RNN_UPB") ? atoi(getenv("BMHRL_RNN_UPB")) : 0;
    if (env_upb > upb) upb = env_upb;
    for (; upb * RNN2_NGRP * 32 <= RNN2_MAXT; ++upb) {
        const size_t smem = sizeof(float) * (2 * (size_t)RNN_BT * H);
        if (smem > 220 * 1024) return 1;
        const int threads = upb * RNN2_NGRP * 32 < 128 ? 128 : upb * RNN2_NGRP * 32;
        int per_sm = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rnn_layer_exchange_kernel<NG>, threads, smem) !=
            cudaSuccess) {
            bmhrl_set_error("rnn exchange: occupancy query failed");
            return BMHRL_ERR_CUDA;
        }
        const int grid = (H + upb - 1) / upb;
        if (per_sm * n_sm < grid) continue;
        float* c = ws;
        unsigned long long* X = reinterpret_cast<unsigned long long*>(ws + rnn_ws_x_offset(B, H));
        if (cudaMemsetAsync(X, 0, sizeof(unsigned long long) * 2 * (size_t)B * H, st) != cudaSuccess) {
            bmhrl_set_error("rnn exchange: memset failed");
            return BMHRL_ERR_CUDA;
        }
        unsigned long long* trace = g_rnn_trace;
        void* args[] = {(void*)&gx, (void*)&W_hh, (void*)&b_hh, (void*)&out, (void*)&c, (void*)&X,
                        (void*)&B,  (void*)&L,    (void*)&H,    (void*)&upb, (void*)&trace};
        cudaError_t e = cudaLaunchCooperativeKernel((const void*)rnn_layer_exchange_kernel<NG>, dim3(grid), dim3(threads),
                                                    args, smem, st);
        if (e != cudaSuccess) {
            bmhrl_set_error(cudaGetErrorString(e));
            return BMHRL_ERR_CUDA;
        }
        bmhrl_count_launch(1);
        return BMHRL_OK;
    }
    return 1;
}

static int launch_rnn_persistent(int cell, const float* gx, const float* W_hh, const float* b_hh, float* out, float* ws,
                                 int B, int L, int H, cudaStream_t st) {
    static const int env_x = getenv("BMHRL_RNN_EXCHANGE") ? atoi(getenv("BMHRL_RNN_EXCHANGE")) : 0;
    if (g_rnn_exchange < 0 ? env_x != 0 : g_rnn_exchange != 0) {
        const int rc = cell == 0 ? launch_rnn_exchange_t<4>(gx, W_hh, b_hh, out, ws, B, L, H, st)
                                 : launch_rnn_exchange_t<3>(gx, W_hh, b_hh, out, ws, B, L, H, st);
        if (rc != 1) return rc;   // 1: not this shape -> the barrier kernel
    }
    return cell == 0 ? launch_rnn_persistent_t<4>(gx, W_hh, b_hh, out, ws, B, L, H, st)
                     : launch_rnn_persistent_t<3>(gx, W_hh, b_hh, out, ws, B, L, H, st);
}

extern "C" long long bmhrl_rnn_ws_floats(int B, int H) { return rnn_ws_x_offset(B, H) + 4LL * B * H; }
extern "C" int bmhrl_rnn_set_units_per_cta(int units) {
    g_rnn_min_upb = units > 0 ? units : 0;
    return BMHRL_OK;
}
extern "C" int bmhrl_rnn_trace(unsigned long long* device_slots) {
    g_rnn_trace = device_slots;
    return BMHRL_OK;
}
extern "C" int bmhrl_rnn_set_exchange(int mode) {
    g_rnn_exchange = mode < 0 ? -1 : (mode ? 1 : 0);
    return BMHRL_OK;
}
extern "C" int bmhrl_rnn_set_persistent(int on) {
    g_rnn_persistent = on ? 1 : 0;
    return BMHRL_OK;
}

// rnn_step_kernel needs more than the default 48 KB of dynamic shared memory for H > 768.
static int rnn_step_smem_attr(int cell, size_t smem) {
    static DeviceOnce attr_set[2];
    if (smem > 48 * 1024 && needs_setup(attr_set[cell])) {
        cudaError_t e = cell == 0 ? cudaFuncSetAttribute(rnn_step_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                         200 * 1024)
                                  : cudaFuncSetAttribute(rnn_step_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                         200 * 1024);
        if (e != cudaSuccess) {
            bmhrl_set_error(cudaGetErrorString(e));
            return BMHRL_ERR_CUDA;
        }
        mark_done(attr_set[cell]);
    }
    return BMHRL_OK;
}

extern "C" int bmhrl_rnn_step_f32(int cell, const float* gx, const float* W_hh, const float* b_hh, float* out,
                                  const float* c_prev, float* c_next, int B, int L, int H, int t, void* stream) {
    BMHRL_CHECK_ARG(cell == 0 || cell == 1, "bmhrl_rnn_step_f32: cell must be 0 (LSTM) or 1 (GRU)");
    BMHRL_CHECK_ARG(gx && W_hh && b_hh && out, "bmhrl_rnn_step_f32: null pointer");
    BMHRL_CHECK_ARG(cell == 1 || (c_next && (t == 0 || c_prev)), "bmhrl_rnn_step_f32: LSTM needs the cell-state buffers");
    BMHRL_CHECK_ARG(B > 0 && L > 0 && H > 0 && t >= 0 && t < L, "bmhrl_rnn_step_f32: bad sizes");
    const size_t smem = (size_t)RNN_BT * H * sizeof(float);
    BMHRL_CHECK_ARG(smem <= 200 * 1024, "bmhrl_rnn_step_f32: hidden size too large");
    if (int rc = rnn_step_smem_attr(cell, smem)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    dim3 grid((H + RNN_WARPS - 1) / RNN_WARPS, (B + RNN_BT - 1) / RNN_BT);
    if (cell == 0) rnn_step_kernel<4><<<grid, RNN_WARPS * 32, smem, st>>>(gx, W_hh, b_hh, out, c_prev, c_next, B, L, H, t);
    else rnn_step_kernel<3><<<grid, RNN_WARPS * 32, smem, st>>>(gx, W_hh, b_hh, out, nullptr, nullptr, B, L, H, t);
    BMHRL_POST_LAUNCH("rnn_step");
    return BMHRL_OK;
}

extern "C" int bmhrl_rnn_layer_f32(int cell, const float* gx, const float* W_hh, const float* b_hh, float* out,
                                   float* ws, int B, int L, int H, void* stream) {
    BMHRL_CHECK_ARG(cell == 0 || cell == 1, "bmhrl_rnn_layer_f32: cell must be 0 (LSTM) or 1 (GRU)");
    BMHRL_CHECK_ARG(gx && W_hh && b_hh && out, "bmhrl_rnn_layer_f32: null pointer");
    BMHRL_CHECK_ARG(cell == 1 || ws, "bmhrl_rnn_layer_f32: LSTM needs the workspace (bmhrl_rnn_ws_floats)");
    BMHRL_CHECK_ARG(B >= 0 && L >= 0 && H > 0, "bmhrl_rnn_layer_f32: bad sizes");
    if (B == 0 || L == 0) return BMHRL_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t smem = (size_t)RNN_BT * H * sizeof(float);
    BMHRL_CHECK_ARG(smem <= 200 * 1024, "bmhrl_rnn_layer_f32: hidden size too large");
    if (int rc = rnn_step_smem_attr(cell, smem)) return rc;
    if (g_rnn_persistent && ws) {
        int rc = launch_rnn_persistent(cell, gx, W_hh, b_hh, out, ws, B, L, H, st);
        if (rc != 1) return rc;  // 1 = shape does not fit the persistent kernel: use the per-step kernels
    }
    dim3 grid((H + RNN_WARPS - 1) / RNN_WARPS, (B + RNN_BT - 1) / RNN_BT);
    for (int t = 0; t < L; ++t) {
        if (cell == 0) rnn_step_kernel<4><<<grid, RNN_WARPS * 32, smem, st>>>(gx, W_hh, b_hh, out, ws, ws, B, L, H, t);
        else rnn_step_kernel<3><<<grid, RNN_WARPS * 32, smem, st>>>(gx, W_hh, b_hh, out, nullptr, nullptr, B, L, H, t);
        BMHRL_POST_LAUNCH("rnn_step");
    }
    return BMHRL_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// bmhrl_rnn_layer_noalloc_f32: the same layer for callers that cannot allocate device memory (CUDA-graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  The workspace of bmhrl_rnn_layer_f32 (LSTM cell state, the grid-
// barrier counter, the tagged exchange buffer) lives in ONE file-scope device array of a fixed, modest size that every
// GPU which loads the library holds until exit.  No counterpart in the reference.
// In host code `g_rnn_noalloc_scratch` is only a stand-in for the symbol, NOT a device address: the address
// comes from cudaGetSymbolAddress, once per device, with that device current.  All calls share the array and it keeps
// the last call's contents: the launcher clears what must start at zero (barrier counter / exchange tags, stream-ordered
// memsets), the cell state is written before it is read (t = 0 reads no state).  Two calls that overlap on the GPU
// (different streams, or concurrent graph branches) RACE on it -- serialise them.
constexpr long long RNN_NOALLOC_MAX_FLOATS = 256 * 1024;   // 1 MB per GPU; B = 16, H = 600 needs 48 032 floats
static __device__ __align__(256) float g_rnn_noalloc_scratch[RNN_NOALLOC_MAX_FLOATS];

static int rnn_noalloc_scratch_ptr(float** out) {
    static std::atomic<float*> ptr[kMaxDevices];
    const int slot = current_device_slot();
    float* p = ptr[slot].load(std::memory_order_acquire);
    if (!p) {
        void* q = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&q, g_rnn_noalloc_scratch);   // the current device's copy of the array
        if (e != cudaSuccess || !q) {
            bmhrl_set_error(e != cudaSuccess ? cudaGetErrorString(e) : "cudaGetSymbolAddress returned a null address");
            return BMHRL_ERR_CUDA;
        }
        p = static_cast<float*>(q);
        ptr[slot].store(p, std::memory_order_release);
    }
    *out = p;
    return BMHRL_OK;
}

extern "C" long long bmhrl_rnn_noalloc_max_floats(void) { return RNN_NOALLOC_MAX_FLOATS; }

extern "C" int bmhrl_rnn_layer_noalloc_f32(int cell, const float* gx, const float* W_hh, const float* b_hh, float* out,
                                           int B, int L, int H, void* stream) {
    BMHRL_CHECK_ARG(B >= 0 && L >= 0 && H > 0, "bmhrl_rnn_layer_noalloc_f32: bad sizes");
    if (bmhrl_rnn_ws_floats(B, H) > RNN_NOALLOC_MAX_FLOATS) {
        bmhrl_set_error("bmhrl_rnn_layer_noalloc_f32: B x H needs more than the built-in scratch (bmhrl_rnn_ws_floats(B, H) > "
                        "bmhrl_rnn_noalloc_max_floats() = 262144 floats): use bmhrl_rnn_layer_f32 with a caller workspace");
        return BMHRL_ERR_ARG;
    }
    float* ws = nullptr;
    if (int rc = rnn_noalloc_scratch_ptr(&ws)) return rc;
    return bmhrl_rnn_layer_f32(cell, gx, W_hh, b_hh, out, ws, B, L, H, stream);
}

extern "C" int bmhrl_arelu_f32(const float* x, const float* alpha, const float* beta, float* y, long long n,
                               void* stream) {
    BMHRL_CHECK_ARG(x && alpha && beta && y && n >= 0, "bmhrl_arelu_f32: bad argument");
    if (n == 0) return BMHRL_OK;
    int blocks = (int)min((long long)148 * 8, (n + 255) / 256);
    arelu_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, alpha, beta, y, n);
    BMHRL_POST_LAUNCH("arelu");
    return BMHRL_OK;
}

extern "C" int bmhrl_segment_labels_f32(const float* x, const float* w, const float* b, float threshold, float* scores,
                                        int* labels, long long rows, int D, void* stream) {
    BMHRL_CHECK_ARG(x && w && b && labels && rows >= 0 && D > 0, "bmhrl_segment_labels_f32: bad argument");
    if (rows == 0) return BMHRL_OK;
    segment_label_kernel<<<(unsigned)((rows + 3) / 4), 128, 0, (cudaStream_t)stream>>>(x, w, b, threshold, scores,
                                                                                       labels, rows, D);
    BMHRL_POST_LAUNCH("segment_labels");
    return BMHRL_OK;
}
