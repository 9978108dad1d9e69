"""Autograd-aware Python wrappers over the C ABI (include/bmhrl_b200.h).

Every function here launches hand-written sm_100a kernels from libbmhrl_b200.so on torch's current CUDA stream and
allocates outputs / workspaces through torch's caching allocator, so whole training steps are CUDA-graph
capturable.  Nothing in this file computes on the CPU and nothing falls back to ATen math: a non-CUDA tensor
raises.
"""
import math

import os

import torch

from ._lib import LIB

_F32 = torch.float32


# ----------------------------------------------------------------------------------------------------------------
# plumbing
# ----------------------------------------------------------------------------------------------------------------
def _stream():
    return torch.cuda.current_stream().cuda_stream


def _p(t):
    return None if t is None else t.data_ptr()


def _chk(t, name, dtype=_F32):
    if not t.is_cuda:
        raise RuntimeError(f"bmhrl_b200: `{name}` must be a CUDA tensor (there is no CPU fallback)")
    if t.dtype != dtype:
        raise RuntimeError(f"bmhrl_b200: `{name}` must be {dtype}, got {t.dtype}")
    return t


def _c(t, name, dtype=_F32):
    return _chk(t, name, dtype).contiguous()


class _Rng:
    """Per-device dropout/sampling state: a device tensor {seed, step} plus the frozen snapshot that the ops of the
    current forward pass (and their backward) read.  `new_epoch()` advances the step ON THE DEVICE (graph-safe)."""

    def __init__(self):
        self.state = {}
        self.active = {}
        self.salt = 0

    def _dev(self, device):
        return torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device()

    def seed(self, seed, device="cuda"):
        d = self._dev(device)
        self.state[d] = torch.tensor([seed, 0], dtype=torch.int64, device=f"cuda:{d}")
        self.active[d] = self.state[d].clone()

    def get_state(self, device):
        d = self._dev(device)
        if d not in self.state:
            self.seed(0x5EED, f"cuda:{d}")
        return self.state[d]

    def new_epoch(self, device):
        st = self.get_state(device)
        LIB.call("bmhrl_tick", None, None, 0, st.data_ptr(), _stream())
        self.active[self._dev(device)] = st.clone()
        return self.active[self._dev(device)]

    def current(self, device):
        self.get_state(device)
        return self.active[self._dev(device)]

    def next_salt(self):
        self.salt += 1
        return self.salt


RNG = _Rng()


# ----------------------------------------------------------------------------------------------------------------
# side streams: independent branches of the agent (frozen critic, worker fusion stack) run concurrently with the main
# stream; fork/join is plain event ordering, so it is captured into the step's CUDA graph as parallel branches.
# ----------------------------------------------------------------------------------------------------------------
_SIDE = {}        # slot key -> stream; keys are paths ("main", i, j, ...) so nested forks get their own streams
_SIDE_KEY = {}    # cuda_stream handle -> slot key
_DIRTY = set()    # side streams that received work since the last join_side_streams()
CONCURRENT = True  # False: every branch runs on the caller's stream (debugging / profiling aid)


def new_stream(device_index):
    """A CUDA stream of our own (bmhrl_stream_create), wrapped for torch.  torch.cuda.Stream() draws from a pool of 32
    handles per device and re-issues them round-robin, so two side streams could be ONE CUDA stream; during graph
    capture (no record_stream) the caching allocator then re-used a block freed on one branch for an unordered branch
    that happened to share the handle.  Every fork target of this module has a unique handle instead."""
    import ctypes
    h = ctypes.c_void_p()
    with torch.cuda.device(device_index):
        LIB.call("bmhrl_stream_create", ctypes.byref(h))
    return torch.cuda.ExternalStream(h.value, device=device_index)


_WARMUP = {}


def warmup_stream(device):
    """One private stream per device for the eager passes that precede a graph capture."""
    idx = torch.device(device).index
    idx = torch.cuda.current_device() if idx is None else idx
    if idx not in _WARMUP:
        _WARMUP[idx] = new_stream(idx)
    return _WARMUP[idx]


def side_stream(device, slot):
    """Side stream number `slot` of the stream that is current on `device` (created on first use)."""
    dev = torch.device(device)
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    cur = torch.cuda.current_stream(idx)
    key = _SIDE_KEY.get((idx, cur.cuda_stream), (idx, "main")) + (slot,)
    if key not in _SIDE:
        _SIDE[key] = new_stream(idx)
        _SIDE_KEY[(idx, _SIDE[key].cuda_stream)] = key
    _DIRTY.add(key)
    return _SIDE[key]


def parallel(device, *thunks):
    """Runs independent branches concurrently: thunk 0 on the current stream, the others on side streams that are
    forked from / joined back into it.  Returns the list of results.  (Same kernels, same arithmetic: only the
    stream each branch is enqueued on differs.)"""
    if not CONCURRENT or len(thunks) == 1:
        return [t() for t in thunks]
    cur = torch.cuda.current_stream(device)
    out = [None] * len(thunks)
    sides = []
    for i in range(1, len(thunks)):
        s = side_stream(device, i)
        s.wait_stream(cur)
        with torch.cuda.stream(s):
            out[i] = thunks[i]()
        sides.append(s)
    out[0] = thunks[0]()
    eager = not torch.cuda.is_current_stream_capturing()
    for i, s in enumerate(sides, 1):
        cur.wait_stream(s)
        if eager:
            for t in (out[i] if isinstance(out[i], (tuple, list)) else (out[i],)):
                if torch.is_tensor(t):
                    t.record_stream(cur)
    return out


_WGRAD_KEEP = []   # operands of weight-gradient GEMMs in flight on the wgrad streams (kept alive until the join)
_WGRAD_RR = [0]
_WGRAD_DEST = {}   # (device, destination data_ptr) -> wgrad stream index: ONE stream per destination, for good


def wgrad_stream(device, dest=None):
    """Weight-gradient GEMMs that accumulate straight into a FusedAdam flat buffer have no consumer before the
    optimiser step: they are forked onto two side streams, off the critical path of backward.  The beta = 1 epilogue
    is a plain read-add-write, so two GEMMs into the SAME destination (a module used twice in one forward, two
    forwards before one backward) must never run concurrently: the stream is chosen by the destination, round-robin
    only over destinations seen for the first time, and work on one stream is ordered."""
    cur = torch.cuda.current_stream(device)
    if dest is None:
        _WGRAD_RR[0] ^= 1
        idx = _WGRAD_RR[0]
    else:
        k = (cur.device.index, dest.data_ptr())
        idx = _WGRAD_DEST.get(k)
        if idx is None:
            _WGRAD_RR[0] ^= 1
            idx = _WGRAD_DEST[k] = _WGRAD_RR[0]
    key = (cur.device.index, "wgrad", idx)
    if key not in _SIDE:
        _SIDE[key] = new_stream(cur.device.index)
    _DIRTY.add(key)
    s = _SIDE[key]
    s.wait_stream(cur)
    return s


def join_side_streams(device=None):
    """Makes the current stream wait for everything enqueued on the side streams since the last join.  Backward
    kernels that accumulate straight into a FusedAdam flat buffer run on the stream of their forward op and hand
    autograd no gradient tensor, so the trainer joins explicitly before the optimiser step."""
    cur = torch.cuda.current_stream(device)
    for key in sorted(_DIRTY, key=lambda k: (len(k), str(k)), reverse=True):
        if key[0] == cur.device.index:
            cur.wait_stream(_SIDE[key])
            _DIRTY.discard(key)
    _WGRAD_KEEP.clear()


def wait_dirty_side_streams(stream, device=None):
    """`stream` waits for everything enqueued so far on the side streams that received work since the last join
    (they stay marked: the trainer's final join still covers what they get afterwards)."""
    idx = stream.device.index
    for key in list(_DIRTY):
        if key[0] == idx and _SIDE[key] is not stream:
            stream.wait_stream(_SIDE[key])


class _GradMarker(torch.autograd.Function):
    """Identity whose backward calls `callback()` first.  Autograd runs ready nodes in reverse creation order, so when
    the marker placed on the encoder outputs fires, the backward kernels of everything created AFTER it (fusion
    stacks, manager / worker heads, loss) have all been enqueued: their gradients can start their all-reduce while
    the encoder's backward is still running."""

    @staticmethod
    def forward(ctx, x, callback):
        ctx.callback = callback
        y = x.view_as(x)
        tw = getattr(x, "_bmhrl16", None)   # bf16 twin emitted by the producing kernel (ops16): same data, keep it
        if tw is not None:
            y._bmhrl16 = tw
        hint = getattr(x, "_bmhrl_bwd_cast", None)   # ops16.hint_backward_cast: an identity keeps it valid
        if hint is not None:
            y._bmhrl_bwd_cast = hint
        return y

    @staticmethod
    def backward(ctx, g):
        ctx.callback()
        return g, None


def grad_marker(x, callback):
    if callback is None or not (torch.is_grad_enabled() and x.requires_grad):
        return x
    return _GradMarker.apply(x, callback)


def manual_seed(seed, device="cuda"):
    RNG.seed(int(seed), device)


def new_rng_epoch(device="cuda"):
    return RNG.new_epoch(device)


def next_salt():
    return RNG.next_salt()


# ----------------------------------------------------------------------------------------------------------------
# contractions
# ----------------------------------------------------------------------------------------------------------------
BF16 = False  # math mode 'bf16': the modules route their blocks through bmhrl_b200.ops16 (bf16 tensor-core operands)


def set_math_mode(mode):
    """'fp32': exact SIMT contractions (default); 'tf32': tcgen05 tensor-core GEMMs (fp32 storage, tf32 multiplies,
    fp32 accumulation in TMEM); 'bf16': the residual stream, statistics, gradients of parameters and optimiser state
    stay fp32, every tensor that is only a tensor-core operand is bf16 (bmhrl_b200/ops16.py: grouped 128x128 tcgen05
    GEMMs, fused one-tile attention forward + backward); whatever has no bf16 kernel (the frozen critic, the value
    heads) runs as in 'tf32'."""
    global BF16
    if mode not in ("fp32", "tf32", "bf16"):
        raise ValueError(f"unknown math mode {mode!r}")
    LIB.call("bmhrl_set_math_mode", 0 if mode == "fp32" else 1)
    BF16 = mode == "bf16"


def get_math_mode():
    return "bf16" if BF16 else ("fp32", "tf32")[LIB.raw("bmhrl_get_math_mode")()]


RECORD = None  # bench.py: when a list, every GEMM launch appends its shape/stride tuple (roofline re-issue)


def gemm(A, sAm, sAk, B, sBk, sBn, C, ldc, M, N, K, batch=1, Hdiv=1, bA=(0, 0), bB=(0, 0), bC=(0, 0), alpha=1.0,
         beta=0.0):
    if RECORD is not None:
        RECORD.append(("gemm", M, N, K, sAm, sAk, sBk, sBn, ldc))
    LIB.call("bmhrl_gemm_f32", _p(A), sAm, sAk, _p(B), sBk, sBn, _p(C), ldc, M, N, K, batch, Hdiv, bA[0], bA[1], bB[0],
             bB[1], bC[0], bC[1], alpha, beta, _stream())


def _gbuf(p):
    """The slice of a FusedAdam flat gradient buffer that belongs to parameter `p` (None outside a trainer, or when
    the parameter is frozen): backward kernels then accumulate into it directly and autograd gets no gradient tensor
    to add (no per-parameter AccumulateGrad kernels)."""
    if p is None or not p.requires_grad:
        return None
    return getattr(p, "_bmhrl_grad", None)


class _Linear(torch.autograd.Function):
    """y = resid + dropout(act(x W^T + b)); reference call sites listed in include/bmhrl_b200.h."""

    @staticmethod
    def forward(ctx, x, weight, bias, resid, act, drop_p, salt, rng, wgrad, bgrad):
        x = _c(x, "x")
        weight = _c(weight, "weight")
        N, K = weight.shape
        if x.shape[-1] != K:
            raise RuntimeError(f"linear: input feature size {x.shape[-1]} != weight in_features {K}")
        M = x.numel() // K
        y = torch.empty(x.shape[:-1] + (N,), device=x.device, dtype=_F32)
        r = None
        if resid is not None:
            r = _c(resid, "resid")
            if r.shape != y.shape:
                raise RuntimeError("linear: residual shape mismatch")
        b = None if bias is None else _c(bias, "bias")
        if RECORD is not None:
            RECORD.append(("linear", M, N, K))
        LIB.call("bmhrl_linear_fwd_f32", _p(x), K, _p(weight), _p(b), _p(r), N, _p(y), N, M, N, K, act, drop_p,
                 _p(rng) if drop_p > 0 else None, salt, _stream())
        ctx.save_for_backward(x, weight, y if act == 1 else None, rng if drop_p > 0 else None)
        ctx.gbufs = (wgrad, bgrad)  # destination slices of the flat gradient buffer: not autograd-saved (the
        # data-parallel all-reduce updates other slices of that buffer in place while backward is still running)
        ctx.cfg = (act, drop_p, salt, M, N, K, bias is not None, resid is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, y, rng = ctx.saved_tensors
        wgrad, bgrad = ctx.gbufs
        act, drop_p, salt, M, N, K, has_bias, has_resid = ctx.cfg
        dy = _c(dy, "dy")
        need_x, need_w, need_b, need_r = ctx.needs_input_grad[0], ctx.needs_input_grad[1], \
            has_bias and ctx.needs_input_grad[2], has_resid and ctx.needs_input_grad[3]
        st = _stream()
        dz = dy
        db = None
        acc_b = 1 if (need_b and bgrad is not None) else 0  # straight into the flat gradient buffer
        if act == 1 or drop_p > 0:
            dz = torch.empty(M, N, device=dy.device, dtype=_F32)
            db = (bgrad if acc_b else torch.empty(N, device=dy.device, dtype=_F32)) if need_b else None
            LIB.call("bmhrl_linear_bwd_prep_f32", _p(dy), N, _p(y), N, _p(dz), _p(db), M, N, act, drop_p, _p(rng), salt,
                     acc_b, st)
        elif need_b:
            db = bgrad if acc_b else torch.empty(N, device=dy.device, dtype=_F32)
            LIB.call("bmhrl_linear_bwd_prep_f32", _p(dy), N, None, N, None, _p(db), M, N, 0, 0.0, None, 0, acc_b, st)
        if acc_b:
            db = None
        dx = dw = None
        if need_x:  # dX = dZ W : (M,N)(N,K)
            dx = torch.empty(x.shape, device=dy.device, dtype=_F32)
            gemm(dz, N, 1, weight, K, 1, dx, K, M, K, N)
        if need_w:  # dW = dZ^T X : (N,M)(M,K)
            if wgrad is not None:  # accumulate into the parameter's slice of the flat gradient buffer (beta = 1)
                if CONCURRENT:
                    _WGRAD_KEEP.append((dz, x))
                    with torch.cuda.stream(wgrad_stream(dy.device, wgrad)):
                        gemm(dz, 1, N, x, K, 1, wgrad, K, N, K, M, beta=1.0)
                else:
                    gemm(dz, 1, N, x, K, 1, wgrad, K, N, K, M, beta=1.0)
            else:
                dw = torch.empty(N, K, device=dy.device, dtype=_F32)
                gemm(dz, 1, N, x, K, 1, dw, K, N, K, M)
        return dx, dw, db, (dy if need_r else None), None, None, None, None, None, None


def linear(x, weight, bias=None, resid=None, act=0, drop_p=0.0, salt=0, rng=None):
    if drop_p > 0 and rng is None:
        rng = RNG.current(x.device)
    return _Linear.apply(x, weight, bias, resid, act, float(drop_p), int(salt), rng, _gbuf(weight), _gbuf(bias))


# ----------------------------------------------------------------------------------------------------------------
# K2
# ----------------------------------------------------------------------------------------------------------------
class _LayerNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, eps, ggrad, bgrad):
        x = _c(x, "x")
        gamma = _c(gamma, "gamma")
        beta = _c(beta, "beta")
        D = x.shape[-1]
        rows = x.numel() // D
        y = torch.empty_like(x)
        mean = torch.empty(rows, device=x.device, dtype=_F32)
        rstd = torch.empty(rows, device=x.device, dtype=_F32)
        LIB.call("bmhrl_layernorm_fwd_f32", _p(x), None, _p(gamma), _p(beta), None, _p(y), _p(mean), _p(rstd), rows, D,
                 eps, _stream())
        ctx.save_for_backward(x, gamma, mean, rstd)
        ctx.gbufs = (ggrad, bgrad)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gamma, mean, rstd = ctx.saved_tensors
        ggrad, bgrad = ctx.gbufs
        dy = _c(dy, "dy")
        D = x.shape[-1]
        rows = x.numel() // D
        need_x = ctx.needs_input_grad[0]
        need_aff = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
        dx = torch.empty_like(x) if need_x else None
        acc = 1 if (need_aff and ggrad is not None and bgrad is not None) else 0
        dg = (ggrad if acc else torch.empty(D, device=x.device, dtype=_F32)) if need_aff else None
        db = (bgrad if acc else torch.empty(D, device=x.device, dtype=_F32)) if need_aff else None
        LIB.call("bmhrl_layernorm_bwd_f32", _p(dy), _p(x), _p(gamma), _p(mean), _p(rstd), None, _p(dx), _p(dg), _p(db),
                 rows, D, acc, _stream())
        if acc:
            dg = db = None
        return dx, dg, db, None, None, None


def layer_norm(x, gamma, beta, eps=1e-5):
    return _LayerNorm.apply(x, gamma, beta, float(eps), _gbuf(gamma), _gbuf(beta))


class _GatedMix(torch.autograd.Function):
    @staticmethod
    def forward(ctx, v, a, gate, ggrad):
        v = _c(v, "v")
        a = _c(a, "a")
        gate = _c(gate, "gate")
        out = torch.empty_like(v)
        LIB.call("bmhrl_gated_mix_fwd_f32", _p(v), _p(a), _p(gate), _p(out), v.numel(), _stream())
        ctx.save_for_backward(v, a, gate)
        ctx.gbufs = (ggrad,)
        return out

    @staticmethod
    def backward(ctx, g):
        v, a, gate = ctx.saved_tensors
        (ggrad,) = ctx.gbufs
        g = _c(g, "g")
        dv = torch.empty_like(v)
        da = torch.empty_like(a)
        acc = 1 if ggrad is not None else 0
        dgate = ggrad if acc else torch.empty(1, device=v.device, dtype=_F32)
        ws = torch.empty(592, device=v.device, dtype=_F32)
        LIB.call("bmhrl_gated_mix_bwd_f32", _p(g), _p(v), _p(a), _p(gate), _p(dv), _p(da), _p(dgate), _p(ws), 592,
                 v.numel(), acc, _stream())
        return dv, da, (None if acc else dgate), None


def gated_mix(v, a, gate):
    return _GatedMix.apply(v, a, gate, _gbuf(gate))


class _DropoutAdd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, r, drop_p, salt, rng):
        x = _c(x, "x")
        r = _c(r, "r")
        out = torch.empty_like(x)
        LIB.call("bmhrl_dropout_add_f32", _p(x), _p(r), _p(out), x.numel(), drop_p, _p(rng) if drop_p > 0 else None,
                 salt, _stream())
        ctx.save_for_backward(rng if drop_p > 0 else None)
        ctx.cfg = (drop_p, salt)
        return out

    @staticmethod
    def backward(ctx, g):
        (rng,) = ctx.saved_tensors
        drop_p, salt = ctx.cfg
        g = _c(g, "g")
        dr = g
        if drop_p > 0 and ctx.needs_input_grad[1]:
            dr = torch.empty_like(g)
            LIB.call("bmhrl_dropout_apply_f32", _p(g), _p(dr), g.numel(), drop_p, _p(rng), salt, _stream())
        return g, dr, None, None, None


def dropout_add(x, r, drop_p=0.0, salt=0, rng=None):
    if drop_p > 0 and rng is None:
        rng = RNG.current(x.device)
    return _DropoutAdd.apply(x, r, float(drop_p), int(salt), rng)


# ----------------------------------------------------------------------------------------------------------------
# K1 attention core (exact-fp32 path)
# ----------------------------------------------------------------------------------------------------------------
def _mask_args(mask, B, Sq, Sk):
    """mask: None | bool (B,1,Sk) / (B,1,1,Sk) key padding | bool (B,Sq,Sk) / (B,1,Sq,Sk) full."""
    if mask is None:
        return None, 0
    if mask.dtype != torch.bool and mask.dtype != torch.uint8:
        mask = mask != 0
    m = mask
    while m.dim() > 3:
        m = m.squeeze(1)
    if m.dim() != 3 or m.shape[0] != B or m.shape[-1] != Sk:
        raise RuntimeError(f"attention mask shape {tuple(mask.shape)} incompatible with B={B}, Sq={Sq}, Sk={Sk}")
    if m.shape[1] == 1 and Sq != 1:
        return m.reshape(B, Sk).contiguous().view(torch.uint8), 1
    if m.shape[1] != Sq:
        raise RuntimeError(f"attention mask shape {tuple(mask.shape)} incompatible with Sq={Sq}")
    return m.contiguous().view(torch.uint8), 2


class _MhaCore(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, mask_u8, mask_kind, causal, H, drop_p, salt, rng):
        q = _c(q, "q")
        k = _c(k, "k")
        v = _c(v, "v")
        B, Sq, D = q.shape
        Sk = k.shape[1]
        dk = D // H
        o = torch.empty(B, Sq, D, device=q.device, dtype=_F32)
        nP = LIB.raw("bmhrl_mha_core_probs_floats")(B, H, Sq, Sk)
        P = torch.empty(nP, device=q.device, dtype=_F32)
        LIB.call("bmhrl_mha_core_fwd_f32", _p(q), D, Sq * D, _p(k), D, Sk * D, _p(v), D, Sk * D, _p(mask_u8), mask_kind,
                 causal, _p(o), D, Sq * D, _p(P), B, H, Sq, Sk, dk, drop_p, _p(rng) if drop_p > 0 else None, salt,
                 _stream())
        ctx.save_for_backward(q, k, v, P, mask_u8, rng if drop_p > 0 else None)
        ctx.cfg = (mask_kind, causal, H, drop_p, salt)
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, P, mask_u8, rng = ctx.saved_tensors
        mask_kind, causal, H, drop_p, salt = ctx.cfg
        do = _c(do, "do")
        B, Sq, D = q.shape
        Sk = k.shape[1]
        dk = D // H
        dq = torch.empty_like(q)
        dkk = torch.empty_like(k)
        dv = torch.empty_like(v)
        ws_dp = torch.empty_like(P)
        ws_do = torch.empty_like(do) if drop_p > 0 else None
        LIB.call("bmhrl_mha_core_bwd_f32", _p(do), D, Sq * D, _p(q), D, Sq * D, _p(k), D, Sk * D, _p(v), D, Sk * D, _p(P),
                 _p(mask_u8), mask_kind, causal, _p(dq), D, Sq * D, _p(dkk), D, Sk * D, _p(dv), D, Sk * D, _p(ws_dp),
                 _p(ws_do), B, H, Sq, Sk, dk, drop_p, _p(rng), salt, _stream())
        return dq, dkk, dv, None, None, None, None, None, None, None


def mha_core(q, k, v, mask, H, causal=False, drop_p=0.0, salt=0, rng=None):
    """q (B,Sq,D), k/v (B,Sk,D) projected tensors; returns the head-merged (B,Sq,D) attention output."""
    B, Sq, D = q.shape
    Sk = k.shape[1]
    if D % H != 0:
        raise RuntimeError("mha_core: d_model must be divisible by H")
    m, kind = _mask_args(mask, B, Sq, Sk)
    if kind == 0:
        causal = False  # the reference only builds the tril when a mask is given (multihead_attention.py:16-21)
    if causal and Sq != Sk:
        raise RuntimeError("mha_core: causal attention needs Sq == Sk")
    if drop_p > 0 and rng is None:
        rng = RNG.current(q.device)
    return _MhaCore.apply(q, k, v, m, kind, int(bool(causal)), H, float(drop_p), int(salt), rng)


def mha_core_bf16(q, k, v, mask, H, causal=False, return_lse=False, use_ws=True):
    """bf16 flash-style attention core, forward only (BASELINE.json config 5, the long-clip encoder attention sweep).
    q (B,Sq,D), k/v (B,Sk,D) bf16 projected tensors -> head-merged (B,Sq,D) bf16 (+ fp32 (B,H,Sq) log-sum-exp).
    Same masking semantics as `mha_core` (model/multihead_attention.py:7-31); no dropout, no autograd."""
    for t, name in ((q, "q"), (k, "k"), (v, "v")):
        if not (torch.is_tensor(t) and t.is_cuda and t.dtype == torch.bfloat16 and t.is_contiguous()):
            raise RuntimeError(f"mha_core_bf16: {name} must be a contiguous CUDA bf16 tensor (there is no CPU fallback)")
    if q.requires_grad or k.requires_grad or v.requires_grad:
        if torch.is_grad_enabled():
            raise RuntimeError("mha_core_bf16 is forward-only: call it under torch.no_grad()")
    B, Sq, D = q.shape
    Sk = k.shape[1]
    if D % H != 0:
        raise RuntimeError("mha_core_bf16: d_model must be divisible by H")
    m, kind = _mask_args(mask, B, Sq, Sk)
    if kind == 0:
        causal = False
    if causal and Sq != Sk:
        raise RuntimeError("mha_core_bf16: causal attention needs Sq == Sk")
    o = torch.empty_like(q)
    lse = torch.empty(B, H, Sq, device=q.device, dtype=_F32) if return_lse else None
    ws = None
    if kind == 1 and not causal and use_ws:  # bit-packed mask + per-row trimming of fully masked tail tiles
        ws = torch.empty(LIB.raw("bmhrl_attn_fwd_bf16_ws_bytes")(B, Sk), device=q.device, dtype=torch.uint8)
    LIB.call("bmhrl_attn_fwd_bf16", _p(q), D, Sq * D, _p(k), D, Sk * D, _p(v), D, Sk * D, _p(m), kind,
             int(bool(causal)), _p(o), D, Sq * D, _p(lse), B, H, Sq, Sk, D // H, _p(ws), _stream())
    return (o, lse) if return_lse else o


def mha_core_bf16_bwd(q, k, v, o, lse, do, mask, H):
    """Backward of `mha_core_bf16` (attn_bf16_bwd.cu): (dq, dk, dv) bf16 from the forward's output `o`, its fp32
    (B,H,Sq) log-sum-exp and the output gradient `do`.  mask: None or a (B,1,Sk) key-padding mask.  The probabilities
    are recomputed from q, k and lse; what autograd derives from model/multihead_attention.py:22-30 in the reference."""
    for t, name in ((q, "q"), (k, "k"), (v, "v"), (o, "o"), (do, "do")):
        if not (torch.is_tensor(t) and t.is_cuda and t.dtype == torch.bfloat16 and t.is_contiguous()):
            raise RuntimeError(f"mha_core_bf16_bwd: {name} must be a contiguous CUDA bf16 tensor (there is no CPU fallback)")
    B, Sq, D = q.shape
    Sk = k.shape[1]
    if D % H != 0:
        raise RuntimeError("mha_core_bf16_bwd: d_model must be divisible by H")
    if not (lse.is_cuda and lse.dtype == _F32 and lse.is_contiguous() and tuple(lse.shape) == (B, H, Sq)):
        raise RuntimeError("mha_core_bf16_bwd: lse must be the contiguous fp32 (B,H,Sq) tensor of the forward")
    m, kind = _mask_args(mask, B, Sq, Sk)
    dq, dk_, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    ws = torch.empty(LIB.raw("bmhrl_attn_bwd_bf16_ws_bytes")(B, H, Sq, Sk), device=q.device, dtype=torch.uint8)
    LIB.call("bmhrl_attn_bwd_bf16", _p(q), D, Sq * D, _p(k), D, Sk * D, _p(v), D, Sk * D, _p(o), D, Sq * D,
             _p(do), D, Sq * D, _p(lse), _p(m), kind, _p(dq), D, Sq * D, _p(dk_), D, Sk * D, _p(dv), D, Sk * D,
             B, H, Sq, Sk, D // H, _p(ws), _stream())
    return dq, dk_, dv


class _MhaCoreBf16(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, mask, H):
        with torch.no_grad():
            o, lse = mha_core_bf16(q, k, v, mask, H, return_lse=True)
        ctx.save_for_backward(q, k, v, o, lse)
        ctx.mask, ctx.H = mask, H
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, o, lse = ctx.saved_tensors
        dq, dk_, dv = mha_core_bf16_bwd(q, k, v, o, lse, do.contiguous(), ctx.mask, ctx.H)
        return dq, dk_, dv, None, None


def mha_core_bf16_train(q, k, v, mask, H):
    """Differentiable bf16 attention core: `mha_core_bf16` forward, `mha_core_bf16_bwd` backward (key-padding masks)."""
    return _MhaCoreBf16.apply(q, k, v, mask, H)


# ----------------------------------------------------------------------------------------------------------------
# K4 / K5
# ----------------------------------------------------------------------------------------------------------------
class _LogSoftmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = _c(x, "x")
        V = x.shape[-1]
        y = torch.empty_like(x)
        LIB.call("bmhrl_log_softmax_fwd_f32", _p(x), _p(y), x.numel() // V, V, _stream())
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        dy = _c(dy, "dy")
        V = y.shape[-1]
        dx = torch.empty_like(y)
        LIB.call("bmhrl_log_softmax_bwd_f32", _p(dy), _p(y), _p(dx), y.numel() // V, V, _stream())
        return dx


def log_softmax(x):
    return _LogSoftmax.apply(x)


class _KlLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, trg, samp, offset, mode, smoothing, pad_idx, rowsum):
        pred = _c(pred, "pred")
        V = pred.shape[-1]
        rows = pred.numel() // V
        trg = _c(trg, "target", torch.int64).reshape(-1)
        if mode == 1:
            samp = _c(samp, "biased_trg", torch.int64).reshape(-1)
            offset = _c(offset, "biased_offset").reshape(-1)
        out = None if rowsum else torch.empty(rows, V, device=pred.device, dtype=_F32)
        rs = torch.empty(rows, device=pred.device, dtype=_F32) if rowsum else None
        LIB.call("bmhrl_kl_loss_fwd_f32", mode, _p(pred), _p(trg), _p(samp), _p(offset), _p(out), _p(rs), rows, V,
                 smoothing, pad_idx, _stream())
        ctx.save_for_backward(pred, trg, samp, offset)
        ctx.cfg = (mode, smoothing, pad_idx, rowsum, rows, V)
        return rs if rowsum else out

    @staticmethod
    def backward(ctx, g):
        pred, trg, samp, offset = ctx.saved_tensors
        mode, smoothing, pad_idx, rowsum, rows, V = ctx.cfg
        g = _c(g, "grad")
        dpred = torch.empty_like(pred) if ctx.needs_input_grad[0] else None
        doff = None
        if mode == 1 and ctx.needs_input_grad[3]:
            doff = torch.empty(rows, device=pred.device, dtype=_F32)
        LIB.call("bmhrl_kl_loss_bwd_f32", mode, _p(pred), _p(trg), _p(samp), _p(offset), None if rowsum else _p(g),
                 _p(g) if rowsum else None, _p(dpred), _p(doff), rows, V, smoothing, pad_idx, _stream())
        return dpred, None, None, doff, None, None, None, None


def label_smoothing_kl(pred, target, smoothing, pad_idx, rowsum=False):
    """loss/label_smoothing.py:12-32 -> (B*L, V) element-wise divergence, or (B*L,) row sums with rowsum=True."""
    return _KlLoss.apply(pred, target, None, None, 0, float(smoothing), int(pad_idx), bool(rowsum))


def biased_kl(pred, trg, biased_trg, biased_offset, smoothing, pad_idx, rowsum=False):
    """loss/biased_kl.py:22-53; differentiable in pred AND biased_offset."""
    return _KlLoss.apply(pred, trg, biased_trg, biased_offset.reshape(-1), 1, float(smoothing), int(pad_idx),
                         bool(rowsum))


class _SampleRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, mode, noise, salt, rng):
        pred = _c(pred, "pred")
        V = pred.shape[-1]
        rows = pred.numel() // V
        tok = torch.empty(pred.shape[:-1], device=pred.device, dtype=torch.int64)
        prob = torch.empty(pred.shape[:-1], device=pred.device, dtype=_F32)
        if mode == 1:
            noise = _c(noise, "exp_noise")
            if noise.numel() != pred.numel():
                raise RuntimeError("sample: exp_noise must have pred's number of elements")
        LIB.call("bmhrl_sample_rows_f32", mode, _p(pred), _p(noise) if mode == 1 else None, _p(tok), _p(prob), rows, V,
                 _p(rng) if mode == 2 else None, salt, _stream())
        ctx.save_for_backward(tok, prob)
        ctx.shape = pred.shape
        ctx.mark_non_differentiable(tok)
        return tok, prob

    @staticmethod
    def backward(ctx, _gtok, gprob):
        tok, prob = ctx.saved_tensors
        gprob = _c(gprob, "gprob")
        V = ctx.shape[-1]
        dpred = torch.zeros(ctx.shape, device=prob.device, dtype=_F32)
        LIB.call("bmhrl_scatter_prob_grad_f32", _p(dpred), _p(tok), _p(prob), _p(gprob), prob.numel(), V, _stream())
        return dpred, None, None, None, None


def sample_tokens(pred, mode="argmax", exp_noise=None, salt=0, rng=None):
    """epoch_loops/captioning_bmrl_loops.py:273-285.  mode: 'argmax' (manager), 'noise' (explicit Exp(1) tensor,
    bit-reproducible) or 'philox' (in-kernel noise).  Returns (tokens int64, exp(pred)[tokens] with grad)."""
    m = {"argmax": 0, "noise": 1, "philox": 2}[mode]
    if m == 2 and rng is None:
        rng = RNG.current(pred.device)
    return _SampleRows.apply(pred, m, exp_noise, int(salt), rng)


class _ReinforceGather(torch.autograd.Function):
    @staticmethod
    def forward(ctx, probs, action, eps):
        probs = _c(probs, "probs")
        V = probs.shape[-1]
        action = _c(action, "action", torch.int64).reshape(-1)
        rows = probs.numel() // V
        out = torch.empty(probs.shape[:-1], device=probs.device, dtype=_F32)
        LIB.call("bmhrl_reinforce_gather_fwd_f32", _p(probs), _p(action), _p(out), rows, V, eps, _stream())
        ctx.save_for_backward(probs, action)
        ctx.eps = eps
        return out

    @staticmethod
    def backward(ctx, g):
        probs, action = ctx.saved_tensors
        g = _c(g, "g")
        V = probs.shape[-1]
        dp = torch.zeros_like(probs)
        LIB.call("bmhrl_reinforce_gather_bwd_f32", _p(probs), _p(action), _p(g), _p(dp), probs.numel() // V, V, ctx.eps,
                 _stream())
        return dp, None, None


def reinforce_log_prob(probs, action, eps=1e-5):
    return _ReinforceGather.apply(probs, action, float(eps))


# ----------------------------------------------------------------------------------------------------------------
# K6 / K7
# ----------------------------------------------------------------------------------------------------------------
_PE_CACHE = {}


def positional_table_f64(seq_len, d_model):
    """model/blocks.py:95-103: fp64 table; column j uses exponent j/d_model itself (even j sin, odd j cos)."""
    pos = torch.arange(seq_len, dtype=torch.float64)[:, None]
    j = torch.arange(d_model, dtype=torch.float64)[None, :]
    ang = pos / torch.pow(torch.tensor(10000.0, dtype=torch.float64), j / d_model)
    even = (torch.arange(d_model) % 2 == 0)[None, :]
    return torch.where(even, torch.sin(ang), torch.cos(ang))


def positional_table(seq_len, d_model, device):
    """fp64 table of model/blocks.py:95-103 cast to fp32 (exactly what `.type_as(x)` yields), cached per device."""
    key = (seq_len, d_model, str(device))
    if key not in _PE_CACHE:
        _PE_CACHE[key] = positional_table_f64(seq_len, d_model).to(_F32).to(device).contiguous()
    return _PE_CACHE[key]


class _EmbedPos(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tokens, W, pe, scale, drop_p, salt, rng):
        tokens = _c(tokens, "tokens", torch.int64)
        W = _c(W, "embedding weight")
        B, L = tokens.shape
        D = W.shape[1]
        emb = torch.empty(B, L, D, device=W.device, dtype=_F32)
        pos = torch.empty(B, L, D, device=W.device, dtype=_F32)
        LIB.call("bmhrl_embed_posenc_f32", _p(tokens), _p(W), _p(pe), _p(emb), _p(pos), B * L, L, D, scale, drop_p,
                 _p(rng) if drop_p > 0 else None, salt, _stream())
        ctx.save_for_backward(tokens, rng if drop_p > 0 else None)
        ctx.cfg = (scale, drop_p, salt, W.shape)
        return emb, pos

    @staticmethod
    def backward(ctx, gemb, gpos):
        tokens, rng = ctx.saved_tensors
        scale, drop_p, salt, wshape = ctx.cfg
        if not ctx.needs_input_grad[1]:
            return None, None, None, None, None, None, None
        dW = torch.zeros(wshape, device=tokens.device, dtype=_F32)
        st = _stream()
        rows, D = tokens.numel(), wshape[1]
        if gemb is not None:
            LIB.call("bmhrl_embed_bwd_f32", _p(tokens), _p(_c(gemb, "gemb")), _p(dW), rows, D, scale, st)
        if gpos is not None:
            g = _c(gpos, "gpos")
            if drop_p > 0:
                g2 = torch.empty_like(g)
                LIB.call("bmhrl_dropout_apply_f32", _p(g), _p(g2), g.numel(), drop_p, _p(rng), salt, st)
                g = g2
            LIB.call("bmhrl_embed_bwd_f32", _p(tokens), _p(g), _p(dW), rows, D, scale, st)
        return None, dW, None, None, None, None, None


def embed_posenc(tokens, W, drop_p=0.0, salt=0, rng=None, max_len=3660):
    """(W[tok]*sqrt(D), dropout(W[tok]*sqrt(D) + PE)) -- model/blocks.py:44-48 and :105-112 in one pass."""
    pe = positional_table(max_len, W.shape[1], W.device)
    if tokens.shape[1] > max_len:
        raise RuntimeError("embed_posenc: sequence longer than the positional table")
    if drop_p > 0 and rng is None:
        rng = RNG.current(W.device)
    return _EmbedPos.apply(tokens, W, pe, math.sqrt(W.shape[1]), float(drop_p), int(salt), rng)


class _Embed(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tokens, W, scale):
        tokens = _c(tokens, "tokens", torch.int64)
        W = _c(W, "embedding weight")
        D = W.shape[1]
        emb = torch.empty(tokens.shape + (D,), device=W.device, dtype=_F32)
        LIB.call("bmhrl_embed_posenc_f32", _p(tokens), _p(W), None, _p(emb), None, tokens.numel(), 1, D, scale, 0.0,
                 None, 0, _stream())
        ctx.save_for_backward(tokens)
        ctx.cfg = (scale, W.shape)
        return emb

    @staticmethod
    def backward(ctx, g):
        (tokens,) = ctx.saved_tensors
        scale, wshape = ctx.cfg
        if not ctx.needs_input_grad[1]:
            return None, None, None
        dW = torch.zeros(wshape, device=tokens.device, dtype=_F32)
        LIB.call("bmhrl_embed_bwd_f32", _p(tokens), _p(_c(g, "g")), _p(dW), tokens.numel(), wshape[1], scale, _stream())
        return None, dW, None


def embed(tokens, W, scale):
    """VocabularyEmbedder model/blocks.py:44-48: W[tokens] * scale."""
    return _Embed.apply(tokens, W, float(scale))


class _Concat2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b):
        a = _c(a, "a")
        b = _c(b, "b")
        Da, Db = a.shape[-1], b.shape[-1]
        if a.shape[:-1] != b.shape[:-1]:
            raise RuntimeError("concat_last: leading dimensions differ")
        out = torch.empty(a.shape[:-1] + (Da + Db,), device=a.device, dtype=_F32)
        LIB.call("bmhrl_concat2_f32", _p(a), Da, _p(b), Db, _p(out), a.numel() // Da, _stream())
        ctx.dims = (Da, Db, a.shape, b.shape)
        return out

    @staticmethod
    def backward(ctx, g):
        Da, Db, sa, sb = ctx.dims
        g = _c(g, "g")
        da = torch.empty(sa, device=g.device, dtype=_F32) if ctx.needs_input_grad[0] else None
        db = torch.empty(sb, device=g.device, dtype=_F32) if ctx.needs_input_grad[1] else None
        if da is not None or db is not None:
            LIB.call("bmhrl_split2_f32", _p(g), _p(da), Da, _p(db), Db, g.numel() // (Da + Db), _stream())
        return da, db


def concat_last(a, b):
    """torch.cat([a, b], dim=-1) (model/bm_hrl_agent.py:483)."""
    return _Concat2.apply(a, b)


def reinforce_loss(logp, value, critic_value):
    """loss/biased_kl.py:76-80 on the (B,S) gathered log-probabilities: -mean(adv.detach * log pi) + mean(adv^2).
    (B*S scalars: plain tensor arithmetic, the (B,S,V) gather is the kernel.)"""
    adv = value - critic_value
    return -torch.mean(adv.clone().detach().squeeze() * logp) + torch.mean(torch.pow(adv, 2))


class _AddPosenc(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, x2, pe, drop_p, salt, rng):
        x = _c(x, "x")
        B, S, D = x.shape
        out = torch.empty_like(x)
        if x2 is not None:
            x2 = _c(x2, "x2")
        LIB.call("bmhrl_add_posenc_f32", _p(x), _p(x2), _p(pe), _p(out), B * S, S, D, drop_p,
                 _p(rng) if drop_p > 0 else None, salt, _stream())
        ctx.save_for_backward(rng if drop_p > 0 else None)
        ctx.cfg = (drop_p, salt, x2 is not None)
        return out

    @staticmethod
    def backward(ctx, g):
        (rng,) = ctx.saved_tensors
        drop_p, salt, has2 = ctx.cfg
        g = _c(g, "g")
        if drop_p > 0:
            g2 = torch.empty_like(g)
            LIB.call("bmhrl_dropout_apply_f32", _p(g), _p(g2), g.numel(), drop_p, _p(rng), salt, _stream())
            g = g2
        return g, (g if has2 else None), None, None, None, None


def add_posenc(x, x2=None, drop_p=0.0, salt=0, rng=None, max_len=3660):
    """dropout(x (+ x2) + PE[:S]) -- PositionalEncoder.forward model/blocks.py:105-112 (x2 fuses rgb + flow)."""
    if x.dim() != 3:
        return x  # the reference leaves non-3D inputs untouched (blocks.py:106)
    if x.shape[1] > max_len:
        raise RuntimeError("add_posenc: sequence longer than the positional table")
    pe = positional_table(max_len, x.shape[-1], x.device)
    if drop_p > 0 and rng is None:
        rng = RNG.current(x.device)
    return _AddPosenc.apply(x, x2, pe, float(drop_p), int(salt), rng)


class _GoalExpand(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, seg, noise):
        x = _c(x, "x")
        seg = _c(seg, "segment labels", torch.int32)
        B, L, D = x.shape
        if seg.shape != (B, L):
            raise RuntimeError(f"expand_goals: segment labels must be (B,L)=({B},{L}), got {tuple(seg.shape)}")
        out = torch.empty_like(x)
        gmap = torch.empty(B, L, device=x.device, dtype=torch.int32)
        if noise is not None:
            noise = _c(noise, "noise")
            if noise.numel() != D:
                raise RuntimeError("expand_goals: noise must have d_goal elements")
        LIB.call("bmhrl_goal_expand_fwd_f32", _p(x), _p(noise), _p(seg), _p(gmap), _p(out), B, L, D, _stream())
        ctx.save_for_backward(gmap)
        ctx.shape = (B, L, D)
        return out

    @staticmethod
    def backward(ctx, g):
        (gmap,) = ctx.saved_tensors
        B, L, D = ctx.shape
        g = _c(g, "g")
        dx = torch.empty_like(g)
        LIB.call("bmhrl_goal_expand_bwd_f32", _p(g), _p(gmap), _p(dx), B, L, D, _stream())
        return dx, None, None


def expand_goals(x, seg, noise=None):
    """Manager.expand_goals model/bm_hrl_agent.py:415-429 as one device-side scan (no host loop, no sync), applied
    to x + noise (the detached exploration vector of model/bm_hrl_agent.py:444-452) when noise is given."""
    return _GoalExpand.apply(x, seg, noise)


def exploration_noise(x, d_goal, eps_normal=None, salt=0, rng=None, return_stats=False):
    """model/bm_hrl_agent.py:444-451: the shared d_goal noise vector N(mean/10, std/5) - 0.5*mean/10 (no grad)."""
    x = _c(x.detach(), "x")
    noise = torch.empty(d_goal, device=x.device, dtype=_F32)
    stats = torch.empty(2, device=x.device, dtype=_F32) if return_stats else None
    if eps_normal is None and rng is None:
        rng = RNG.current(x.device)
    if eps_normal is not None:
        eps_normal = _c(eps_normal, "eps_normal")
    LIB.call("bmhrl_exploration_noise_f32", _p(x), x.numel(), d_goal, _p(eps_normal), _p(noise), _p(stats),
             _p(rng), int(salt), _stream())
    return (noise, stats) if return_stats else noise


# ----------------------------------------------------------------------------------------------------------------
# frozen SegmentCritic (forward only)
# ----------------------------------------------------------------------------------------------------------------
def rnn_layer(cell, x, w_ih, w_hh, b_ih, b_hh):
    """One batch_first nn.LSTM ('lstm') / nn.GRU ('gru') layer with zero initial state, no autograd."""
    x = _c(x, "x")
    B, L, _ = x.shape
    H = w_hh.shape[1]
    with torch.no_grad():
        gx = linear(x, w_ih, b_ih)
        out = torch.empty(B, L, H, device=x.device, dtype=_F32)
        c = torch.empty(LIB.raw("bmhrl_rnn_ws_floats")(B, H), device=x.device, dtype=_F32)
        LIB.call("bmhrl_rnn_layer_f32", 0 if cell == "lstm" else 1, _p(gx), _p(_c(w_hh, "w_hh")), _p(_c(b_hh, "b_hh")),
                 _p(out), _p(c), B, L, H, _stream())
    return out


NODE_PRIORITY = False   # train.StepRunner(node_priority=True) sets it while it captures: see launch_priority


class launch_priority:
    """with ops.launch_priority(p): launches of the chain kernels inside carry cudaLaunchAttributePriority p (0 = lowest,
    negative = placed first).  No-op unless ops.NODE_PRIORITY (a graph instantiated with node priorities is being captured)."""

    def __init__(self, priority):
        self.p = int(priority)

    def __enter__(self):
        if NODE_PRIORITY:
            self.prev = _PRIO[0]
            _PRIO[0] = self.p
            LIB.call("bmhrl_set_launch_priority", 1, self.p)
        return self

    def __exit__(self, *exc):
        if NODE_PRIORITY:
            _PRIO[0] = self.prev
            LIB.call("bmhrl_set_launch_priority", 1, self.prev)
        return False


_PRIO = [0]


class adam_grid_limit:
    """with ops.adam_grid_limit(296): opt.step(...) -- Adam launches inside use at most that many CTAs (grid-stride loop over
    the 8192-element chunks).  For an update that runs BESIDE other work: the unlimited grid fills every thread slot of
    every SM and every later launch waits until it has drained (140 us without one kernel start behind the first stage's
    update of the training step); 2 CTAs per SM keep the update near HBM speed and leave room for the step's kernels."""

    def __init__(self, max_ctas):
        self.n = int(max_ctas)

    def __enter__(self):
        self.prev = LIB.raw("bmhrl_adam_get_grid_limit")()
        LIB.call("bmhrl_adam_set_grid_limit", self.n)
        return self

    def __exit__(self, *exc):
        LIB.call("bmhrl_adam_set_grid_limit", self.prev)
        return False


class rnn_units_per_cta:
    """with ops.rnn_units_per_cta(16): ... -- recurrent layers launched inside use at least that many hidden units per
    CTA (H / 16 CTAs instead of one per SM).  For a layer that runs BESIDE other kernels (the frozen critic of the next
    batch under the current training step): a time step is a latency chain, so the layer is barely slower on a quarter
    of the SMs, and the rest of the GPU stays free (measured: 3.01 -> 2.73 ms per warm-start step)."""

    def __init__(self, units):
        self.units = int(units)

    def __enter__(self):
        LIB.call("bmhrl_rnn_set_units_per_cta", self.units)
        return self

    def __exit__(self, *exc):
        LIB.call("bmhrl_rnn_set_units_per_cta", 0)
        return False


def rnn_step(cell, gx, w_hh, b_hh, out, c, t):
    """Step t of an nn.LSTM / nn.GRU layer IN PLACE on caller-kept buffers gx (B,Lmax,NG*H), out (B,Lmax,H) and, for
    the LSTM, c (Lmax,B,H) (the cell state after every step, so a step can be repeated): the incremental form of
    `rnn_layer` for token-by-token decoding."""
    B, L, H = out.shape
    NG = 4 if cell == "lstm" else 3
    if gx.shape != (B, L, NG * H) or not gx.is_contiguous() or not out.is_contiguous():
        raise RuntimeError("rnn_step: gx / out must be contiguous (B,Lmax,NG*H) / (B,Lmax,H) buffers")
    _chk(gx, "gx")
    _chk(out, "out")
    cp = cn = None
    if cell == "lstm":
        if c is None or c.shape != (L, B, H) or not c.is_contiguous():
            raise RuntimeError("rnn_step: the LSTM needs a contiguous (Lmax,B,H) cell-state buffer")
        _chk(c, "c")
        cp, cn = c[max(t - 1, 0)], c[t]
    LIB.call("bmhrl_rnn_step_f32", 0 if cell == "lstm" else 1, _p(gx), _p(_c(w_hh, "w_hh")), _p(_c(b_hh, "b_hh")),
             _p(out), _p(cp), _p(cn), B, L, H, int(t), _stream())


def critic_wavefront(emb, lstm, gru, alpha, beta):
    """All recurrent layers of the SegmentCritic (nn.LSTM stack -> AReLU -> nn.GRU stack, model/bm_hrl_agent.py:204-211)
    in ONE cooperative launch (rnn.cu: critic_wavefront_kernel).  Returns the last layer's (B,L,H) hidden states, or
    None when the grid cannot be co-resident (the caller then runs layer by layer)."""
    import ctypes
    emb = _c(emb, "embedded tokens")
    B, L, E = emb.shape
    H = lstm.hidden_size
    nl, ng = lstm.num_layers, gru.num_layers
    if gru.hidden_size != H or H % 4 or E % 4:
        return None
    names = [(lstm, l) for l in range(nl)] + [(gru, l) for l in range(ng)]
    arr = lambda fmt: (ctypes.c_void_p * (nl + ng))(*[_c(getattr(m, fmt.format(l)), fmt).data_ptr() for m, l in names])
    w_ih, w_hh, b_ih, b_hh = arr("weight_ih_l{}"), arr("weight_hh_l{}"), arr("bias_ih_l{}"), arr("bias_hh_l{}")
    n = LIB.raw("bmhrl_critic_wavefront_ws_floats")(nl + ng, nl, B, L, H)
    ws = torch.empty(n, device=emb.device, dtype=_F32)
    rc = LIB.raw("bmhrl_critic_wavefront_f32")(_p(emb), E, w_ih, w_hh, b_ih, b_hh, nl, ng, _p(_c(alpha, "alpha")),
                                               _p(_c(beta, "beta")), _p(ws), B, L, H, _stream())
    if rc == 1:
        return None
    if rc != 0:
        raise RuntimeError(f"bmhrl_critic_wavefront_f32 failed ({rc}): {LIB.raw('bmhrl_last_error')().decode()}")
    blk = B * L * H
    return ws[(nl + ng - 1) * blk:(nl + ng) * blk].view(B, L, H)


def arelu(x, alpha, beta):
    x = _c(x, "x")
    y = torch.empty_like(x)
    LIB.call("bmhrl_arelu_f32", _p(x), _p(_c(alpha, "alpha")), _p(_c(beta, "beta")), _p(y), x.numel(), _stream())
    return y


def segment_labels(x, w, b, threshold):
    """Linear(D->1) + sigmoid > threshold (model/bm_hrl_agent.py:213,638-640) -> (scores (B,L,1), labels int32 (B,L))."""
    x = _c(x, "x")
    D = x.shape[-1]
    rows = x.numel() // D
    scores = torch.empty(x.shape[:-1] + (1,), device=x.device, dtype=_F32)
    labels = torch.empty(x.shape[:-1], device=x.device, dtype=torch.int32)
    LIB.call("bmhrl_segment_labels_f32", _p(x), _p(_c(w, "w")), _p(_c(b, "b")), float(threshold), _p(scores),
             _p(labels), rows, D, _stream())
    return scores, labels


class _SegmentScan(torch.autograd.Function):
    @staticmethod
    def forward(ctx, probs, expected, seg):
        probs = _c(probs, "probs")
        expected = _c(expected, "expected")
        seg = _c(seg, "segments", torch.int32)
        B, L = probs.shape
        sp = torch.empty_like(probs)
        eo = torch.empty_like(expected)
        LIB.call("bmhrl_segment_scan_fwd_f32", _p(probs), _p(expected), _p(seg), _p(sp), _p(eo), B, L, _stream())
        ctx.save_for_backward(probs, seg)
        ctx.mark_non_differentiable(eo)
        return sp, eo

    @staticmethod
    def backward(ctx, gsp, _geo):
        probs, seg = ctx.saved_tensors
        B, L = probs.shape
        gsp = _c(gsp, "g")
        dp = torch.empty_like(probs)
        LIB.call("bmhrl_segment_scan_bwd_f32", _p(probs), _p(seg), _p(gsp), _p(dp), B, L, _stream())
        return dp, None, None


def segment_scan(probs, expected, seg):
    """Manager branch of biased_kl() (epoch_loops/captioning_bmrl_loops.py:299-317): per-segment product of the
    sampled-token probabilities and per-segment sum of the value baseline."""
    return _SegmentScan.apply(probs, expected, seg)


def make_masks_fused(rgb, audio, cap_in, pad_idx, targets=None):
    """model/masking.py:28-53 (modality audio_video) in one launch: {'V_mask' (B,1,Sv), 'A_mask' (B,1,Sa), 'C_mask'
    (B,L,L)} bool tensors, plus the int64 count of non-pad `targets` (the loss normaliser) when targets are given."""
    for t, nm in ((rgb, "rgb"), (audio, "audio")):
        if not (torch.is_tensor(t) and t.is_cuda and t.dtype == _F32 and t.dim() == 3):
            raise RuntimeError(f"make_masks_fused: `{nm}` must be a CUDA fp32 (B, S, D) tensor (there is no CPU fallback)")
    B, Sv, Sa = rgb.shape[0], rgb.shape[1], audio.shape[1]
    dev = rgb.device
    vm = torch.empty(B, 1, Sv, device=dev, dtype=torch.bool)
    am = torch.empty(B, 1, Sa, device=dev, dtype=torch.bool)
    cm, L = None, 0
    if cap_in is not None:
        cap_in = cap_in.contiguous()
        L = cap_in.shape[-1]
        cm = torch.empty(B, L, L, device=dev, dtype=torch.bool)
    nt = None
    if targets is not None:
        targets = targets.contiguous()
        nt = torch.empty((), device=dev, dtype=torch.int64)
    LIB.call("bmhrl_make_masks", rgb.data_ptr(), rgb.stride(0), rgb.stride(1), Sv, audio.data_ptr(), audio.stride(0),
             audio.stride(1), Sa, _p(cap_in), L, int(pad_idx), B, vm.data_ptr(), am.data_ptr(), _p(cm), _p(targets),
             0 if targets is None else targets.numel(), _p(nt), _stream())
    masks = {"V_mask": vm, "A_mask": am}
    if cm is not None:
        masks["C_mask"] = cm
    return masks, nt


# ----------------------------------------------------------------------------------------------------------------
# optimiser
# ----------------------------------------------------------------------------------------------------------------
class FusedAdam:
    """torch.optim.Adam(lr, betas=(0.9,0.999), eps=1e-8, weight_decay=0) (scripts/train_rl_captioning_module.py:81-83)
    as ONE kernel launch over all currently trainable parameters.  State (moments, one step count per tensor like
    torch) lives on the device and the counters are advanced by a kernel, so `step()` is graph-capturable.  Gradients
    accumulate straight into one persistent flat buffer (`p.grad` are views of it): pointer tables stay valid across
    steps and the flat buffer is what the data-parallel all-reduce buckets slice.

    Phase switching (BMHrlAgent.teach_worker / teach_manager flip requires_grad): call `refresh_active()` afterwards;
    frozen tensors are then skipped exactly like torch skips parameters whose grad is None."""

    def __init__(self, params, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, grad_scale=1.0, peer=None):
        """peer: a peer.PeerGroup (not yet shared) -- the flat gradient buffer and the parameters themselves (re-pointed
        at slices of one flat buffer, values kept) are then carved from its IPC-shareable arenas, which is what
        `step_peer` (gradient sum over the ranks + Adam + parameter broadcast in one kernel) works on."""
        self.params = list(params)
        self.peer = peer
        # step_peer: keep the fp32 master of weights that have a bf16 shadow on the owning rank only (bf16 math mode reads
        # such weights through the shadow); `gather_params()` completes every rank's fp32 copies again
        self.owner_only_master = False
        self._masters_stale = False
        if not self.params:
            raise RuntimeError("FusedAdam: no parameters")
        dev = self.params[0].device
        if dev.type != "cuda":
            raise RuntimeError("FusedAdam: parameters must live on a CUDA device (there is no CPU fallback)")
        self.lr, self.betas, self.eps, self.grad_scale = lr, betas, eps, grad_scale
        # every tensor's slice starts on a 16-byte boundary (128-bit accesses in adam_multi_kernel); the <= 3 padding
        # elements between slices stay zero
        total = sum((p.numel() + 3) & ~3 for p in self.params)
        self.flat_param = None
        if peer is not None:
            self.flat_grad = peer.arena(peer.GRAD, total * 4 + 4096).take(total, _F32)
            self.flat_param = peer.arena(peer.PARAM, total * 4 + 4096).take(total, _F32)
            off = 0
            with torch.no_grad():
                for p in self.params:   # the parameters move into the shared flat buffer (same values, same objects)
                    n = p.numel()
                    fp = self.flat_param[off:off + n].view_as(p)
                    fp.copy_(p.data)
                    p.data = fp
                    off += (n + 3) & ~3
        else:
            self.flat_grad = torch.zeros(total, device=dev, dtype=_F32)
        self.exp_avg = torch.zeros(total, device=dev, dtype=_F32)
        self.exp_avg_sq = torch.zeros(total, device=dev, dtype=_F32)
        self.step_counts = torch.zeros(len(self.params), device=dev, dtype=_F32)
        self.chunk = LIB.raw("bmhrl_adam_chunk_elems")()
        pp, gp, mp, vp, ne = [], [], [], [], []
        off = 0
        self.grad_views, self.offsets = [], []
        for p in self.params:
            n = p.numel()
            gv = self.flat_grad[off:off + n].view_as(p)
            self.grad_views.append(gv)
            self.offsets.append(off)
            pp.append(p.data_ptr())
            gp.append(gv.data_ptr())
            mp.append(self.exp_avg[off:off + n].data_ptr())
            vp.append(self.exp_avg_sq[off:off + n].data_ptr())
            ne.append(n)
            off += (n + 3) & ~3
        i64 = dict(device=dev, dtype=torch.int64)
        self._pp, self._gp = torch.tensor(pp, **i64), torch.tensor(gp, **i64)
        self._mp, self._vp = torch.tensor(mp, **i64), torch.tensor(vp, **i64)
        self._ne = torch.tensor(ne, **i64)
        self._shadow_gen = -1      # generation of the bf16 weight-shadow registry (ops16) the tables below were built for
        self._shadow_tabs = None
        self._shadow_ptrs = []
        self.refresh_active()

    def refresh_active(self):
        dev = self.flat_grad.device
        act, ct, cs = [], [], []
        for i, p in enumerate(self.params):
            if not p.requires_grad:
                p.grad = None
                p._bmhrl_grad = None
                continue
            p.grad = self.grad_views[i]  # autograd accumulates in place into the flat buffer ...
            p._bmhrl_grad = self.grad_views[i]  # ... and the library's backward kernels write into it directly
            act.append(i)
            for s in range(0, p.numel(), self.chunk):
                ct.append(i)
                cs.append(s)
        self.active = act
        self._act = torch.tensor(act, device=dev, dtype=torch.int32)
        self._ct = torch.tensor(ct, device=dev, dtype=torch.int32)
        self._cs = torch.tensor(cs, device=dev, dtype=torch.int32)
        self.n_chunks = len(ct)
        self._subsets = {}  # frozenset of parameter indices -> (act, ct, cs, n_act, n_chunks) device tables
        self._peer_subsets = {}

    def _subset_tables(self, only):
        """Chunk tables of the active parameters whose index is in `only` (built on first use: call once eagerly
        before capturing a CUDA graph -- StepRunner's warm-up steps do)."""
        key = only if isinstance(only, frozenset) else frozenset(only)
        t = self._subsets.get(key)
        if t is None:
            dev = self.flat_grad.device
            act = [i for i in self.active if i in key]
            ct, cs = [], []
            for i in act:
                for s in range(0, self.params[i].numel(), self.chunk):
                    ct.append(i)
                    cs.append(s)
            i32 = dict(device=dev, dtype=torch.int32)
            t = (torch.tensor(act, **i32), torch.tensor(ct, **i32), torch.tensor(cs, **i32), len(act), len(ct))
            self._subsets[key] = t
        return t

    def zero_grad(self, set_to_none=False):
        self.flat_grad.zero_()
        for i in self.active:
            self.params[i].grad = self.grad_views[i]

    def step(self, only=None):
        """One Adam update of every active parameter, or of those whose index is in `only` (the data-parallel trainer
        updates the parameters whose gradients are already reduced while backward is still running)."""
        st = _stream()
        act, ct, cs, n_act, n_chunks = (self._act, self._ct, self._cs, len(self.active), self.n_chunks) \
            if only is None else self._subset_tables(only)
        if n_chunks == 0:
            return
        LIB.call("bmhrl_tick", _p(self.step_counts), _p(act), n_act, None, st)
        sh = self._shadow_tables()
        if sh is None:
            LIB.call("bmhrl_adam_multi_f32", _p(self._pp), _p(self._gp), _p(self._mp), _p(self._vp), _p(self._ne),
                     _p(ct), _p(cs), n_chunks, _p(self.step_counts), self.lr, self.betas[0], self.betas[1],
                     self.eps, self.grad_scale, st)
        else:
            LIB.call("bmhrl_adam_multi_shadow_f32", _p(self._pp), _p(self._gp), _p(self._mp), _p(self._vp), _p(self._ne),
                     _p(ct), _p(cs), n_chunks, _p(self.step_counts), self.lr, self.betas[0], self.betas[1],
                     self.eps, self.grad_scale, _p(sh[0]), _p(sh[1]), _p(sh[2]), _p(sh[3]), st)

    def _peer_tables(self, only):
        """Chunk tables of the chunks THIS rank owns among the active parameters (in `only`): every N-th chunk."""
        key = None if only is None else (only if isinstance(only, frozenset) else frozenset(only))
        t = self._peer_subsets.get(key)
        if t is None:
            dev = self.flat_grad.device
            act = [i for i in self.active if key is None or i in key]
            # the owner of a chunk holds its Adam moments, so ownership must not depend on which subset a step updates:
            # owner = (index of the chunk among ALL chunks of ALL parameters) mod N
            if not hasattr(self, "_chunk_base"):
                self._chunk_base, k = [], 0
                for p in self.params:
                    self._chunk_base.append(k)
                    k += (p.numel() + self.chunk - 1) // self.chunk
            ct, cs = [], []
            for i in act:
                for s in range(0, self.params[i].numel(), self.chunk):
                    if (self._chunk_base[i] + s // self.chunk) % self.peer.world == self.peer.rank:
                        ct.append(i)
                        cs.append(s)
            i32 = dict(device=dev, dtype=torch.int32)
            t = (torch.tensor(act, **i32), torch.tensor(ct, **i32), torch.tensor(cs, **i32), len(act), len(ct))
            self._peer_subsets[key] = t
        return t

    def step_peer(self, slot, only=None):
        """Data-parallel step over NVLink peer memory (csrc/peer.cu): cross-GPU signal "my gradients of these parameters
        are final", wait for everyone's, then for the chunks this rank owns: gradient sum over all ranks (peer loads),
        Adam, new parameter and bf16 shadow stored into every rank's buffers.  Stream-ordered, CUDA-graph capturable.
        The caller ends the step with `peer.barrier(end_slot)` before any rank touches gradients or parameters again."""
        if self._peer_tables(only)[3] == 0:
            return
        self.peer.signal(slot)
        self.step_peer_finish(slot, only)

    def step_peer_finish(self, slot, only=None):
        """step_peer behind an already issued `peer.signal(slot)` (a single-process simulation of several ranks must
        issue every rank's signal before any rank's wait: a kernel that is launched for the first time is loaded lazily,
        which waits for running kernels -- such as another simulated rank's spinning wait)."""
        peer = self.peer
        st = _stream()
        act, ct, cs, n_act, n_chunks = self._peer_tables(only)
        if n_act == 0:
            return
        peer.wait(slot)
        LIB.call("bmhrl_tick", _p(self.step_counts), _p(act), n_act, None, st)
        if n_chunks == 0:
            return
        sh = self._shadow_tables()
        if sh is not None:
            ptrs = [p for p in self._shadow_ptrs if p]
            if not ptrs:
                sh = None   # the registry holds shadows, but none of THIS optimiser's parameters
            else:
                arena = peer.arenas.get(peer.SHADOW)
                if arena is None or not all(arena.contains(p) for p in ptrs):
                    raise RuntimeError("FusedAdam.step_peer: the weight shadows must live in the peer group's shadow arena "
                                       "(ops16.SHADOWS.arena) so that every rank's copy can be written")
        if self.owner_only_master and sh is not None:
            from . import ops16
            if not BF16:
                raise RuntimeError("FusedAdam.step_peer(owner_only_master): only in bf16 math mode (other modes read the fp32 "
                                   "weights, which then exist on their owner only)")
            if any(g["uncovered"] for g in ops16.SHADOWS.groups.values()):
                raise RuntimeError("FusedAdam.step_peer(owner_only_master): a weight sits in two shadow groups (the second "
                                   "is re-cast from the fp32 master on every use): keep the masters replicated")
        go, po = peer.arena(peer.GRAD).offsets(), peer.arena(peer.PARAM).offsets()
        so = peer.arena(peer.SHADOW).offsets() if sh is not None else None
        LIB.call("bmhrl_adam_peer_f32", _p(self._pp), _p(self._gp), _p(self._mp), _p(self._vp), _p(self._ne), _p(ct), _p(cs),
                 n_chunks, _p(self.step_counts), self.lr, self.betas[0], self.betas[1], self.eps, self.grad_scale,
                 None if sh is None else _p(sh[0]), None if sh is None else _p(sh[1]), None if sh is None else _p(sh[2]),
                 None if sh is None else _p(sh[3]), go, po, so, peer.world,
                 1 if (self.owner_only_master and sh is not None) else 0, st)
        if self.owner_only_master and sh is not None:
            self._masters_stale = True

    def gather_params(self):
        """After steps with `owner_only_master`: every owner pushes the fp32 values of its chunks to all ranks, so that
        every rank's parameters are complete again (state_dict / checkpoints, leaving bf16 math mode).  Collective over
        the peer group: every rank calls it; returns when all ranks' pushes have landed."""
        if self.peer is None or not getattr(self, "_masters_stale", False):
            return
        peer = self.peer
        st = _stream()
        act, ct, cs, n_act, n_chunks = self._peer_tables(None)
        peer.barrier(peer.n_slots - 2)      # nobody is still inside a step
        LIB.call("bmhrl_peer_bcast_params_f32", _p(self._pp), _p(self._ne), _p(ct), _p(cs), n_chunks,
                 peer.arena(peer.PARAM).offsets(), peer.world, st)
        peer.barrier(peer.n_slots - 2)
        torch.cuda.current_stream().synchronize()
        self._masters_stale = False

    def _shadow_tables(self):
        """Device tables of the bf16 / fp32 shadows (ops16.SHADOWS) of this optimiser's parameters: the Adam kernel
        writes every updated weight into its shadow as well, so the tensor-core operand copies never go stale and no
        separate conversion pass exists.  Rebuilt (eagerly, never inside a graph capture: the warm-up steps that precede
        a capture create every shadow) when the registry has grown."""
        from . import ops16
        reg = ops16.SHADOWS
        if reg.generation == 0:
            return None
        if reg.generation != self._shadow_gen:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("FusedAdam: weight shadows were created during a CUDA-graph capture; run the step "
                                   "eagerly once before capturing")
            dev = self.flat_grad.device
            ptr, cols, ld, kind = [], [], [], []
            for p in self.params:
                e = reg.entry(p)
                ptr.append(e[0] if e else 0)
                cols.append(e[1] if e else 1)
                ld.append(e[2] if e else 1)
                kind.append(e[3] if e else 0)
            self._shadow_ptrs = list(ptr)
            self._shadow_tabs = (torch.tensor(ptr, device=dev, dtype=torch.int64),
                                 torch.tensor(cols, device=dev, dtype=torch.int32),
                                 torch.tensor(ld, device=dev, dtype=torch.int32),
                                 torch.tensor(kind, device=dev, dtype=torch.int32))
            self._shadow_gen = reg.generation
        return self._shadow_tabs
