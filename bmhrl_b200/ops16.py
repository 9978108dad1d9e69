"""bf16 math mode: Python wrappers over the bf16 entry points of the C ABI (include/bmhrl_b200.h).

Data flow of this mode: the residual stream, LayerNorm statistics, accumulators, gradients of
parameters and optimiser state stay fp32; every tensor that is ONLY a tensor-core operand (LayerNorm outputs, Q / K / V,
attention outputs, FFN hidden activations, dZ of every linear layer, weight shadows) is bf16, emitted by the epilogue of
the kernel that produces it.  bf16 matrices are 2-D views with unit column stride and a row pitch that is a multiple of
8 elements (TMA needs 16-byte pitches: 300 -> 304, 364 -> 368, 10172 -> 10176).  Nothing here computes on the CPU.
"""
import ctypes

import torch

from ._lib import LIB

_F32, _BF16 = torch.float32, torch.bfloat16


class GemmDesc(ctypes.Structure):
    """Mirror of `bmhrl_gemm_desc` (include/bmhrl_b200.h); checked against bmhrl_gemm_desc_bytes() at first use."""
    _fields_ = [("A", ctypes.c_void_p), ("lda", ctypes.c_longlong), ("a_mn", ctypes.c_int),
                ("B", ctypes.c_void_p), ("ldb", ctypes.c_longlong), ("b_mn", ctypes.c_int),
                ("M", ctypes.c_int), ("N", ctypes.c_int), ("K", ctypes.c_int), ("alpha", ctypes.c_float),
                ("bias", ctypes.c_void_p), ("act", ctypes.c_int),
                ("gate", ctypes.c_void_p), ("ldg", ctypes.c_longlong), ("gate_scale", ctypes.c_float),
                ("drop_p", ctypes.c_float), ("rng_state", ctypes.c_void_p), ("salt", ctypes.c_uint),
                ("resid", ctypes.c_void_p), ("ldr", ctypes.c_longlong), ("beta", ctypes.c_float),
                ("C32", ctypes.c_void_p), ("ldc32", ctypes.c_longlong),
                ("C16", ctypes.c_void_p), ("ldc16", ctypes.c_longlong),
                ("seg_rows", ctypes.c_int), ("C32_seg1", ctypes.c_void_p), ("C32_seg2", ctypes.c_void_p),
                ("b_static", ctypes.c_int)]


_DESC_CHECKED = [False]
RECORD = None  # bench.py: when a list, every LAUNCH appends the list of its problems' (M, N, K, a_mn, b_mn)


def _stream():
    return torch.cuda.current_stream().cuda_stream


def pad8(n):
    return (int(n) + 7) & ~7


def empty16(rows, cols, device):
    """(rows, cols) bf16 matrix with a TMA-legal row pitch."""
    return torch.empty(rows, pad8(cols), device=device, dtype=_BF16)[:, :cols]


def zeros16(rows, cols, device):
    return torch.zeros(rows, pad8(cols), device=device, dtype=_BF16)[:, :cols]


def _mat16(t, name):
    if not (torch.is_tensor(t) and t.is_cuda and t.dtype == _BF16 and t.dim() == 2 and t.stride(1) == 1 and
            t.stride(0) % 8 == 0 and t.data_ptr() % 16 == 0):
        raise RuntimeError(f"bmhrl_b200 (bf16 mode): `{name}` must be a 2-D CUDA bf16 matrix with unit column stride, a "
                           f"row pitch that is a multiple of 8 and a 16-byte aligned base (there is no CPU fallback)")
    return t


def _f32mat(t, name, rows, cols):
    if t is None:
        return None
    if not (t.is_cuda and t.dtype == _F32 and t.stride(-1) == 1 and t.numel() == rows * cols):
        raise RuntimeError(f"bmhrl_b200 (bf16 mode): `{name}` must be a CUDA fp32 ({rows},{cols}) matrix")
    return t


def gemm_problem(A, a_mn, B, b_mn, M, N, K, *, alpha=1.0, bias=None, act=0, gate=None, gate_scale=1.0, drop_p=0.0,
                 rng=None, salt=0, resid=None, beta=0.0, C32=None, C16=None, seg=None, b_static=False):
    """One `bmhrl_gemm_desc`.  A / B: bf16 matrices as stored (see the header for the two majors); C32 (fp32, 2-D,
    contiguous rows) / C16 (bf16 matrix) outputs; seg = (rows_per_segment, [seg1, seg2]) splits C32 by rows."""
    _mat16(A, "A")
    _mat16(B, "B")
    d = GemmDesc()
    d.A, d.lda, d.a_mn = A.data_ptr(), A.stride(0), int(a_mn)
    d.B, d.ldb, d.b_mn = B.data_ptr(), B.stride(0), int(b_mn)
    d.M, d.N, d.K, d.alpha = int(M), int(N), int(K), float(alpha)
    d.bias = None if bias is None else bias.data_ptr()
    d.act = int(act)
    if gate is not None:
        _mat16(gate, "gate")
        d.gate, d.ldg, d.gate_scale = gate.data_ptr(), gate.stride(0), float(gate_scale)
    d.drop_p = float(drop_p)
    if drop_p > 0:
        d.rng_state, d.salt = rng.data_ptr(), int(salt)
    if resid is not None:
        d.resid, d.ldr = resid.data_ptr(), resid.stride(-2) if resid.dim() >= 2 else N
    d.beta = float(beta)
    if C32 is not None:
        d.C32, d.ldc32 = C32.data_ptr(), (C32.stride(-2) if C32.dim() >= 2 else N)
    if C16 is not None:
        _mat16(C16, "C16")
        d.C16, d.ldc16 = C16.data_ptr(), C16.stride(0)
    if seg is not None:
        rows, others = seg
        d.seg_rows = int(rows)
        d.C32_seg1 = others[0].data_ptr()
        d.C32_seg2 = others[1].data_ptr() if len(others) > 1 else None
    if b_static and getattr(B, "_bmhrl_refreshed", False):
        b_static = False              # re-cast on this stream a moment ago: this first use must wait for that kernel
        B._bmhrl_refreshed = False
    d.b_static = 1 if b_static else 0   # B is a weight shadow: may be fetched before the grid dependency resolves (PDL)
    d._keep = (A, B, bias, gate, rng, resid, C32, C16, seg)  # operands stay alive until the call returns
    return d


def gemm_group(problems):
    """1..4 problems -> ONE launch of gemm_bf16_tc_kernel on the current stream."""
    if not _DESC_CHECKED[0]:
        if LIB.raw("bmhrl_gemm_desc_bytes")() != ctypes.sizeof(GemmDesc):
            raise RuntimeError("bmhrl_b200: GemmDesc does not match bmhrl_gemm_desc of the loaded library")
        _DESC_CHECKED[0] = True
    n = len(problems)
    if RECORD is not None:
        RECORD.append([(p.M, p.N, p.K, p.a_mn, p.b_mn) for p in problems])
    arr = (GemmDesc * n)(*problems)
    LIB.call("bmhrl_gemm_bf16_group", ctypes.cast(arr, ctypes.c_void_p), n, _stream())


# ----------------------------------------------------------------------------------------------------------------
# one-tile fused attention (attn16.cu)
# ----------------------------------------------------------------------------------------------------------------
def attn16_supported(Sq, Sk, dk):
    return (Sq <= 128 and Sk <= 128 and dk in (64, 128, 256, 512) and
            LIB.raw("bmhrl_attn16_smem_bytes")(1, Sq, Sk, dk) <= 220 * 1024)


def attn16_fwd(q, k, v, mask_u8, mask_kind, causal, B, H, Sq, Sk, dk, o, lse2, drop_p=0.0, rng=None, salt=0):
    """q (B*Sq, .), k / v (B*Sk, .) bf16 matrices (column slices of the projection output are fine) -> o (B*Sq, H*dk)
    bf16 and lse2 (B,H,Sq) fp32, both preallocated."""
    for t, n in ((q, "q"), (k, "k"), (v, "v"), (o, "o")):
        _mat16(t, n)
    LIB.call("bmhrl_attn16_fwd", q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0), v.data_ptr(), v.stride(0),
             None if mask_u8 is None else mask_u8.data_ptr(), int(mask_kind), int(bool(causal)), o.data_ptr(), o.stride(0),
             lse2.data_ptr(), B, H, Sq, Sk, dk, float(drop_p), rng.data_ptr() if drop_p > 0 else None, int(salt), _stream())


def attn16_bwd(q, k, v, do, lse2, mask_u8, mask_kind, causal, B, H, Sq, Sk, dk, dq, dk_, dv):
    """dq / dk_ / dv: preallocated bf16 matrices (or column slices of one N-concatenated dZ buffer)."""
    for t, n in ((q, "q"), (k, "k"), (v, "v"), (do, "do"), (dq, "dq"), (dk_, "dk"), (dv, "dv")):
        _mat16(t, n)
    LIB.call("bmhrl_attn16_bwd", q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0), v.data_ptr(), v.stride(0),
             do.data_ptr(), do.stride(0), lse2.data_ptr(), None if mask_u8 is None else mask_u8.data_ptr(),
             int(mask_kind), int(bool(causal)), dq.data_ptr(), dq.stride(0), dk_.data_ptr(), dk_.stride(0), dv.data_ptr(),
             dv.stride(0), B, H, Sq, Sk, dk, _stream())


# ----------------------------------------------------------------------------------------------------------------
# bf16 twins of fp32 activations and bf16 shadows of fp32 weights
# ----------------------------------------------------------------------------------------------------------------
def cast16(x2d, out=None, drop_p=0.0, rng=None, salt=0, db=None, accumulate_db=False):
    """bf16(dropout(x2d)) for a contiguous-row fp32 (rows, cols) matrix; db (fp32 (cols,)): bias gradient, fixed order."""
    rows, cols = x2d.shape
    if out is None:
        out = empty16(rows, cols, x2d.device)
    LIB.call("bmhrl_cast_bf16", x2d.data_ptr(), x2d.stride(0), out.data_ptr(), out.stride(0), rows, cols, float(drop_p),
             rng.data_ptr() if drop_p > 0 else None, int(salt), None if db is None else db.data_ptr(),
             1 if accumulate_db else 0, _stream())
    return out


def twin16(t):
    """The bf16 (rows, D) matrix of an fp32 (..., D) activation.  Kernels that produce an activation which a later GEMM
    consumes emit the twin in their epilogue and attach it (`attach_twin`); otherwise it is cast here, once."""
    tw = getattr(t, "_bmhrl16", None)
    if tw is not None and tw[1] == t._version:
        if tw[2] is not None:
            # cast lazily by ANOTHER consumer, possibly on another stream (the worker / manager fusion stacks share the
            # encoder memory): order this stream behind that cast
            torch.cuda.current_stream(t.device).wait_event(tw[2])
        return tw[0]
    x = t.detach()
    if not x.is_contiguous():
        x = x.contiguous()
    y = cast16(x.reshape(-1, x.shape[-1]))
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(t.device))
    t._bmhrl16 = (y, t._version, ev)
    return y


def attach_twin(t, y16):
    """y16 was written by the SAME kernel as t: whoever is ordered behind t is ordered behind y16 (no event needed)."""
    t._bmhrl16 = (y16, t._version, None)


class _Shadows:
    """bf16 operand copies of fp32 weights (and fp32 N-concatenated biases).  A group = the weights of linear layers
    that read the same input, stacked along N (Q|K|V of a self-attention, K|V of a cross-attention): ONE GEMM.
    Freshness: ops.FusedAdam writes the shadow of every weight it updates from inside its kernel (bmhrl_adam_multi_
    shadow_f32), which is what keeps a captured CUDA graph self-contained; any other in-place change of a parameter
    (load_state_dict, torch.optim.Adam, .copy_) bumps its `_version` and the group is re-cast on its next use."""

    def __init__(self):
        self.groups = {}      # tuple(id(weight)) -> dict
        self.by_param = {}    # id(param) -> (data_ptr, cols, ld, kind) of the FIRST group that holds it
        self.generation = 0
        # peer.Arena the shadows are carved from (data parallel over peer memory: FusedAdam.step_peer stores the
        # updated shadows into every rank's arena, so all ranks must create them in the same order); None: torch memory
        self.arena = None

    def entry(self, p):
        return self.by_param.get(id(p))

    def clear(self):
        self.groups.clear()
        self.by_param.clear()
        self.generation += 1

    def get(self, weights, biases=None):
        key = tuple(id(w) for w in weights)
        g = self.groups.get(key)
        if g is None:
            g = self._create(key, weights, biases)
        ver = sum(w._version for w in weights) + (sum(b._version for b in biases if b is not None) if biases else 0)
        if ver != g["ver"] or g["uncovered"]:
            self._refresh(g, weights, biases)
            g["ver"] = ver
        return g["w16"], g["b32"]

    def _create(self, key, weights, biases):
        import weakref
        dev = weights[0].device
        K = weights[0].shape[1]
        if any(w.dim() != 2 or w.shape[1] != K for w in weights):
            raise RuntimeError("weight shadows: the stacked weights must share in_features")
        n_tot = sum(w.shape[0] for w in weights)
        if self.arena is not None:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("weight shadows (peer arena) must exist before a CUDA-graph capture: run the step eagerly once")
            w16 = self.arena.take(n_tot * pad8(K), _BF16).view(n_tot, pad8(K))[:, :K]
        else:
            w16 = empty16(n_tot, K, dev)
        b32 = None
        if biases is not None and any(b is not None for b in biases):
            if len(weights) == 1:
                b32 = biases[0]          # a single layer reads its own bias parameter
            elif self.arena is not None:
                b32 = self.arena.take(n_tot, _F32)
            else:
                b32 = torch.zeros(n_tot, device=dev, dtype=_F32)
        g = {"w16": w16, "b32": b32, "ver": -1, "uncovered": False}
        r0 = 0
        for i, w in enumerate(weights):
            n = w.shape[0]
            if id(w) in self.by_param:
                g["uncovered"] = True    # Adam keeps only the first shadow of a parameter fresh: re-cast this one per use
            else:
                self.by_param[id(w)] = (w16[r0:r0 + n].data_ptr(), K, w16.stride(0), 0)
            if len(weights) > 1 and biases is not None and biases[i] is not None:
                if id(biases[i]) in self.by_param:
                    g["uncovered"] = True
                else:
                    self.by_param[id(biases[i])] = (b32[r0:r0 + n].data_ptr(), n, n, 1)
            r0 += n
        self.groups[key] = g
        self.generation += 1
        ids = [id(w) for w in weights] + [id(b) for b in (biases or []) if b is not None]

        def _drop(reg=weakref.ref(self), key=key, ids=ids, w16ptr=w16.data_ptr()):
            reg = reg()
            if reg is None:
                return
            reg.groups.pop(key, None)
            for i in ids:
                reg.by_param.pop(i, None)
            reg.generation += 1
        weakref.finalize(weights[0], _drop)
        return g

    def _refresh(self, g, weights, biases):
        g["w16"]._bmhrl_refreshed = True   # see gemm_problem: no early (pre-dependency) fetch of a shadow just re-cast
        r0 = 0
        for i, w in enumerate(weights):
            n = w.shape[0]
            wd = w.detach()
            LIB.call("bmhrl_cast_bf16", wd.data_ptr(), wd.stride(0), g["w16"][r0:r0 + n].data_ptr(), g["w16"].stride(0), n,
                     w.shape[1], 0.0, None, 0, None, 0, _stream())
            if len(weights) > 1 and biases is not None and biases[i] is not None:
                g["b32"][r0:r0 + n].copy_(biases[i].detach())
            r0 += n


SHADOWS = _Shadows()


# ----------------------------------------------------------------------------------------------------------------
# raw (non-autograd) building blocks
# ----------------------------------------------------------------------------------------------------------------
def ln_fwd16(x2d, gamma, beta, eps):
    rows, D = x2d.shape
    y = empty16(rows, D, x2d.device)
    mean = torch.empty(rows, device=x2d.device, dtype=_F32)
    rstd = torch.empty(rows, device=x2d.device, dtype=_F32)
    LIB.call("bmhrl_layernorm_fwd_bf16", x2d.data_ptr(), gamma.data_ptr(), beta.data_ptr(), y.data_ptr(), y.stride(0),
             mean.data_ptr(), rstd.data_ptr(), rows, D, float(eps), _stream())
    return y, mean, rstd


def ln_bwd(dy2d, x2d, gamma, mean, rstd, dres2d, need_dx, ggrad, bgrad, need_affine, twin=None):
    """dx = LN'(dy) (+ dres); dgamma / dbeta into the flat-buffer slices (accumulating) or fresh tensors.  Deterministic.
    When the affine gradients go straight into a FusedAdam slice nothing downstream waits for them: they are computed
    by a second launch on a weight-gradient side stream, and the critical path only carries the dx rows."""
    from . import ops
    rows, D = x2d.shape
    dx = torch.empty_like(x2d) if need_dx else None
    acc = need_affine and ggrad is not None and bgrad is not None
    dg = (ggrad if acc else torch.empty(D, device=x2d.device, dtype=_F32)) if need_affine else None
    db = (bgrad if acc else torch.empty(D, device=x2d.device, dtype=_F32)) if need_affine else None

    def call(dxp, dgp, dbp):
        LIB.call("bmhrl_layernorm_bwd_f32", dy2d.data_ptr(), x2d.data_ptr(), gamma.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                 None if dres2d is None else dres2d.data_ptr(), None if dxp is None else dxp.data_ptr(),
                 None if dgp is None else dgp.data_ptr(), None if dbp is None else dbp.data_ptr(), rows, D, 1 if acc else 0,
                 _stream())
    dx16 = None
    if twin is not None and need_dx and D % 4 == 0:
        # the dx rows by the variant that also writes bf16(dropout(dx)) for the producer of x (see hint_backward_cast)
        p, rng, salt = twin
        dx16 = empty16(rows, D, x2d.device)
        ops._WGRAD_KEEP.append(dx16)   # read by another node, possibly on another stream: alive until the step ends

        def call_dx():
            LIB.call("bmhrl_layernorm_bwd_twin_f32", dy2d.data_ptr(), x2d.data_ptr(), gamma.data_ptr(), mean.data_ptr(),
                     rstd.data_ptr(), None if dres2d is None else dres2d.data_ptr(), dx.data_ptr(), dx16.data_ptr(),
                     dx16.stride(0), float(p), rng.data_ptr() if p > 0 else None, int(salt), rows, D, _stream())
    else:
        def call_dx():
            call(dx, None, None)
    if acc and ops.CONCURRENT:
        ops._WGRAD_KEEP.append((dy2d, x2d, mean, rstd))
        with torch.cuda.stream(ops.wgrad_stream(x2d.device, dg)), ops.launch_priority(0):
            call(None, dg, db)
        if need_dx:
            call_dx()
    elif dx16 is not None:
        if need_affine:
            call(None, dg, db)
        call_dx()
    else:
        call(dx, dg, db)
    if dx16 is not None:
        dx._bmhrl16m_pending = (dx16, float(twin[0]), int(twin[2]))
    return dx, (None if (acc or not need_affine) else dg), (None if (acc or not need_affine) else db)


def hint_backward_cast(out, drop_p, rng, salt):
    """Marks a block's fp32 output: the block's backward will start from bf16(dropout-mask o d_out) with these dropout
    parameters.  A block that consumes `out` as its residual-stream input reads the hint and makes its LayerNorm backward
    emit that matrix next to dx (`grad16_of` finds it) -- one launch fewer at the head of every block's backward."""
    out._bmhrl_bwd_cast = (float(drop_p), rng, int(salt))


def grad16_of(dy, rows, cols, drop_p, salt):
    """The masked bf16 copy of the incoming gradient `dy`, if the node that produced dy emitted one for exactly these
    dropout parameters.  The attribute travels with the tensor object through autograd as long as the gradient has a single
    producer (a summed gradient is a new tensor without it); the version check rejects an in-place accumulation."""
    tw = getattr(dy, "_bmhrl16m", None)
    if tw is None:
        return None
    t16, p, sl, ver = tw
    if p == float(drop_p) and sl == int(salt) and ver == dy._version and tuple(t16.shape) == (rows, cols):
        return t16
    return None


def colsum16(z16, outs, seg_cols=0, accumulate=False):
    rows, cols = z16.shape
    o = list(outs) + [None] * (3 - len(outs))
    LIB.call("bmhrl_colsum_bf16", z16.data_ptr(), z16.stride(0), rows, cols, o[0].data_ptr(),
             None if o[1] is None else o[1].data_ptr(), None if o[2] is None else o[2].data_ptr(), int(seg_cols),
             1 if accumulate else 0, _stream())


def _wgrad(dz16, x16, weights, n_each):
    """dW (+)= dZ^T X for the stacked weights of a group, off the critical path (wgrad streams); returns the list of
    gradient tensors autograd still has to see (None where the kernel accumulated into a FusedAdam slice)."""
    from . import ops
    live = [w for w in weights if w.requires_grad]
    if not live:
        return [None] * len(weights)
    M = dz16.shape[0]
    K = x16.shape[1]
    dsts = [ops._gbuf(w) if w.requires_grad else None for w in weights]
    outs = [None] * len(weights)
    all_flat = all(d is not None for d in dsts) and len(live) == len(weights)
    if all_flat:
        prob = gemm_problem(dz16, 1, x16, 1, dz16.shape[1], K, M, C32=dsts[0], beta=1.0,
                            seg=(n_each, dsts[1:]) if len(weights) > 1 else None)
        if ops.CONCURRENT:
            ops._WGRAD_KEEP.append((dz16, x16))
            with torch.cuda.stream(ops.wgrad_stream(dz16.device, dsts[0])), ops.launch_priority(0):
                gemm_group([prob])
        else:
            gemm_group([prob])
        return outs
    # mixed / no flat buffer: one problem per live weight, fresh or accumulating destinations
    probs = []
    r0 = 0
    for i, w in enumerate(weights):
        n = w.shape[0]
        if w.requires_grad:
            if dsts[i] is not None:
                probs.append(gemm_problem(dz16[:, r0:r0 + n], 1, x16, 1, n, K, M, C32=dsts[i], beta=1.0))
            else:
                outs[i] = torch.empty(n, K, device=dz16.device, dtype=_F32)
                probs.append(gemm_problem(dz16[:, r0:r0 + n], 1, x16, 1, n, K, M, C32=outs[i]))
        r0 += n
    gemm_group(probs)
    return outs


_ONES16 = {}   # device index -> (1, L) bf16 row of ones: the B operand that turns a bias gradient into a GEMM problem


def _ones_row(device, n):
    dev = torch.device(device)
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    t = _ONES16.get(idx)
    if t is None or t.shape[1] < n:
        if torch.cuda.is_current_stream_capturing():
            return None     # never allocate a process-wide buffer from inside a graph's private pool
        t = torch.ones(1, pad8(max(n, 4096)), device=dev, dtype=_BF16)
        _ONES16[idx] = t
    return t


FUSE_BGRAD = True   # bias gradients as one more problem of the weight-gradient launch (switch for experiments / tests)


def _wbgrad(dz16, x16, weights, biases, n_each):
    """Weight AND bias gradients of the (stacked) linear layers that produced dz16 from x16: (gw, gb), same convention as
    _wgrad / _bgrad16.  When everything accumulates into FusedAdam slices the bias gradient db = dZ^T 1 rides along as a
    second problem of the SAME tensor-core launch (A = dZ^T as for dW, B = a row of ones, N = 1): no column-sum kernel,
    fixed summation order (one CTA per 128 outputs)."""
    from . import ops
    M = dz16.shape[0]
    K = x16.shape[1]
    ok = (FUSE_BGRAD and biases is not None and len(biases) == len(weights)
          and all(w.requires_grad and ops._gbuf(w) is not None for w in weights)
          and all(b is not None and b.requires_grad and ops._gbuf(b) is not None for b in biases))
    ones = _ones_row(dz16.device, M) if ok else None
    if ones is None:
        return _wgrad(dz16, x16, weights, n_each), (_bgrad16(dz16, biases, n_each) if biases is not None else [None] * len(weights))
    wd = [ops._gbuf(w) for w in weights]
    bd = [ops._gbuf(b) for b in biases]
    n_tot = dz16.shape[1]
    probs = [gemm_problem(dz16, 1, x16, 1, n_tot, K, M, C32=wd[0], beta=1.0, seg=(n_each, wd[1:]) if len(weights) > 1 else None),
             gemm_problem(dz16, 1, ones, 0, n_tot, 1, M, C32=bd[0], beta=1.0, seg=(n_each, bd[1:]) if len(biases) > 1 else None)]
    if ops.CONCURRENT:
        ops._WGRAD_KEEP.append((dz16, x16))
        with torch.cuda.stream(ops.wgrad_stream(dz16.device, wd[0])), ops.launch_priority(0):
            gemm_group(probs)
    else:
        gemm_group(probs)
    return [None] * len(weights), [None] * len(biases)


def _bgrad16(dz16, biases, n_each):
    """bias gradients = column sums of a bf16 dZ (deterministic); same return convention as _wgrad."""
    from . import ops
    outs = [None] * len(biases)
    live = [i for i, b in enumerate(biases) if b is not None and b.requires_grad]
    if not live:
        return outs
    dsts = []
    acc = all(ops._gbuf(biases[i]) is not None for i in live) and len(live) == len(biases)
    if acc:
        dsts = [ops._gbuf(b) for b in biases]
        if ops.CONCURRENT:   # only the optimiser reads these: off the critical path
            ops._WGRAD_KEEP.append((dz16,))
            with torch.cuda.stream(ops.wgrad_stream(dz16.device, dsts[0])), ops.launch_priority(0):
                colsum16(dz16, dsts, seg_cols=n_each if len(biases) > 1 else 0, accumulate=True)
        else:
            colsum16(dz16, dsts, seg_cols=n_each if len(biases) > 1 else 0, accumulate=True)
        return outs
    r0 = 0
    for i, b in enumerate(biases):
        n = n_each
        if b is not None and b.requires_grad:
            g = ops._gbuf(b)
            if g is not None:
                colsum16(dz16[:, r0:r0 + n], [g], accumulate=True)
            else:
                outs[i] = torch.empty(n, device=dz16.device, dtype=_F32)
                colsum16(dz16[:, r0:r0 + n], [outs[i]])
        r0 += n
    return outs


# ----------------------------------------------------------------------------------------------------------------
# block-level autograd nodes: one node per ResidualConnection block, every interior tensor is bf16 and invisible to
# autograd; the residual-stream gradient adds are fused (LayerNorm backward takes the skip gradient as `dres`)
# ----------------------------------------------------------------------------------------------------------------
def _finish_dx(dx2d, shape):
    """The (B, S, D) gradient a block returns; a masked bf16 copy emitted by ln_bwd moves onto the returned object."""
    out = dx2d.view(shape)
    pend = getattr(dx2d, "_bmhrl16m_pending", None)
    if pend is not None:
        out._bmhrl16m = (pend[0], pend[1], pend[2], out._version)
    return out


class _AttnBlock(torch.autograd.Function):
    """y = [x +] dropout_r( W_o dropout_a( attention( W_q n, W_k m, W_v m ) ) + b_o ),  n = LayerNorm(x) or x,
    m = n (self-attention) or the fp32 tensor `kv` (cross-attention: the un-normed other modality / encoder memory).
    model/blocks.py:135-144 + model/multihead_attention.py:60-92 as ONE autograd node:
      forward  LN -> Q|K|V GEMM (one N-concatenated problem; cross: Q and K|V as two problems of one launch) ->
               fused attention -> out-projection GEMM with bias + dropout + residual in its epilogue
      backward cast (+ dropout mask, + db_o) -> dO GEMM (attention dropout mask in its epilogue) | dW_o -> fused attention
               backward -> dX GEMM(s) | dW, db -> LayerNorm backward with the skip gradient folded in."""

    @staticmethod
    def forward(ctx, x, kv, mask_u8, cfg, ln_g, ln_b, Wq, bq, Wk, bk, Wv, bv, Wo, bo):
        from . import ops
        B, Sq, Dq = x.shape
        M = B * Sq
        D = Wq.shape[0]
        H = cfg["H"]
        dk = D // H
        x2 = x.detach().reshape(M, Dq)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        dev = x.device
        if ln_g is not None:
            n16, mean, rstd = ln_fwd16(x2, ln_g.detach(), ln_b.detach(), cfg["eps"])
        else:
            n16, mean, rstd = twin16(x), None, None
        if kv is None:
            Sk, Mk = Sq, M
            w16, b32 = SHADOWS.get([Wq, Wk, Wv], [bq, bk, bv])
            qkv = empty16(M, 3 * D, dev)
            gemm_group([gemm_problem(n16, 0, w16, 0, M, 3 * D, Dq, b_static=True, bias=b32, C16=qkv)])
            q, k, v = qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:]
            m16 = None
        else:
            Sk = kv.shape[1]
            Mk = B * Sk
            Dkv = kv.shape[2]
            m16 = twin16(kv)
            wq16, _ = SHADOWS.get([Wq], [bq])
            wkv16, bkv32 = SHADOWS.get([Wk, Wv], [bk, bv])
            q = empty16(M, D, dev)
            kvp = empty16(Mk, 2 * D, dev)
            gemm_group([gemm_problem(n16, 0, wq16, 0, M, D, Dq, b_static=True, bias=None if bq is None else bq.detach(), C16=q),
                        gemm_problem(m16, 0, wkv16, 0, Mk, 2 * D, Dkv, b_static=True, bias=bkv32, C16=kvp)])
            k, v = kvp[:, :D], kvp[:, D:]
            qkv = None
        o16 = empty16(M, D, dev)
        lse2 = torch.empty(B, H, Sq, 2, device=dev, dtype=_F32)
        p_a = cfg["p_attn"]
        attn16_fwd(q, k, v, mask_u8, cfg["mask_kind"], cfg["causal"], B, H, Sq, Sk, dk, o16, lse2, p_a, cfg["rng"],
                   cfg["s_core"])
        wo16, _ = SHADOWS.get([Wo], [bo])
        out = torch.empty(B, Sq, Dq, device=dev, dtype=_F32)
        out16 = empty16(M, Dq, dev) if cfg.get("emit_twin") else None
        gemm_group([gemm_problem(o16, 0, wo16, 0, M, Dq, D, b_static=True, bias=None if bo is None else bo.detach(), drop_p=cfg["p_res"],
                                 rng=cfg["rng"], salt=cfg["s_out"], resid=x2 if cfg["resid"] else None,
                                 C32=out.view(M, Dq), C16=out16)])
        if out16 is not None:
            attach_twin(out, out16)
        hint_backward_cast(out, cfg["p_res"], cfg["rng"], cfg["s_out"])
        ctx.in_hint = getattr(x, "_bmhrl_bwd_cast", None) if ln_g is not None else None
        ctx.save_for_backward(x2, mean, rstd, n16, q, k, v, m16, o16, lse2, mask_u8, ln_g, Wq, Wk, Wv, Wo)
        ctx.params = (ln_g, ln_b, Wq, bq, Wk, bk, Wv, bv, Wo, bo)
        ctx.cfg = cfg
        ctx.dims = (B, Sq, Sk, Dq, D, H, dk, kv is None, None if kv is None else kv.shape[2])
        ctx.bufs = (qkv,)
        return out

    @staticmethod
    def backward(ctx, dy):
        from . import ops
        x2, mean, rstd, n16, q, k, v, m16, o16, lse2, mask_u8, ln_g_s, *_ = ctx.saved_tensors
        ln_g, ln_b, Wq, bq, Wk, bk, Wv, bv, Wo, bo = ctx.params
        cfg = ctx.cfg
        B, Sq, Sk, Dq, D, H, dk, self_att, Dkv = ctx.dims
        M, Mk = B * Sq, B * Sk
        dev = dy.device
        dy2 = dy.reshape(M, Dq)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        need_x, need_kv = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        # ---- out-projection: dZ_o = dropout-mask o dy (bf16), db_o
        dzo = grad16_of(dy, M, Dq, cfg["p_res"], cfg["s_out"])
        if dzo is None:
            dzo = cast16(dy2, drop_p=cfg["p_res"], rng=cfg["rng"], salt=cfg["s_out"])
        wo16, _ = SHADOWS.get([Wo], [bo])
        do16 = empty16(M, D, dev)
        gemm_group([gemm_problem(dzo, 0, wo16, 1, M, D, Dq, b_static=True, drop_p=cfg["p_attn"], rng=cfg["rng"], salt=cfg["s_core"], C16=do16)])
        gwo, gbo = _wbgrad(dzo, o16, [Wo], [bo], Dq)
        dWo, dbo = gwo[0], gbo[0]
        # ---- attention core
        if self_att:
            dz = empty16(M, 3 * D, dev)
            dq, dk_, dv = dz[:, :D], dz[:, D:2 * D], dz[:, 2 * D:]
        else:
            dq = empty16(M, D, dev)
            dzkv = empty16(Mk, 2 * D, dev)
            dk_, dv = dzkv[:, :D], dzkv[:, D:]
        attn16_bwd(q, k, v, do16, lse2, mask_u8, cfg["mask_kind"], cfg["causal"], B, H, Sq, Sk, dk, dq, dk_, dv)
        # ---- projections
        grads_w = {}
        dkv = None
        if self_att:
            w16, _ = SHADOWS.get([Wq, Wk, Wv], [bq, bk, bv])
            dn = torch.empty(M, Dq, device=dev, dtype=_F32) if (need_x or ln_g is not None) else None
            if dn is not None:
                gemm_group([gemm_problem(dz, 0, w16, 1, M, Dq, 3 * D, b_static=True, C32=dn)])
            gw, gb = _wbgrad(dz, n16, [Wq, Wk, Wv], [bq, bk, bv], D)
        else:
            wq16, _ = SHADOWS.get([Wq], [bq])
            wkv16, _ = SHADOWS.get([Wk, Wv], [bk, bv])
            probs = []
            dn = torch.empty(M, Dq, device=dev, dtype=_F32) if (need_x or ln_g is not None) else None
            if dn is not None:
                probs.append(gemm_problem(dq, 0, wq16, 1, M, Dq, D, b_static=True, C32=dn))
            if need_kv:
                dkv = torch.empty(B, Sk, Dkv, device=dev, dtype=_F32)
                probs.append(gemm_problem(dzkv, 0, wkv16, 1, Mk, Dkv, 2 * D, b_static=True, C32=dkv.view(Mk, Dkv)))
            if probs:
                gemm_group(probs)
            gw1, gb1 = _wbgrad(dq, n16, [Wq], [bq], D)
            gw2, gb2 = _wbgrad(dzkv, m16, [Wk, Wv], [bk, bv], D)
            gw, gb = gw1 + gw2, gb1 + gb2
        # ---- LayerNorm backward with the skip gradient folded in
        dx = dg = dbeta = None
        if ln_g is not None:
            need_aff = ln_g.requires_grad
            if need_x or need_aff:
                dx, dg, dbeta = ln_bwd(dn, x2, ln_g.detach(), mean, rstd, dy2 if cfg["resid"] else None, need_x,
                                       ops._gbuf(ln_g), ops._gbuf(ln_b), need_aff, twin=ctx.in_hint)
        elif need_x:
            dx = dn + dy2 if cfg["resid"] else dn
        if dx is not None:
            dx = _finish_dx(dx, (B, Sq, Dq))
        return (dx, dkv, None, None, dg, dbeta, gw[0], gb[0], gw[1], gb[1], gw[2], gb[2], dWo, dbo)


class _FfnBlock(torch.autograd.Function):
    """y = x + dropout_r( W_2 dropout_1( relu( W_1 LayerNorm(x) + b_1 ) ) + b_2 )   (model/blocks.py:135-144,175-187)."""

    @staticmethod
    def forward(ctx, x, cfg, ln_g, ln_b, W1, b1, W2, b2):
        B, S, D = x.shape
        M = B * S
        F_ = W1.shape[0]
        dev = x.device
        x2 = x.detach().reshape(M, D)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        if ln_g is not None:
            n16, mean, rstd = ln_fwd16(x2, ln_g.detach(), ln_b.detach(), cfg["eps"])
        else:
            n16, mean, rstd = twin16(x), None, None
        w1, _ = SHADOWS.get([W1], [b1])
        w2, _ = SHADOWS.get([W2], [b2])
        h16 = empty16(M, F_, dev)
        gemm_group([gemm_problem(n16, 0, w1, 0, M, F_, D, b_static=True, bias=b1.detach(), act=1, drop_p=cfg["p1"], rng=cfg["rng"],
                                 salt=cfg["s1"], C16=h16)])
        out = torch.empty(B, S, D, device=dev, dtype=_F32)
        out16 = empty16(M, D, dev) if cfg.get("emit_twin") else None
        gemm_group([gemm_problem(h16, 0, w2, 0, M, D, F_, b_static=True, bias=b2.detach(), drop_p=cfg["p_res"], rng=cfg["rng"], salt=cfg["s2"],
                                 resid=x2 if cfg["resid"] else None, C32=out.view(M, D), C16=out16)])
        if out16 is not None:
            attach_twin(out, out16)
        hint_backward_cast(out, cfg["p_res"], cfg["rng"], cfg["s2"])
        ctx.in_hint = getattr(x, "_bmhrl_bwd_cast", None) if ln_g is not None else None
        ctx.save_for_backward(x2, mean, rstd, n16, h16, ln_g, W1, W2)
        ctx.params = (ln_g, ln_b, W1, b1, W2, b2)
        ctx.cfg = cfg
        ctx.dims = (B, S, D, F_)
        return out

    @staticmethod
    def backward(ctx, dy):
        from . import ops
        x2, mean, rstd, n16, h16, *_ = ctx.saved_tensors
        ln_g, ln_b, W1, b1, W2, b2 = ctx.params
        cfg = ctx.cfg
        B, S, D, F_ = ctx.dims
        M = B * S
        dev = dy.device
        dy2 = dy.reshape(M, D)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        need_x = ctx.needs_input_grad[0]
        dz2 = grad16_of(dy, M, D, cfg["p_res"], cfg["s2"])
        if dz2 is None:
            dz2 = cast16(dy2, drop_p=cfg["p_res"], rng=cfg["rng"], salt=cfg["s2"])
        w1, _ = SHADOWS.get([W1], [b1])
        w2, _ = SHADOWS.get([W2], [b2])
        dz1 = empty16(M, F_, dev)   # dZ_1 = (dZ_2 W_2) o [h > 0] / (1 - p1): relu and dropout gates in the GEMM epilogue
        gs = 1.0 / (1.0 - cfg["p1"]) if cfg["p1"] > 0 else 1.0
        gemm_group([gemm_problem(dz2, 0, w2, 1, M, F_, D, b_static=True, gate=h16, gate_scale=gs, C16=dz1)])
        g2w, g2b = _wbgrad(dz2, h16, [W2], [b2], D)
        dW2, db2 = g2w[0], g2b[0]
        dn = torch.empty(M, D, device=dev, dtype=_F32)
        gemm_group([gemm_problem(dz1, 0, w1, 1, M, D, F_, b_static=True, C32=dn)])
        g1w, g1b = _wbgrad(dz1, n16, [W1], [b1], F_)
        dW1, db1 = g1w[0], g1b[0]
        dx = dg = dbeta = None
        if ln_g is not None:
            need_aff = ln_g.requires_grad
            if need_x or need_aff:
                dx, dg, dbeta = ln_bwd(dn, x2, ln_g.detach(), mean, rstd, dy2 if cfg["resid"] else None, need_x,
                                       ops._gbuf(ln_g), ops._gbuf(ln_b), need_aff, twin=ctx.in_hint)
        elif need_x:
            dx = dn + dy2 if cfg["resid"] else dn
        if dx is not None:
            dx = _finish_dx(dx, (B, S, D))
        return dx, None, dg, dbeta, dW1, db1, dW2, db2


class _Linear16(torch.autograd.Function):
    """y = dropout(x W^T + b) for an fp32 activation x (its bf16 twin is the operand); fp32 output."""

    @staticmethod
    def forward(ctx, x, W, b, cfg):
        K = x.shape[-1]
        N = W.shape[0]
        M = x.numel() // K
        x16 = twin16(x)
        w16, _ = SHADOWS.get([W], [b])
        out = torch.empty(x.shape[:-1] + (N,), device=x.device, dtype=_F32)
        out16 = empty16(M, N, x.device) if cfg.get("emit_twin") else None
        gemm_group([gemm_problem(x16, 0, w16, 0, M, N, K, b_static=True, bias=None if b is None else b.detach(), drop_p=cfg["p"],
                                 rng=cfg["rng"], salt=cfg["salt"], C32=out.view(M, N), C16=out16)])
        if out16 is not None:
            attach_twin(out, out16)
        ctx.save_for_backward(x16, W)
        ctx.params = (W, b)
        ctx.cfg = cfg
        ctx.dims = (M, N, K, x.shape)
        return out

    @staticmethod
    def backward(ctx, dy):
        from . import ops
        x16, _ = ctx.saved_tensors
        W, b = ctx.params
        cfg = ctx.cfg
        M, N, K, xshape = ctx.dims
        dev = dy.device
        dy2 = dy.reshape(M, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dz = cast16(dy2, drop_p=cfg["p"], rng=cfg["rng"], salt=cfg["salt"])
        dx = None
        if ctx.needs_input_grad[0]:
            w16, _ = SHADOWS.get([W], [b])
            dx = torch.empty(xshape, device=dev, dtype=_F32)
            gemm_group([gemm_problem(dz, 0, w16, 1, M, K, N, b_static=True, C32=dx.view(M, K))])
        if b is not None:
            gw, gb = _wbgrad(dz, x16, [W], [b], N)
            dW, db = gw[0], gb[0]
        else:
            dW, db = _wgrad(dz, x16, [W], N)[0], None
        return dx, dW, db, None


def linear16(x, W, b=None, drop_p=0.0, salt=0, rng=None, emit_twin=False):
    from . import ops
    if drop_p > 0 and rng is None:
        rng = ops.RNG.current(x.device)
    if W.shape[0] % 4 != 0 or W.shape[1] % 4 != 0:
        raise RuntimeError("linear16: feature sizes must be multiples of 4")
    return _Linear16.apply(x, W, b, {"p": float(drop_p), "salt": int(salt), "rng": rng, "emit_twin": emit_twin})


def attention_block(x, kv, mask, att, norm=None, resid=False, p_res=0.0, causal=False, emit_twin=False):
    """The ResidualConnection + MultiheadedAttention block (or the bare attention when norm / resid are off) as one
    autograd node.  `att`: a bmhrl_b200 MultiheadedAttention; kv: None (self-attention on the normed input) or the fp32
    (B,Sk,Dkv) tensor both keys and values are projected from."""
    from . import ops
    B, Sq, _ = x.shape
    Sk = Sq if kv is None else kv.shape[1]
    m_u8, kind = ops._mask_args(mask, B, Sq, Sk)
    if kind == 0:
        causal = False
    s_core, s_out = att._call_salts()
    p_a = att.dropout.p if att.training else 0.0
    rng = ops.RNG.current(x.device) if (p_a > 0 or p_res > 0) else None
    cfg = {"H": att.H, "mask_kind": kind, "causal": bool(causal), "p_attn": float(p_a), "p_res": float(p_res), "s_core": s_core,
           "s_out": s_out, "resid": bool(resid), "eps": norm.eps if norm is not None else 0.0, "rng": rng,
           "emit_twin": emit_twin}
    return _AttnBlock.apply(x, kv, m_u8, cfg, None if norm is None else norm.weight, None if norm is None else norm.bias,
                            att.linear_Q2d.weight, att.linear_Q2d.bias, att.linear_K2d.weight, att.linear_K2d.bias,
                            att.linear_V2d.weight, att.linear_V2d.bias, att.linear_d2Q.weight, att.linear_d2Q.bias)


def attention_block_supported(att, x, kv):
    D = att.d_model
    Sq = x.shape[1]
    Sk = Sq if kv is None else kv.shape[1]
    dims_ok = all(d % 8 == 0 or d % 4 == 0 for d in (x.shape[-1], D)) and x.shape[-1] % 4 == 0 and \
        (kv is None or kv.shape[-1] % 4 == 0)
    return dims_ok and D % att.H == 0 and attn16_supported(Sq, Sk, D // att.H)


def ffn_block(x, ff, norm=None, resid=False, p_res=0.0, emit_twin=False):
    from . import ops
    if ff._salts is None:
        ff._salts = (ops.next_salt(), ops.next_salt())
    p1 = ff.dropout.p if ff.training else 0.0
    rng = ops.RNG.current(x.device) if (p1 > 0 or p_res > 0) else None
    cfg = {"p1": float(p1), "p_res": float(p_res), "s1": ff._salts[0], "s2": ff._salts[1], "resid": bool(resid),
           "eps": norm.eps if norm is not None else 0.0, "rng": rng, "emit_twin": emit_twin}
    return _FfnBlock.apply(x, cfg, None if norm is None else norm.weight, None if norm is None else norm.bias, ff.fc1.weight,
                           ff.fc1.bias, ff.fc2.weight, ff.fc2.bias)


# ----------------------------------------------------------------------------------------------------------------
# K4 fused: vocabulary projection + log-softmax + KL-loss statistics (gemm_bf16.cu, MODE 1 / 2)
# ----------------------------------------------------------------------------------------------------------------
class _VocabStats(torch.autograd.Function):
    """(M,K) features -> the five per-row statistics {lse, sum logit, logit[target], logit[pad], logit[token]} of the
    vocabulary logits x W^T + b and the sampled / arg-max token, without the (M,V) logits ever reaching memory.
    Backward: d logits (bf16) from the five statistic gradients (recomputing GEMM), then the ordinary dX / dW GEMMs."""

    @staticmethod
    def forward(ctx, x, W, b, trg, pad_idx, smode, noise, rng, salt):
        K = x.shape[-1]
        V = W.shape[0]
        M = x.numel() // K
        x16 = twin16(x)
        w16, _ = SHADOWS.get([W], [b])
        trg = trg.reshape(-1).contiguous()
        dev = x.device
        part = torch.empty(LIB.raw("bmhrl_vocab_stats_ws_floats")(M, V), device=dev, dtype=_F32)
        stats = torch.empty(M, 5, device=dev, dtype=_F32)
        tok = torch.empty(M, device=dev, dtype=torch.int64)
        if smode == 1:
            noise = noise.reshape(M, V).contiguous()
        LIB.call("bmhrl_vocab_stats_bf16", x16.data_ptr(), x16.stride(0), w16.data_ptr(), w16.stride(0),
                 None if b is None else b.data_ptr(), M, V, K, trg.data_ptr(), int(pad_idx), int(smode),
                 noise.data_ptr() if smode == 1 else None, rng.data_ptr() if smode == 2 else None, int(salt), part.data_ptr(),
                 stats.data_ptr(), tok.data_ptr(), _stream())
        ctx.save_for_backward(x16, stats, trg, tok)
        ctx.params = (W, b)
        ctx.cfg = (M, V, K, int(pad_idx), x.shape)
        ctx.mark_non_differentiable(tok)
        return stats, tok

    @staticmethod
    def backward(ctx, dstats, _dtok):
        x16, stats, trg, tok = ctx.saved_tensors
        W, b = ctx.params
        M, V, K, pad_idx, xshape = ctx.cfg
        dev = dstats.device
        dstats = dstats.contiguous()
        w16, _ = SHADOWS.get([W], [b])
        dl = empty16(M, V, dev)
        LIB.call("bmhrl_vocab_dlogits_bf16", x16.data_ptr(), x16.stride(0), w16.data_ptr(), w16.stride(0),
                 None if b is None else b.data_ptr(), M, V, K, trg.data_ptr(), tok.data_ptr(), pad_idx, stats.data_ptr(),
                 dstats.data_ptr(), dl.data_ptr(), dl.stride(0), _stream())
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(xshape, device=dev, dtype=_F32)
            gemm_group([gemm_problem(dl, 0, w16, 1, M, K, V, b_static=True, C32=dx.view(M, K))])
        if b is not None:
            gw, gb = _wbgrad(dl, x16, [W], [b], V)
            dW, db = gw[0], gb[0]
        else:
            dW, db = _wgrad(dl, x16, [W], V)[0], None
        return dx, dW, db, None, None, None, None, None, None


def vocab_stats(x, W, b, trg, pad_idx, mode="argmax", exp_noise=None, salt=0, rng=None):
    """stats (M,5) fp32 (differentiable in x, W, b) and tokens (M,) int64; see bmhrl_vocab_stats_bf16."""
    from . import ops
    m = {"argmax": 0, "noise": 1, "philox": 2}[mode]
    if m == 2 and rng is None:
        rng = ops.RNG.current(x.device)
    return _VocabStats.apply(x, W, b, trg, pad_idx, m, exp_noise, rng, salt)


class _KlStats(torch.autograd.Function):
    @staticmethod
    def forward(ctx, stats, trg, samp, offset, mode, smoothing, pad_idx, V):
        stats = stats.contiguous()
        rows = stats.shape[0]
        trg = trg.reshape(-1).contiguous()
        if mode == 1:
            samp = samp.reshape(-1).contiguous()
            offset = offset.reshape(-1).contiguous()
        rs = torch.empty(rows, device=stats.device, dtype=_F32)
        LIB.call("bmhrl_kl_stats_fwd_f32", mode, stats.data_ptr(), trg.data_ptr(), samp.data_ptr() if mode == 1 else None,
                 offset.data_ptr() if mode == 1 else None, rs.data_ptr(), rows, V, smoothing, pad_idx, _stream())
        ctx.save_for_backward(stats, trg, samp if mode == 1 else None, offset if mode == 1 else None)
        ctx.cfg = (mode, smoothing, pad_idx, V, rows)
        return rs

    @staticmethod
    def backward(ctx, g):
        stats, trg, samp, offset = ctx.saved_tensors
        mode, smoothing, pad_idx, V, rows = ctx.cfg
        g = g.contiguous()
        ds = torch.empty(rows, 5, device=g.device, dtype=_F32)
        doff = torch.empty(rows, device=g.device, dtype=_F32) if (mode == 1 and ctx.needs_input_grad[3]) else None
        LIB.call("bmhrl_kl_stats_bwd_f32", mode, stats.data_ptr(), trg.data_ptr(), samp.data_ptr() if mode == 1 else None,
                 offset.data_ptr() if mode == 1 else None, g.data_ptr(), ds.data_ptr(), None if doff is None else doff.data_ptr(),
                 rows, V, smoothing, pad_idx, _stream())
        return ds, None, None, doff, None, None, None, None


def kl_rowsum_from_stats(stats, trg, smoothing, pad_idx, V, biased_trg=None, biased_offset=None):
    """Row sums of the LabelSmoothing (biased_trg None; loss/label_smoothing.py:12-32) or BiasedKL
    (loss/biased_kl.py:22-53) divergence, sum_v [xlogy(d_v) - d_v log p_v], in closed form from the five statistics
    (loss.cu: kl_stats_fwd / bwd kernels).  Differentiable in stats and biased_offset."""
    if biased_trg is None:
        return _KlStats.apply(stats, trg, None, None, 0, float(smoothing), int(pad_idx), int(V))
    return _KlStats.apply(stats, trg, biased_trg, biased_offset.reshape(-1), 1, float(smoothing), int(pad_idx), int(V))


class VocabHead:
    """What BMHrlAgent returns in place of the (B,L,V) log-probability tensor when `fused_vocab` is on (the trainer's
    bf16 path): the 364-d features and the projection, from which the losses are evaluated without the logits."""

    def __init__(self, x, proj):
        self.x, self.proj = x, proj
        self.shape = tuple(x.shape[:-1]) + (proj.out_features,)

    def stats(self, trg, pad_idx, mode="argmax", exp_noise=None, salt=0):
        st, tok = vocab_stats(self.x, self.proj.weight, self.proj.bias, trg, pad_idx, mode, exp_noise, salt)
        return st, tok.view(self.x.shape[:-1])

    def log_probs(self):
        from . import ops
        return ops.log_softmax(linear16(self.x, self.proj.weight, self.proj.bias))
