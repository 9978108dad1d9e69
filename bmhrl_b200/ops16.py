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
                ("seg_rows", ctypes.c_int), ("C32_seg1", ctypes.c_void_p), ("C32_seg2", ctypes.c_void_p)]


_DESC_CHECKED = [False]
RECORD = None  # bench.py: when a list, every problem appends (M, N, K, a_mn, b_mn)


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
                 rng=None, salt=0, resid=None, beta=0.0, C32=None, C16=None, seg=None):
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
        RECORD.extend((p.M, p.N, p.K, p.a_mn, p.b_mn) for p in problems)
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
