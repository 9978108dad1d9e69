"""GPU: the bf16 flash-style attention backward (attn_bf16_bwd.cu, BASELINE.json config 5 "fwd+bwd") against autograd
through the oracle's attention_core (model/multihead_attention.py:7-31) in fp64 on the SAME bf16-rounded inputs.
Tolerance: 2e-2 relative (north_star's bf16 bound); asserted at 1e-2 so that a layout bug cannot hide inside it."""
import pytest
import torch

from conftest import rel

pytestmark = pytest.mark.gpu
BF16_TOL = 1e-2


def _inputs(B, H, Sq, Sk, dk, seed):
    g = torch.Generator().manual_seed(seed)
    D = H * dk
    q = torch.randn(B, Sq, D, generator=g).bfloat16()
    k = torch.randn(B, Sk, D, generator=g).bfloat16()
    v = torch.randn(B, Sk, D, generator=g).bfloat16()
    do = torch.randn(B, Sq, D, generator=g).bfloat16()
    return q, k, v, do


def _oracle_grads(q, k, v, do, mask4, H):
    from oracle import bmhrl_oracle as O  # checker
    B, Sq, D = q.shape
    Sk = k.shape[1]
    dk = D // H
    qd, kd, vd = (t.double().requires_grad_(True) for t in (q, k, v))
    split = lambda t, S: t.view(B, S, H, dk).transpose(1, 2)
    o = O.attention_core(split(qd, Sq), split(kd, Sk), split(vd, Sk), mask4, False)
    o = o.transpose(1, 2).reshape(B, Sq, D)
    o.backward(do.double())
    return qd.grad, kd.grad, vd.grad


def _run(q, k, v, do, mask, H):
    from bmhrl_b200 import ops
    qc, kc, vc = q.cuda(), k.cuda(), v.cuda()
    with torch.no_grad():
        o, lse = ops.mha_core_bf16(qc, kc, vc, None if mask is None else mask.cuda(), H, return_lse=True)
        dq, dk_, dv = ops.mha_core_bf16_bwd(qc, kc, vc, o, lse, do.cuda(), None if mask is None else mask.cuda(), H)
    torch.cuda.synchronize()
    return dq.float().cpu(), dk_.float().cpu(), dv.float().cpu()


def _check(got, ref, what=""):
    for name, g, r in zip(("dq", "dk", "dv"), got, ref):
        assert torch.isfinite(g).all(), (what, name)
        # 1e-4 per element of absolute slack: gradients that are exactly 0 in exact arithmetic (one key: dS = P (dP - delta)
        # = 0) come out as the bf16 rounding of O inside delta
        err = float((g.double() - r.double()).norm())
        assert err < BF16_TOL * float(r.double().norm()) + 1e-4 * g.numel() ** 0.5, (what, name, rel(g, r))


@pytest.mark.parametrize("B,H,Sq,Sk,dk", [(2, 4, 30, 30, 256), (2, 4, 80, 80, 256), (2, 2, 300, 300, 256),
                                          (1, 2, 300, 800, 256), (1, 2, 800, 300, 256), (3, 2, 129, 257, 128),
                                          (2, 3, 128, 128, 64), (1, 1, 1, 1, 256), (2, 4, 20, 80, 256),
                                          (5, 4, 64, 65, 256), (40, 4, 150, 150, 256)])
def test_no_mask(gpu_lib, B, H, Sq, Sk, dk):
    q, k, v, do = _inputs(B, H, Sq, Sk, dk, seed=Sq + 3 * Sk + dk)
    _check(_run(q, k, v, do, None, H), _oracle_grads(q, k, v, do, None, H))


@pytest.mark.parametrize("Sq,Sk", [(30, 30), (300, 300), (300, 800), (200, 130)])
def test_key_padding_mask(gpu_lib, Sq, Sk):
    """(B,1,Sk) masks: none masked, ~50 % tail padding, one FULLY masked batch row (uniform weights: dV = mean of dO,
    dQ = dK = 0 like the reference, whose masked_fill cuts the gradient), and holes."""
    B, H, dk = 4, 4, 256
    q, k, v, do = _inputs(B, H, Sq, Sk, dk, seed=7 * Sq + Sk)
    mask = torch.ones(B, 1, Sk, dtype=torch.bool)
    mask[1, :, Sk // 2:] = False
    mask[2, :, :] = False
    mask[3, :, 1::2] = False
    got = _run(q, k, v, do, mask, H)
    ref = _oracle_grads(q, k, v, do, mask[:, None], H)
    _check(got, ref)
    for b in range(B):
        _check([g[b] for g in got], [r[b] for r in ref], what=f"batch row {b}")
    assert float(got[0][2].abs().max()) == 0.0 and float(got[1][2].abs().max()) == 0.0  # fully masked row
    assert float(got[1][1, Sk // 2:].abs().max()) == 0.0  # masked keys get no dK / dV
    assert float(got[2][1, Sk // 2:].abs().max()) == 0.0


def test_autograd_wrapper(gpu_lib):
    from bmhrl_b200 import ops
    B, H, Sq, Sk, dk = 3, 4, 140, 200, 256
    q, k, v, do = _inputs(B, H, Sq, Sk, dk, seed=17)
    mask = torch.ones(B, 1, Sk, dtype=torch.bool)
    mask[0, :, 77:] = False
    qc, kc, vc = (t.cuda().requires_grad_(True) for t in (q, k, v))
    o = ops.mha_core_bf16_train(qc, kc, vc, mask.cuda(), H)
    o.backward(do.cuda())
    torch.cuda.synchronize()
    _check([t.grad.float().cpu() for t in (qc, kc, vc)], _oracle_grads(q, k, v, do, mask[:, None], H))


def test_rejects_bad_inputs(gpu_lib):
    from bmhrl_b200 import ops
    q, k, v, do = _inputs(1, 2, 8, 8, 256, seed=0)
    qc, kc, vc, dc = q.cuda(), k.cuda(), v.cuda(), do.cuda()
    with torch.no_grad():
        o, lse = ops.mha_core_bf16(qc, kc, vc, None, 2, return_lse=True)
    with pytest.raises(RuntimeError):
        ops.mha_core_bf16_bwd(q, k, v, o.cpu(), lse.cpu(), do, None, 2)  # CPU tensors: no fallback
    full = torch.ones(1, 8, 8, dtype=torch.bool, device="cuda")
    with pytest.raises(RuntimeError):
        ops.mha_core_bf16_bwd(qc, kc, vc, o, lse, dc, full, 2)  # full (B,Sq,Sk) masks: fp32 / tf32 kernels only
    with pytest.raises(RuntimeError):
        ops.mha_core_bf16_bwd(qc, kc, vc, o, lse[:, :1].contiguous(), dc, None, 2)


def test_full_size_against_training_kernels(gpu_lib):
    """BASELINE config 5 size (B=64, H=4, S=300, d_k=256): the bf16 gradients against autograd through the fp32-operand
    training kernels (ops.mha_core, exact-fp32 mode) on the same values, and linearity in dO."""
    from bmhrl_b200 import ops
    B, H, S, dk = 64, 4, 300, 256
    D = H * dk
    g = torch.Generator(device="cuda").manual_seed(5)
    q, k, v, do = (torch.randn(B, S, D, device="cuda", generator=g).bfloat16() for _ in range(4))
    mask = torch.ones(B, 1, S, dtype=torch.bool, device="cuda")
    mask[::3, :, 180:] = False
    with torch.no_grad():
        o, lse = ops.mha_core_bf16(q, k, v, mask, H, return_lse=True)
        got = ops.mha_core_bf16_bwd(q, k, v, o, lse, do, mask, H)
        got2 = ops.mha_core_bf16_bwd(q, k, v, o, lse, (do.float() * 2).bfloat16(), mask, H)
    for a, b2 in zip(got, got2):  # delta and dP are linear in dO: doubling dO doubles every gradient (bf16 exact x2)
        assert rel(b2.float(), 2 * a.float()) < 1e-3
    ops.set_math_mode("fp32")
    qf, kf, vf = (t.float().requires_grad_(True) for t in (q, k, v))
    of = ops.mha_core(qf, kf, vf, mask, H)
    of.backward(do.float())
    torch.cuda.synchronize()
    for name, a, r in zip(("dq", "dk", "dv"), got, (qf.grad, kf.grad, vf.grad)):
        assert rel(a.float(), r) < BF16_TOL, (name, rel(a.float(), r))
