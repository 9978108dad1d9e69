"""GPU: the one-tile fused bf16 attention kernels of the training path (attn16.cu) against the fp32 oracle
(oracle.attention_core == model/multihead_attention.py:7-31) evaluated on the same bf16-rounded inputs.  The kernel
rounds the probabilities / dS to bf16 before the second contraction: bound 2e-2 (north_star bf16 bound), observed < 1e-2."""
import math

import pytest
import torch

from oracle import bmhrl_oracle as O
from conftest import rel

pytestmark = pytest.mark.gpu
TOL = 2e-2


def _inputs(B, H, Sq, Sk, dk, kind, seed, concat):
    from bmhrl_b200 import ops16
    g = torch.Generator().manual_seed(seed)
    D = H * dk
    if concat and Sq == Sk:  # Q|K|V as column slices of ONE (B*S, 3D) matrix, like the fused projection emits them
        buf = ops16.empty16(B * Sq, 3 * D, "cuda")
        buf.copy_(torch.randn(B * Sq, 3 * D, generator=g))
        q, k, v = buf[:, :D], buf[:, D:2 * D], buf[:, 2 * D:]
    else:
        q = ops16.empty16(B * Sq, D, "cuda"); q.copy_(torch.randn(B * Sq, D, generator=g))
        k = ops16.empty16(B * Sk, D, "cuda"); k.copy_(torch.randn(B * Sk, D, generator=g))
        v = ops16.empty16(B * Sk, D, "cuda"); v.copy_(torch.randn(B * Sk, D, generator=g))
    if kind == 0:
        mask = None
    elif kind == 1:
        lens = torch.randint(1, Sk + 1, (B,), generator=g)
        lens[0] = Sk
        mask = (torch.arange(Sk)[None, :] < lens[:, None])[:, None, :]
        if B > 1:
            mask[1] = False  # fully masked clip: uniform attention over every key (reference semantics)
    else:
        mask = torch.tril(torch.ones(Sq, Sk, dtype=torch.bool))[None].repeat(B, 1, 1)
        mask[0, :, Sk // 2:] = False
        mask[-1, Sq - 1, :] = False  # one fully masked query row
    do = ops16.empty16(B * Sq, D, "cuda"); do.copy_(torch.randn(B * Sq, D, generator=g))
    return q, k, v, mask, do


def _oracle(q, k, v, mask, do, B, H, Sq, Sk, dk, causal=False):
    split = lambda t, S: t.float().cpu().reshape(B, S, H, dk).transpose(1, 2).clone().requires_grad_(True)
    qq, kk, vv = split(q, Sq), split(k, Sk), split(v, Sk)
    m = None if mask is None else (mask[:, None] if mask.dim() == 3 else mask)
    o = O.attention_core(qq, kk, vv, m, causal)
    o.backward(do.float().cpu().reshape(B, Sq, H, dk).transpose(1, 2))
    merge = lambda t, S: t.transpose(1, 2).reshape(B * S, H * dk)
    return merge(o.detach(), Sq), merge(qq.grad, Sq), merge(kk.grad, Sk), merge(vv.grad, Sk)


@pytest.mark.parametrize("B,H,Sq,Sk,dk,kind,concat", [
    (3, 4, 30, 30, 256, 1, True), (2, 4, 20, 80, 256, 1, False), (2, 2, 20, 20, 512, 2, True), (2, 4, 80, 80, 256, 1, True),
    (2, 4, 80, 30, 256, 1, False), (16, 4, 80, 80, 256, 0, True), (2, 4, 33, 47, 128, 1, False), (1, 4, 5, 7, 64, 2, False),
    (2, 4, 128, 64, 256, 1, False), (2, 4, 30, 30, 256, 2, True)])
def test_attn16_forward_backward_vs_oracle(gpu_lib, B, H, Sq, Sk, dk, kind, concat):
    from bmhrl_b200 import ops16, ops
    q, k, v, mask, do = _inputs(B, H, Sq, Sk, dk, kind, B * 1000 + Sq * 10 + Sk, concat)
    assert ops16.attn16_supported(Sq, Sk, dk)
    m_u8, mk = ops._mask_args(None if mask is None else mask.cuda(), B, Sq, Sk)
    D = H * dk
    o = ops16.empty16(B * Sq, D, "cuda")
    lse2 = torch.empty(B, H, Sq, 2, device="cuda")
    ops16.attn16_fwd(q, k, v, m_u8, mk, False, B, H, Sq, Sk, dk, o, lse2)
    ro, rdq, rdk, rdv = _oracle(q, k, v, mask, do, B, H, Sq, Sk, dk)
    assert rel(o.float(), ro) < TOL, ("o", rel(o.float(), ro))
    dq, dk_, dv = (ops16.zeros16(B * S, D, "cuda") for S in (Sq, Sk, Sk))
    ops16.attn16_bwd(q, k, v, do, lse2, m_u8, mk, False, B, H, Sq, Sk, dk, dq, dk_, dv)
    for name, a, r in (("dq", dq, rdq), ("dk", dk_, rdk), ("dv", dv, rdv)):
        assert torch.isfinite(a.float()).all(), name
        assert rel(a.float(), r) < TOL, (name, rel(a.float(), r))


def test_attn16_causal_and_dropout(gpu_lib):
    from bmhrl_b200 import ops16, ops
    B, H, Sq, Sk, dk = 2, 4, 24, 24, 256
    q, k, v, mask, do = _inputs(B, H, Sq, Sk, dk, 1, 77, True)
    mask = torch.ones(B, 1, Sk, dtype=torch.bool)
    m_u8, mk = ops._mask_args(mask.cuda(), B, Sq, Sk)
    D = H * dk
    o = ops16.empty16(B * Sq, D, "cuda")
    lse2 = torch.empty(B, H, Sq, 2, device="cuda")
    ops16.attn16_fwd(q, k, v, m_u8, mk, True, B, H, Sq, Sk, dk, o, lse2)   # causal (DETR decoder self-attention)
    ro, rdq, rdk, rdv = _oracle(q, k, v, mask, do, B, H, Sq, Sk, dk, causal=True)
    assert rel(o.float(), ro) < TOL
    dq, dk_, dv = (ops16.zeros16(B * S, D, "cuda") for S in (Sq, Sk, Sk))
    ops16.attn16_bwd(q, k, v, do, lse2, m_u8, mk, True, B, H, Sq, Sk, dk, dq, dk_, dv)
    assert rel(dq.float(), rdq) < TOL and rel(dk_.float(), rdk) < TOL and rel(dv.float(), rdv) < TOL
    # dropout on the product: keep rate 1 - p, survivors scaled by 1 / (1 - p), same mask as the GEMM epilogue convention
    ops.manual_seed(3)
    rng = ops.new_rng_epoch("cuda")
    od = ops16.empty16(B * Sq, D, "cuda")
    ops16.attn16_fwd(q, k, v, m_u8, mk, True, B, H, Sq, Sk, dk, od, lse2, drop_p=0.25, rng=rng, salt=21)
    keep = od.float() != 0
    assert abs(float(keep.float().mean()) - 0.75) < 0.02
    assert rel(od.float()[keep], (o.float() / 0.75)[keep]) < 1e-2
    # the mask is the one a GEMM epilogue regenerates for element index row * D + column
    ones16 = ops16.empty16(B * Sq, 64, "cuda"); ones16.fill_(1.0)
    eye = ops16.zeros16(D, 64, "cuda")      # C = ones (M,64) x eye^T (D,64) -> all-ones (M, D) for column < 64 ...
    eye[:64, :64] = torch.eye(64, device="cuda", dtype=torch.bfloat16)
    probe = torch.zeros(B * Sq, D, device="cuda")
    ops16.gemm_group([ops16.gemm_problem(ones16, 0, eye, 0, B * Sq, D, 64, drop_p=0.25, rng=rng, salt=21, C32=probe)])
    assert torch.equal(probe[:, :64] != 0, keep[:, :64])
