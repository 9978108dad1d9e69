"""GPU: the bf16 flash-style attention forward (attn_bf16.cu, BASELINE.json config 5) against the oracle's
attention_core (model/multihead_attention.py:7-31) evaluated in fp64 on the SAME bf16-rounded inputs.
Tolerance: 2e-2 relative (north_star's bf16 bound) on the output; asserted tighter (1e-2) so that a descriptor /
layout bug cannot hide inside it.  Log-sum-exp: 1e-3 absolute + relative."""
import math

import pytest
import torch

from conftest import rel

pytestmark = pytest.mark.gpu
BF16_TOL = 1e-2


@pytest.fixture(autouse=True, params=["one_cta", "cta_pair"])
def kernel_variant(request, gpu_lib):
    """Every test runs on both forms of the kernel: one CTA per query tile, and CTA pairs (cta_group::2 MMAs with the
    K / V tiles split between the two shared memories); d_k = 64 always takes the one-CTA form."""
    gpu_lib.call("bmhrl_attn_bf16_tuning", 1 if request.param == "one_cta" else 2)
    yield request.param
    gpu_lib.call("bmhrl_attn_bf16_tuning", 0)


def _inputs(B, H, Sq, Sk, dk, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    D = H * dk
    q = (torch.randn(B, Sq, D, generator=g) * scale).bfloat16()
    k = (torch.randn(B, Sk, D, generator=g) * scale).bfloat16()
    v = torch.randn(B, Sk, D, generator=g).bfloat16()
    return q, k, v


def _oracle(q, k, v, mask4, H, causal=False):
    from oracle import bmhrl_oracle as O  # checker
    B, Sq, D = q.shape
    Sk = k.shape[1]
    dk = D // H
    split = lambda t, S: t.double().view(B, S, H, dk).transpose(1, 2)
    o = O.attention_core(split(q, Sq), split(k, Sk), split(v, Sk), mask4, causal)
    s = split(q, Sq) @ split(k, Sk).transpose(-1, -2) / math.sqrt(dk)
    if mask4 is not None:
        if causal:
            s = s.masked_fill(~torch.tril(torch.ones(Sk, Sk, dtype=torch.bool))[None, None], -1e9)
        s = s.masked_fill(~mask4, -1e9)
    return o.transpose(1, 2).reshape(B, Sq, D), torch.logsumexp(s, -1)


def _run(gpu_lib, q, k, v, mask, H, causal=False, use_ws=True):
    from bmhrl_b200 import ops
    with torch.no_grad():
        o, lse = ops.mha_core_bf16(q.cuda(), k.cuda(), v.cuda(), None if mask is None else mask.cuda(), H, causal,
                                   return_lse=True, use_ws=use_ws)
    torch.cuda.synchronize()
    return o.float().cpu(), lse.cpu()


@pytest.mark.parametrize("B,H,Sq,Sk,dk", [(2, 4, 30, 30, 256), (2, 4, 80, 80, 256), (1, 4, 300, 300, 256),
                                          (2, 2, 300, 800, 256), (1, 2, 800, 300, 256), (3, 2, 129, 257, 128),
                                          (2, 3, 128, 128, 64), (1, 1, 1, 1, 256), (2, 4, 20, 80, 256),
                                          (5, 4, 300, 300, 256)])
def test_no_mask(gpu_lib, B, H, Sq, Sk, dk):
    q, k, v = _inputs(B, H, Sq, Sk, dk, seed=Sq + Sk + dk)
    o, lse = _run(gpu_lib, q, k, v, None, H)
    ro, rl = _oracle(q, k, v, None, H)
    assert rel(o, ro) < BF16_TOL, rel(o, ro)
    assert torch.allclose(lse.double(), rl, rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("Sq,Sk", [(30, 30), (300, 300), (300, 800), (200, 130)])
def test_key_padding_mask(gpu_lib, Sq, Sk):
    """(B,1,Sk) masks with 0 %, ~50 % tail padding and one FULLY masked batch row (uniform over all keys)."""
    B, H, dk = 4, 4, 256
    q, k, v = _inputs(B, H, Sq, Sk, dk, seed=7 * Sq + Sk)
    mask = torch.ones(B, 1, Sk, dtype=torch.bool)
    mask[1, :, Sk // 2:] = False
    mask[2, :, :] = False
    mask[3, :, 1::2] = False  # holes, first tile partially visible
    o, lse = _run(gpu_lib, q, k, v, mask, H)
    ro, rl = _oracle(q, k, v, mask[:, None], H)
    assert rel(o, ro) < BF16_TOL, rel(o, ro)
    assert rel(o[2], ro[2]) < BF16_TOL  # the fully masked row: mean of V over ALL keys, like the reference
    assert torch.allclose(lse.double(), rl, rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("use_ws", [True, False])
def test_trimmed_tail_tiles(gpu_lib, use_ws):
    """Per batch row the kernel skips the key tiles behind the last visible key (workspace given); rows with NO visible
    key keep all Sk keys (uniform attention like the reference); ragged lengths across the batch; ws = NULL path."""
    B, H, Sq, Sk, dk = 6, 2, 200, 700, 256
    q, k, v = _inputs(B, H, Sq, Sk, dk, seed=41)
    mask = torch.zeros(B, 1, Sk, dtype=torch.bool)
    for b, n in enumerate([700, 1, 128, 129, 0, 385]):
        mask[b, :, :n] = True
    mask[3, :, 5:100] = False  # holes before the last visible key
    o, lse = _run(gpu_lib, q, k, v, mask, H, use_ws=use_ws)
    ro, rl = _oracle(q, k, v, mask[:, None], H)
    for b in range(B):
        assert rel(o[b], ro[b]) < BF16_TOL, (b, rel(o[b], ro[b]))
    assert torch.allclose(lse.double(), rl, rtol=1e-3, atol=1e-3)


def test_late_visible_keys_force_rescale(gpu_lib):
    """Only keys of the LAST key tile are visible, and scores are large: the running maximum must be raised after
    P V of earlier tiles has been accumulated (the accumulator-rescale path)."""
    B, H, Sq, Sk, dk = 2, 2, 140, 300, 256
    q, k, v = _inputs(B, H, Sq, Sk, dk, seed=99, scale=2.0)
    mask = torch.zeros(B, 1, Sk, dtype=torch.bool)
    mask[:, :, 270:] = True
    o, _ = _run(gpu_lib, q, k, v, mask, H)
    ro, _ = _oracle(q, k, v, mask[:, None], H)
    assert rel(o, ro) < BF16_TOL, rel(o, ro)
    # growing scores without a mask: later tiles dominate by far more than the lazy threshold
    k2 = k.clone().float()
    k2[:, 128:] *= 3
    k2[:, 256:] *= 3
    k2 = k2.bfloat16()
    o, lse = _run(gpu_lib, q, k2, v, None, H)
    ro, rl = _oracle(q, k2, v, None, H)
    assert rel(o, ro) < BF16_TOL, rel(o, ro)
    assert torch.allclose(lse.double(), rl, rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("L", [20, 150])
def test_full_and_causal_mask(gpu_lib, L):
    """(B,L,L) caption mask = pad & tril (model/masking.py:13-25), and the causal flag (multihead_attention.py:18-21)."""
    B, H, dk = 3, 2, 128
    q, k, v = _inputs(B, H, L, L, dk, seed=L)
    pad = torch.ones(B, L, dtype=torch.bool)
    pad[1, L - 5:] = False
    full = pad[:, None, :] & torch.tril(torch.ones(L, L, dtype=torch.bool))[None]
    o, _ = _run(gpu_lib, q, k, v, full, H)
    ro, _ = _oracle(q, k, v, full[:, None], H)
    assert rel(o, ro) < BF16_TOL, rel(o, ro)
    km = pad[:, None, :]
    o, _ = _run(gpu_lib, q, k, v, km, H, causal=True)
    ro, _ = _oracle(q, k, v, km[:, None], H, causal=True)
    assert rel(o, ro) < BF16_TOL, rel(o, ro)


def test_matches_training_path(gpu_lib):
    """Same inputs through the fp32/tf32 training kernel (ops.mha_core) and the bf16 kernel."""
    from bmhrl_b200 import ops
    B, H, Sq, Sk, dk = 4, 4, 300, 300, 256
    q, k, v = _inputs(B, H, Sq, Sk, dk, seed=3)
    mask = torch.ones(B, 1, Sk, dtype=torch.bool)
    mask[0, :, 100:] = False
    o, _ = _run(gpu_lib, q, k, v, mask, H)
    ops.set_math_mode("fp32")
    of = ops.mha_core(q.float().cuda(), k.float().cuda(), v.float().cuda(), mask.cuda(), H)
    assert rel(o, of) < BF16_TOL


def test_rejects_bad_inputs(gpu_lib):
    from bmhrl_b200 import ops
    q, k, v = _inputs(1, 2, 8, 8, 256, seed=0)
    with pytest.raises(RuntimeError):
        ops.mha_core_bf16(q, k, v, None, 2)  # CPU tensors: no fallback
    with pytest.raises(RuntimeError):
        ops.mha_core_bf16(q.float().cuda(), k.float().cuda(), v.float().cuda(), None, 2)
    q, k, v = _inputs(1, 2, 8, 8, 96, seed=0)
    with pytest.raises(RuntimeError):
        ops.mha_core_bf16(q.cuda(), k.cuda(), v.cuda(), None, 2)  # d_k = 96 unsupported


def test_full_size_properties(gpu_lib):
    """BASELINE config 5 sizes (B=64, H=4, S=300, d_k=256), size-independent properties instead of the oracle:
    (1) V = 1 -> every output is exactly 1 (the probabilities of a row sum to 1 against the same normaliser);
    (2) masked key padding does not change the un-padded result, and neither does the workspace (tile trimming);
    (3) permuting the keys (and values) leaves the output unchanged up to bf16 summation order;
    (4) log-sum-exp of a row with identical scores = score + log(Sk)."""
    from bmhrl_b200 import ops
    B, H, S, dk = 64, 4, 300, 256
    D = H * dk
    g = torch.Generator(device="cuda").manual_seed(5)
    q = torch.randn(B, S, D, device="cuda", generator=g).bfloat16()
    k = torch.randn(B, S, D, device="cuda", generator=g).bfloat16()
    v = torch.randn(B, S, D, device="cuda", generator=g).bfloat16()
    with torch.no_grad():
        o1 = ops.mha_core_bf16(q, k, torch.ones_like(v), None, H)
        assert float((o1.float() - 1).abs().max()) <= 2 ** -7  # bf16 rounding of P and of the output only
        o = ops.mha_core_bf16(q, k, v, None, H)
        # (2) 300 -> 500 keys, the tail masked
        kp = torch.cat([k, torch.randn(B, 200, D, device="cuda", generator=g).bfloat16()], 1)
        vp = torch.cat([v, torch.randn(B, 200, D, device="cuda", generator=g).bfloat16()], 1)
        mask = torch.zeros(B, 1, 500, dtype=torch.bool, device="cuda")
        mask[:, :, :S] = True
        for use_ws in (True, False):
            op = ops.mha_core_bf16(q, kp, vp, mask, H, use_ws=use_ws)
            assert rel(op, o) < 2e-3, (use_ws, rel(op, o))
        # (3) key permutation
        perm = torch.randperm(S, device="cuda", generator=g)
        opm = ops.mha_core_bf16(q, k[:, perm].contiguous(), v[:, perm].contiguous(), None, H)
        assert rel(opm, o) < 5e-3, rel(opm, o)
        # (4) constant scores
        _, lse = ops.mha_core_bf16(torch.zeros_like(q), k, v, None, H, return_lse=True)
        assert torch.allclose(lse, torch.full_like(lse, math.log(S)), atol=1e-4)
    torch.cuda.synchronize()
