"""GPU: the grouped bf16 tcgen05 GEMM (bmhrl_gemm_bf16_group) against fp32 torch on the SAME bf16-rounded operands:
the only differences are the fp32 accumulation order and the final bf16 rounding of C16, so the bound is tight (1e-5 for
fp32 outputs, 2^-8 for bf16 outputs)."""
import math

import pytest
import torch

from conftest import rel

pytestmark = pytest.mark.gpu


@pytest.fixture()
def g16(gpu_lib):
    from bmhrl_b200 import ops16
    yield ops16
    gpu_lib.call("bmhrl_gemm_bf16_tuning", 0, 0, 0)


def _g(seed):
    return torch.Generator().manual_seed(seed)


def _mat(rows, cols, gen, scale=1.0):
    """random bf16 matrix with a padded row pitch; returns (device bf16 view, the same values as fp32 on the device)"""
    from bmhrl_b200 import ops16
    v = (torch.randn(rows, cols, generator=gen) * scale).to(torch.bfloat16)
    d = ops16.empty16(rows, cols, "cuda")
    d.copy_(v)
    return d, v.float().cuda()


SHAPES = [(128, 128, 64), (480, 1024, 1024), (320, 300, 1024), (320, 1024, 300), (1280, 128, 3072), (37, 300, 64),
          (320, 10172, 364), (5, 64, 1024), (129, 520, 200), (480, 3072, 1024)]


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("bn,split", [(0, 0), (64, 1), (128, 1), (128, 2), (128, 4), (64, 2)])
def test_forward_major_kk(g16, gpu_lib, M, N, K, bn, split):
    """A K-major (activations), B K-major (nn.Linear weight): y = x W^T + b, fp32 and bf16 outputs."""
    gpu_lib.call("bmhrl_gemm_bf16_tuning", bn, split, 0)
    gen = _g(M * 7 + N * 3 + K)
    x, xf = _mat(M, K, gen)
    w, wf = _mat(N, K, gen, 1 / math.sqrt(K))
    b = torch.randn(N, generator=gen).cuda()
    ref = xf @ wf.t() + b
    c32 = torch.empty(M, N, device="cuda")
    c16 = g16.empty16(M, N, "cuda")
    g16.gemm_group([g16.gemm_problem(x, 0, w, 0, M, N, K, bias=b, C32=c32, C16=c16)])
    assert rel(c32, ref) < 1e-5
    assert rel(c16.float(), ref) < 4e-3
    assert float((c16.float() - c32).abs().max()) <= float(ref.abs().max()) * 2 ** -8


@pytest.mark.parametrize("M,N,K", [(480, 1024, 1024), (320, 300, 1024), (320, 364, 10172), (1280, 128, 512), (37, 64, 300)])
@pytest.mark.parametrize("split", [0, 1, 2, 4])
def test_dx_and_dw_majors(g16, gpu_lib, M, N, K, split):
    """dX = dZ W (A K-major, B MN-major) and dW = dZ^T X (A MN-major, B MN-major) of a linear layer x (M,K) -> (M,N)."""
    gpu_lib.call("bmhrl_gemm_bf16_tuning", 0, split, 0)
    gen = _g(M + N + K)
    x, xf = _mat(M, K, gen)
    w, wf = _mat(N, K, gen, 1 / math.sqrt(K))
    dz, dzf = _mat(M, N, gen)
    dx = torch.empty(M, K, device="cuda")
    g16.gemm_group([g16.gemm_problem(dz, 0, w, 1, M, K, N, C32=dx)])
    assert rel(dx, dzf @ wf) < 1e-5
    dw = torch.full((N, K), 0.5, device="cuda")
    g16.gemm_group([g16.gemm_problem(dz, 1, x, 1, N, K, M, C32=dw, beta=1.0)])
    assert rel(dw, dzf.t() @ xf + 0.5) < 1e-5


def test_epilogue_relu_dropout_residual_gate(g16, gpu_lib):
    from bmhrl_b200 import ops
    gen = _g(3)
    M, N, K = 200, 304, 96
    x, xf = _mat(M, K, gen)
    w, wf = _mat(N, K, gen, 1 / math.sqrt(K))
    b = torch.randn(N, generator=gen).cuda()
    r = torch.randn(M, N, generator=gen).cuda()
    ops.manual_seed(5)
    rng = ops.new_rng_epoch("cuda")
    # forward of fc1: h = dropout(relu(x W^T + b)); same mask as the fp32 path's epilogue (element index m*N + n)
    h16 = g16.empty16(M, N, "cuda")
    h32 = torch.empty(M, N, device="cuda")
    g16.gemm_group([g16.gemm_problem(x, 0, w, 0, M, N, K, bias=b, act=1, drop_p=0.3, rng=rng, salt=9, C32=h32, C16=h16)])
    ops.set_math_mode("fp32")
    y0 = ops.linear(xf, wf, b, act=1, drop_p=0.3, salt=9, rng=rng)
    assert torch.equal(h32 == 0, y0 == 0)
    assert rel(h32, y0) < 1e-5
    keep = float((h32 != 0).float().mean()) / float((torch.relu(xf @ wf.t() + b) != 0).float().mean())
    assert abs(keep - 0.7) < 0.02  # keep rate 1 - p ...
    nz = h32 != 0
    assert rel(h32[nz], (torch.relu(xf @ wf.t() + b) / 0.7)[nz]) < 1e-5  # ... and survivors scaled by 1 / (1 - p)
    # residual + beta
    out = torch.ones(M, N, device="cuda")
    g16.gemm_group([g16.gemm_problem(x, 0, w, 0, M, N, K, bias=b, resid=r, beta=2.0, C32=out)])
    assert rel(out, xf @ wf.t() + b + r + 2.0) < 1e-5
    # gate: the gradient through dropout(relu(.)) is dY * 1/(1-p) where the saved output is > 0
    dy, dyf = _mat(M, 64, gen)
    w2, w2f = _mat(64, N, gen, 1 / math.sqrt(N))   # fc2 weight (64, N): dH = dY W2
    dh = g16.empty16(M, N, "cuda")
    g16.gemm_group([g16.gemm_problem(dy, 0, w2, 1, M, N, 64, gate=h16, gate_scale=1 / 0.7, C16=dh)])
    ref = torch.where(h16.float() > 0, (dyf @ w2f) / 0.7, torch.zeros((), device="cuda"))
    assert rel(dh.float(), ref) < 4e-3


def test_row_segments_and_grouped_problems(g16, gpu_lib):
    """dW of N-concatenated Q|K|V weights lands in three separate slices; four problems of different shapes in ONE launch."""
    from bmhrl_b200._lib import launch_count, reset_launch_count
    gen = _g(11)
    M, D, K = 480, 256, 300
    x, xf = _mat(M, K, gen)
    dz, dzf = _mat(M, 3 * D, gen)
    gq, gk, gv = (torch.zeros(D, K, device="cuda") for _ in range(3))
    g16.gemm_group([g16.gemm_problem(dz, 1, x, 1, 3 * D, K, M, C32=gq, beta=1.0, seg=(D, [gk, gv]))])
    ref = dzf.t() @ xf
    assert rel(gq, ref[:D]) < 1e-5 and rel(gk, ref[D:2 * D]) < 1e-5 and rel(gv, ref[2 * D:]) < 1e-5
    probs, refs, outs = [], [], []
    for (m, n, k) in [(480, 1024, 1024), (1280, 128, 128), (320, 300, 300), (37, 64, 520)]:
        a, af = _mat(m, k, gen)
        b, bf = _mat(n, k, gen, 1 / math.sqrt(k))
        c = torch.empty(m, n, device="cuda")
        probs.append(g16.gemm_problem(a, 0, b, 0, m, n, k, C32=c))
        refs.append(af @ bf.t())
        outs.append(c)
    reset_launch_count()
    g16.gemm_group(probs)
    assert launch_count() == 1
    for c, r in zip(outs, refs):
        assert rel(c, r) < 1e-5
