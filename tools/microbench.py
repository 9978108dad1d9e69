"""Device-side timings of single ops of libbmhrl_b200 (CUDA events around a CUDA-graph of `reps` back-to-back calls).
usage: python tools/microbench.py [critic] [gemm] [ln] [attn] ..."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from bmhrl_b200 import ops  # noqa: E402

dev = torch.device("cuda")


def timeit(fn, reps=20, iters=5, name=""):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(iters):
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / reps)
    print(f"{name:60s} {best:9.2f} us")
    return best


def critic():
    from types import SimpleNamespace as NS
    from bmhrl_b200.model.bm_hrl_agent import SegmentCritic
    c = SegmentCritic(NS(d_model_caps=300, rl_critic_path=None)).to(dev)
    x = torch.randn(16, 20, 300, device=dev)
    for mode in ("fp32", "tf32"):
        ops.set_math_mode(mode)
        timeit(lambda: c.scores_and_labels(x, 0.25), name=f"SegmentCritic B=16 L=20 ({mode} input GEMMs)")
    from bmhrl_b200._lib import LIB
    LIB.raw("bmhrl_rnn_set_persistent")(0)
    timeit(lambda: c.scores_and_labels(x, 0.25), name="SegmentCritic, per-step kernels")
    LIB.raw("bmhrl_rnn_set_persistent")(1)


def gemm():
    ops.set_math_mode("tf32")
    for (M, N, K) in [(480, 1024, 1024), (1280, 1024, 128), (1280, 128, 1024), (320, 1024, 300), (320, 300, 1024),
                      (320, 10172, 364), (1280, 512, 128), (480, 1024, 128), (1280, 1024, 1024), (4096, 4096, 4096)]:
        x = torch.randn(M, K, device=dev)
        w = torch.randn(N, K, device=dev)
        b = torch.randn(N, device=dev)
        t = timeit(lambda: ops.linear(x, w, b), name=f"linear fwd M={M} N={N} K={K}")
        print(f"{'':60s} {2.0 * M * N * K / t / 1e6:9.1f} TFLOP/s")
        dy = torch.randn(M, N, device=dev)
        dx = torch.empty(M, K, device=dev)
        dw = torch.empty(N, K, device=dev)
        timeit(lambda: ops.gemm(dy, N, 1, w, K, 1, dx, K, M, K, N), name="   dX = dY W")
        timeit(lambda: ops.gemm(dy, 1, N, x, K, 1, dw, K, N, K, M), name="   dW = dY^T X")


def ln():
    for (rows, D) in [(480, 1024), (1280, 128), (320, 300)]:
        x = torch.randn(rows, D, device=dev, requires_grad=True)
        g = torch.randn(D, device=dev, requires_grad=True)
        b = torch.randn(D, device=dev, requires_grad=True)
        timeit(lambda: ops.layer_norm(x, g, b), name=f"layer_norm fwd rows={rows} D={D}")
        dy = torch.randn(rows, D, device=dev)
        timeit(lambda: torch.autograd.grad(ops.layer_norm(x, g, b), (x, g, b), dy), name="   fwd + bwd (dx, dgamma, dbeta)")


def attn(mode="tf32"):
    ops.set_math_mode(mode)
    print(f"-- attention core, math mode {mode}")
    for (B, H, Sq, Sk, D) in [(16, 4, 30, 30, 1024), (16, 4, 80, 80, 1024), (16, 4, 20, 80, 1024), (16, 2, 20, 20, 1024),
                              (64, 4, 300, 300, 1024)]:
        q = torch.randn(B, Sq, D, device=dev, requires_grad=True)
        k = torch.randn(B, Sk, D, device=dev, requires_grad=True)
        v = torch.randn(B, Sk, D, device=dev, requires_grad=True)
        mask = torch.ones(B, 1, Sk, dtype=torch.bool, device=dev)
        t = timeit(lambda: ops.mha_core(q, k, v, mask, H), name=f"mha_core fwd B={B} H={H} Sq={Sq} Sk={Sk} dk={D // H}")
        print(f"{'':60s} {4.0 * B * Sq * Sk * D / t / 1e6:9.2f} TFLOP/s")
        do = torch.randn(B, Sq, D, device=dev)
        timeit(lambda: torch.autograd.grad(ops.mha_core(q, k, v, mask, H), (q, k, v), do), name="   fwd + bwd")


if __name__ == "__main__":
    which = sys.argv[1:] or ["critic", "gemm", "ln", "attn"]
    for w in which:
        globals()[w]()
