#!/usr/bin/env python
"""Microbenchmark: bf16 grouped tcgen05 GEMM (gemm_bf16_tc_kernel) vs the tf32 kernel on the training step's shapes.
Weights rotate through > 126 MB of copies (cold: every launch streams W from HBM, as in the step); launches are
captured in one CUDA graph so that the number is kernel time + graph-node latency, not Python."""
import json
import math
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bmhrl_b200 import ops, ops16
from bmhrl_b200._lib import LIB

SHAPES = [  # (M, N, K, what)
    (480, 3072, 1024, "video self-att Q|K|V"), (480, 1024, 1024, "video out-proj / FFN"),
    (1280, 3072, 128, "audio self-att Q|K|V"), (1280, 128, 1024, "audio out-proj"),
    (1280, 2048, 128, "K|V of audio memory"), (480, 2048, 1024, "K|V of video memory"),
    (320, 3072, 300, "caption self-att Q|K|V"), (320, 300, 1024, "caption out-proj"),
    (320, 10172, 364, "vocabulary projection"), (4096, 4096, 4096, "large")]


def timed(fn, iters):
    g = torch.cuda.CUDAGraph()
    fn(0)
    torch.cuda.synchronize()
    with torch.cuda.graph(g):
        for i in range(iters):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def main():
    LIB.load()
    out = []
    for M, N, K, what in SHAPES:
        ncopy = max(2, min(64, int(160e6 / (N * K * 2)) + 1))
        iters = ncopy * 2
        x32 = torch.randn(M, K, device="cuda")
        w32 = [torch.randn(N, K, device="cuda") / math.sqrt(K) for _ in range(ncopy)]
        x16 = ops16.empty16(M, K, "cuda"); x16.copy_(x32)
        w16 = []
        for w in w32:
            t = ops16.empty16(N, K, "cuda"); t.copy_(w); w16.append(t)
        c32 = torch.empty(M, N, device="cuda")
        c16 = ops16.empty16(M, N, "cuda")
        dz16 = ops16.empty16(M, N, "cuda"); dz16.copy_(torch.randn(M, N, device="cuda"))
        dx = torch.empty(M, K, device="cuda")
        dw = torch.zeros(N, K, device="cuda")
        rec = {"M": M, "N": N, "K": K, "what": what, "flops": 2.0 * M * N * K}
        ops.set_math_mode("tf32")
        with torch.no_grad():
            rec["tf32_fwd_us"] = timed(lambda i: ops.linear(x32, w32[i % ncopy]), iters)
        rec["bf16_fwd_us"] = timed(lambda i: ops16.gemm_group([ops16.gemm_problem(x16, 0, w16[i % ncopy], 0, M, N, K, C16=c16)]), iters)
        rec["bf16_fwd_hotW_us"] = timed(lambda i: ops16.gemm_group([ops16.gemm_problem(x16, 0, w16[0], 0, M, N, K, C16=c16)]), iters)
        rec["bf16_dx_hotW_us"] = timed(lambda i: ops16.gemm_group([ops16.gemm_problem(dz16, 0, w16[0], 1, M, K, N, C32=dx)]), iters)
        rec["bf16_fwd32_us"] = timed(lambda i: ops16.gemm_group([ops16.gemm_problem(x16, 0, w16[i % ncopy], 0, M, N, K, C32=c32)]), iters)
        rec["bf16_dx_us"] = timed(lambda i: ops16.gemm_group([ops16.gemm_problem(dz16, 0, w16[i % ncopy], 1, M, K, N, C32=dx)]), iters)
        rec["bf16_dw_us"] = timed(lambda i: ops16.gemm_group([ops16.gemm_problem(dz16, 1, x16, 1, N, K, M, C32=dw, beta=1.0)]), iters)
        for bn, sp in ((128, 1), (64, 1), (64, 2)):
            LIB.call("bmhrl_gemm_bf16_tuning", bn, sp, 0)
            rec[f"bf16_fwd_bn{bn}_s{sp}_us"] = timed(
                lambda i: ops16.gemm_group([ops16.gemm_problem(x16, 0, w16[i % ncopy], 0, M, N, K, C16=c16)]), iters)
        LIB.call("bmhrl_gemm_bf16_tuning", 0, 0, 0)
        rec["bf16_fwd_tflops"] = rec["flops"] / rec["bf16_fwd_us"] / 1e6
        rec["tf32_fwd_tflops"] = rec["flops"] / rec["tf32_fwd_us"] / 1e6
        print(json.dumps({k: (round(v, 2) if isinstance(v, float) else v) for k, v in rec.items()}), flush=True)
        out.append(rec)
        del w32, w16
    ops.set_math_mode("fp32")


if __name__ == "__main__":
    main()
