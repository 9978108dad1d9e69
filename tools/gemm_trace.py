#!/usr/bin/env python
"""Where one CTA of gemm_bf16_tc_kernel spends its time: CTA 0 stamps clock64() at its hand-over points
(bmhrl_gemm_bf16_trace) and this prints the intervals in nanoseconds for the step's shapes, weights cold (rotated through
> 126 MB) and hot, launches back to back in one CUDA graph (so the grid dependency of the previous launch is in the picture).
usage: python tools/gemm_trace.py > gpurun_out/gemm_trace.txt"""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bmhrl_b200 import ops16
from bmhrl_b200._lib import LIB

SHAPES = [(480, 3072, 1024, "video Q|K|V"), (480, 1024, 1024, "video out-proj"), (1280, 3072, 128, "audio Q|K|V"),
          (1280, 128, 1024, "audio out-proj"), (320, 3072, 300, "caption Q|K|V"), (320, 300, 1024, "caption out-proj")]
NAMES = ["entry", "setup", "dep", "ring", "land0", "mma_end", "acc_vis", "epi_end", "exit"]


def run(M, N, K, what, hot, out_kind):
    ncopy = 1 if hot else max(2, min(64, int(160e6 / (N * K * 2)) + 1))
    x16 = ops16.empty16(M, K, "cuda"); x16.copy_(torch.randn(M, K, device="cuda"))
    w16 = []
    for _ in range(ncopy):
        t = ops16.empty16(N, K, "cuda"); t.copy_(torch.randn(N, K, device="cuda") / math.sqrt(K)); w16.append(t)
    c16 = ops16.empty16(M, N, "cuda")
    c32 = torch.empty(M, N, device="cuda")
    slots = torch.zeros(64, dtype=torch.int64, device="cuda")
    kw = {"C16": c16} if out_kind == "16" else {"C32": c32}
    iters = max(ncopy, 8)

    def fn(i):
        ops16.gemm_group([ops16.gemm_problem(x16, 0, w16[i % ncopy], 0, M, N, K, b_static=True, **kw)])
    fn(0)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    LIB.call("bmhrl_gemm_bf16_trace", slots.data_ptr())
    with torch.cuda.graph(g):
        for i in range(iters):
            fn(i)
    LIB.call("bmhrl_gemm_bf16_trace", 0)
    acc = None
    reps = 5
    for _ in range(reps):      # the stamps of the LAST launch of the graph survive
        g.replay()
        torch.cuda.synchronize()
        s = slots.cpu().tolist()
        v = [s[i] - s[0] for i in range(9)] + [s[10] - s[9]] + [s[i] - s[0] for i in (11, 12, 13)] + \
            [s[16 + i] - s[0] for i in range(min(48, (K + 63) // 64))]
        acc = v if acc is None else [a + b for a, b in zip(acc, v)]
    v = [a / reps for a in acc]
    clk_per_ns = v[8] / max(1.0, v[9])
    ns = [x / clk_per_ns for x in v[:9]]
    epi = [x / clk_per_ns for x in v[10:13]]
    lands = [x / clk_per_ns for x in v[13:]]
    print(f"{what:18s} M={M} N={N} K={K} W {'hot' if hot else 'cold'} out bf{out_kind}: CTA 0 lives {v[9]:.0f} ns "
          f"({clk_per_ns:.2f} clk/ns)")
    print("   " + "  ".join(f"{n}={t:.0f}" for n, t in zip(NAMES, ns)))
    print(f"   epilogue: first chunk in registers {epi[0]:.0f}, first pass staged {epi[1]:.0f}, first pass stored {epi[2]:.0f}")
    print("   stage landed (ns after entry): " + " ".join(f"{t:.0f}" for t in lands))


def main():
    LIB.load()
    print(f"# PDL {'on' if LIB.raw("bmhrl_get_pdl")() else 'off'}; tuning from the environment "
          f"(BMHRL_GB_BN={os.environ.get('BMHRL_GB_BN')}, BMHRL_GB_STAGES={os.environ.get('BMHRL_GB_STAGES')})")
    for M, N, K, what in SHAPES:
        for hot in (False, True):
            run(M, N, K, what, hot, "16")
    run(480, 1024, 1024, "video out-proj", False, "32")


if __name__ == "__main__":
    main()
