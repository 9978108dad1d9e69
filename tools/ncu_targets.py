#!/usr/bin/env python
"""A short program for ncu: launches each bf16-mode kernel of the training step a few times at its step shape
(weights / optimiser state larger than L2 where the step streams them from HBM).
    ncu --set full --clock-control none --import-source on -k regex:<kernel> -s 2 -c 1 -o gpurun_out/<name> python tools/ncu_targets.py <what>"""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bmhrl_b200 import ops, ops16
from bmhrl_b200._lib import LIB


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    LIB.load()
    dev = "cuda"
    reps = 4
    if what in ("gemm", "all"):
        M, N, K = 480, 3072, 1024    # video self-attention Q|K|V
        x = ops16.empty16(M, K, dev); x.normal_()
        ws = []
        for _ in range(24):          # 24 x 6.3 MB of weights: every launch streams W from HBM, like the step
            w = ops16.empty16(N, K, dev); w.normal_(); ws.append(w)
        c = ops16.empty16(M, N, dev)
        for i in range(reps):
            ops16.gemm_group([ops16.gemm_problem(x, 0, ws[i], 0, M, N, K, C16=c)])
        dz = ops16.empty16(M, N, dev); dz.normal_()
        dw = torch.zeros(N, K, device=dev)
        for i in range(reps):
            ops16.gemm_group([ops16.gemm_problem(dz, 1, x, 1, N, K, M, C32=dw, beta=1.0)])
    if what in ("attn", "all"):
        B, H, S, dk = 16, 4, 80, 256
        D = H * dk
        q, k, v, do = (ops16.empty16(B * S, D, dev) for _ in range(4))
        for t in (q, k, v, do):
            t.normal_()
        o = ops16.empty16(B * S, D, dev)
        lse = torch.empty(B, H, S, 2, device=dev)
        mask = torch.ones(B, S, dtype=torch.uint8, device=dev)
        dq, dk_, dv = (ops16.empty16(B * S, D, dev) for _ in range(3))
        for _ in range(reps):
            ops16.attn16_fwd(q, k, v, mask, 1, False, B, H, S, S, dk, o, lse)
        for _ in range(reps):
            ops16.attn16_bwd(q, k, v, do, lse, mask, 1, False, B, H, S, S, dk, dq, dk_, dv)
    if what in ("norm", "all"):
        rows, D = 480, 1024
        x = torch.randn(rows, D, device=dev)
        g_, b_ = torch.ones(D, device=dev), torch.zeros(D, device=dev)
        dy = torch.randn(rows, D, device=dev)
        for _ in range(reps):
            y, mean, rstd = ops16.ln_fwd16(x, g_, b_, 1e-5)
        for _ in range(reps):
            ops16.ln_bwd(dy, x, g_, mean, rstd, dy, True, None, None, True)
        for _ in range(reps):
            z = ops16.cast16(dy)
        out = torch.zeros(D, device=dev)
        for _ in range(reps):
            ops16.colsum16(z, [out])
    if what in ("loss", "all"):
        M, K, V = 320, 364, 10172
        import torch.nn as nn
        lin = nn.Linear(K, V).to(dev)
        x = torch.randn(M, K, device=dev, requires_grad=True)
        trg = torch.randint(2, V, (M,), device=dev)
        for _ in range(reps):
            st, tok = ops16.vocab_stats(x, lin.weight, lin.bias, trg, 1, "argmax")
            rows = ops16.kl_rowsum_from_stats(st, trg, 0.7, 1, V)
            rows.sum().backward()
            ops.join_side_streams(dev)
    if what in ("adam", "all"):
        ps = [torch.nn.Parameter(torch.randn(1024, 1024, device=dev)) for _ in range(48)]   # 50 M parameters
        opt = ops.FusedAdam(ps)
        opt.flat_grad.normal_()
        for _ in range(reps):
            opt.step()
    torch.cuda.synchronize()
    print("ncu_targets done:", what)


if __name__ == "__main__":
    main()
