#!/usr/bin/env python
"""Isolated timings (CUDA graph of 20 back-to-back launches, CUDA events) of the bf16-mode kernels at the training step's
shapes: one-tile attention forward / backward, LayerNorm forward (bf16 out) / backward, cast, column sums."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bmhrl_b200 import ops, ops16
from bmhrl_b200._lib import LIB


def timed(fn, iters=20):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters):
            fn()
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
    dev = "cuda"
    for (B, H, Sq, Sk, dk, what) in [(16, 4, 30, 30, 256, "video self"), (16, 4, 80, 80, 256, "audio self"),
                                     (16, 4, 30, 80, 256, "video<-audio"), (16, 4, 80, 30, 256, "audio<-video"),
                                     (16, 4, 20, 20, 256, "caption self"), (16, 4, 20, 30, 256, "caption<-video"),
                                     (16, 4, 20, 80, 256, "caption<-audio"), (16, 2, 20, 20, 512, "goal attention")]:
        D = H * dk
        q = ops16.empty16(B * Sq, D, dev); q.normal_()
        k = ops16.empty16(B * Sk, D, dev); k.normal_()
        v = ops16.empty16(B * Sk, D, dev); v.normal_()
        do = ops16.empty16(B * Sq, D, dev); do.normal_()
        o = ops16.empty16(B * Sq, D, dev)
        lse = torch.empty(B, H, Sq, 2, device=dev)
        mask = torch.ones(B, Sk, dtype=torch.uint8, device=dev)
        dq, dk_, dv = ops16.empty16(B * Sq, D, dev), ops16.empty16(B * Sk, D, dev), ops16.empty16(B * Sk, D, dev)
        f = timed(lambda: ops16.attn16_fwd(q, k, v, mask, 1, False, B, H, Sq, Sk, dk, o, lse))
        b = timed(lambda: ops16.attn16_bwd(q, k, v, do, lse, mask, 1, False, B, H, Sq, Sk, dk, dq, dk_, dv))
        print(json.dumps({"kernel": "attn16", "what": what, "B": B, "H": H, "Sq": Sq, "Sk": Sk, "dk": dk, "fwd_us": round(f, 2),
                          "bwd_us": round(b, 2), "fwd_bwd_us": round(f + b, 2)}), flush=True)
    for rows, D in [(480, 1024), (1280, 128), (320, 300)]:
        x = torch.randn(rows, D, device=dev)
        g_, b_ = torch.ones(D, device=dev), torch.zeros(D, device=dev)
        f = timed(lambda: ops16.ln_fwd16(x, g_, b_, 1e-5))
        y, mean, rstd = ops16.ln_fwd16(x, g_, b_, 1e-5)
        dy = torch.randn(rows, D, device=dev)
        bw = timed(lambda: ops16.ln_bwd(dy, x, g_, mean, rstd, dy, True, None, None, False))
        ba = timed(lambda: ops16.ln_bwd(dy, x, g_, mean, rstd, None, False, None, None, True))
        c = timed(lambda: ops16.cast16(dy))
        z = ops16.cast16(dy)
        out = torch.zeros(D, device=dev)
        cs = timed(lambda: ops16.colsum16(z, [out]))
        print(json.dumps({"kernel": "ln/cast/colsum", "rows": rows, "D": D, "ln_fwd16_us": round(f, 2), "ln_bwd_dx_us": round(bw, 2),
                          "ln_bwd_affine_us": round(ba, 2), "cast16_us": round(c, 2), "colsum16_us": round(cs, 2)}), flush=True)


if __name__ == "__main__":
    main()
