#!/usr/bin/env python
"""One recurrent layer of the frozen critic (B=16, L=20, H=600), the three variants, alone on the GPU: us per layer."""
import math
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bmhrl_b200 import ops
from bmhrl_b200._lib import LIB


def main():
    LIB.load()
    ops.set_math_mode("tf32")
    g = torch.Generator().manual_seed(1)
    for cell, NG in (("lstm", 4), ("gru", 3)):
        H, E, B = 600, 600, 16
        w_ih = (torch.randn(NG * H, E, generator=g) / math.sqrt(E)).cuda()
        w_hh = (torch.randn(NG * H, H, generator=g) / math.sqrt(H)).cuda()
        b_ih = torch.zeros(NG * H).cuda(); b_hh = torch.zeros(NG * H).cuda()
        for L in (20, 40):
            x = torch.randn(B, L, E, generator=g).cuda()
            for name, (p, e) in {"exchange": (1, 1), "barrier": (1, 0), "per-step": (0, 0)}.items():
                LIB.call("bmhrl_rnn_set_persistent", p); LIB.call("bmhrl_rnn_set_exchange", e)
                for _ in range(3):
                    ops.rnn_layer(cell, x, w_ih, w_hh, b_ih, b_hh)
                torch.cuda.synchronize()
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr):
                    for _ in range(6):
                        ops.rnn_layer(cell, x, w_ih, w_hh, b_ih, b_hh)
                gr.replay(); torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); gr.replay(); gr.replay(); e1.record(); torch.cuda.synchronize()
                print(f"{cell} B={B} L={L} H={H} {name:9s}: {e0.elapsed_time(e1) * 1e3 / 12:8.1f} us per layer (input GEMM included)", flush=True)
            # where a time step of the exchange kernel goes (CTA 0, clock64 stamps, averaged over steps 2 .. L-1)
            LIB.call("bmhrl_rnn_set_persistent", 1); LIB.call("bmhrl_rnn_set_exchange", 1)
            slots = torch.zeros(480, dtype=torch.int64, device="cuda")
            LIB.call("bmhrl_rnn_trace", slots.data_ptr())
            ops.rnn_layer(cell, x, w_ih, w_hh, b_ih, b_hh)
            torch.cuda.synchronize()
            LIB.call("bmhrl_rnn_trace", 0)
            s = slots.cpu().view(60, 8).tolist()
            n = min(L, 60)
            names = ["hint seen", "tile loaded", "barrier", "dot products", "reduce+gates+store"]
            acc = [0.0] * 5
            for t in range(2, n):
                for i in range(5):
                    acc[i] += (s[t][i + 1] - s[t][i]) / (n - 2)
            per_step = (s[n - 1][0] - s[2][0]) / (n - 3)
            print(f"   exchange kernel, CTA 0, clocks per step {per_step:.0f} (~{per_step / 1.9e3:.2f} us at 1.9 GHz): "
                  + ", ".join(f"{nm} {a:.0f}" for nm, a in zip(names, acc)), flush=True)
    LIB.call("bmhrl_rnn_set_persistent", 1); LIB.call("bmhrl_rnn_set_exchange", -1)


if __name__ == "__main__":
    main()
