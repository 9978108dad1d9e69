"""Skinny linear layer (M <= 32 rows, tensor-core math mode) against HBM bandwidth: the kernel reads every weight once,
so its algorithmic bytes are (N*K + M*K + M*N) * 4 per launch.  Weights rotate through enough copies to exceed the
126 MB L2, so every launch streams its W from HBM.  The same calls with the kernel switched off (tcgen05 path) are timed
next to it.  The launches of a shape are captured into one CUDA graph (no Python issue cost in the timing); CUDA events on the
launching stream around one replay, after a warm-up replay.  Prints one JSON line.

    python tools/bench_skinny.py
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SHAPES = [(1, 364, 10172, "vocabulary projection, last position, B=1"),
          (16, 364, 10172, "vocabulary projection, last position, B=16"),
          (30, 1024, 1024, "video-width projection over a 30-token prefix / memory K,V"),
          (1, 300, 2400, "critic LSTM input contribution, one position"),
          (1, 600, 2400, "critic LSTM layers 2-4 input contribution, one position"),
          (30, 300, 1024, "caption Q/K/V projection over a 30-token prefix")]


def main():
    from bmhrl_b200 import ops
    from bmhrl_b200._lib import LIB
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = None
    for k, v in (peaks.items() if isinstance(peaks, dict) else []):
        if "hbm" in k.lower() and isinstance(v, (int, float)):
            hbm = float(v) if hbm is None else max(hbm, float(v))
    ops.set_math_mode("tf32")
    out = {"metric": "skinny_linear_hbm_fraction", "peak_hbm_gbs": hbm, "shapes": []}
    g = torch.Generator().manual_seed(0)
    for M, K, N, what in SHAPES:
        copies = max(2, int(200e6 // (N * K * 4)) + 1)
        ws = [torch.randn(N, K, generator=g).cuda() for _ in range(copies)]
        b = torch.randn(N, generator=g).cuda()
        x = torch.randn(M, K, generator=g).cuda()
        row = {"M": M, "K": K, "N": N, "what": what, "weight_copies": copies,
               "algorithmic_bytes": (N * K + M * K + M * N) * 4}
        for name, on in (("skinny", 1), ("tcgen05", 0)):
            LIB.call("bmhrl_set_skinny_linear", on)
            iters = 3 * copies
            with torch.no_grad():
                for w in ws[:3]:
                    ops.linear(x, w, b)
                torch.cuda.synchronize()
                # the launches are captured into ONE CUDA graph: the Python / ctypes issue cost stays out of the timing
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    for i in range(iters):
                        ops.linear(x, ws[i % copies], b)
            graph.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            graph.replay()
            e1.record()
            torch.cuda.synchronize()
            del graph
            us = 1e3 * e0.elapsed_time(e1) / iters
            row[name] = {"us_per_launch": us, "gbs": row["algorithmic_bytes"] / us / 1e3,
                         "hbm_frac": (row["algorithmic_bytes"] / us / 1e3 / hbm) if hbm else None}
        out["shapes"].append(row)
        del ws
    LIB.call("bmhrl_set_skinny_linear", 1)
    ops.set_math_mode("fp32")
    out["note"] = ("per-launch time = one graph replay of 3 x weight_copies serial launches / their number: it includes "
                   "the ~1-2 us gap between dependent graph nodes, which bounds the small shapes")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
