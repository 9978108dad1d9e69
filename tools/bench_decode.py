"""Config 1 (single-clip greedy captioning, B=1, Sv=14, Sa=36, 30 generated tokens, random-init BMHRL): tokens/s of the
full-recompute decoder (the reference's algorithm, captioning_bmrl_loops.py:127-152), of the encode-once decoder and of
the graphed decoder (encoder + memory K/V once per clip, one CUDA-graph replay per token),
timed with CUDA events on the current stream.  Prints one JSON line.

    python tools/bench_decode.py [--B 1] [--tokens 30] [--reps 5]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=1)
    ap.add_argument("--tokens", type=int, default=30)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--only", default="", help="comma list of decoders to time (full_recompute,encode_once,graphed)")
    ap.add_argument("--modes", default="fp32,tf32")
    ap.add_argument("--profile", action="store_true",
                    help="bracket ONE graphed caption with cudaProfilerStart/Stop (ncu --profile-from-start off)")
    a = ap.parse_args()
    from bmhrl_b200 import ops, synth
    from bmhrl_b200.decode import GraphedGreedyDecoder, greedy_decoder, greedy_decoder_cached
    from conftest import PAD, VOC, build_agent
    golden = torch.load(os.path.join(ROOT, "tests", "golden", "reference_golden.pt"), weights_only=False)
    sd = synth.fill_state_dict(golden["agent_shapes"], seed=1234)
    model, _ = build_agent(sd)
    model.eval()
    model.set_inference_mode(True)
    b = synth.synthetic_batch(B=a.B, Sv=14, Sa=36, L=6, V=VOC, seed=3)
    fs = {k: b[k].cuda() for k in ("rgb", "flow", "audio")}
    out = {"metric": "greedy_decode_tokens_per_sec", "unit": "tokens/s",
           "config": {"workload": f"config 1 shape: B={a.B}, Sv=14, Sa=36, {a.tokens} generated tokens, V={VOC}, "
                                  f"random-init BMHRL, end token disabled so every run generates {a.tokens} tokens"}}
    only = set(filter(None, a.only.split(",")))
    for mode in a.modes.split(","):
        ops.set_math_mode(mode)
        res = {}
        toks = {}
        # graphed: static buffers a little larger than the clip (Sv 16, Sa 40), one completeness read-back per caption
        dec = GraphedGreedyDecoder(model, a.B, 16, 40, a.tokens, 2, -1, PAD, check_every=a.tokens)
        for name, fn in (("full_recompute", greedy_decoder), ("encode_once", greedy_decoder_cached),
                         ("graphed", None)):
            if only and name not in only:
                continue
            run = (lambda: dec(fs)) if fn is None else (lambda: fn(model, fs, a.tokens, 2, -1, PAD, "audio_video"))
            for _ in range(2):
                toks[name] = run()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.reps):
                run()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / a.reps
            res[name] = {"ms_per_caption": ms, "tokens_per_s": 1e3 * a.B * a.tokens / ms}
        if a.profile:
            torch.cuda.synchronize()
            torch.cuda.profiler.start()
            dec(fs)
            torch.cuda.synchronize()
            torch.cuda.profiler.stop()
        if not only:
            res["same_tokens"] = bool(torch.equal(toks["full_recompute"], toks["encode_once"]))
            res["same_tokens_graphed"] = bool(torch.equal(toks["full_recompute"], toks["graphed"]))
            res["speedup"] = res["full_recompute"]["ms_per_caption"] / res["encode_once"]["ms_per_caption"]
            res["speedup_graphed"] = res["full_recompute"]["ms_per_caption"] / res["graphed"]["ms_per_caption"]
            res["graph_replays_per_caption"] = dec.replays // (2 + a.reps)
        out[mode] = res
    ops.set_math_mode("fp32")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
