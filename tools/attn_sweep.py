"""BASELINE.json config 5: long-clip encoder attention sweep (B=64, H=4, d_k=256, bf16) on one B200.
Times the bf16 flash-style forward (`ops.mha_core_bf16`, attn_bf16.cu) with CUDA events around a CUDA graph of `reps`
back-to-back launches (inputs 39-105 MB per call; the 64-clip working set cycles through L2 between launches), next to
the fp32-operand training kernel (`ops.mha_core`, tf32 tensor cores, writes the probabilities for backward).
Algorithmic FLOPs = 4*B*H*Sq*Sk*d_k; peak = MEASURED_PEAKS.json bf16_tflops (burst: the
kernel is timed alone).  usage: python tools/attn_sweep.py [--quick] > gpurun_out/attn_sweep.jsonl"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from bmhrl_b200 import ops  # noqa: E402

dev = torch.device("cuda")


def timeit(fn, reps=10, iters=5):
    fn()
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
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
        best = min(best, e0.elapsed_time(e1) * 1e-3 / reps)
    return best


def main():
    quick = "--quick" in sys.argv
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]
        src = "measured bf16_tflops (burst)"
    except Exception:
        pk, src = 1590.0, "fallback"
    B, H, dk = 64, 4, 256
    if os.environ.get("ATTN_PAIR"):  # 1 = one CTA per query tile, 2 = CTA pairs (cta_group::2)
        from bmhrl_b200._lib import LIB
        LIB.call("bmhrl_attn_bf16_tuning", int(os.environ["ATTN_PAIR"]))
    D = H * dk
    shapes = [(300, 300), (800, 800), (300, 800), (800, 300)] if not quick else [(300, 300)]
    if os.environ.get("ATTN_SHAPES"):  # e.g. ATTN_SHAPES=800x800,300x300
        shapes = [tuple(int(v) for v in t.split("x")) for t in os.environ["ATTN_SHAPES"].split(",")]
    for (Sq, Sk) in shapes:
        g = torch.Generator(device="cuda").manual_seed(Sq * 1000 + Sk)
        q = torch.randn(B, Sq, D, device=dev, generator=g).bfloat16()
        k = torch.randn(B, Sk, D, device=dev, generator=g).bfloat16()
        v = torch.randn(B, Sk, D, device=dev, generator=g).bfloat16()
        for mname in ("none", "pad50", "full_row_masked"):
            mask = torch.ones(B, 1, Sk, dtype=torch.bool, device=dev)
            if mname == "pad50":
                mask[:, :, Sk // 2:] = False
            elif mname == "full_row_masked":
                mask[0] = False
            m = None if mname == "none" else mask
            with torch.no_grad():
                t = timeit(lambda: ops.mha_core_bf16(q, k, v, m, H))
            fl = 4.0 * B * H * Sq * Sk * dk
            line = {"kernel": "attn_fwd_bf16_kernel", "B": B, "H": H, "Sq": Sq, "Sk": Sk, "dk": dk, "mask": mname,
                    "us": t * 1e6, "tflops": fl / t / 1e12, "peak_tflops": pk, "frac": fl / t / 1e12 / pk,
                    "peak_source": src, "dtype": "bf16", "pass": "fwd",
                    "flops": "nominal 4*B*H*Sq*Sk*dk" + (" (key tiles behind the last visible key are skipped: their "
                                                          "probabilities are exactly 0)" if mname == "pad50" else "")}
            print(json.dumps(line), flush=True)
        if "--bwd" in sys.argv:  # backward (attn_bf16_bwd.cu): rowsum(dO o O) + the dQ, dK, dV passes; nominal 2.5 x forward FLOPs
            from bmhrl_b200._lib import LIB
            do = torch.randn(B, Sq, D, device=dev, generator=g).bfloat16()
            for mname in ("none", "pad50"):
                m = None
                if mname == "pad50":
                    m = torch.ones(B, 1, Sk, dtype=torch.bool, device=dev)
                    m[:, :, Sk // 2:] = False
                with torch.no_grad():
                    o, lse = ops.mha_core_bf16(q, k, v, m, H, return_lse=True)
                    fl = 10.0 * B * H * Sq * Sk * dk
                    for passes, pname in ((7, "bwd (delta + dQ + dK + dV)"), (1, "delta + dQ"), (2, "delta + dK"), (4, "delta + dV")):
                        LIB.call("bmhrl_attn_bwd_bf16_tuning", passes)
                        t = timeit(lambda: ops.mha_core_bf16_bwd(q, k, v, o, lse, do, m, H), reps=4, iters=3)
                        line = {"kernel": "attn_bwd_bf16_kernel", "B": B, "H": H, "Sq": Sq, "Sk": Sk, "dk": dk, "mask": mname,
                                "pass": pname, "us": t * 1e6, "dtype": "bf16"}
                        if passes == 7:
                            line.update({"tflops": fl / t / 1e12, "peak_tflops": pk, "frac": fl / t / 1e12 / pk,
                                         "flops": "nominal 10*B*H*Sq*Sk*dk (executed: 16, S is recomputed in every pass)"})
                        print(json.dumps(line), flush=True)
                    LIB.call("bmhrl_attn_bwd_bf16_tuning", 7)
        if not quick:
            ops.set_math_mode("tf32")
            qf, kf, vf = q.float(), k.float(), v.float()
            with torch.no_grad():
                t = timeit(lambda: ops.mha_core(qf, kf, vf, None, H), reps=4, iters=3)
            fl = 4.0 * B * H * Sq * Sk * dk
            print(json.dumps({"kernel": "attn_fwd_tc_kernel (fp32 operands, tf32 MMA, stores P)", "B": B, "H": H,
                              "Sq": Sq, "Sk": Sk, "dk": dk, "mask": "none", "us": t * 1e6, "tflops": fl / t / 1e12,
                              "peak_tflops": pk, "frac": fl / t / 1e12 / pk, "dtype": "tf32", "pass": "fwd"}),
                  flush=True)
            del qf, kf, vf


if __name__ == "__main__":
    main()
