#!/usr/bin/env python
"""bench.py -- BMHRL train-step throughput (clips/s, B=16 per GPU) on N B200s, or the CPU reference arm.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload warmstart|worker|manager] [--impl reference]

N > 1 is launched by torch.distributed.run (one rank per GPU, NCCL).  Prints ONE JSON line (rank 0).
  value     device-timed whole-job clips/s, inputs resident in HBM, one CUDA-graph replay per step
  e2e       same step through the public API (StepRunner.step) from pinned HOST buffers: H2D copies of the batch and
            the D2H read of the loss inside the timed region
  roofline  the dominant kernel class (projection / FFN / vocab / gradient GEMMs): the step's GEMM launches are
            re-issued with their exact shapes and timed with CUDA events; achieved = algorithmic FLOPs / time
  cpu_baseline  the oracle port of the reference step timed on this box's host cores (bounded sample)
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

VOC, PAD, SMOOTH = 10172, 1, 0.7
SHAPE = dict(B=16, Sv=30, Sa=80, L=20)  # SURVEY.md section 8 "C16"
METRIC = "bmhrl_train_step_clips_per_sec"


def cfg_ns(device):
    from types import SimpleNamespace as NS
    return NS(d_vid=1024, d_aud=128, d_model_video=1024, d_model_audio=128, d_model=1024, d_model_caps=300,
              rl_projection_d=512, rl_att_heads=4, rl_att_layers=2, rl_goal_d=64, rl_ff_v=1024, rl_ff_a=512,
              rl_ff_c=2048, dout_p=0.1, rl_critic_score_threshhold=0.25, unfreeze_word_emb=False, device=device,
              device_ids=[torch.cuda.current_device()] if device == "cuda" else [0], rl_critic_path=None)


def agent_shapes():
    g = torch.load(os.path.join(ROOT, "tests", "golden", "reference_golden.pt"), weights_only=False)
    return g["agent_shapes"], g["value_fn_shapes"]


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


# ------------------------------------------------------------------------------------------------ CPU reference arm
def cpu_step_fn(workload, threads):
    """The oracle port of the reference step (fwd + loss + bwd + Adam) on the host cores.  Checker code, timed only
    as the reported CPU baseline."""
    from oracle import bmhrl_oracle as O
    from bmhrl_b200 import synth
    O.DROPOUT_P = 0.1  # the benchmark's dropout, at the reference's dropout sites (timing only)
    torch.set_num_threads(threads)
    shapes, vshapes = agent_shapes()
    sd = synth.fill_state_dict(shapes, seed=1234)
    b = synth.synthetic_batch(V=VOC, seed=1234, **SHAPE)
    cin, cy = b["caption"][:, :-1], b["caption"][:, 1:]
    masks = O.make_masks(b["rgb"], b["audio"], cin, PAD)
    frozen_w = ("bm_manager_fus.", "manager.")
    frozen_m = ("bm_enc.", "bm_worker_fus.", "worker.")
    state = {"params": {k: v.clone() for k, v in sd.items()}, "adam": {}}
    vsd = synth.fill_state_dict(vshapes, seed=77)
    vstate = {"params": {k: v.clone() for k, v in vsd.items()}, "adam": {}}

    def trainable(k):
        if k.startswith("critic.") or k.startswith("emb_C."):
            return False
        if workload == "worker":
            return not k.startswith(frozen_w)
        if workload == "manager":
            return not k.startswith(frozen_m)
        return True

    cin0 = cin

    def step():
        p = {k: (v.clone().requires_grad_(True) if v.dtype.is_floating_point and trainable(k) else v)
             for k, v in state["params"].items()}
        # feature_getter(random_synonyms=0.3): the reference corrupts the input tokens of every step (Python loop)
        cin = O.generate_synonyms(cin0, 0.3, VOC, torch.rand(cin0.shape), torch.rand(cin0.shape),
                                  torch.randint(2, VOC, cin0.shape))
        masks = O.make_masks(b["rgb"], b["audio"], cin, PAD)
        pred, wf, mf, gf, seg = O.agent_forward(p, b["rgb"] + b["flow"], b["audio"], cin, masks)
        if workload == "warmstart":
            loss = O.label_smoothing_kl(pred, cy, SMOOTH, PAD).sum() / (cy != PAD).sum()
            loss.backward()
        else:
            tw = workload == "worker"
            pv = {k: v.clone().requires_grad_(True) for k, v in vstate["params"].items()}
            ev = O.value_function(pv, (wf if tw else mf).detach()).squeeze(-1)
            q = torch.empty(pred.shape).exponential_(1) if tw else None
            lm = cy != PAD
            div, score, tok, oexp = O.biased_kl_rl(pred, cy, b["score"].clone(), ev.detach(), lm, seg, tw, False,
                                                   SMOOTH, PAD, q)
            loss = div.sum() / (lm.sum() * 0.2)
            loss.backward()
            ev_used = ev + (oexp - ev).detach()
            vl = (((ev_used - score.float()) ** 2) * (lm if tw else seg).float()).mean()
            vl.backward()
            vg = {k: v.grad for k, v in pv.items() if v.grad is not None}
            vstate["params"], vstate["adam"] = O.adam_step({k: v.detach() for k, v in pv.items()}, vg, vstate["adam"])
        grads = {k: v.grad for k, v in p.items() if torch.is_tensor(v) and v.requires_grad and v.grad is not None}
        state["params"], state["adam"] = O.adam_step({k: v.detach() for k, v in p.items()}, grads, state["adam"])
        return float(loss.detach())
    return step


def time_cpu(workload, steps, warmup, threads):
    step = cpu_step_fn(workload, threads)
    for _ in range(warmup):
        step()
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        step()
        ts.append(time.perf_counter() - t0)
    ts.sort()
    return ts[len(ts) // 2]


CPU_PORT_NOTE = ("CPU arm = oracle/bmhrl_oracle.py, a functional-PyTorch restatement of the reference step (kind 'port') with "
                 "dropout 0.1 at the reference's dropout sites: no nn.Module overhead, no debug prints, no per-call "
                 "fp64 positional table copy -- it is FASTER than the reference's own modules on the same cores (survey "
                 "probe: 15-23 clips/s on 8 threads), so ratios against it are conservative")


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    med = time_cpu(args.workload, max(1, args.steps), max(1, args.warmup), threads)
    val = SHAPE["B"] / med
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "clips/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": med * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload_config(args.workload), note=CPU_PORT_NOTE),
            "cpu_baseline": {"value": val, "unit": "clips/s", "cores": threads, "kind": "port",
                             "sample": f"{args.steps} full B=16 steps (median) of the oracle port of the reference "
                                       f"step: fwd + loss + bwd + Adam, fp32, torch CPU, {threads} threads, "
                                       f"dropout 0.1 at the reference's dropout sites"},
            "e2e": {"value": val, "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def workload_config(workload, dropout=0.1):
    names = {"warmstart": "warm-start KL (label-smoothing) caption train step",
             "worker": "BMHRL RL step, worker phase (REINFORCE-biased KL + value fn)",
             "manager": "BMHRL RL step, manager phase (REINFORCE-biased KL + value fn)"}
    return {"workload": f"{names[workload]}, B=16/GPU, Sv=30, Sa=80, L=20, V=10172, d_model=1024, H=4, N=2, "
                        f"synthetic I3D(1024 rgb+flow)+VGGish(128), random-init BMHRL, dropout {dropout}, Adam lr 1e-4",
            "phase": workload, "batch_per_gpu": SHAPE["B"], "dropout": dropout,
            "l2": "no explicit flush: parameters + Adam state + gradients touched every step (~1 GB) exceed the 126 MB L2"}


# ------------------------------------------------------------------------------------------------ GPU arm
class Clocks:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-i", str(index), "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [l.strip().split(", ") for l in open(self.f.name) if l.strip()]
        os.unlink(self.f.name)
        sm, mx, reasons = [], None, set()
        for r in rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.strip().lower() == "active":
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def gemm_roofline(recorded, iters=20):
    """Re-issues every GEMM launch of one training step (exact shapes/strides, scratch operands) and times the
    sequence with CUDA events on the launching stream.  Returns (flops per launch, seconds per launch, n)."""
    from bmhrl_b200 import ops
    if not recorded:
        return 0.0, 0.0, 0
    dev = torch.device("cuda")
    calls = []
    flops = 0.0
    span = lambda n, s: 1 + (n - 1) * s
    for rec in recorded:
        kind, M, N, K = rec[:4]
        if min(M, N, K) <= 0:
            continue
        flops += 2.0 * M * N * K
        if kind == "linear":
            a, w, c = torch.randn(M, K, device=dev), torch.randn(N, K, device=dev), torch.empty(M, N, device=dev)
            calls.append((a, K, 1, w, 1, K, c, N, M, N, K))
        else:
            sAm, sAk, sBk, sBn, ldc = rec[4:]
            a = torch.randn(span(M, sAm) + span(K, sAk), device=dev)
            w = torch.randn(span(K, sBk) + span(N, sBn), device=dev)
            c = torch.empty(M * ldc, device=dev)
            calls.append((a, sAm, sAk, w, sBk, sBn, c, ldc, M, N, K))
    torch.cuda.synchronize()

    def run():
        for a, sAm, sAk, w, sBk, sBn, c, ldc, M, N, K in calls:
            ops.gemm(a, sAm, sAk, w, sBk, sBn, c, ldc, M, N, K)
    with torch.no_grad():
        for _ in range(3):
            run()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            run()
        e1.record()
        torch.cuda.synchronize()
    sec = e0.elapsed_time(e1) / 1e3 / iters
    n = len(calls)
    return flops / n, sec / n, n


TRAFFIC_SOURCE = ("profiles/r02_gemm_traffic.json (committed ncu capture of the same command, dram__bytes_read.sum + "
                  "dram__bytes_write.sum per GEMM launch; not re-measured in this run: ncu cannot run inside a timed bench)")


def gemm_roofline16(recorded, iters=20):
    """bf16 math mode: re-issues the step's grouped GEMM launches as they ran (the same problems per launch: exact M, N, K
    and operand majors, scratch operands; a weight-gradient launch carries its bias-gradient problem along) and times the
    sequence with CUDA events on the launching stream.  Returns (FLOPs per launch, seconds per launch, launches)."""
    from bmhrl_b200 import ops16
    if not recorded:
        return 0.0, 0.0, 0
    dev = torch.device("cuda")
    groups, flops = [], 0.0
    for launch in recorded:
        probs = []
        for (M, N, K, a_mn, b_mn) in launch:
            flops += 2.0 * M * N * K
            A = ops16.zeros16(K, M, dev) if a_mn else ops16.zeros16(M, K, dev)
            B = ops16.zeros16(K, N, dev) if b_mn else ops16.zeros16(N, K, dev)
            A.normal_()
            B.normal_()
            if N % 8:   # bias-gradient problem (N = 1): fp32 output like in the step
                probs.append(ops16.gemm_problem(A, a_mn, B, b_mn, M, N, K, C32=torch.zeros(M, N, device=dev)))
            else:
                probs.append(ops16.gemm_problem(A, a_mn, B, b_mn, M, N, K, C16=ops16.empty16(M, N, dev)))
        groups.append(probs)
    torch.cuda.synchronize()

    def run():
        for g in groups:
            ops16.gemm_group(g)
    for _ in range(3):
        run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        run()
    e1.record()
    torch.cuda.synchronize()
    sec = e0.elapsed_time(e1) / 1e3 / iters
    n = len(groups)
    return flops / n, sec / n, n


def gemm_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per GEMM launch (average over the GEMM launches of one training step),
    from the committed ncu capture profiles/r02_gemm_traffic.json (see its `command`; written by
    profiles/gemm_traffic_from_launches.py from the launch list of the same command); None when absent."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "r02_gemm_traffic.json")))["avg_dram_bytes_per_launch"]
    except Exception:
        return None


def attention_roofline(pk):
    """BASELINE.json config 5 (bf16 long-clip encoder attention, B=64, H=4, d_k=256) timed live: the fused tcgen05
    forward kernel attn_fwd_bf16_kernel, CUDA events around 10 back-to-back launches (68-105 MB of Q/K/V/O per call)."""
    from bmhrl_b200 import ops
    dev = torch.device("cuda")
    out = []
    for S in (800, 300):
        B, H, dk = 64, 4, 256
        q, k, v = (torch.randn(B, S, H * dk, device=dev).bfloat16() for _ in range(3))
        with torch.no_grad():
            for _ in range(3):
                ops.mha_core_bf16(q, k, v, None, H)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(10):
                ops.mha_core_bf16(q, k, v, None, H)
            e1.record()
            torch.cuda.synchronize()
        sec = e0.elapsed_time(e1) / 1e3 / 10
        fl = 4.0 * B * H * S * S * dk
        out.append({"kernel": "attn_fwd_bf16_kernel", "shape": f"B=64 H=4 Sq=Sk={S} d_k=256 bf16 fwd", "bound": "tensor",
                    "achieved": fl / sec / 1e12, "peak": pk["bf16_tflops"], "unit": "TFLOP/s",
                    "frac": fl / sec / 1e12 / pk["bf16_tflops"], "us": sec * 1e6})
        # backward (attn_bf16_bwd.cu): rowsum(dO o O) + the dQ, dK and dV passes; nominal FLOPs = 2.5 x forward
        do = torch.randn(B, S, H * dk, device=dev).bfloat16()
        with torch.no_grad():
            o, lse = ops.mha_core_bf16(q, k, v, None, H, return_lse=True)
            for _ in range(2):
                ops.mha_core_bf16_bwd(q, k, v, o, lse, do, None, H)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(5):
                ops.mha_core_bf16_bwd(q, k, v, o, lse, do, None, H)
            e1.record()
            torch.cuda.synchronize()
        bsec = e0.elapsed_time(e1) / 1e3 / 5
        out.append({"kernel": "attn_bwd_bf16_kernel x3 + delta", "shape": f"B=64 H=4 Sq=Sk={S} d_k=256 bf16 bwd",
                    "bound": "tensor", "achieved": 2.5 * fl / bsec / 1e12, "peak": pk["bf16_tflops"], "unit": "TFLOP/s",
                    "frac": 2.5 * fl / bsec / 1e12 / pk["bf16_tflops"], "us": bsec * 1e6})
        del q, k, v, do, o, lse
    return out


def run_gpu(args):
    import torch.distributed as dist
    from bmhrl_b200 import ops, synth, _lib
    from bmhrl_b200.ddp import GradSync
    from bmhrl_b200.model.bm_hrl_agent import BMHrlAgent, BMManagerValueFunction, BMWorkerValueFunction
    from bmhrl_b200.train import BMHRLTrainer, StepRunner
    from types import SimpleNamespace as NS

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    _lib.LIB.load()
    assert _lib.LIB.raw("bmhrl_check_device")() == 0, _lib.LIB.raw("bmhrl_last_error")().decode()

    shapes, vshapes = agent_shapes()
    sd = synth.fill_state_dict(shapes, seed=1234)  # same weights on every rank (replicated, never broadcast)
    cfg = cfg_ns("cuda")
    ds = NS(trg_voc_size=VOC, train_vocab=NS(vectors=sd["emb_C.embedder.weight"].clone()), pad_idx=PAD, start_idx=2,
            end_idx=3)
    model = BMHrlAgent(cfg, ds)
    model.load_state_dict(sd)
    model = model.to(dev)
    wv, mv = BMWorkerValueFunction(cfg), BMManagerValueFunction(cfg)
    wv.load_state_dict(synth.fill_state_dict(vshapes, seed=77))
    mv.load_state_dict(synth.fill_state_dict(vshapes, seed=78))
    sync = GradSync() if world > 1 else None
    # data parallel: --dp peer (default) = gradient sum + Adam + parameter broadcast as one kernel over NVLink peer memory
    # (bmhrl_b200/peer.py, csrc/peer.cu); --dp nccl = bucketed NCCL all-reduce + the same Adam on every rank (ddp.GradSync)
    peer, tr = None, None
    if world > 1 and args.dp == "peer":
        from bmhrl_b200.peer import PeerGroup
        try:
            peer = PeerGroup(dev)
            # random_synonyms 0.3: the reference's feature_getter corrupts the input tokens of every training step
            tr = BMHRLTrainer(model, wv.to(dev), mv.to(dev), pad_idx=PAD, smoothing=SMOOTH, grad_sync=sync,
                              random_synonyms=0.3, peer=peer,
                              owner_only_master=(args.math == "bf16" and not args.replicate_masters))
        except RuntimeError as e:   # raised on EVERY rank (PeerGroup.share agrees on the outcome): no CUDA IPC here
            if rank == 0:
                print(f"bench.py: peer-memory data parallel unavailable ({e}); using --dp nccl", file=sys.stderr)
            from bmhrl_b200 import ops16
            ops16.SHADOWS.arena = None
            peer, tr = None, None
    if tr is None:
        tr = BMHRLTrainer(model, wv.to(dev), mv.to(dev), pad_idx=PAD, smoothing=SMOOTH, grad_sync=sync, random_synonyms=0.3)
    {"warmstart": tr.teach_warmstart, "worker": tr.teach_worker, "manager": tr.teach_manager}[args.workload]()
    ops.manual_seed(1234 + rank, dev)
    ops.set_math_mode(args.math)
    if os.environ.get("BMHRL_ATTN_FUSED") is not None:  # experiment switches (default: fused tcgen05 attention, on)
        _lib.LIB.call("bmhrl_attn_tuning", int(os.environ["BMHRL_ATTN_FUSED"]))
    if os.environ.get("BMHRL_CONCURRENT") is not None:
        ops.CONCURRENT = bool(int(os.environ["BMHRL_CONCURRENT"]))

    # pinned host batches (a different one per rank and per step slot), H2D'd inside the e2e region
    n_host = 4
    host = []
    for i in range(n_host):
        b = synth.synthetic_batch(V=VOC, seed=1234 + rank + 1000 * i, **SHAPE)
        host.append({k: b[k].pin_memory() for k in ("rgb", "flow", "audio", "caption", "score")})
    h2d = sum(t.numel() * t.element_size() for t in host[0].values())

    # shapes of the GEMM launches of one eager step (for the roofline section)
    eager = StepRunner(tr, args.workload, host[0], use_graph=False)
    eager.replay()
    from bmhrl_b200 import ops16
    ops.RECORD, ops16.RECORD = [], []
    eager.replay()
    per_step_gemms, ops.RECORD = ops.RECORD, None
    per_step_gemms16, ops16.RECORD = ops16.RECORD, None
    del eager
    if args.scorer:
        # the REAL reward scorer in the loop: synthetic vocabulary, a 37k-caption Zipf
        # corpus for the document frequencies, the batch's own captions as references; two graphs around the host call
        import random
        from bmhrl_b200.metrics import BleuScorer, CiderScorer
        from bmhrl_b200.train import ScoredStepRunner
        if args.workload == "warmstart":
            raise SystemExit("--scorer needs --workload worker|manager (the warm-start step has no rewards)")
        itos = ["<unk>", "<pad>", "<s>", "</s>"] + [f"w{i}" for i in range(4, VOC)]
        rng = random.Random(99)
        corpus = [[itos[4 + min(int(rng.paretovariate(1.1)) - 1, VOC - 5)] for _ in range(rng.randint(6, 22))]
                  for _ in range(37421)]
        for b in host:
            b["trg_caption"] = [" ".join(itos[int(t)] for t in row[1:] if int(t) > 3) for row in b["caption"]]
            b["trg_caption"] = [" ".join(c.split()[:SHAPE["L"] - 2]) for c in b["trg_caption"]]
        tr.scorer = (CiderScorer(NS(itos=itos), corpus, "cpu", 0.9, 0.9) if args.scorer == "cider"
                     else BleuScorer(NS(itos=itos), "cpu", 0.9, 0.9))
        runner = ScoredStepRunner(tr, args.workload, host[0], use_graph=not args.no_graph, warmup=3,
                                  pipeline_critic=not args.no_pipeline)
    else:
        runner = StepRunner(tr, args.workload, host[0], use_graph=not args.no_graph, warmup=3,
                            pipeline_critic=not args.no_pipeline)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing (value)
    for _ in range(args.warmup):
        runner.replay()
    barrier()
    _lib.reset_launch_count()
    clocks = Clocks(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        runner.replay()
    e1.record()
    barrier()
    sec = e0.elapsed_time(e1) / 1e3
    # ---- end-to-end timing through the public step API with host buffers
    def estep(r, i, asynchronous=False):
        # the public step API: this batch from pinned host memory (+ the captions of the batch that follows, which a
        # prefetching loader has at hand, for the pipelined frozen critic)
        f = r.step_async if asynchronous else r.step
        if getattr(r, "pipeline_critic", False):
            return f(host[i % n_host], host[(i + 1) % n_host])
        return f(host[i % n_host])
    for i in range(args.warmup):
        float(estep(runner, i)[0])
    barrier()
    t0 = time.perf_counter()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    last = 0.0
    if args.sync_loss or not hasattr(runner, "step_async"):
        for i in range(args.steps):
            last = float(estep(runner, i)[0])  # D2H read of the step's loss (synchronises every step)
    else:
        # every step's loss is copied to pinned host memory and READ, one step late: the host issues step i + 1 before it
        # waits for the loss of step i (StepRunner.step_async), like any training loop that logs its loss
        pending = None
        for i in range(args.steps):
            h = estep(runner, i, asynchronous=True)
            if pending is not None:
                last = pending.value()
            pending = h
        last = pending.value()
    e3.record()
    barrier()
    sec_e2e = max(e2.elapsed_time(e3) / 1e3, 0.0)
    wall_e2e = time.perf_counter() - t0
    clk = clocks.stop() if clocks else None
    launches_eager = _lib.launch_count()  # launches issued by python after capture (0 with graphs) ...
    t = torch.tensor([sec, sec_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    sec, sec_e2e = float(t[0]), float(t[1])

    # ---- BASELINE config 3: the worker and manager RL steps, same model / process, their own graphs (all ranks)
    phases = {}
    if args.workload == "warmstart" and not args.scorer and not args.no_phases:
        main_runner = runner
        for ph in ("worker", "manager"):
            {"worker": tr.teach_worker, "manager": tr.teach_manager}[ph]()
            r2 = StepRunner(tr, ph, host[0], use_graph=not args.no_graph, warmup=3, pipeline_critic=not args.no_pipeline)
            for _ in range(args.warmup):
                r2.replay()
            barrier()
            p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            p0.record()
            for _ in range(args.steps):
                r2.replay()
            p1.record()
            barrier()
            psec = p0.elapsed_time(p1) / 1e3
            for i in range(args.warmup):
                float(estep(r2, i)[0])
            barrier()
            p0.record()
            pending = None
            for i in range(args.steps):
                if args.sync_loss:
                    float(estep(r2, i)[0])
                else:
                    h = estep(r2, i, asynchronous=True)
                    if pending is not None:
                        pending.value()
                    pending = h
            if pending is not None:
                pending.value()
            p1.record()
            barrier()
            psec_e2e = p0.elapsed_time(p1) / 1e3
            tt = torch.tensor([psec, psec_e2e], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            psec, psec_e2e = float(tt[0]), float(tt[1])
            nclips = SHAPE["B"] * world * args.steps
            phases[ph] = {"ms_per_step": psec / args.steps * 1e3, "value": nclips / psec, "unit": "clips/s",
                          "e2e": {"value": nclips / psec_e2e, "ms_per_step": psec_e2e / args.steps * 1e3},
                          "launches_per_step": getattr(r2, "captured_launches", None),
                          "config": workload_config(ph)["workload"]}
            del r2
        tr.teach_warmstart()
        runner = main_runner

    if rank == 0:
        pk, pk_src = peaks()
        graph_launches = getattr(runner, "captured_launches", None)
        if args.math == "bf16":
            fl, sl, ng = gemm_roofline16(per_step_gemms16)
        else:
            fl, sl, ng = gemm_roofline(per_step_gemms)
        ach = fl / sl / 1e12 if sl > 0 else 0.0
        peak = pk["bf16_tflops_sustained"]
        cpu_threads = os.cpu_count() or 1
        # N > 1: torchrun pins OMP_NUM_THREADS=1 and the other ranks would idle in the barrier for minutes: the CPU
        # baseline is an N=1 number (rank 0 of a single-GPU run), as the bench contract says
        do_cpu = not args.no_cpu and world == 1
        cpu_med = time_cpu(args.workload, 5, 1, cpu_threads) if do_cpu else None
        clips = SHAPE["B"] * world * args.steps
        line = {
            "metric": METRIC, "value": clips / sec, "unit": "clips/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": {"bf16": "bf16", "tf32": "tf32", "fp32": "f32"}[args.math], "data": "synthetic",
            "config": dict(workload_config(args.workload),
                           parallelism=("single GPU" if world == 1 else
                                        f"dp{world}, one process per GPU, " +
                                        ("gradient sum + Adam + parameter broadcast fused in one kernel over NVLink peer "
                                         "memory (CUDA IPC), Adam state sharded" +
                                         (", fp32 masters of GEMM weights on their owner only (bf16 shadows broadcast)"
                                          if tr.cap_opt.owner_only_master else "") if peer is not None else
                                         "NCCL all-reduce of the flat gradient buffer overlapped with backward, replicated Adam"))),
            "e2e": {"value": clips / sec_e2e, "unit": "clips/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                    "loss_read": ("float(loss) after every step" if args.sync_loss else
                                  "every step's loss copied to pinned host memory and read one step late (step_async)"),
                    "ms_per_step": sec_e2e / args.steps * 1e3, "wall_ms_per_step": wall_e2e / args.steps * 1e3},
            "gpu_launches": graph_launches * (2 * args.steps + args.warmup) if graph_launches else launches_eager,
            "launches_per_step": graph_launches,
            "clocks": clk,
            "roofline": {"bound": "tensor",
                         "kernel": {"bf16": "projection/FFN/vocab/gradient GEMMs (gemm_bf16_tc_kernel: grouped, TMA + "
                                            "tcgen05.mma kind::f16 bf16 operands, 128x128/64 tiles, TMEM accumulators, "
                                            "cluster split-K over DSMEM)",
                                    "tf32": "projection/FFN/vocab/gradient GEMMs (gemm_tf32_tc_kernel: TMA + tcgen05.mma "
                                            "kind::tf32, TMEM accumulators)",
                                    "fp32": "projection/FFN/vocab/gradient GEMMs (gemm_simt_kernel, fp32 SIMT)"}[args.math],
                         "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak if peak else None,
                         "traffic": gemm_traffic(), "traffic_source": TRAFFIC_SOURCE, "peak_source": f"{pk_src} bf16_tflops_sustained",
                         "launches_per_step": ng, "avg_launch_us": sl * 1e6, "flops_per_launch": fl},
            "step_hbm_roofline": {"algorithmic_bytes": step_bytes(args.workload), "peak_gbs": pk["hbm_gbs"],
                                  "frac": step_bytes(args.workload) / (sec / args.steps) / 1e9 / pk["hbm_gbs"]},
            "attention_roofline": attention_roofline(pk),
            "loss": last,
        }
        if args.scorer:
            line["config"]["scorer"] = (f"{args.scorer} reward scorer in the loop (libbmhrl_reward.so on the host between "
                                        f"graph A forward+sampling and graph B loss+backward+Adam); tokens D2H, scores H2D")
            line["host_part_ms_last_step"] = runner.scorer_seconds * 1e3  # D2H wait for graph A + scorer + H2D enqueue
            line["e2e"]["d2h_bytes_per_step"] = 4 + 8 * SHAPE["B"] * SHAPE["L"] + 4 * SHAPE["B"] * SHAPE["L"]
        if cpu_med is not None:
            line["cpu_baseline"] = {"value": SHAPE["B"] / cpu_med, "unit": "clips/s", "cores": cpu_threads,
                                    "kind": "port", "dropout": 0.1, "note": CPU_PORT_NOTE,
                                    "sample": "5 full B=16 steps (median, 1 warm-up) of the oracle port of the "
                                              "reference step on the host cores: fwd + loss + bwd + Adam, fp32"}
        for ph, rec in phases.items():
            rec["step_hbm_roofline"] = {"algorithmic_bytes": step_bytes(ph), "peak_gbs": pk["hbm_gbs"],
                                        "frac": step_bytes(ph) / (rec["ms_per_step"] / 1e3) / 1e9 / pk["hbm_gbs"]}
            if do_cpu:
                med = time_cpu(ph, 3, 1, cpu_threads)
                rec["cpu_port"] = {"value": SHAPE["B"] / med, "unit": "clips/s", "cores": cpu_threads, "dropout": 0.1,
                                   "sample": "3 full B=16 steps (median, 1 warm-up) of the oracle port of this phase"}
        if phases:
            line["phases"] = phases
        if world == 1 and not args.no_eager:
            line["eager_module_swap"] = eager_module_swap(args, cfg, ds, sd, host, dev)
        print(json.dumps(line), flush=True)
    if world > 1:
        # The step graph holds captured NCCL work: tearing the communicator down (destroy_process_group, or the
        # interpreter's own teardown order) with it alive hangs.  Everything is measured and printed; leave together.
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def eager_module_swap(args, cfg, ds, sd, host, dev, steps=10):
    """INTEGRATION.md section 1 as a maintainer of the reference gets it: the reference's own warm-start loop body
    (epoch_loops/captioning_bmrl_loops.py:1148-1160) over the swapped-in modules -- eager launches through
    Python + ctypes, the materialising LabelSmoothing.forward, torch.optim.Adam, no StepRunner, no CUDA graph."""
    from bmhrl_b200 import ops
    from bmhrl_b200.loss.label_smoothing import LabelSmoothing
    from bmhrl_b200.model.bm_hrl_agent import BMHrlAgent
    from bmhrl_b200.model.masking import make_masks
    model = BMHrlAgent(cfg, ds)
    model.load_state_dict(sd)
    model = model.to(dev).train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-4)
    crit = LabelSmoothing(SMOOTH, PAD)
    b = {k: v.to(dev) for k, v in host[0].items()}

    def step():
        ops.new_rng_epoch(dev)
        opt.zero_grad()
        cap = b["caption"]
        cin, y = cap[:, :-1], cap[:, 1:]
        masks = make_masks({"rgb": b["rgb"], "flow": b["flow"], "audio": b["audio"]}, cin, "audio_video", PAD)
        pred = model((b["rgb"] + b["flow"], b["audio"]), cin, masks)[0]
        n_tokens = (y != PAD).sum()
        loss = torch.sum(crit(pred, y)) / n_tokens
        loss.backward()
        ops.join_side_streams(dev)
        opt.step()
        return loss
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        loss = step()
    last = float(loss)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / steps * 1e3
    del model, opt
    return {"ms_per_step": ms, "value": SHAPE["B"] / ms * 1e3, "unit": "clips/s", "steps": steps, "loss": last,
            "what": "reference loop body over the swapped-in bmhrl_b200 modules: eager, torch.optim.Adam, materialised "
                    "(B*L,V) loss, wall clock incl. Python/ctypes overhead (math mode " + args.math + ")"}


def step_bytes(workload):
    """SURVEY.md section 8(d) algorithmic bytes per step (weights fwd + 36 B per updated param + logits)."""
    p_fwd = 62.33e6
    p_upd = {"warmstart": 47.18e6, "worker": 37.50e6, "manager": 9.68e6}[workload]
    return 4 * p_fwd + 36 * p_upd + 2 * 4 * SHAPE["B"] * SHAPE["L"] * VOC


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="warmstart", choices=["warmstart", "worker", "manager"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--math", default="bf16", choices=["bf16", "tf32", "fp32"],
                    help="bf16: bf16 tensor-core operands, grouped tcgen05 GEMMs + fused attention fwd/bwd (default); "
                         "tf32: fp32 storage, tcgen05 kind::tf32 GEMMs; fp32: exact SIMT contractions")
    ap.add_argument("--dp", default="peer", choices=["peer", "nccl"],
                    help="N > 1: peer = gradient sum + Adam + parameter broadcast in one kernel over NVLink peer memory "
                         "(default); nccl = bucketed all-reduce + replicated Adam")
    ap.add_argument("--replicate-masters", action="store_true",
                    help="--dp peer: also broadcast the fp32 master of GEMM weights every step (default in bf16 mode: the "
                         "owner of a chunk keeps it, replicas compute with the broadcast bf16 shadow)")
    ap.add_argument("--sync-loss", action="store_true",
                    help="e2e leg: float(loss) right after every step (host stalls the GPU between steps) instead of reading "
                         "every step's loss one step late from pinned memory (StepRunner.step_async)")
    ap.add_argument("--no-graph", action="store_true", help="eager launches instead of one CUDA graph per step")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-pipeline", action="store_true",
                    help="evaluate the frozen critic inside the step instead of one step ahead on a side stream")
    ap.add_argument("--no-phases", action="store_true", help="skip the worker / manager RL-step phases (config 3)")
    ap.add_argument("--no-eager", action="store_true", help="skip the eager module-swap measurement")
    ap.add_argument("--scorer", default=None, choices=["cider", "bleu"],
                    help="RL workloads: score the sampled tokens with the real host reward scorer inside the step")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
