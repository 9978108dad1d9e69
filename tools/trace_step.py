"""Kernel timeline of one CUDA-graph replay of the training step (torch.profiler / CUPTI): per-kernel totals, union
busy time per stream and overall, and the gaps.  usage: python tools/trace_step.py [workload] > gpurun_out/trace.txt"""
import json
import os
import sys
import collections

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
from types import SimpleNamespace as NS  # noqa: E402


def main():
    workload = sys.argv[1] if len(sys.argv) > 1 else "warmstart"
    from bmhrl_b200 import ops, synth, _lib
    from bmhrl_b200.model.bm_hrl_agent import BMHrlAgent, BMManagerValueFunction, BMWorkerValueFunction
    from bmhrl_b200.train import BMHRLTrainer, StepRunner
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    sync = None
    if world > 1:   # under torchrun: the data-parallel step (rank 0 prints its own timeline, NCCL kernels included)
        import torch.distributed as dist
        from bmhrl_b200.ddp import GradSync
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        sync = GradSync(bucket_bytes=int(os.environ.get("BMHRL_BUCKET_MB", "128")) << 20)
    shapes, vshapes = bench.agent_shapes()
    sd = synth.fill_state_dict(shapes, seed=1234)
    cfg = bench.cfg_ns("cuda")
    ds = NS(trg_voc_size=bench.VOC, train_vocab=NS(vectors=sd["emb_C.embedder.weight"].clone()), pad_idx=1, start_idx=2,
            end_idx=3)
    model = BMHrlAgent(cfg, ds)
    model.load_state_dict(sd)
    model = model.to(dev)
    wv, mv = BMWorkerValueFunction(cfg), BMManagerValueFunction(cfg)
    peer = None
    if world > 1 and os.environ.get("BMHRL_DP", "peer") == "peer":
        from bmhrl_b200.peer import PeerGroup
        peer = PeerGroup(dev)
    tr = BMHRLTrainer(model, wv.to(dev), mv.to(dev), pad_idx=1, smoothing=0.7, grad_sync=sync, peer=peer,
                      owner_only_master=os.environ.get("BMHRL_OWNER_MASTER", "1") != "0")
    {"warmstart": tr.teach_warmstart, "worker": tr.teach_worker, "manager": tr.teach_manager}[workload]()
    ops.manual_seed(1234 + rank, dev)
    ops.set_math_mode(os.environ.get("BMHRL_MATH", "bf16"))
    if os.environ.get("BMHRL_CONCURRENT") is not None:
        ops.CONCURRENT = bool(int(os.environ["BMHRL_CONCURRENT"]))
    b = synth.synthetic_batch(V=bench.VOC, seed=1234 + rank, **bench.SHAPE)
    runner = StepRunner(tr, workload, b, use_graph=True, warmup=3, pipeline_critic=os.environ.get("BMHRL_PIPELINE", "1") != "0")
    if os.environ.get("BMHRL_DIAG_NO_CRITIC"):   # diagnosis only: what the step costs with the critic taken out altogether
        lab = runner.labels_cur.clone()
        model.segment_labels_of = lambda trg: lab
        runner = StepRunner(tr, workload, b, use_graph=True, warmup=3, pipeline_critic=True)
    for _ in range(5):
        runner.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 40
    e0.record()
    for _ in range(reps):
        runner.replay()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    if rank == 0:
        print(f"# world size {world}")
    if rank == 0:
        print(f"# back-to-back replays without the profiler: {e0.elapsed_time(e1) / reps * 1e3:.1f} us per step "
              f"(CUDA_DEVICE_MAX_CONNECTIONS={os.environ.get('CUDA_DEVICE_MAX_CONNECTIONS', 'default')})")
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        runner.replay()   # exactly ONE replay inside the profiled region: every kernel event belongs to it
        torch.cuda.synchronize()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        if rank != 0:
            os._exit(0)
    path = os.path.join(ROOT, "gpurun_out", "trace.json")
    prof.export_chrome_trace(path)
    ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") == "kernel"]
    ev.sort(key=lambda e: e["ts"])
    step = ev
    t0 = step[0]["ts"]
    t1 = max(e["ts"] + e["dur"] for e in step)
    print(f"# one graph replay, workload {workload}: {len(step)} kernels, span {t1 - t0:.1f} us")
    by = collections.defaultdict(lambda: [0, 0.0])
    streams = collections.defaultdict(list)
    for e in step:
        n = e["name"].split("(")[0][:70]
        by[n][0] += 1
        by[n][1] += e["dur"]
        streams[e["args"].get("stream")].append((e["ts"], e["ts"] + e["dur"]))

    def union(iv):
        iv = sorted(iv)
        tot, cs, ce = 0.0, None, None
        for s, e in iv:
            if cs is None:
                cs, ce = s, e
            elif s <= ce:
                ce = max(ce, e)
            else:
                tot += ce - cs
                cs, ce = s, e
        return tot + (ce - cs if cs is not None else 0)
    allv = [iv for v in streams.values() for iv in v]
    print(f"# GPU busy (union over streams): {union(allv):.1f} us; sum of kernel durations {sum(e['dur'] for e in step):.1f} us")
    for sid, v in sorted(streams.items(), key=lambda kv: -len(kv[1])):
        print(f"#   stream {sid}: {len(v)} kernels, busy {union(v):.1f} us, first {min(s for s, _ in v) - t0:.1f}, last {max(e for _, e in v) - t0:.1f}")
    print(f"{'total_us':>10} {'count':>5} {'avg_us':>8}  kernel")
    for n, (c, t) in sorted(by.items(), key=lambda kv: -kv[1][1]):
        print(f"{t:10.1f} {c:5d} {t / c:8.2f}  {n}")
    if os.environ.get("BMHRL_TIMELINE"):
        print("# timeline: start_us dur_us stream kernel")
        for e in step:
            print(f"{e['ts'] - t0:9.1f} {e['dur']:7.1f} {e['args'].get('stream')} {e['name'].split('(')[0][:60]}")
    os.remove(path)
    if world > 1:
        sys.stdout.flush()
        os._exit(0)   # tearing the communicator down with the captured graph alive hangs


if __name__ == "__main__":
    main()
