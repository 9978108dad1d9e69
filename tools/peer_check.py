#!/usr/bin/env python
"""torchrun -n N: the data-parallel warm-start step through NCCL (ddp.GradSync: all-reduce + replicated Adam) and through
peer memory (peer.PeerGroup: gradient sum + Adam + parameter broadcast in one kernel) from the same initial state and the
same batches: parameters after K steps must agree (N = 2: bit for bit, a two-term sum has one order), every rank must
hold the same parameters, and the bf16 shadows must equal the cast of the fp32 parameters.  Then both as CUDA graphs, timed."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import bench  # noqa: E402
from types import SimpleNamespace as NS  # noqa: E402


def build(mode, dev, rank):
    from bmhrl_b200 import ops, ops16, synth
    from bmhrl_b200.ddp import GradSync
    from bmhrl_b200.peer import PeerGroup
    from bmhrl_b200.model.bm_hrl_agent import BMHrlAgent, BMManagerValueFunction, BMWorkerValueFunction
    from bmhrl_b200.train import BMHRLTrainer
    shapes, vshapes = bench.agent_shapes()
    sd = synth.fill_state_dict(shapes, seed=1234)
    cfg = bench.cfg_ns("cuda")
    ds = NS(trg_voc_size=bench.VOC, train_vocab=NS(vectors=sd["emb_C.embedder.weight"].clone()), pad_idx=1, start_idx=2, end_idx=3)
    model = BMHrlAgent(cfg, ds)
    model.load_state_dict(sd)
    model = model.to(dev)
    wv, mv = BMWorkerValueFunction(cfg), BMManagerValueFunction(cfg)
    ops16.SHADOWS.clear()
    ops16.SHADOWS.arena = None
    ops.RNG.salt = 0   # same dropout salts (handed out per call site on first use) in both builds
    peer = PeerGroup(dev) if mode.startswith("peer") else None
    tr = BMHRLTrainer(model, wv.to(dev), mv.to(dev), pad_idx=1, smoothing=0.7, grad_sync=GradSync(), peer=peer,
                      owner_only_master=(mode == "peer-owner"))
    tr.teach_warmstart()
    ops.manual_seed(1234 + rank, dev)
    ops.set_math_mode("bf16")
    return tr, model


def main():
    from bmhrl_b200 import synth, ops16
    from bmhrl_b200.train import StepRunner
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
    batches = [synth.synthetic_batch(V=bench.VOC, seed=77 + rank + 100 * i, **bench.SHAPE) for i in range(3)]
    batches = [{k: v.to(dev) for k, v in b.items() if torch.is_tensor(v)} for b in batches]
    out = {}
    for mode in ("nccl", "peer", "peer-owner"):
        tr, model = build(mode, dev, rank)
        losses = []
        for b in batches:
            losses.append(float(tr.warmstart_step(b)))
        torch.cuda.synchronize()
        if mode == "peer-owner":
            stale = sum(0 if torch.equal(p.detach(), out["peer"][0][n]) else 1 for n, p in model.named_parameters())
            tr.cap_opt.gather_params()
            torch.cuda.synchronize()
            if rank == 0:
                print(f"peer-owner: {stale} tensors of rank 0 differ from the replicated run before gather_params()")
        out[mode] = ({n: p.detach().clone() for n, p in model.named_parameters()}, losses)
        if mode.startswith("peer"):
            # every rank holds the same parameters; shadows = cast of the parameters
            chk = torch.stack([p.detach().double().sum() for p in model.parameters()]).sum().reshape(1)
            allc = [torch.zeros_like(chk) for _ in range(world)]
            dist.all_gather(allc, chk)
            same = all(float(c) == float(allc[0]) for c in allc)
            bad = 0
            for p in model.parameters():
                e = ops16.SHADOWS.entry(p)
                if e is None or e[3] != 0 or p.dim() != 2:
                    continue
                g = [g for g in ops16.SHADOWS.groups.values() if g["w16"].data_ptr() <= e[0] < g["w16"].data_ptr() + g["w16"].numel() * 4]
                rows = p.shape[0]
                w16 = None
                for gg in g:
                    r0 = (e[0] - gg["w16"].data_ptr()) // (gg["w16"].stride(0) * 2)
                    w16 = gg["w16"][r0:r0 + rows]
                if w16 is not None and not torch.equal(w16, p.detach().to(torch.bfloat16)):
                    bad += 1
            if rank == 0:
                print(f"{mode}: all {world} ranks hold the same parameters: {same}; stale bf16 shadows: {bad}")
        if os.environ.get("PEER_CHECK_NO_TIMING"):
            continue
        # timing as a CUDA graph
        host = {k: v for k, v in batches[0].items()}
        runner = StepRunner(tr, "warmstart", host, use_graph=True, warmup=3, pipeline_critic=True)
        for _ in range(5):
            runner.replay()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(30):
            runner.replay()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 30], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"{mode}: {float(t) * 1e3:.1f} us per step on {world} GPUs (max over ranks), losses {losses}")
        del runner, tr, model
    for other in ("peer", "peer-owner"):
        pa, pb = out["nccl"][0], out[other][0]
        worst = max(float((pa[n] - pb[n]).abs().max() / (pa[n].abs().max() + 1e-12)) for n in pa)
        nd = sum(0 if torch.equal(pa[n], pb[n]) else 1 for n in pa)
        if rank == 0:
            print(f"nccl vs {other} after 3 steps: {nd} of {len(pa)} tensors differ, worst relative difference {worst:.3e}; "
                  f"losses nccl {out['nccl'][1]} {other} {out[other][1]}")
    sys.stdout.flush()
    dist.barrier()
    os._exit(0)


if __name__ == "__main__":
    main()
