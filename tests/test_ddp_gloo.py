"""World-size-2 gloo tests (CPU) of the data-parallel host logic in bmhrl_b200/ddp.py: bucket ranges over the flat
gradient buffer, the frozen / never-used spans staying untouched, and the reference's DataParallel semantics
(scripts/train_rl_captioning_module.py:97-99 + captioning_bmrl_loops.py:1156-1158: ONE loss over the gathered global
batch) = sum over ranks of d(local row sum / GLOBAL token count)."""
import os
import socket
from types import SimpleNamespace as NS

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from bmhrl_b200.ddp import GradSync


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class _FlatOpt:
    """The attributes of ops.FusedAdam that GradSync reads, on CPU tensors."""

    def __init__(self, shapes, frozen):
        self.params = [torch.zeros(s) for s in shapes]
        self.offsets, off = [], 0
        for p in self.params:
            self.offsets.append(off)
            off += p.numel()
        self.flat_grad = torch.zeros(off)
        self.active = [i for i in range(len(shapes)) if i not in frozen]


def _toy_problem(world, V=23, D=8, pad=1):
    g = torch.Generator().manual_seed(5)
    W = torch.randn(V, D, generator=g) * 0.3
    xs = [torch.randn(3, 5, D, generator=g) for _ in range(world)]
    ys = []
    for r in range(world):
        y = torch.randint(4, V, (3, 5), generator=g)
        y[0, 3 + r:] = pad  # ragged: ranks hold different token counts
        ys.append(y)
    return W, xs, ys


def _worker(rank, world, port, q):
    from oracle import bmhrl_oracle as O  # checker
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        res = {}
        # ---- 1. bucketed sum over the active spans only
        shapes = [(7, 3), (5,), (4, 4), (11,), (2, 9)]
        opt = _FlatOpt(shapes, frozen={1, 3})
        sync = GradSync(bucket_bytes=4 * 6)  # 6-element buckets: spans split unevenly
        opt.flat_grad.copy_(torch.arange(opt.flat_grad.numel(), dtype=torch.float32) * (rank + 1))
        before = opt.flat_grad.clone()
        ranges = sync.ranges(opt)
        sync(opt)
        tot = sum(r + 1 for r in range(world))
        exp = before.clone()
        for i in opt.active:
            s, e = opt.offsets[i], opt.offsets[i] + opt.params[i].numel()
            exp[s:e] = before[s:e] / (rank + 1) * tot
        res["bucket_ok"] = bool(torch.equal(opt.flat_grad, exp))
        res["calls"] = sync.calls
        res["n_ranges"] = len(ranges)
        res["max_bucket"] = max(e - s for s, e in ranges)
        covered = sorted(ranges)
        res["disjoint"] = all(covered[i][1] <= covered[i + 1][0] for i in range(len(covered) - 1))
        # mean=True (value-function loss is a .mean() over the global batch)
        opt.flat_grad.fill_(float(rank + 1))
        sync(opt, mean=True)
        s0 = opt.offsets[0]
        res["mean_ok"] = abs(float(opt.flat_grad[s0]) - tot / world) < 1e-6
        # ---- 2. DataParallel loss semantics with the label-smoothing criterion
        W, xs, ys = _toy_problem(world)
        Wr = W.clone().requires_grad_(True)
        pred = torch.log_softmax(xs[rank] @ Wr.t(), -1)
        n_local = (ys[rank] != 1).sum()
        n_glob = sync.all_reduce_scalar(n_local)
        res["n_glob"] = int(n_glob)
        res["n_local_unchanged"] = int(n_local)
        (O.label_smoothing_kl(pred, ys[rank], 0.7, 1).sum() / n_glob).backward()
        opt2 = _FlatOpt([tuple(W.shape)], frozen=set())
        opt2.flat_grad.copy_(Wr.grad.reshape(-1))
        sync(opt2)
        res["grad"] = opt2.flat_grad.tolist()  # plain python: the queue outlives this process
        if rank == 0:
            q.put(res)
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_grad_sync_world2():
    from oracle import bmhrl_oracle as O
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    mp.spawn(_worker, args=(world, _free_port(), q), nprocs=world, join=True)
    res = q.get()
    assert res["bucket_ok"] and res["mean_ok"] and res["disjoint"]
    assert res["calls"] >= res["n_ranges"] and res["max_bucket"] <= 6
    # single-process global batch: what nn.DataParallel + one criterion call computes
    W, xs, ys = _toy_problem(world)
    Wg = W.clone().requires_grad_(True)
    x, y = torch.cat(xs), torch.cat(ys)
    n = (y != 1).sum()
    assert res["n_glob"] == int(n) and res["n_local_unchanged"] == int((ys[0] != 1).sum())
    (O.label_smoothing_kl(torch.log_softmax(x @ Wg.t(), -1), y, 0.7, 1).sum() / n).backward()
    assert torch.allclose(torch.tensor(res["grad"]).view_as(W), Wg.grad, rtol=1e-5, atol=1e-7)


def test_ranges_single_process():
    opt = _FlatOpt([(10,), (3, 3), (6,)], frozen={1})
    s = GradSync.__new__(GradSync)
    s.bucket_elems, s.world, s.group, s.calls = 4, 1, None, 0
    assert s.ranges(opt) == [(0, 4), (4, 8), (8, 10), (19, 23), (23, 25)]
    # adjacent active tensors merge into one span before bucketing
    opt = _FlatOpt([(3,), (3,), (2,)], frozen=set())
    assert s.ranges(opt) == [(0, 4), (4, 8)]
