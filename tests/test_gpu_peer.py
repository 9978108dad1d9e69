"""Data parallel over peer memory (csrc/peer.cu, bmhrl_b200/peer.py): gradient sum over the ranks + Adam + parameter
broadcast in one kernel.  Here on ONE GPU: N simulated ranks in one process whose "peer" arenas are each other's local
arenas, every rank on its own stream (the cross-GPU signal / wait kernels are the real ones).  The result must equal
-- bit for bit -- one Adam step on the summed gradients, which is what nn.DataParallel + torch.optim.Adam compute
(scripts/train_rl_captioning_module.py:81-83,97-99).  The 2-GPU run over CUDA IPC is tools/peer_check.py."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _params(seed, dev):
    g = torch.Generator().manual_seed(seed)
    shapes = [(300, 1024), (1024,), (10172, 364), (7,), (3, 5, 11), (40000,)]
    return [torch.nn.Parameter(torch.randn(*s, generator=g).to(dev)) for s in shapes]


@pytest.mark.parametrize("world", [2, 3])
def test_step_peer_equals_adam_on_summed_gradients(gpu_lib, world):
    from bmhrl_b200 import ops
    from bmhrl_b200.peer import PeerGroup
    dev = torch.device("cuda", 0)
    groups = PeerGroup.simulate(dev, world)
    ranks = [_params(5, dev) for _ in range(world)]
    opts = [ops.FusedAdam(ranks[r], lr=1e-2, peer=groups[r]) for r in range(world)]
    PeerGroup.share_simulated(groups)
    ref_p = _params(5, dev)
    ref = ops.FusedAdam(ref_p, lr=1e-2)
    streams = [torch.cuda.Stream(device=dev) for _ in range(world)]
    gen = torch.Generator().manual_seed(11)
    for step in range(3):
        only = None if step != 1 else frozenset({0, 2, 5})
        ref.zero_grad()
        grads = [[torch.randn(p.shape, generator=gen).to(dev) for p in ref_p] for _ in range(world)]
        for r in range(world):
            opts[r].zero_grad()
            for i, p in enumerate(ranks[r]):
                p.grad.copy_(grads[r][i])
        for i, p in enumerate(ref_p):
            for r in range(world):          # rank order 0 .. N-1, like the kernel
                p.grad.add_(grads[r][i])
        torch.cuda.synchronize()
        # one process plays all ranks: every rank's signal is issued before any rank's wait (see step_peer_finish)
        for r in range(world):
            streams[r].wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(streams[r]):
                groups[r].signal(step)
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                opts[r].step_peer_finish(step, only=only)
                groups[r].signal(groups[r].n_slots - 1)
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                groups[r].wait(groups[r].n_slots - 1)
        for s in streams:
            torch.cuda.current_stream(dev).wait_stream(s)
        ref.step(only=only)
        torch.cuda.synchronize()
        for r in range(world):
            for i, (a, b) in enumerate(zip(ranks[r], ref_p)):
                d = (a.detach() - b.detach()).abs().flatten()
                assert torch.equal(a.detach(), b.detach()), (step, r, i, float(d.max()), int(d.argmax()), int((d > 0).sum()))
            assert torch.equal(opts[r].step_counts, ref.step_counts)
    # every chunk has exactly one owner, and the same one whatever subset a step updates (the owner keeps the moments)
    owned = [set(zip(t[1].tolist(), t[2].tolist())) for t in (o._peer_tables(None) for o in opts)]
    assert sum(len(s) for s in owned) == ref.n_chunks and len(set().union(*owned)) == ref.n_chunks
    sub = frozenset({0, 2, 5})
    for r, o in enumerate(opts):
        t = o._peer_tables(sub)
        assert set(zip(t[1].tolist(), t[2].tolist())) == {c for c in owned[r] if c[0] in sub}


def test_parameters_keep_identity_and_values_in_the_arena(gpu_lib):
    from bmhrl_b200 import ops
    from bmhrl_b200.peer import PeerGroup
    dev = torch.device("cuda", 0)
    g = PeerGroup.simulate(dev, 1)[0]
    ps = _params(9, dev)
    before = [p.detach().clone() for p in ps]
    opt = ops.FusedAdam(ps, peer=g)
    PeerGroup.share_simulated([g])
    arena = g.arena(g.PARAM)
    for p, b in zip(ps, before):
        assert torch.equal(p.detach(), b) and arena.contains(p.data_ptr()) and p.grad is not None
        assert g.arena(g.GRAD).contains(p.grad.data_ptr())
    assert opt.flat_param.data_ptr() == ps[0].data_ptr()
