"""Data parallel over NVLink peer memory: gradient sum + Adam + parameter broadcast as ONE kernel (csrc/peer.cu).

The reference wraps the agent in single-process nn.DataParallel and steps torch.optim.Adam on the full model
(scripts/train_rl_captioning_module.py:81-83,97-99).  `ddp.GradSync` is the NCCL form of that (all-reduce of the flat
gradient buffer, then the same Adam update on every rank).  `PeerGroup` is the B200 form for one process per GPU on one
NVLink / NVSwitch node: the flat gradient buffer, the parameters and the bf16 weight shadows of every rank live in
cudaMalloc'd arenas that all ranks map through CUDA IPC, laid out identically, and rank r owns every N-th 8192-element
chunk -- it sums that chunk's gradients over the ranks with peer loads, runs Adam on its own moments and stores the new
parameter (fp32 + bf16 shadow) into every rank's arenas (`ops.FusedAdam.step_peer`).  Same result as DataParallel's
sum (fixed summation order 0 .. N-1, each element summed by exactly one rank), Adam state traffic divided by N, no
collective library on the path.  The loss normaliser (a scalar) and the value-function optimisers stay on
torch.distributed.

No CPU fallback; needs CUDA IPC between the ranks (same node, peer access)."""
import ctypes

import torch
import torch.distributed as dist

from ._lib import LIB

_ALIGN = 256


class _CAI:
    """Minimal __cuda_array_interface__ carrier: lets torch alias memory it did not allocate."""

    def __init__(self, ptr, nelem, typestr):
        self.__cuda_array_interface__ = {"shape": (int(nelem),), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class Arena:
    """One cudaMalloc'd (IPC-shareable, zero-filled) region with bump allocation.  Every rank must request the same
    tensors in the same order: a peer's copy of a tensor is then at the same offset of the peer's arena."""

    def __init__(self, nbytes, device):
        self.device = torch.device(device)
        self.nbytes = int(nbytes)
        out = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            LIB.call("bmhrl_peer_alloc", self.nbytes, ctypes.byref(out))
        self.base = int(out.value)
        self.used = 0
        self.peer_bases = None   # per rank, after PeerGroup.share()
        self._keep = []

    def take(self, nelem, dtype):
        item = torch.empty((), dtype=dtype).element_size()
        off = (self.used + _ALIGN - 1) // _ALIGN * _ALIGN
        if off + nelem * item > self.nbytes:
            raise RuntimeError(f"peer arena exhausted ({self.nbytes} bytes): raise its size")
        self.used = off + nelem * item
        typestr, view = {torch.float32: ("<f4", None), torch.bfloat16: ("<i2", torch.bfloat16), torch.int64: ("<i8", None),
                         torch.int32: ("<i4", None)}[dtype]
        t = torch.as_tensor(_CAI(self.base + off, nelem, typestr), device=self.device)
        if view is not None:
            t = t.view(view)
        return t

    def contains(self, ptr):
        return self.base <= ptr < self.base + self.nbytes

    def offsets(self):
        """Byte offset from a local address to the same address in rank r's arena, per rank (ctypes long long array)."""
        if self.peer_bases is None:
            raise RuntimeError("peer arena not shared yet (PeerGroup.share)")
        return (ctypes.c_longlong * len(self.peer_bases))(*[b - self.base for b in self.peer_bases])


class PeerGroup:
    """The ranks of `group` (default: world) on one node, their shared arenas and the cross-GPU signals."""

    GRAD, PARAM, SHADOW = "grad", "param", "shadow"

    def __init__(self, device, group=None, shadow_bytes=None):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        if self.world > LIB.raw("bmhrl_peer_max_ranks")():
            raise RuntimeError(f"PeerGroup: at most {LIB.raw('bmhrl_peer_max_ranks')()} ranks")
        self.device = torch.device(device)
        self.arenas = {}
        self.n_slots = LIB.raw("bmhrl_peer_slots")()
        self.flags = None
        self._shared = False

    # ---- set-up --------------------------------------------------------------------------------------------------
    def arena(self, name, nbytes=None):
        a = self.arenas.get(name)
        if a is None:
            if nbytes is None:
                raise RuntimeError(f"peer arena {name!r} does not exist yet")
            if self._shared:
                raise RuntimeError("PeerGroup: arenas must be created before share()")
            a = self.arenas[name] = Arena(nbytes, self.device)
        return a

    def share(self):
        """Exchanges the IPC handles of all arenas (collective over the group) and maps the peers' copies."""
        if self._shared:
            return
        fl = self.arena("_flags", LIB.raw("bmhrl_peer_flag_bytes")() + _ALIGN)
        self.flags = fl.take(LIB.raw("bmhrl_peer_flag_bytes")() // 8, torch.int64)
        hb = LIB.raw("bmhrl_peer_handle_bytes")()
        names = sorted(self.arenas)
        mine = {}
        for n in names:
            buf = ctypes.create_string_buffer(hb)
            LIB.call("bmhrl_peer_get_handle", self.arenas[n].base, buf)
            mine[n] = (bytes(buf.raw), self.arenas[n].nbytes)
        if self.world > 1:
            everyone = [None] * self.world
            dist.all_gather_object(everyone, mine, group=self.group)
        else:
            everyone = [mine]
        err = None
        try:
            for n in names:
                bases = []
                for r in range(self.world):
                    if r == self.rank:
                        bases.append(self.arenas[n].base)
                        continue
                    h, nb = everyone[r][n]
                    if nb != self.arenas[n].nbytes:
                        raise RuntimeError(f"peer arena {n!r}: rank {r} allocated {nb} bytes, this rank {self.arenas[n].nbytes}")
                    out = ctypes.c_void_p()
                    with torch.cuda.device(self.device):
                        LIB.call("bmhrl_peer_open_handle", ctypes.create_string_buffer(h, hb), ctypes.byref(out))
                    bases.append(int(out.value))
                self.arenas[n].peer_bases = bases
        except Exception as e:   # noqa: BLE001 -- every rank must learn about it, or the others wait forever
            err = e
        if self.world > 1:
            ok = torch.tensor([0 if err is not None else 1], device=self.device, dtype=torch.int32)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
            if int(ok) == 0:
                raise RuntimeError(f"PeerGroup.share: mapping the peers' arenas failed on at least one rank "
                                   f"(this rank: {err!r}); CUDA IPC / peer access between all GPUs of the node is required")
        elif err is not None:
            raise err
        self._shared = True

    @classmethod
    def simulate(cls, device, world):
        """`world` groups inside ONE process on one device (tests): same kernels, "peer" arenas are each other's local
        arenas.  Call `share_simulated(groups)` once all arenas exist; drive every rank on a stream of its own."""
        out = []
        for r in range(world):
            g = cls.__new__(cls)
            g.group, g.rank, g.world, g.device = None, r, world, torch.device(device)
            g.arenas, g.n_slots, g.flags, g._shared = {}, LIB.raw("bmhrl_peer_slots")(), None, False
            out.append(g)
        return out

    @staticmethod
    def share_simulated(groups):
        for g in groups:
            fl = g.arena("_flags", LIB.raw("bmhrl_peer_flag_bytes")() + _ALIGN)
            g.flags = fl.take(LIB.raw("bmhrl_peer_flag_bytes")() // 8, torch.int64)
        for g in groups:
            for n, a in g.arenas.items():
                a.peer_bases = [h.arenas[n].base for h in groups]
            g._shared = True

    # ---- stream-ordered cross-GPU ordering (CUDA-graph capturable) ---------------------------------------------------
    def signal(self, slot):
        from .ops import _stream
        LIB.call("bmhrl_peer_signal", self.flags.data_ptr(), self.arenas["_flags"].offsets(), self.world, self.rank,
                 int(slot), _stream())

    def wait(self, slot):
        from .ops import _stream
        LIB.call("bmhrl_peer_wait", self.flags.data_ptr(), self.world, int(slot), _stream())

    def barrier(self, slot):
        """Every rank's stream passes this point only when every rank's stream has reached it."""
        self.signal(slot)
        self.wait(slot)
