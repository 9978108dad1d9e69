import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bmhrl_b200._lib import LIB
from bmhrl_b200.peer import PeerGroup, Arena
LIB.load()
dev = torch.device("cuda", 0)
a = Arena(1 << 20, dev)
t = a.take(1000, torch.float32)
print("alias:", t.data_ptr() == a.base, hex(t.data_ptr()), hex(a.base), t.device, t.dtype)
t.fill_(3.0); torch.cuda.synchronize()
t2 = a.take(16, torch.int64); print("i64", t2.data_ptr() - a.base)
gs = PeerGroup.simulate(dev, 2)
for g in gs:
    g.arena("grad", 1 << 16)
PeerGroup.share_simulated(gs)
print("flags ptrs", [hex(g.flags.data_ptr()) for g in gs], [list(g.arenas["_flags"].offsets()) for g in gs])
s = [torch.cuda.Stream() for _ in gs]
for r, g in enumerate(gs):
    with torch.cuda.stream(s[r]):
        g.signal(3)
torch.cuda.synchronize(); print("signals ok", gs[0].flags[3 * 8:3 * 8 + 2].tolist(), gs[1].flags[3 * 8:3 * 8 + 2].tolist())
for r, g in enumerate(gs):
    with torch.cuda.stream(s[r]):
        g.wait(3)
torch.cuda.synchronize(); print("waits ok")
for r, g in enumerate(gs):
    with torch.cuda.stream(s[r]):
        g.barrier(5)
torch.cuda.synchronize(); print("barrier ok")
