import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bmhrl_b200 import ops
from bmhrl_b200._lib import LIB
LIB.load()
ops.manual_seed(5)
P = 0.3
def mask(shape, salt):
    z = torch.zeros(shape, device="cuda")
    return ops.dropout_add(z, torch.ones_like(z), P, salt)
g = torch.Generator().manual_seed(1)
for (M, K, N) in ((36, 64, 128), (36, 64, 1024), (60, 128, 512)):
    x = torch.randn(M, K, generator=g).cuda(); w = torch.randn(N, K, generator=g).cuda(); b = torch.randn(N, generator=g).cuda()
    y0 = ops.linear(x, w, b)
    y = ops.linear(x, w, b, drop_p=P, salt=9)
    m = mask((M, N), 9)
    print("linear", (M, K, N), "err", float((y - y0 * m).abs().max()), "keep", float((y != 0).float().mean()), float((m != 0).float().mean()))
    r = torch.randn(M, N, generator=g).cuda()
    y = ops.linear(x, w, b, resid=r, drop_p=P, salt=9)
    print("  +resid err", float((y - (r + y0 * m)).abs().max()))
    y = ops.linear(x, w, b, act=1, drop_p=P, salt=9)
    print("  relu err", float((y - torch.relu(y0) * m).abs().max()))
B, H, S, dk = 3, 4, 12, 256
q = torch.randn(B, S, H * dk, generator=g).cuda(); k = torch.randn(B, S, H * dk, generator=g).cuda(); v = torch.randn(B, S, H * dk, generator=g).cuda()
o0 = ops.mha_core(q, k, v, None, H)
o = ops.mha_core(q, k, v, None, H, drop_p=P, salt=11)
m = mask((B, S, H * dk), 11)
print("mha_core err", float((o - o0 * m).abs().max()), "keep", float((o != 0).float().mean()))
