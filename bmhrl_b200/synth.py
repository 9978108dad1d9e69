"""Deterministic synthetic weights and ActivityNet-shaped inputs (CPU, torch only).

There is no dataset and no checkpoint on the boxes this runs on, so the benchmark, the parity tests and
the golden fixtures all draw from these generators (SURVEY.md section 8(d) "Synthetic inputs").  Tensors are
keyed by NAME, so the same (name, shape, seed) gives the same values in the build container (where the
reference is importable) and on the GPU box (where it is not).
"""
import zlib

import torch


def _gen(name, seed):
    g = torch.Generator(device="cpu")
    g.manual_seed((zlib.crc32(name.encode()) ^ (seed * 0x9E3779B1)) & 0x7FFFFFFF)
    return g


def fill_tensor(name, shape, seed, dtype=torch.float32):
    """Value rule by parameter name (reference key set, model/bm_hrl_agent.py)."""
    g = _gen(name, seed)
    shape = tuple(shape)
    leaf = name.split(".")[-1]
    if leaf in ("alpha",):
        return torch.full(shape, 0.9, dtype=dtype)
    if leaf in ("beta",):
        return torch.full(shape, 2.0, dtype=dtype)
    if leaf == "a_v_constant":
        return (0.5 * torch.randn(shape, generator=g)).to(dtype)
    if "norm" in name and leaf == "weight":
        return (1.0 + 0.1 * torch.randn(shape, generator=g)).to(dtype)
    if "norm" in name and leaf == "bias":
        return (0.05 * torch.randn(shape, generator=g)).to(dtype)
    if name.endswith("embedder.weight"):
        return torch.randn(shape, generator=g).to(dtype)
    if name == "critic.lin.bias":
        return torch.full(shape, -2.5, dtype=dtype)
    if name == "critic.lin.weight":
        return (0.1 * torch.randn(shape, generator=g)).to(dtype)
    if name.startswith("critic.") and len(shape) == 2:  # strong recurrences -> varied segment labels
        k = 3.0 / (shape[-1] ** 0.5)
        return ((torch.rand(shape, generator=g) * 2 - 1) * k).to(dtype)
    if len(shape) >= 2:
        k = 1.0 / (shape[-1] ** 0.5)
        return ((torch.rand(shape, generator=g) * 2 - 1) * k).to(dtype)
    return (0.02 * torch.randn(shape, generator=g)).to(dtype)


def fill_state_dict(shapes, seed=1234, dtype=torch.float32):
    """shapes: mapping name -> shape (e.g. {k: v.shape for k, v in model.state_dict().items()})."""
    sd = {}
    for name, shape in shapes.items():
        sd[name] = fill_tensor(name, shape, seed, dtype)
    # one module registered under two names in the reference (manager_core == manager.core)
    for a, b in (("manager_core.linear.weight", "manager.core.linear.weight"),
                 ("manager_core.linear.bias", "manager.core.linear.bias")):
        if a in sd and b in sd:
            sd[b] = sd[a]
    return sd


def synthetic_batch(B=16, Sv=30, Sa=80, L=20, V=10172, seed=1234, ragged=False, d_vid=1024, d_aud=128,
                    pad_idx=1, start_idx=2, end_idx=3, min_cap=6):
    """One training batch of the reference's shapes (epoch_loops/captioning_bmrl_loops.py:488-500):
    rgb/flow (B,Sv,d_vid), audio (B,Sa,d_aud) fp32 with zero padding frames and component 0 of every valid
    frame forced positive (so model/masking.py reproduces the length mask), caption (B,L+1) int64 =
    <s> w.. </s> pad.., plus a synthetic reward in [0, 0.3) standing in for the host scorer."""
    g = _gen("batch", seed)
    if ragged:
        len_v = torch.randint(4, Sv + 1, (B,), generator=g)
        len_a = torch.clamp(torch.round(len_v.float() * (Sa / max(Sv, 1))).long(), 1, Sa)
        len_v[0], len_a[0] = Sv, Sa
    else:
        len_v = torch.full((B,), Sv, dtype=torch.long)
        len_a = torch.full((B,), Sa, dtype=torch.long)

    def feats(S, D, lens):
        x = torch.relu(0.25 * torch.randn(B, S, D, generator=g) + 0.1)
        x[:, :, 0] = 0.05 + torch.rand(B, S, generator=g)
        valid = torch.arange(S)[None, :] < lens[:, None]
        return x * valid[:, :, None].float()

    rgb = feats(Sv, d_vid, len_v)
    flow = feats(Sv, d_vid, len_v)
    audio = feats(Sa, d_aud, len_a)
    cap = torch.full((B, L + 1), pad_idx, dtype=torch.long)
    lo = min(min_cap, L)
    lens_c = torch.randint(lo, L + 1, (B,), generator=g)
    lens_c[0] = L
    for b in range(B):
        n = int(lens_c[b])
        cap[b, 0] = start_idx
        cap[b, 1:n] = torch.randint(4, V, (n - 1,), generator=g)
        cap[b, n] = end_idx
    score = 0.3 * torch.rand(B, L, generator=g)
    return {"rgb": rgb, "flow": flow, "audio": audio, "caption": cap, "score": score,
            "len_v": len_v, "len_a": len_a, "len_c": lens_c}
