"""CPU: the C-ABI library loads and exports every symbol include/bmhrl_b200.h declares; the Python module surface
keeps the reference's state_dict layout; host-side glue (masks, synthetic data) behaves like the reference."""
import ctypes
import os

import pytest
import torch

from bmhrl_b200 import _lib, synth
from conftest import PAD, ROOT, VOC, make_cfg, make_dataset


def test_header_symbols_exported():
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as ge
        ge.build()
    protos = _lib.parse_header()
    assert len(protos) >= 40
    dll = ctypes.CDLL(_lib.LIB_PATH)
    for name in protos:
        assert hasattr(dll, name), f"{name} declared in include/bmhrl_b200.h but not exported"
    dll.bmhrl_abi_version.restype = ctypes.c_int
    assert dll.bmhrl_abi_version() == 1


def test_no_cpu_fallback():
    from bmhrl_b200 import ops
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.layer_norm(torch.randn(2, 8), torch.ones(8), torch.zeros(8))
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.linear(torch.randn(2, 8), torch.randn(4, 8))


def test_product_never_imports_oracle():
    import re
    for dp, _, files in os.walk(os.path.join(ROOT, "bmhrl_b200")):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f"{f} imports the oracle"


def test_state_dict_layout(golden):
    from bmhrl_b200.model.bm_hrl_agent import BMHrlAgent, BMManagerValueFunction, BMWorkerValueFunction
    cfg = make_cfg(device="cpu")
    m = BMHrlAgent(cfg, make_dataset(torch.randn(VOC, 300)))
    sd = {k: tuple(v.shape) for k, v in m.state_dict().items()}
    assert list(sd.items()) == list(golden["agent_shapes"].items())
    assert sum(p.numel() for p in m.parameters()) == 70322365
    for cls in (BMWorkerValueFunction, BMManagerValueFunction):
        assert {k: tuple(v.shape) for k, v in cls(cfg).state_dict().items()} == golden["value_fn_shapes"]
    # phase switching: trainable parameter counts of the reference
    count = lambda: sum(p.numel() for p in m.parameters() if p.requires_grad)
    assert count() == 52123760
    m.teach_worker()
    assert count() == 39962910 and m.manager.exploration is False
    m.teach_manager()
    assert count() == 12160850 and m.manager.exploration is True
    m.teach_warmstart()  # F10: the encoder stays frozen after a manager epoch
    assert count() == 28744560  # SURVEY.md F10 probe: 28.7 M


def test_masks_match_oracle():
    from oracle import bmhrl_oracle as O
    from bmhrl_b200.model.masking import make_masks
    b = synth.synthetic_batch(B=4, Sv=9, Sa=13, L=7, V=VOC, seed=3, ragged=True)
    cin = b["caption"][:, :-1]
    m = make_masks({"rgb": b["rgb"], "flow": b["flow"], "audio": b["audio"]}, cin, "audio_video", PAD)
    o = O.make_masks(b["rgb"], b["audio"], cin, PAD)
    for k in o:
        assert torch.equal(m[k], o[k])
    assert m["V_mask"].sum(-1).flatten().tolist() == b["len_v"].tolist()


def test_used_parameters_skip_never_called():
    from bmhrl_b200.model.bm_hrl_agent import BMHrlAgent
    from bmhrl_b200.train import used_parameters
    m = BMHrlAgent(make_cfg(device="cpu"), make_dataset(torch.randn(VOC, 300)))
    used = sum(p.numel() for p in used_parameters(m) if p.requires_grad)
    assert used == 52123760 - 4943856
