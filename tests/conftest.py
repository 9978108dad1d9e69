"""Shared fixtures.  `-m "not gpu"`: oracle vs committed golden vectors, host logic, C-ABI symbol checks (no GPU).
`-m gpu`: parity of the CUDA path (through the C ABI) against the oracle / golden vectors on a B200."""
import os
import sys
from types import SimpleNamespace as NS

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

VOC = 10172
PAD, START, END = 1, 2, 3
SMOOTH = 0.7


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a) and the built libbmhrl_b200.so")


@pytest.fixture(scope="session")
def golden():
    return torch.load(os.path.join(ROOT, "tests", "golden", "reference_golden.pt"), weights_only=False)


def make_cfg(**over):
    cfg = dict(d_vid=1024, d_aud=128, d_model_video=1024, d_model_audio=128, d_model=1024, d_model_caps=300,
               rl_projection_d=512, rl_att_heads=4, rl_att_layers=2, rl_goal_d=64, rl_ff_v=1024, rl_ff_a=512,
               rl_ff_c=2048, dout_p=0.0, rl_critic_score_threshhold=0.25, unfreeze_word_emb=False,
               device="cuda", device_ids=[0], rl_critic_path=None)
    cfg.update(over)
    return NS(**cfg)


def make_dataset(vectors, voc=VOC):
    return NS(trg_voc_size=voc, train_vocab=NS(vectors=vectors), pad_idx=PAD, start_idx=START, end_idx=END)


def rel(a, b, floor=1e-30):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / (b.norm() + floor))


@pytest.fixture(scope="session")
def agent_sd(golden):
    from bmhrl_b200 import synth
    return synth.fill_state_dict(golden["agent_shapes"], seed=1234)


@pytest.fixture(scope="session")
def gpu_lib():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from bmhrl_b200._lib import LIB
    LIB.load()
    assert LIB.raw("bmhrl_check_device")() == 0, LIB.raw("bmhrl_last_error")().decode()
    return LIB


def build_agent(agent_sd, device="cuda", **cfg_over):
    """bmhrl_b200 BMHrlAgent with the deterministic synthetic weights (reference key set)."""
    from bmhrl_b200.model.bm_hrl_agent import BMHrlAgent
    cfg = make_cfg(**cfg_over)
    model = BMHrlAgent(cfg, make_dataset(agent_sd["emb_C.embedder.weight"].clone()))
    model.load_state_dict(agent_sd, strict=True)
    return model.to(device), cfg
