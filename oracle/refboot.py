"""Bootstrap for importing the UNMODIFIED reference (/root/reference) on CPU.

TEST INFRASTRUCTURE ONLY.  Used by oracle/make_golden.py and oracle/pin_oracle.py inside the
build container to (a) pin the CPU restatement in oracle/bmhrl_oracle.py against the real
reference modules and (b) generate the golden fixtures committed under tests/golden/.
/root/reference does not exist on the GPU box, so nothing under tests/ -m gpu, smoke() or
bench.py imports this file.

Wrapper-level deviations (none of them edits the reference):
  (i)   nltk stub in sys.modules (nltk is not installed; imported at module scope by
        epoch_loops/captioning_bmrl_loops.py:19-20 and metrics/batched_meteor.py:1-5)
  (ii)  SimpleNamespace cfg + fake dataset with the fields model/bm_hrl_agent.py:495-507 reads
  (iii) synthetic SegmentCritic checkpoint (keys of model/bm_hrl_agent.py:186-201)
  (iv)  stdout/stderr sunk while the reference runs (it prints whole tensors every forward)
"""
import contextlib
import io
import os
import sys
import types
from types import SimpleNamespace as NS

import torch
import torch.nn as nn

REF_ROOT = os.environ.get("BMHRL_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "model"))


def _stub_nltk():
    names = ["nltk", "nltk.corpus", "nltk.translate", "nltk.translate.meteor_score", "nltk.tokenize",
             "nltk.tokenize.treebank"]
    for name in names:
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["nltk.corpus"].wordnet = None
    ms = sys.modules["nltk.translate.meteor_score"]
    ms.meteor_score = ms.single_meteor_score = lambda *a, **k: 0.0
    sys.modules["nltk.translate"].meteor = lambda *a, **k: 0.0
    sys.modules["nltk.tokenize.treebank"].TreebankWordDetokenizer = type("TreebankWordDetokenizer", (), {})


@contextlib.contextmanager
def sink():
    with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
        yield


def import_reference():
    """Returns a namespace with the reference's hot-path python modules."""
    if not available():
        raise RuntimeError(f"reference tree not found at {REF_ROOT}")
    _stub_nltk()
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    with sink():
        from model import bm_hrl_agent, blocks, masking, multihead_attention, encoder, decoder
        from loss import label_smoothing, biased_kl
        from epoch_loops import captioning_bmrl_loops as loops
    return NS(agent=bm_hrl_agent, blocks=blocks, masking=masking, mha=multihead_attention,
              encoder=encoder, decoder=decoder, label_smoothing=label_smoothing, biased_kl=biased_kl,
              loops=loops)


def make_cfg(**over):
    cfg = dict(d_vid=1024, d_aud=128, d_model_video=1024, d_model_audio=128, d_model=1024, d_model_caps=300,
               rl_projection_d=512, rl_att_heads=4, rl_att_layers=2, rl_goal_d=64, rl_ff_v=1024, rl_ff_a=512,
               rl_ff_c=2048, dout_p=0.0, rl_critic_score_threshhold=0.25, unfreeze_word_emb=False,
               device="cpu", device_ids=[0], rl_critic_path="/tmp/bmhrl_oracle_critic.cp")
    cfg.update(over)
    return NS(**cfg)


def make_dataset(voc_size, d_caps, vectors):
    itos = ["<unk>", "<blank>", "<s>", "</s>"] + [f"w{i}" for i in range(voc_size - 4)]
    return NS(trg_voc_size=voc_size, train_vocab=NS(vectors=vectors, itos=itos), pad_idx=1, start_idx=2, end_idx=3)


def build_reference_agent(ref, cfg, voc_size, state_dict):
    """Constructs the reference BMHrlAgent and loads `state_dict` (reference key set) into it."""
    d = cfg.d_model_caps
    critic_sd = {k[len("critic."):]: v.clone() for k, v in state_dict.items() if k.startswith("critic.")}
    torch.save(critic_sd, cfg.rl_critic_path)
    ds = make_dataset(voc_size, d, state_dict["emb_C.embedder.weight"].clone())
    with sink():
        model = ref.agent.BMHrlAgent(cfg, ds)
    missing, unexpected = model.load_state_dict(state_dict, strict=True)
    return model, ds
