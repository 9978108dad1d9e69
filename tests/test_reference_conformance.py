"""CPU, build container only: the module surface of bmhrl_b200 against the UNMODIFIED reference itself.

SURVEY.md section 8(b) lists the Python signatures that ARE the drop-in boundary (the reference has no FFI).  For every
class named there the constructor and `forward` (and the BMHrlAgent entry points the epoch loops call,
epoch_loops/captioning_bmrl_loops.py:844,1155,72,147, scripts/train_rl_captioning_module.py:55,63-67,132) must take the
same parameters, in the same order, with the same defaults, as the class of the same name imported from
/root/reference.  Skipped where the reference tree is absent (the GPU box)."""
import inspect

import pytest

from oracle import refboot

pytestmark = pytest.mark.skipif(not refboot.available(), reason="reference tree not present (build container only)")

# (reference module attribute of refboot.import_reference(), bmhrl_b200 module, class, methods)
SURFACE = [
    ("mha", "bmhrl_b200.model.multihead_attention", "MultiheadedAttention", ["__init__", "forward"]),
    ("blocks", "bmhrl_b200.model.blocks", "ResidualConnection", ["__init__", "forward"]),
    ("blocks", "bmhrl_b200.model.blocks", "PositionwiseFeedForward", ["__init__", "forward"]),
    ("blocks", "bmhrl_b200.model.blocks", "PositionalEncoder", ["__init__", "forward"]),
    ("blocks", "bmhrl_b200.model.blocks", "VocabularyEmbedder", ["__init__", "forward", "init_word_embeddings"]),
    ("blocks", "bmhrl_b200.model.blocks", "BridgeConnection", ["__init__", "forward"]),
    ("blocks", "bmhrl_b200.model.blocks", "FeatureEmbedder", ["__init__", "forward"]),
    ("blocks", "bmhrl_b200.model.blocks", "LayerStack", ["__init__", "forward"]),
    ("agent", "bmhrl_b200.model.bm_hrl_agent", "BMEncoder", ["__init__", "forward"]),
    ("agent", "bmhrl_b200.model.bm_hrl_agent", "BMEncoderLayer", ["__init__", "forward"]),
    ("agent", "bmhrl_b200.model.bm_hrl_agent", "BMFusion", ["__init__", "forward"]),
    ("agent", "bmhrl_b200.model.bm_hrl_agent", "BMFusionLayer", ["__init__", "forward"]),
    ("agent", "bmhrl_b200.model.bm_hrl_agent", "SegmentCritic", ["__init__", "forward"]),
    ("agent", "bmhrl_b200.model.bm_hrl_agent", "Manager", ["__init__", "forward", "expand_goals"]),
    ("agent", "bmhrl_b200.model.bm_hrl_agent", "Worker", ["__init__", "forward"]),
    ("agent", "bmhrl_b200.model.bm_hrl_agent", "WorkerCore", ["__init__", "forward"]),
    ("agent", "bmhrl_b200.model.bm_hrl_agent", "BMWorkerValueFunction", ["__init__", "forward"]),
    ("agent", "bmhrl_b200.model.bm_hrl_agent", "BMManagerValueFunction", ["__init__", "forward"]),
    ("agent", "bmhrl_b200.model.bm_hrl_agent", "BMHrlAgent",
     ["__init__", "forward", "inference", "prediction", "warmstart", "teach_warmstart", "teach_worker", "teach_manager",
      "set_inference_mode", "save_model", "load_model", "mixed_prediction", "predict_with_features"]),
    ("encoder", "bmhrl_b200.model.encoder", "TransformerEncoder", ["__init__", "forward"]),
    ("encoder", "bmhrl_b200.model.encoder", "TransformerEncoderLayer", ["__init__", "forward"]),
    ("decoder", "bmhrl_b200.model.decoder", "TransformerDecoder", ["__init__", "forward"]),
    ("decoder", "bmhrl_b200.model.decoder", "TransformerDecoderLayer", ["__init__", "forward"]),
    ("label_smoothing", "bmhrl_b200.loss.label_smoothing", "LabelSmoothing", ["__init__", "forward"]),
    ("biased_kl", "bmhrl_b200.loss.biased_kl", "BiasedKL", ["__init__", "forward"]),
    ("biased_kl", "bmhrl_b200.loss.biased_kl", "Reinforce", ["__init__", "forward"]),
]

FUNCTIONS = [
    ("masking", "bmhrl_b200.model.masking", ["subsequent_mask", "mask", "make_masks"]),
    ("mha", "bmhrl_b200.model.multihead_attention", ["attention"]),
]


@pytest.fixture(scope="module")
def ref():
    return refboot.import_reference()


def _params(fn):
    """[(name, kind, default)] of the parameters the REFERENCE declares."""
    return [(p.name, p.kind, p.default) for p in inspect.signature(fn).parameters.values()]


def _compatible(ours, theirs):
    """Every reference parameter exists at the same position with the same default; ours may only APPEND optional
    (defaulted) parameters -- extensions such as `resid=` never break a reference call site."""
    if len(ours) < len(theirs):
        return False
    for o, t in zip(ours, theirs):
        if o[0] != t[0] or o[1] != t[1] or o[2] != t[2]:
            return False
    return all(o[2] is not inspect.Parameter.empty or o[1] in (inspect.Parameter.VAR_KEYWORD,
                                                                  inspect.Parameter.VAR_POSITIONAL)
               for o in ours[len(theirs):])


@pytest.mark.parametrize("refmod,ourmod,cls,methods", SURFACE, ids=[s[2] for s in SURFACE])
def test_class_signatures_match_reference(ref, refmod, ourmod, cls, methods):
    import importlib
    theirs_cls = getattr(getattr(ref, refmod), cls)
    ours_cls = getattr(importlib.import_module(ourmod), cls)
    for m in methods:
        assert hasattr(theirs_cls, m), f"reference {cls} has no {m} (stale test table)"
        assert hasattr(ours_cls, m), f"bmhrl_b200 {cls} lacks {m}"
        t, o = _params(getattr(theirs_cls, m)), _params(getattr(ours_cls, m))
        assert _compatible(o, t), f"{cls}.{m}: reference {t} vs ours {o}"


@pytest.mark.parametrize("refmod,ourmod,names", FUNCTIONS, ids=[f[0] for f in FUNCTIONS])
def test_function_signatures_match_reference(ref, refmod, ourmod, names):
    import importlib
    for n in names:
        t = _params(getattr(getattr(ref, refmod), n))
        o = _params(getattr(importlib.import_module(ourmod), n))
        assert _compatible(o, t), f"{n}: reference {t} vs ours {o}"


def test_state_dict_keys_match_reference(ref, golden):
    """Same parameter / buffer names, order and shapes as the reference BMHrlAgent (checkpoints interchange)."""
    import torch
    from conftest import VOC, make_cfg, make_dataset
    from bmhrl_b200.model.bm_hrl_agent import BMHrlAgent
    ours = BMHrlAgent(make_cfg(device="cpu"), make_dataset(torch.randn(VOC, 300)))
    assert [(k, tuple(v.shape)) for k, v in ours.state_dict().items()] == list(golden["agent_shapes"].items())
