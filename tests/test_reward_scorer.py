"""CPU: the host reward scorer (libbmhrl_reward.so behind bmhrl_b200.metrics).

Three-way parity: reference golden (oracle/pin_reward.py ran the UNMODIFIED metrics/cider.py) == oracle restatement
== product.  Rewards are floating point: tolerance 1e-6 absolute on fp32 outputs (fp32 accumulation order inside
torch.sum may differ), 1e-12 on the fp64 per-prefix scores (the incremental update re-associates the sums); the
`sections` edit is integer work and must be bit-exact.
"""
import ctypes
import os
import random

import numpy as np
import pytest
import torch

from conftest import ROOT
from oracle import reward_oracle as RO

GOLDEN = os.path.join(ROOT, "tests", "golden", "reward_golden.pt")
F32_TOL = 1e-6
F64_TOL = 1e-12


def f32ok(a, b):
    """fp32 outputs: 1e-6 absolute, scaled by the magnitude of the reference values (fp32 sums of discounted rewards
    reach a few units; torch.sum's accumulation order differs from a sequential loop in the last bit)."""
    a = a.detach().cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
    b = b.detach().cpu().numpy() if torch.is_tensor(b) else np.asarray(b)
    return a.shape == b.shape and (a.size == 0 or float(np.abs(a - b).max()) <= F32_TOL * max(1.0, float(np.abs(b).max())))


@pytest.fixture(scope="module")
def rgold():
    return torch.load(GOLDEN, weights_only=False)


@pytest.fixture(scope="module")
def built():
    from bmhrl_b200.metrics import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as ge
        ge.build()
    return _lib


def _vocab(itos):
    return type("V", (), {"itos": itos})()


def test_header_symbols_exported(built):
    from bmhrl_b200._lib import parse_header
    protos = parse_header(built.HEADER)
    assert len(protos) == 11
    dll = ctypes.CDLL(built.LIB_PATH)
    for name in protos:
        assert hasattr(dll, name), f"{name} declared in include/bmhrl_reward.h but not exported"


def test_oracle_matches_reference_golden(rgold):
    df = RO.precook_corpus(rgold["corpus"])
    assert len(df) == rgold["df_size"]
    for k, v in rgold["df_samples"]:
        assert df[k] == v
    for c in rgold["cases"]:
        score, rewards = RO.worker_reward(rgold["itos"], df, c["pred"].numpy(), c["targets"], c["gamma"])
        assert np.abs(rewards - c["worker_rewards"].numpy()).max() <= F64_TOL
        assert f32ok(score, c["worker_score"].numpy())
        for name in ("bool", "long"):
            sec = c[f"sections_in_{name}"].numpy().copy()
            m = RO.manager_reward(rgold["itos"], df, c["pred"].numpy(), c["targets"], sec, c["gamma"])
            assert np.array_equal(sec, c[f"sections_out_{name}"].numpy())
            assert f32ok(m, c[f"manager_score_{name}"].numpy())
        bw, br = RO.bleu_worker_reward(rgold["itos"], c["pred"].numpy(), c["targets"], c["gamma"])
        assert f32ok(br, c["bleu_rewards"].numpy())
        assert f32ok(bw, c["bleu_worker_score"].numpy())
        bm = RO.bleu_manager_reward(rgold["itos"], c["pred"].numpy(), c["targets"], c["sections_in_bool"].numpy(),
                                    c["gamma"])
        assert f32ok(bm, c["bleu_manager_score"].numpy())
    for e in rgold["scan_cases"]:
        x, sec = e["x"].numpy(), e["sections"].numpy()
        assert f32ok(RO.discount(x, e["gamma"]), e["discount"].numpy())
        assert f32ok(RO.discount(x, e["gamma"], 3), e["discount_n3"].numpy())
        assert f32ok(RO.segment_sums(x, sec), e["segment_reward"].numpy())
        y = x.copy()
        out = RO.discount_segments(y, sec, e["gamma"])
        assert f32ok(out, e["discount_segments"].numpy())
        assert f32ok(y, e["discount_segments_inplace"].numpy())


def test_product_matches_reference_golden(built, rgold):
    from bmhrl_b200.metrics import CiderScorer
    s0 = CiderScorer(_vocab(rgold["itos"]), rgold["corpus"], "cpu", 0.9, 0.9)
    assert s0.df_size() == rgold["df_size"]
    for k, v in rgold["df_samples"]:
        assert s0.document_frequency(k) == v
    assert s0.document_frequency(("no-such-word",)) == 0
    for c in rgold["cases"]:
        s = CiderScorer(_vocab(rgold["itos"]), rgold["corpus"], "cpu", c["gamma"], c["gamma"])
        assert s.type == "CIDER"
        score, rewards = s.delta_cider_worker(c["pred"].clone(), list(c["targets"]))
        assert score.dtype == torch.float32 and rewards.dtype == torch.float64
        assert (rewards - c["worker_rewards"]).abs().max() <= F64_TOL
        assert f32ok(score, c["worker_score"])
        for name in ("bool", "long"):
            sec = c[f"sections_in_{name}"].clone()
            m, none = s.delta_cider_manager(c["pred"].clone(), list(c["targets"]), None, sec)
            assert none is None and m.dtype == torch.float32
            assert torch.equal(sec, c[f"sections_out_{name}"]), "sections must be closed in place, dtype kept"
            assert f32ok(m, c[f"manager_score_{name}"])
        assert s.counter == 3


def test_bleu_matches_reference_golden_and_oracle(built, rgold):
    from bmhrl_b200.metrics import BleuScorer
    for c in rgold["cases"]:
        s = BleuScorer(_vocab(rgold["itos"]), "cpu", c["gamma"], c["gamma"])
        assert s.type == "BLEU"
        w, r = s.delta_bleu_worker(c["pred"].clone(), list(c["targets"]))
        assert w.dtype == torch.float32 and r.dtype == torch.float32
        assert f32ok(r, c["bleu_rewards"])
        assert f32ok(w, c["bleu_worker_score"])
        sec = c["sections_in_bool"].clone()
        m, none = s.delta_bleu_manager(c["pred"].clone(), list(c["targets"]), None, sec)
        assert none is None and torch.equal(sec, c["sections_in_bool"]), "the BLEU scorer leaves sections alone"
        assert f32ok(m, c["bleu_manager_score"])
    # fresh case: n = 2, threads, a token that lower()s onto another word, an empty reference
    itos = ["<unk>", "<pad>", "<s>", "</s>", "a", "b", "c", "A", "x y", ""]
    pred = torch.randint(0, len(itos), (24, 15), generator=torch.Generator().manual_seed(3))
    targets = [" ".join(itos[4 + (i + j) % 3] for j in range(i % 7)) for i in range(24)]
    s = BleuScorer(_vocab(itos), "cpu", 0.8, 0.8, n=2, n_threads=4)
    w, r = s.delta_bleu_worker(pred, targets)
    ow, orw = RO.bleu_worker_reward(itos, pred.numpy(), targets, 0.8, n=2)
    assert f32ok(r.numpy(), orw) and f32ok(w.numpy(), ow)
    # a hypothesis equal to its reference scores 1 at the last word
    s4 = BleuScorer(_vocab(itos), "cpu", 1.0, 1.0)
    _, r = s4.delta_bleu_worker(torch.tensor([[4, 5, 6, 4, 5]]), ["a b c a b"])
    assert abs(float(r[0, -1]) - 1.0) < 1e-6


def test_scans_match_reference_golden(built, rgold):
    from bmhrl_b200.metrics import discontinue_reward, segment_reward
    for e in rgold["scan_cases"]:
        x, sec, g = e["x"], e["sections"], e["gamma"]
        assert (discontinue_reward(x.clone(), g) - e["discount"]).abs().max() <= F32_TOL
        assert (discontinue_reward(x.clone(), g, n_step=3) - e["discount_n3"]).abs().max() <= F32_TOL
        sr, idx = segment_reward(x.clone(), sec)
        assert f32ok(sr, e["segment_reward"])
        assert torch.equal(idx, torch.nonzero(sec))
        y = x.clone()
        out = discontinue_reward(y, g, segments=sec)
        assert f32ok(out, e["discount_segments"])
        assert f32ok(y, e["discount_segments_inplace"]), "in-place semantics of the reference"
        assert out.data_ptr() != y.data_ptr()


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_product_matches_oracle_random(built, seed):
    """Fresh seeded cases the golden file does not hold: long rows, n < 4, multi-threaded rows, out-of-vocabulary
    reference words, tokens that split into two words."""
    from bmhrl_b200.metrics import CiderScorer
    rng = random.Random(seed)
    words = [f"w{i}" for i in range(40)]
    itos = ["<unk>", "<pad>", "<s>", "</s>"] + words + ["", "two words", "\t"]
    draw = lambda n: [words[min(int(rng.expovariate(0.15)), 39)] for _ in range(n)]
    corpus = [draw(rng.randint(3, 25)) for _ in range(400)]
    B, L = (40, 64) if seed == 0 else (9, 17)
    n = 4 if seed < 2 else 2
    targets = [" ".join(draw(rng.randint(1, L - 2)) + (["OOV-word"] if b % 5 == 0 else [])) for b in range(B)]
    pred = torch.tensor([[rng.randrange(0, len(itos)) if rng.random() < 0.97 else 3 for _ in range(L)] for _ in range(B)])
    pred[1, 0] = 3
    df = RO.precook_corpus(corpus, n)
    want_score, want_rewards = RO.worker_reward(itos, df, pred.numpy(), targets, 0.9, n=n)
    s = CiderScorer(_vocab(itos), corpus, "cpu", 0.9, 0.9, n=n, n_threads=3 if seed == 0 else 1)
    score, rewards = s.delta_cider_worker(pred, targets)
    assert np.abs(rewards.numpy() - want_rewards).max() <= F64_TOL
    assert f32ok(score.numpy(), want_score)
    assert float(rewards[1, 0]) == float(np.float32(-0.1)) and bool((rewards[1] == rewards[1, 0]).all())
    sec = torch.rand(B, L, generator=torch.Generator().manual_seed(seed)) < 0.2
    sec_o = sec.numpy().copy()
    want_m = RO.manager_reward(itos, df, pred.numpy(), targets, sec_o, 0.9, n=n)
    m, _ = s.delta_cider_manager(pred, targets, None, sec)
    assert np.array_equal(sec.numpy(), sec_o)
    assert f32ok(m.numpy(), want_m)


def test_properties_and_errors(built):
    from bmhrl_b200.metrics import CiderScorer, discontinue_reward
    words = "a b c d e f g h".split()
    itos = ["<unk>", "<pad>", "<s>", "</s>"] + words
    corpus = [words, words, words[::-1], words[::-1], words[2:6], words[2:6]]
    s = CiderScorer(_vocab(itos), corpus, "cpu", 1.0, 1.0)
    ids = [itos.index(w) for w in words]
    # a hypothesis equal to the reference scores 1 at its last word (cosine 1 for every order, no length penalty)
    pred = torch.tensor([ids + [3, 1, 1]])
    _, rewards = s.delta_cider_worker(pred, [" ".join(words)])
    assert abs(float(rewards[0, len(words) - 1]) - 1.0) < 1e-12
    assert bool((rewards[0, len(words):] == rewards[0, len(words) - 1]).all())
    # gamma = 1: the discounted sum of the deltas telescopes to (final score - previous prefix score)
    score, rewards = s.delta_cider_worker(pred, [" ".join(words)])
    assert abs(float(score[0, 0]) - float(rewards[0, -1])) < 1e-5
    # empty batch row list / bad token id / reference longer than the caption tensor
    with pytest.raises(RuntimeError, match="vocabulary"):
        s.delta_cider_worker(torch.tensor([[len(itos)]]), ["a"])
    with pytest.raises(IndexError):
        s.delta_cider_manager(pred[:, :4], ["a b c d e"], None, torch.zeros(1, 4, dtype=torch.bool))
    with pytest.raises(ValueError):
        discontinue_reward(torch.zeros(3), 0.9)
    out = discontinue_reward(torch.zeros(0, 5), 0.9)
    assert tuple(out.shape) == (0, 5)
