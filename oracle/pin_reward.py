"""Generates tests/golden/reward_golden.pt by running the UNMODIFIED reference reward code
(/root/reference/metrics/{cider,util,batched_meteor}.py) on seeded synthetic captions.

TEST INFRASTRUCTURE ONLY; runs in the build container (the reference tree does not exist on the GPU box).
    python -m oracle.pin_reward
Deviations: the nltk stub of oracle/refboot.py (metrics/batched_meteor.py imports nltk at module scope); stdout sunk
(precook_corpus prints).
"""
import os
import random
import sys

import torch

from . import refboot

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden", "reward_golden.pt")

WORDS = ("a man is jumping over the long bar and lands in sand woman runs down track people are watching him she "
         "then stands up walks away crowd cheers with her arms raised another person measures distance on field while "
         "camera pans to scoreboard several athletes wait their turn near line").split()


def make_vocab():
    itos = ["<unk>", "<pad>", "<s>", "</s>"] + sorted(set(WORDS)) + [" ", "n't", "Sand"]
    return itos


def make_caption(rng, lo=4, hi=14):
    # a Zipf-ish draw so that n-grams repeat across the corpus
    n = rng.randint(lo, hi)
    out, i = [], rng.randrange(len(WORDS))
    for _ in range(n):
        out.append(WORDS[i])
        i = (i + 1) % len(WORDS) if rng.random() < 0.7 else rng.randrange(len(WORDS))
    return out


def main():
    if not refboot.available():
        raise SystemExit("reference tree not found")
    refboot._stub_nltk()
    if refboot.REF_ROOT not in sys.path:
        sys.path.insert(0, refboot.REF_ROOT)
    with refboot.sink():
        from metrics import cider as rcider, util as rutil, batched_meteor as rbm, bleu as rbleu
    rng = random.Random(20261104)
    torch.manual_seed(7)
    itos = make_vocab()
    stoi = {w: i for i, w in enumerate(itos)}
    vocab = type("V", (), {"itos": itos})()
    corpus = [make_caption(rng) for _ in range(300)]
    corpus[5] = corpus[5] + [" "]          # a raw whitespace corpus token (never matches a split() word)
    g = {"itos": itos, "corpus": corpus, "cases": [], "scan_cases": []}
    for case, (B, L, gamma) in enumerate([(6, 12, 0.9), (16, 20, 0.95), (4, 8, 0.5), (3, 30, 1.0)]):
        with refboot.sink():
            ref = rcider.CiderScorer(vocab, corpus, "cpu", gamma, gamma)
        targets, pred = [], torch.zeros(B, L, dtype=torch.long)
        for b in range(B):
            cap = make_caption(rng, 3, L - 2)
            # mixed case in the reference string: the scorer lower()s it
            targets.append(" ".join(cap).capitalize() if b % 2 else " ".join(cap))
            kind = (b + case) % 6
            toks = [stoi[w] for w in cap]
            if kind == 0:      # the reference caption itself, ended
                row = toks + [3]
            elif kind == 1:    # noisy copy, no end token
                row = [t if rng.random() < 0.7 else rng.randrange(4, len(itos)) for t in toks]
            elif kind == 2:    # end token first
                row = [3] + toks
            elif kind == 3:    # random tokens incl. specials and the whitespace token
                row = [rng.randrange(0, len(itos)) for _ in range(L)]
                row = [t if t != 3 else 0 for t in row]
            elif kind == 4:    # starts with whitespace token and <s>, repeated words
                row = [stoi[" "], 2] + toks[:3] * 3 + [3]
            else:              # shifted copy
                row = toks[1:] + toks[:1] + [3]
            row = (row + [rng.randrange(4, len(itos)) for _ in range(L)])[:L]
            pred[b] = torch.tensor(row)
        with refboot.sink():
            w_score, w_rewards = ref.delta_cider_worker(pred.clone(), list(targets))
        entry = {"B": B, "L": L, "gamma": gamma, "pred": pred, "targets": targets,
                 "worker_score": w_score.clone(), "worker_rewards": w_rewards.clone().double()}
        for name, dtype in (("bool", torch.bool), ("long", torch.long)):
            sections = (torch.rand(B, L) < 0.25)
            if case == 0:
                sections[0] = False       # first row without a segment end before the closing edit
            sections = sections.to(dtype)
            sec_in = sections.clone()
            with refboot.sink():
                m_score, _ = ref.delta_cider_manager(pred.clone(), list(targets), None, sections)
            entry[f"sections_in_{name}"] = sec_in
            entry[f"sections_out_{name}"] = sections.clone()
            entry[f"manager_score_{name}"] = m_score.clone()
        # BLEU scorer on the same predictions (metrics/bleu.py); "Mixed Case" vocabulary entries are lower()ed there
        with refboot.sink():
            bl = rbleu.BleuScorer(vocab, "cpu", gamma, gamma)
            bw, br = bl.delta_bleu_worker(pred.clone(), list(targets))
            sec_b = entry["sections_in_bool"].clone()
            bm, _ = bl.delta_bleu_manager(pred.clone(), list(targets), None, sec_b)
        assert torch.equal(sec_b, entry["sections_in_bool"])
        entry.update({"bleu_worker_score": bw.clone(), "bleu_rewards": br.clone(), "bleu_manager_score": bm.clone()})
        g["cases"].append(entry)
        # df table spot checks
        if case == 0:
            df = ref.dictionary
            keys = list(df.keys())
            rng.shuffle(keys)
            g["df_size"] = len([k for k, v in df.items() if v > 1])
            g["df_samples"] = [(k, int(df[k])) for k in keys[:50] if df[k] > 1]
    # the scans on arbitrary fp32 inputs (quirk coverage: rows without segment ends, leading empty rows, > 2 ends)
    for B, L, gamma, p in [(5, 9, 0.9, 0.3), (8, 20, 0.95, 0.15), (3, 6, 0.5, 0.6), (4, 12, 0.99, 0.0), (1, 7, 0.8, 0.4)]:
        x = torch.randn(B, L)
        sec = torch.rand(B, L) < p
        if B > 2:
            sec[0] = False
            sec[2] = False
        e = {"B": B, "L": L, "gamma": gamma, "x": x.clone(), "sections": sec.clone()}
        e["discount"] = rutil.discontinue_reward(x.clone(), gamma).clone()
        e["discount_n3"] = rutil.discontinue_reward(x.clone(), gamma, n_step=3).clone()
        sr, idx = rbm.segment_reward(x.clone(), sec)
        e["segment_reward"] = sr.clone()
        y = x.clone()
        e["discount_segments"] = rutil.discontinue_reward(y, gamma, segments=sec).clone()
        e["discount_segments_inplace"] = y.clone()
        g["scan_cases"].append(e)
    torch.save(g, OUT)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
