"""Pins the input-pipeline restatements (oracle.crop_bounds / pad_batch / generate_synonyms) against the UNMODIFIED
reference and writes tests/golden/pipeline_golden.pt.  TEST INFRASTRUCTURE; build container only.

    python -m oracle.pin_pipeline --write

generate_synonyms draws from Python's global `random` in a data-dependent order.  The reference function is run on a
recorded stream (random.random / random.randint patched to log every draw); a shadow walk of the same rules attributes
each draw to its (row, position) so that the per-position tables the oracle / the kernel consume can be built -- and the
reference's own output is the golden."""
import argparse
import os
import random
import sys
from types import SimpleNamespace as NS

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import bmhrl_oracle as O  # noqa: E402
from oracle import refboot  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden", "pipeline_golden.pt")


def pin_crop(gold):
    sys.path.insert(0, refboot.REF_ROOT)
    from captioning_datasets.load_features import crop_a_segment, pad_segment  # the reference's own functions
    rng = np.random.RandomState(3)
    cases = []
    for _ in range(400):
        S = int(rng.randint(1, 120))
        dur = float(rng.uniform(1.0, 300.0))
        a, b = sorted(rng.uniform(0, dur, 2))
        if rng.rand() < 0.1:
            b = a
        if rng.rand() < 0.05:
            a = b = dur
        feat = torch.arange(S, dtype=torch.float32)[:, None].repeat(1, 2)
        r = crop_a_segment(feat, a, b, dur)
        s, e = O.crop_bounds(S, a, b, dur)
        ref_rows = [] if r is None else [int(v) for v in r[:, 0]]
        assert ref_rows == list(range(s, min(e, S))), (S, a, b, dur, ref_rows, s, e)
        cases.append((S, a, b, dur, s, e, len(ref_rows)))
    gold["crop"] = cases
    # pad_sequence semantics used by the dataset (rgb / audio padded with pad_idx, flow with 0)
    from torch.nn.utils.rnn import pad_sequence
    stacks = [torch.randn(int(n), 8, generator=torch.Generator().manual_seed(int(n))) for n in (3, 7, 1, 5)]
    for pv in (0, 1):
        assert torch.equal(pad_sequence(stacks, batch_first=True, padding_value=pv), O.pad_batch(stacks, pv))
    assert torch.equal(pad_segment(stacks[0], 6, 1), O.pad_batch([stacks[0], torch.zeros(6, 8)], 1)[0])
    print(f"[crop / pad] {len(cases)} segments agree with load_features.crop_a_segment; pad_sequence agrees")


def pin_synonyms(ref, gold):
    getter = ref.loops.feature_getter(True, False, random_synonyms=0.3)
    # generate_synonyms is a closure of feature_getter: reach it through the returned function's closure cells
    cells = {c.cell_contents.__name__: c.cell_contents for c in getter.__closure__ if callable(c.cell_contents)}
    gen = cells["generate_synonyms"]
    V = 50
    loader = NS(dataset=NS(train_vocab=NS(itos=list(range(V))), pad_idx=1))
    g = torch.Generator().manual_seed(17)
    B, L = 9, 12
    cap = torch.randint(4, V, (B, L), generator=g)
    cap[:, 0] = 2
    for i, n in enumerate([5, 11, 3, 8, 12, 1, 7, 9, 6]):   # end token at different places; row 4 has none
        if n < L:
            cap[i, n] = 3
            cap[i, n + 1:] = 1
    cases = []
    for p, seed in ((0.3, 1), (0.9, 2), (0.0, 3), (1.0, 4)):
        log = []
        rr, ri = random.random, random.randint
        state = random.Random(seed)

        def fake_random():
            v = state.random()
            log.append(("u", v))
            return v

        def fake_randint(a, b):
            v = state.randint(a, b)
            log.append(("r", v, a, b))
            return v
        random.random, random.randint = fake_random, fake_randint
        try:
            ref_out = gen(cap, loader, p)
        finally:
            random.random, random.randint = rr, ri
        # attribute the draws to positions with a shadow walk of the same rules
        u1 = torch.ones(B, L)
        u2 = torch.ones(B, L)
        rw = torch.zeros(B, L, dtype=torch.int64)
        k = 0
        for i in range(B):
            for j in range(L):
                if int(cap[i, j]) == 3:
                    break
                u1[i, j] = log[k][1]; k += 1
                if u1[i, j] < p:
                    u2[i, j] = log[k][1]; k += 1
                    if u2[i, j] >= 0.9:
                        assert log[k][0] == "r" and log[k][2:] == (2, V - 1)
                        rw[i, j] = log[k][1]; k += 1
        assert k == len(log), (k, len(log))
        # float32 tables must decide like the float64 draws did
        u1f, u2f = u1.float(), u2.float()
        o = O.generate_synonyms(cap, p, V, u1f, u2f, rw)
        assert torch.equal(o, ref_out), p
        cases.append({"p": p, "cap": cap.clone(), "u1": u1f, "u2": u2f, "rw": rw, "out": ref_out.clone(), "V": V})
        print(f"[generate_synonyms] p={p}: {len(log)} draws, {int((ref_out != cap).sum())} tokens changed -- oracle == reference")
    gold["synonyms"] = cases


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--write", action="store_true")
    args = ap.parse_args()
    ref = refboot.import_reference()
    gold = {}
    pin_crop(gold)
    pin_synonyms(ref, gold)
    if args.write:
        torch.save(gold, GOLD)
        print("wrote", GOLD, os.path.getsize(GOLD) // 1024, "KiB")
    print("PIPELINE PINNED: crop / pad / generate_synonyms agree with the reference")


if __name__ == "__main__":
    main()
