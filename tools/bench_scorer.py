"""Times the CIDEr reward scorer of one RL step (B=16 sampled captions, L=20, V=10172, a 37k-caption corpus of
ActivityNet shape) on the host: the product (libbmhrl_reward.so through bmhrl_b200.metrics.CiderScorer) next to the
oracle port of the reference's Python scorer (oracle/reward_oracle.py, the checker, timed here as `cpu_baseline`).
Prints one JSON line.  Host-only: no GPU is used.

    python tools/bench_scorer.py [--B 16] [--L 20] [--steps 200] [--oracle-steps 5]
"""
import argparse
import json
import os
import random
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=16)
    ap.add_argument("--L", type=int, default=20)
    ap.add_argument("--V", type=int, default=10172)
    ap.add_argument("--corpus", type=int, default=37421)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--oracle-steps", type=int, default=5)
    ap.add_argument("--threads", type=int, default=1)
    a = ap.parse_args()
    from bmhrl_b200.metrics import CiderScorer
    from oracle import reward_oracle as RO  # checker / cpu baseline only

    rng = random.Random(1234)
    itos = ["<unk>", "<pad>", "<s>", "</s>"] + [f"w{i}" for i in range(a.V - 4)]
    zipf = lambda: itos[4 + min(int(rng.paretovariate(1.1)) - 1, a.V - 5)]

    def caption(lo, hi):
        return [zipf() for _ in range(rng.randint(lo, hi))]

    corpus = [caption(6, 22) for _ in range(a.corpus)]
    t0 = time.perf_counter()
    scorer = CiderScorer(type("V", (), {"itos": itos})(), corpus, "cpu", 0.9, 0.9, n_threads=a.threads)
    t_build = time.perf_counter() - t0
    t0 = time.perf_counter()
    df = RO.precook_corpus(corpus)
    t_build_oracle = time.perf_counter() - t0
    assert scorer.df_size() == len(df)

    batches = []
    for _ in range(8):
        targets = [" ".join(caption(6, a.L - 2)) for _ in range(a.B)]
        pred = torch.tensor([[itos.index(w) if rng.random() < 0.6 else rng.randrange(4, a.V) for w in
                              (t.split() + [itos[3]] * a.L)[:a.L]] for t in targets])
        sec = torch.rand(a.B, a.L) < 0.2
        batches.append((pred, targets, sec))

    def run(fn, steps):
        ts = []
        for i in range(steps):
            pred, targets, sec = batches[i % len(batches)]
            t0 = time.perf_counter()
            fn(pred, targets, sec.clone())
            ts.append(time.perf_counter() - t0)
        return float(np.median(ts))

    def product(pred, targets, sec):
        w, _ = scorer.delta_cider_worker(pred, targets)
        m, _ = scorer.delta_cider_manager(pred, targets, None, sec)
        return w, m

    def oracle(pred, targets, sec):
        w, _ = RO.worker_reward(itos, df, pred.numpy(), targets, 0.9)
        m = RO.manager_reward(itos, df, pred.numpy(), targets, sec.numpy(), 0.9)
        return w, m

    # parity on the timed inputs before timing anything
    for pred, targets, sec in batches[:2]:
        w, m = product(pred, targets, sec.clone())
        ow, om = oracle(pred, targets, sec.clone())
        assert np.abs(w.numpy() - ow).max() <= 1e-6 and np.abs(m.numpy() - om).max() <= 1e-6
    run(product, 20)
    t_prod = run(product, a.steps)
    t_orac = run(oracle, a.oracle_steps)
    print(json.dumps({
        "metric": "cider_reward_captions_per_sec", "unit": "captions/s",
        "value": 2 * a.B / t_prod, "ms_per_step": 1e3 * t_prod,
        "config": {"workload": f"worker + manager CIDEr reward of one RL step: B={a.B} sampled captions, L={a.L}, "
                               f"V={a.V}, corpus {a.corpus} captions ({scorer.df_size()} n-grams kept), n=4, gamma 0.9",
                   "threads": a.threads},
        "corpus_build_s": t_build,
        "cpu_baseline": {"value": 2 * a.B / t_orac, "unit": "captions/s", "ms_per_step": 1e3 * t_orac, "cores": 1,
                         "kind": "port", "corpus_build_s": t_build_oracle,
                         "sample": f"{a.oracle_steps} steps (median) of oracle/reward_oracle.py, the restatement of "
                                   f"metrics/cider.py pinned against the reference"},
        "speedup": t_orac / t_prod, "host": {"cpu_count": os.cpu_count()}}))


if __name__ == "__main__":
    main()
