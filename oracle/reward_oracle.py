"""CPU restatement of the reference's CIDEr reward path (TEST INFRASTRUCTURE ONLY: imported by tests/ and by the
cpu_baseline leg of tools/bench_scorer.py, never by the product).

Follows, function by function:
  precook_corpus            metrics/cider.py:112-120
  ngram_counts              metrics/util.py:16-33 (precook)
  cider_pair                metrics/cider.py:153-239 (compute_cider for ONE hypothesis / ONE reference, ref_len = log 1)
  cider_prefix_rewards      metrics/cider.py:28-70 (_cider_diff)
  discount                  metrics/util.py:56-66, :90 (discontinue_reward without segments)
  segment_sums              metrics/batched_meteor.py:19-36 (segment_reward)
  discount_segments         metrics/util.py:67-88 (discontinue_reward with segments, in place, quirks kept)
  close_sections            metrics/cider.py:74-77 (delta_cider_manager's in-place edit of `sections`)
  worker_reward / manager_reward   metrics/cider.py:72-109
  bleu_pair                 metrics/util.py:6-15, :93-139 + metrics/bleu.py:216-279 (ONE hypothesis / ONE reference)
  bleu_prefix_rewards       metrics/bleu.py:26-59 (_bleu_diff; no stop at "</s>", fp32 rewards)
  bleu_worker_reward / bleu_manager_reward   metrics/bleu.py:61-92 (the segment variant discounts WITHOUT segments)

Pinned against the UNMODIFIED reference by oracle/pin_reward.py -> tests/golden/reward_golden.pt
(tests/test_reward_scorer.py::test_oracle_matches_reference_golden).
Plain Python / numpy: every prefix is recounted from scratch (the O(L^2) cost the product removes).
"""
import math
from collections import Counter

import numpy as np


def ngram_counts(words, n=4):
    c = Counter()
    for k in range(1, n + 1):
        for i in range(len(words) - k + 1):
            c[tuple(words[i:i + k])] += 1
    return c


def precook_corpus(caps, n=4):
    total = Counter()
    for cap in caps:
        total.update(ngram_counts(list(cap), n))
    return {g: v for g, v in total.items() if v > 1}


def _tfidf(counts, df, n):
    vec = [dict() for _ in range(n)]
    norm2 = [0.0] * n
    length = 0
    for g, tf in counts.items():
        idf = float(np.log(max(1.0, df.get(g, 0))))
        k = len(g) - 1
        vec[k][g] = float(tf) * (0.0 - idf)
        norm2[k] += pow(vec[k][g], 2)
        if k == 1:
            length += tf
    return vec, [float(np.sqrt(x)) for x in norm2], length


def cider_pair(hyp_words, ref_words, df, n=4, sigma=6.0):
    vh, nh, lh = _tfidf(ngram_counts(hyp_words, n), df, n)
    vr, nr, lr = _tfidf(ngram_counts(ref_words, n), df, n)
    delta = float(lh - lr)
    val = np.zeros(n)
    for k in range(n):
        for g, w in vh[k].items():
            r = vr[k].get(g, 0.0)
            val[k] += min(w, r) * r
        if nh[k] != 0 and nr[k] != 0:
            val[k] /= (nh[k] * nr[k])
        val[k] *= np.e ** (-(delta ** 2) / (2 * sigma ** 2))
    return float(np.mean(val))


def cider_prefix_rewards(itos, df, pred, targets, n=4, sigma=6.0):
    """pred (B, L) ints, targets list of B strings.  Returns (delta fp32 (B, L), rewards fp64 (B, L))."""
    pred = np.asarray(pred)
    B, L = pred.shape
    rewards = np.zeros((B, L), dtype=np.float64)
    for b in range(B):
        hypo = [itos[int(t)] for t in pred[b]]
        ref = targets[b].lower().split()
        scores = []
        for l in range(L):
            if hypo[l] == "</s>":
                if not scores:
                    scores.append(float(np.float32(-0.1)))
                break
            scores.append(cider_pair(" ".join(hypo[:l + 1]).split(), ref, df, n, sigma))
        rewards[b, :len(scores)] = scores
        rewards[b, len(scores):] = scores[-1]
    delta = np.concatenate([rewards[:, :1], rewards[:, 1:] - rewards[:, :-1]], axis=1).astype(np.float32)
    return delta, rewards


def discount(diff, gamma, n_step=100):
    diff = np.asarray(diff, dtype=np.float32)
    out = np.zeros_like(diff)
    B, L = diff.shape
    for b in range(B):
        for l in range(L):
            acc = np.float32(0)
            for i in range(min(n_step, L - l)):
                acc = np.float32(acc + np.float32(np.float32(gamma ** i) * diff[b, l + i]))
            out[b, l] = acc
    return out


def segment_sums(reward, sections):
    reward = np.asarray(reward, dtype=np.float32)
    sec = np.asarray(sections) != 0
    out = np.zeros_like(reward)
    for b in range(reward.shape[0]):
        start = 0
        for l in np.nonzero(sec[b])[0]:
            out[b, start:l + 1] = reward[b, start:l + 1].sum(dtype=np.float32)
            start = l + 1
    return out


def discount_segments(diff, sections, gamma):
    """In place on `diff` (np.float32 array); returns a copy."""
    ends = [(int(b), int(l)) for b, l in zip(*np.nonzero(np.asarray(sections) != 0))]
    prev_b, start = 0, 0
    for i, (b, l) in enumerate(ends):
        if prev_b != b:
            diff[prev_b, start:] = 0
            prev_b, start = b, 0
        if i < 2:  # util.py:79 compares the position in the list with len((b, l)) == 2
            for k, (b2, l2) in enumerate(ends[i + 1:]):
                if b2 != b:
                    break
                diff[b, l] = np.float32(diff[b, l] + np.float32(np.float32(gamma ** (k + 1)) * diff[b2, l2]))
        diff[b, start:l + 1] = diff[b, l]
        start = l + 1
    return diff.copy()


def close_sections(sections, targets):
    for b, t in enumerate(targets):
        e = len(t.split())
        sections[b][e] = 1
        sections[b][e + 1:] = 0
    return sections


def worker_reward(itos, df, pred, targets, gamma, n=4, sigma=6.0):
    delta, rewards = cider_prefix_rewards(itos, df, pred, targets, n, sigma)
    return discount(delta, gamma), rewards


def manager_reward(itos, df, pred, targets, sections, gamma, n=4, sigma=6.0):
    """`sections` (np array) is edited in place like the reference does."""
    close_sections(sections, targets)
    step, _ = worker_reward(itos, df, pred, targets, gamma, n, sigma)
    seg = segment_sums(step, sections)
    return discount_segments(seg, sections, gamma)


def bleu_pair(hyp_words, ref_words, n=4):
    ref = ngram_counts(ref_words, n)
    hyp = ngram_counts(hyp_words, n)
    testlen, reflen = len(hyp_words), float(len(ref_words)) / 1
    correct = [0] * n
    for g, c in hyp.items():
        correct[len(g) - 1] += min(ref.get(g, 0), c)
    small, tiny = 1e-9, 1e-15
    bleus, bleu = [], 1.
    for k in range(n):
        bleu *= float(correct[k] + tiny) / (max(0, testlen - k) + small)
        bleus.append(bleu ** (1. / (k + 1)))
    ratio = (testlen + tiny) / (reflen + small)
    if ratio < 1:
        bleus = [b * math.exp(1 - 1 / ratio) for b in bleus]
    acc = np.float32(0)
    for b in bleus:  # torch.sum(torch.tensor(bleus) * torch.tensor(1 / n)) in fp32
        acc = np.float32(acc + np.float32(np.float32(b) * np.float32(1 / n)))
    return acc


def bleu_prefix_rewards(itos, pred, targets, n=4):
    pred = np.asarray(pred)
    B, L = pred.shape
    rewards = np.zeros((B, L), dtype=np.float32)
    for b in range(B):
        hypo = [itos[int(t)] for t in pred[b]]
        ref = targets[b].lower().split()
        for l in range(L):
            rewards[b, l] = bleu_pair(" ".join(hypo[:l + 1]).lower().split(), ref, n)
    delta = np.concatenate([rewards[:, :1], rewards[:, 1:] - rewards[:, :-1]], axis=1).astype(np.float32)
    return delta, rewards


def bleu_worker_reward(itos, pred, targets, gamma, n=4):
    delta, rewards = bleu_prefix_rewards(itos, pred, targets, n)
    return discount(delta, gamma), rewards


def bleu_manager_reward(itos, pred, targets, sections, gamma, n=4):
    step, _ = bleu_worker_reward(itos, pred, targets, gamma, n)
    return discount(segment_sums(step, sections), gamma)
