"""`BleuScorer` with the reference's interface (metrics/bleu.py:13-92): per-prefix BLEU rewards of the sampled
captions.  N-gram clipping and scoring run in libbmhrl_reward.so (bmhrl_bleu_prefix_rewards)."""
import ctypes

import numpy as np
import torch

from ._lib import RLIB
from .util import discontinue_reward, segment_reward


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class BleuScorer:
    def __init__(self, vocab, device, gamma, gamma_manager, n=4, sigma=6.0, n_threads=0):
        assert (n <= 4 and n > 0)
        self.counter = 0
        self.vocab = vocab
        self.device = device
        self._n = n
        self.type = "BLEU"
        self.gamma = gamma
        self.gamma_m = gamma_manager
        self.n_threads = n_threads
        self._word_id = {}
        self._ref_cache = {}
        # token id -> the words it contributes to " ".join(hypo).lower().split() (bleu.py:39, util.py:10)
        tok_words, tok_off = [], [0]
        for s in vocab.itos:
            tok_words.extend(self._wid(w) for w in s.lower().split())
            tok_off.append(len(tok_words))
        self._tok_words = np.asarray(tok_words if tok_words else [0], dtype=np.int32)
        self._tok_off = np.asarray(tok_off, dtype=np.int32)

    def _wid(self, w):
        i = self._word_id.get(w)
        if i is None:
            i = self._word_id[w] = len(self._word_id)
        return i

    def _ref_ids(self, caption):
        ids = self._ref_cache.get(caption)
        if ids is None:
            if len(self._ref_cache) >= 1 << 18:
                self._ref_cache.clear()
            ids = self._ref_cache[caption] = np.asarray([self._wid(w) for w in caption.lower().split()], dtype=np.int32)
        return ids

    def _bleu_diff(self, pred, target):
        """bleu.py:26-59: returns (delta (B, L) fp32, rewards (B, L) fp32), both on self.device."""
        p = pred.detach().to("cpu", torch.int64).contiguous()
        B, L = p.shape
        if len(target) < B:
            raise IndexError("fewer reference captions than predictions")
        rows = [self._ref_ids(target[b]) for b in range(B)]
        ro = np.zeros(B + 1, dtype=np.int64)
        np.cumsum([len(r) for r in rows], out=ro[1:])
        rw = np.concatenate(rows + [np.zeros(1, dtype=np.int32)])
        rewards = torch.empty(B, L, dtype=torch.float32)
        delta = torch.empty(B, L, dtype=torch.float32)
        RLIB.call("bmhrl_bleu_prefix_rewards", p.data_ptr(), B, L, _ptr(self._tok_off), _ptr(self._tok_words),
                  len(self._tok_off) - 1, _ptr(rw), _ptr(ro), self._n, rewards.data_ptr(), delta.data_ptr(),
                  int(self.n_threads))
        self.counter += 1
        return delta.to(self.device), rewards.to(self.device)

    def delta_bleu_manager(self, pred, trg, mask, sections):
        manager_segment_score, rewards = self.delta_bleu(pred, trg, mask, sections)
        return manager_segment_score.clone().detach().float(), None

    def delta_bleu_worker(self, pred, trg):
        delta_bleu_step_reward, rewards = self.delta_bleu_step(pred, trg, self.gamma)
        return delta_bleu_step_reward.clone().detach().float(), rewards

    def delta_bleu(self, pred, trg, mask, sections):
        delta_bleu_step_reward, rewards = self.delta_bleu_step(pred, trg, self.gamma)
        delta_bleu_section_reward, segment_idx = self.delta_bleu_segment(
            delta_bleu_step_reward.clone().detach(), sections, self.gamma)
        return delta_bleu_section_reward, rewards

    def delta_bleu_segment(self, delta_bleu_step_reward, sections, gamma):
        """bleu.py:80-83: unlike the CIDEr scorer the discount here is the plain per-position one (no segments)."""
        segment_bleu_dif, segment_reward_index = segment_reward(delta_bleu_step_reward, sections)
        discounted_segment_reward = discontinue_reward(segment_bleu_dif, gamma)
        return discounted_segment_reward, segment_reward_index

    def delta_bleu_step(self, pred, tar, gamma):
        bleu_diff, rewards = self._bleu_diff(pred, tar)
        result = discontinue_reward(bleu_diff, gamma)
        return result, rewards
