"""`CiderScorer` with the reference's interface (metrics/cider.py:12-109): per-prefix CIDEr rewards of the sampled
captions for the worker / manager REINFORCE steps.  Strings are mapped to int32 word ids once (vocabulary and corpus
at construction, reference captions per call); the n-gram counting and scoring run in libbmhrl_reward.so.
"""
import ctypes

import numpy as np
import torch

from ._lib import RLIB
from .util import discontinue_reward, segment_reward


def word_from_vector(vocab, indices):
    """metrics/batched_meteor.py:52-53"""
    return np.array([vocab.itos[int(i)] for i in indices])


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class CiderScorer:
    """Same constructor and methods as the reference class.  `vocab` needs `.itos`; `dictionary` is the iterable of
    tokenised training captions (`train_dataset.word_counter`, captioning_dataset.py:365)."""

    def __init__(self, vocab, dictionary, device, gamma, gamma_manager, n=4, sigma=6.0, n_threads=0):
        assert (n <= 4 and n > 0)
        self.counter = 0
        self.vocab = vocab
        self.device = device
        self._n = n
        self._sigma = sigma
        self.type = "CIDER"
        self.gamma = gamma
        self.gamma_m = gamma_manager
        self.n_threads = n_threads
        self._word_id = {}
        self._ref_cache = {}
        # token id -> the words it contributes to " ".join(hypo).split() (cider.py:42, util.py:26)
        tok_words, tok_off = [], [0]
        for s in vocab.itos:
            tok_words.extend(self._wid(w) for w in s.split())
            tok_off.append(len(tok_words))
        self._tok_words = np.asarray(tok_words if tok_words else [0], dtype=np.int32)
        self._tok_off = np.asarray(tok_off, dtype=np.int32)
        self._tok_is_end = np.asarray([s == "</s>" for s in vocab.itos], dtype=np.uint8)
        # precook_corpus (cider.py:112-120): n-grams are tuples of the RAW corpus tokens
        words, off = [], [0]
        for cap in dictionary:
            words.extend(self._wid(w) for w in cap)
            off.append(len(words))
        cw = np.asarray(words if words else [0], dtype=np.int32)
        co = np.asarray(off, dtype=np.int64)
        h = ctypes.c_void_p()
        RLIB.call("bmhrl_cider_create", _ptr(cw), _ptr(co), len(off) - 1, n, float(sigma), ctypes.byref(h))
        self._h = h

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                RLIB.raw("bmhrl_cider_destroy")(h)
            except Exception:
                pass

    def _wid(self, w):
        i = self._word_id.get(w)
        if i is None:
            i = self._word_id[w] = len(self._word_id)
        return i

    def _ref_ids(self, caption):
        """target[b].lower().split() as word ids (cider.py:40); captions recur every epoch, so the mapping is kept."""
        ids = self._ref_cache.get(caption)
        if ids is None:
            if len(self._ref_cache) >= 1 << 18:
                self._ref_cache.clear()
            ids = self._ref_cache[caption] = np.asarray([self._wid(w) for w in caption.lower().split()], dtype=np.int32)
        return ids

    def document_frequency(self, ngram):
        """Count of `ngram` (tuple of strings) in the kept document-frequency table, 0 if absent."""
        ids = [self._word_id.get(w, -1) for w in ngram]
        if min(ids) < 0:
            return 0
        a = np.asarray(ids, dtype=np.int32)
        return int(RLIB.raw("bmhrl_cider_df_lookup")(self._h, _ptr(a), len(ids)))

    def df_size(self):
        return int(RLIB.raw("bmhrl_cider_df_size")(self._h))

    def _cider_diff(self, pred, target):
        """cider.py:28-70: returns (delta (B, L) fp32 on self.device, rewards (B, L) fp64 on self.device)."""
        p = pred.detach().to("cpu", torch.int64).contiguous()
        B, L = p.shape
        if len(target) < B:
            raise IndexError("fewer reference captions than predictions")
        rows = [self._ref_ids(target[b]) for b in range(B)]
        ro = np.zeros(B + 1, dtype=np.int64)
        np.cumsum([len(r) for r in rows], out=ro[1:])
        rw = np.concatenate(rows + [np.zeros(1, dtype=np.int32)])
        rewards = torch.empty(B, L, dtype=torch.float64)
        delta = torch.empty(B, L, dtype=torch.float32)
        RLIB.call("bmhrl_cider_prefix_rewards", self._h, p.data_ptr(), B, L, _ptr(self._tok_off), _ptr(self._tok_words),
                  _ptr(self._tok_is_end), len(self._tok_is_end), _ptr(rw), _ptr(ro), rewards.data_ptr(),
                  delta.data_ptr(), int(self.n_threads))
        self.counter += 1
        return delta.to(self.device), rewards.to(self.device)

    def delta_cider_manager(self, pred, trg, mask, sections):
        """cider.py:72-80; like the reference it closes `sections` at the reference caption's length IN PLACE."""
        B, L = pred.shape[0], sections.shape[1]
        ends = [len(trg[i].split()) for i in range(B)]
        if B and max(ends) >= L:
            raise IndexError(f"index {max(ends)} is out of bounds for dimension 0 with size {L}")
        e = torch.tensor(ends, dtype=torch.long, device=sections.device)
        pos = torch.arange(L, device=sections.device)
        head = sections[:B]
        head.masked_fill_(pos[None, :] == e[:, None], 1)
        head.masked_fill_(pos[None, :] > e[:, None], 0)
        manager_segment_score, rewards = self.delta_cider(pred, trg, mask, sections)
        return manager_segment_score.clone().detach().float(), None

    def delta_cider_worker(self, pred, trg):
        delta_cider_step_reward, rewards = self.delta_cider_step(pred, trg, self.gamma)
        return delta_cider_step_reward.clone().detach().float(), rewards

    def delta_cider(self, pred, trg, mask, sections):
        delta_cider_step_reward, rewards = self.delta_cider_step(pred, trg, self.gamma)
        delta_cider_section_reward, segment_idx = self.delta_cider_segment(
            delta_cider_step_reward.clone().detach(), sections, self.gamma)
        return delta_cider_section_reward, rewards

    def delta_cider_segment(self, delta_cider_step_reward, sections, gamma):
        segment_cider_dif, segment_reward_index = segment_reward(delta_cider_step_reward, sections)
        discounted_segment_reward = discontinue_reward(segment_cider_dif, gamma, segments=sections)
        return discounted_segment_reward, segment_reward_index

    def delta_cider_step(self, pred, tar, gamma):
        cider_diff, rewards = self._cider_diff(pred, tar)
        result = discontinue_reward(cider_diff, gamma)
        return result, rewards
