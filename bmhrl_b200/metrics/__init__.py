"""Reward scorers of the RL step (reference package `metrics/`): CIDEr / BLEU prefix rewards and the discount / segment
scans, computed by the host library libbmhrl_reward.so (include/bmhrl_reward.h).  SURVEY.md section 8(f) rank 1."""
from .bleu import BleuScorer
from .cider import CiderScorer
from .util import discontinue_reward, segment_reward

__all__ = ["BleuScorer", "CiderScorer", "discontinue_reward", "segment_reward"]
