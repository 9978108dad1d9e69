"""loss/biased_kl.py of the reference: BiasedKL and Reinforce on fused sm_100a kernels."""
import torch
import torch.nn as nn

from .. import ops


class BiasedKL(nn.Module):
    """loss/biased_kl.py:11-53.  The target distribution (prior + (1-ls)(1-a) at the ground truth + (1-ls) a at the
    sampled token, + 1e-8 everywhere) is evaluated inside the kernel; gradients flow to `pred` AND to
    `biased_offset` (torch >= 2 differentiates kl_div w.r.t. its target)."""

    def __init__(self, label_smoothing, pad_idx):
        super(BiasedKL, self).__init__()
        self.pad_idx = pad_idx
        self.ls = label_smoothing
        self.trg_factor = 1 - self.ls
        self.eps = 1e-6

    def forward(self, pred, trg, biased_trg, biased_offset, segments=None):
        return ops.biased_kl(pred, trg.long(), biased_trg.long(), biased_offset, self.ls, self.pad_idx)

    def row_sums(self, pred, trg, biased_trg, biased_offset, segments=None):
        return ops.biased_kl(pred, trg.long(), biased_trg.long(), biased_offset, self.ls, self.pad_idx, rowsum=True)


class Reinforce(nn.Module):
    """loss/biased_kl.py:61-81: -mean(adv.detach() * log clamp(pi(a))) + mean(adv^2)."""

    def __init__(self):
        super(Reinforce, self).__init__()
        self.pad_idx = 0
        self.value_const = 0.5
        self.entropie_const = 0.005
        self.eps = 1e-5

    def forward(self, pred, action, value, critic_value):
        B, S, V = pred.shape
        logp = ops.reinforce_log_prob(pred, action.reshape(B, S).long(), self.eps)
        return ops.reinforce_loss(logp, value, critic_value)
