"""loss/label_smoothing.py of the reference as one fused kernel (no (B*L, V) target distribution is built)."""
import torch.nn as nn

from .. import ops


class LabelSmoothing(nn.Module):

    def __init__(self, smoothing, pad_idx):
        super(LabelSmoothing, self).__init__()
        self.smoothing = smoothing
        self.pad_idx = pad_idx

    def forward(self, pred, target):  # pred (B, S, V) log-probs, target (B, S)
        """-> (B*S, V) element-wise F.kl_div(pred, smoothed_target, reduction='none') (loss/label_smoothing.py:12-32)."""
        return ops.label_smoothing_kl(pred, target.long(), self.smoothing, self.pad_idx)

    def row_sums(self, pred, target):
        """(B*S,) sums over the vocabulary of forward()'s result, without materialising it."""
        return ops.label_smoothing_kl(pred, target.long(), self.smoothing, self.pad_idx, rowsum=True)
