"""model/encoder.py of the reference (DETR-style post-LN encoder) on sm_100a kernels."""
import torch
from torch import nn

from .. import ops
from .multihead_attention import MultiheadedAttention
from .utils import _get_clones, _get_activation_fn


class TransformerEncoder(nn.Module):
    """model/encoder.py:10-36 (including the double application of `norm` to the last intermediate, :28-33)."""

    def __init__(self, encoder_layer, num_layers, norm=None, cfg=None, return_intermediate=True):
        super().__init__()
        self.cfg = cfg
        self.layers = _get_clones(encoder_layer, num_layers)
        self.num_layers = num_layers
        self.norm = norm
        self.return_intermediate = return_intermediate

    def _norm(self, x):
        if isinstance(self.norm, nn.LayerNorm):
            return ops.layer_norm(x, self.norm.weight, self.norm.bias, self.norm.eps)
        return self.norm(x)

    def forward(self, src, mask, pos_enc):
        output = src
        intermediate = []
        for layer in self.layers:
            output = layer(output, src_mask=mask, pos=pos_enc)
            if self.return_intermediate:
                intermediate.append(output)
        if self.norm is not None:
            output = self._norm(output)
            if self.return_intermediate:
                intermediate.pop()
                intermediate.append(self._norm(output))
        if self.return_intermediate:
            return torch.stack(intermediate)
        return output


class TransformerEncoderLayer(nn.Module):
    """model/encoder.py:39-75."""

    def __init__(self, d_model, nhead, dim_feedforward=2048, dropout=0.1,
                 activation="relu", normalize_before=True, embed_size=300):
        super().__init__()
        self.self_attn = MultiheadedAttention(d_model, d_model, d_model, nhead, dropout, d_model)
        self.linear1 = nn.Linear(d_model, dim_feedforward)
        self.dropout = nn.Dropout(dropout)
        self.linear2 = nn.Linear(dim_feedforward, d_model)
        self.embed = nn.Linear(d_model, embed_size)
        self.norm1 = nn.LayerNorm(d_model)
        self.norm2 = nn.LayerNorm(d_model)
        self.dropout1 = nn.Dropout(dropout)
        self.dropout2 = nn.Dropout(dropout)

        self.activation = _get_activation_fn(activation)
        self.normalize_before = normalize_before
        self._salts = None

    def _p(self, d):
        return d.p if self.training else 0.0

    def forward_post(self, src, mask, pos):
        if self._salts is None:
            self._salts = (ops.next_salt(), ops.next_salt())
        q = k = pos(src)
        src = self.self_attn(q, k, src, mask, resid=src, resid_drop_p=self._p(self.dropout1))
        src = ops.layer_norm(src, self.norm1.weight, self.norm1.bias, self.norm1.eps)
        h = ops.linear(src, self.linear1.weight, self.linear1.bias, act=1, drop_p=self._p(self.dropout),
                       salt=self._salts[0])
        src = ops.linear(h, self.linear2.weight, self.linear2.bias, resid=src, drop_p=self._p(self.dropout2),
                         salt=self._salts[1])
        return ops.layer_norm(src, self.norm2.weight, self.norm2.bias, self.norm2.eps)

    def forward(self, src, src_mask, pos):
        return self.forward_post(src, src_mask, pos)
