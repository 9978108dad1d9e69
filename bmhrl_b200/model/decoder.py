"""model/decoder.py of the reference (DETR-style decoder with goal / detected-object attention) on sm_100a kernels."""
import torch
from torch import nn

from .. import ops
from .multihead_attention import MultiheadedAttention
from .utils import _get_clones, _get_activation_fn


class TransformerDecoder(nn.Module):
    """model/decoder.py:7-36."""

    def __init__(self, decoder_layer, num_layers, norm=None, return_intermediate=True):
        super().__init__()
        self.layers = _get_clones(decoder_layer, num_layers)
        self.num_layers = num_layers
        self.norm = norm
        self.return_intermediate = return_intermediate

    def _norm(self, x):
        if isinstance(self.norm, nn.LayerNorm):
            return ops.layer_norm(x, self.norm.weight, self.norm.bias, self.norm.eps)
        return self.norm(x)

    def forward(self, tgt, memory, mask, pos, query_pos, query_mask, goal, goal_mask, goal_pos=None, add_pos=False,
                detected_objects=None, obj_mask=None,):
        output = tgt
        intermediate = []
        for layer in self.layers:
            output = layer(output, memory, mask, pos, query_pos, query_mask, goal, goal_mask, goal_pos, add_pos,
                           detected_objects, obj_mask)
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


class TransformerDecoderLayer(nn.Module):
    """model/decoder.py:39-107, post-LN variant; norm1 is applied BEFORE the first residual add (:77-78)."""

    def __init__(self, d_model, nhead, d_model_C, d_goal, dim_feedforward=2048, dropout=0.1,
                 activation="relu", normalize_before=False):
        super().__init__()
        self.self_attn = MultiheadedAttention(d_model_C, d_model_C, d_model_C, nhead, dropout, d_model)
        self.multihead_attn = MultiheadedAttention(d_model_C, d_model, d_model, nhead, dropout, d_model)
        self.detected_attention = MultiheadedAttention(d_model_C, 256, 256, nhead, dropout, d_model)
        self.linear1 = nn.Linear(d_model_C, dim_feedforward)
        self.dropout = nn.Dropout(dropout)
        self.linear2 = nn.Linear(dim_feedforward, d_model_C)
        self.norm1 = nn.LayerNorm(d_model_C)
        self.norm2 = nn.LayerNorm(d_model_C)
        self.norm3 = nn.LayerNorm(d_model_C)
        self.norm4 = nn.LayerNorm(d_model_C)
        self.norm5 = nn.LayerNorm(d_model_C)
        self.dropout1 = nn.Dropout(dropout)
        self.dropout2 = nn.Dropout(dropout)
        self.dropout3 = nn.Dropout(dropout)
        self.dropout4 = nn.Dropout(dropout)
        self.dropout5 = nn.Dropout(dropout)
        self.goal_attention = MultiheadedAttention(d_model_C, d_goal, d_goal, nhead, dropout, d_model)
        self.activation = _get_activation_fn(activation)
        self.normalize_before = normalize_before
        self.positional_encoding = nn.Parameter()
        self._salts = None

    def _p(self, d):
        return d.p if self.training else 0.0

    @staticmethod
    def _ln(norm, x):
        return ops.layer_norm(x, norm.weight, norm.bias, norm.eps)

    def forward_post(self, tgt, memory, memory_mask, pos, query_pos, query_mask, goal, goal_mask, goal_pos,
                     detected_objects=None, add_pos=False, obj_mask=None):
        if self._salts is None:
            self._salts = (ops.next_salt(), ops.next_salt())
        if not add_pos:
            causal = True
            q = k = query_pos(tgt)
        else:
            q = k = ops.dropout_add(tgt, query_pos.expand_as(tgt), 0.0)
            causal = False
        tgt = self.self_attn(q, k, tgt, query_mask, causal=causal, resid=self._ln(self.norm1, tgt),
                             resid_drop_p=self._p(self.dropout1))
        tgt = self.multihead_attn(q, pos(memory), memory, memory_mask, resid=tgt, resid_drop_p=self._p(self.dropout2))
        tgt = self._ln(self.norm2, tgt)
        if goal is not None:
            tgt = self.goal_attention(query_pos(tgt), goal_pos(goal), goal, goal_mask, resid=tgt,
                                      resid_drop_p=self._p(self.dropout4))
            tgt = self._ln(self.norm4, tgt)
        if detected_objects is not None:
            tgt = self.detected_attention(q, detected_objects, detected_objects, obj_mask, resid=tgt,
                                          resid_drop_p=self._p(self.dropout5))
            tgt = self._ln(self.norm5, tgt)
        h = ops.linear(tgt, self.linear1.weight, self.linear1.bias, act=1, drop_p=self._p(self.dropout),
                       salt=self._salts[0])
        tgt = ops.linear(h, self.linear2.weight, self.linear2.bias, resid=tgt, drop_p=self._p(self.dropout3),
                         salt=self._salts[1])
        return self._ln(self.norm3, tgt)

    def forward(self, tgt, memory, memory_mask, pos, query_pos, query_mask, goal, goal_mask, goal_pos, add_pos=False,
                detected_objects=None, obj_mask=None):
        # the reference drops the caller's obj_mask here (model/decoder.py:106-107)
        return self.forward_post(tgt, memory, memory_mask, pos, query_pos, query_mask, goal, goal_mask, goal_pos,
                                 add_pos=add_pos, detected_objects=detected_objects, obj_mask=None)
