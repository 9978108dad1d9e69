"""model/multihead_attention.py of the reference on sm_100a kernels.

`MultiheadedAttention` keeps the reference's constructor, forward signature, attribute names and state_dict keys
(linear_Q2d / linear_K2d / linear_V2d / linear_d2Q, each an nn.Linear used as a parameter container); its forward
runs the projection GEMMs and the fused masked attention core of libbmhrl_b200 -- no ATen math, no CPU fallback.
"""
import torch.nn as nn

from .. import ops


def attention(Q, K, V, mask, dropout=None, causal=False):
    """model/multihead_attention.py:7-31 for already head-split tensors Q (B,H,Sq,dk), K/V (B,H,Sk,dk) and a mask
    (B,1,1,Sk) / (B,1,Sq,Sk).  `dropout`: None, an nn.Dropout (its p is used while it is in training mode) or a
    float.  Returns (B,H,Sq,dk)."""
    B, H, Sq, dk = Q.shape
    p = 0.0
    if isinstance(dropout, nn.Dropout):
        p = dropout.p if dropout.training else 0.0
    elif dropout is not None:
        p = float(dropout)
    merge = lambda t: t.transpose(1, 2).reshape(B, t.shape[2], H * dk)
    out = ops.mha_core(merge(Q), merge(K), merge(V), mask, H, causal=causal, drop_p=p, salt=ops.next_salt())
    return out.view(B, Sq, H, dk).transpose(1, 2)


class MultiheadedAttention(nn.Module):

    def __init__(self, d_model_Q, d_model_K, d_model_V, H, dout_p=0.0, d_model=None):
        super(MultiheadedAttention, self).__init__()
        self.d_model_Q = d_model_Q
        self.d_model_K = d_model_K
        self.d_model_V = d_model_V
        self.H = H
        self.d_model = d_model
        self.dout_p = dout_p

        if self.d_model is None:
            self.d_model = self.d_model_Q

        self.d_k = self.d_model // H

        self.linear_Q2d = nn.Linear(self.d_model_Q, self.d_model)
        self.linear_K2d = nn.Linear(self.d_model_K, self.d_model)
        self.linear_V2d = nn.Linear(self.d_model_V, self.d_model)
        self.linear_d2Q = nn.Linear(self.d_model, self.d_model_Q)

        self.dropout = nn.Dropout(self.dout_p)

        assert self.d_model % H == 0
        self._salts = None
        # inference extension (bmhrl_b200.decode.GraphedGreedyDecoder): (k, v) = the K / V projections of a fixed
        # memory, computed once per clip; while it is set, forward() skips linear_K2d / linear_V2d
        self.memory_kv = None

    def project_memory(self, K, V):
        """linear_K2d(K), linear_V2d(V): what forward() computes from its K / V arguments before the core."""
        return (ops.linear(K, self.linear_K2d.weight, self.linear_K2d.bias),
                ops.linear(V, self.linear_V2d.weight, self.linear_V2d.bias))

    def _call_salts(self):
        # dropout call-site ids: fixed per module instance so a CUDA-graph replay and an eager run agree
        if self._salts is None:
            self._salts = (ops.next_salt(), ops.next_salt())
        return self._salts

    def forward(self, Q, K, V, mask, causal=False, resid=None, resid_drop_p=0.0):
        '''
            Q, K, V: (B, Sq, Dq), (B, Sk, Dk), (B, Sv, Dv); mask: (B, 1, Sk) or (B, Sq, Sk); Sk = Sv.
            resid / resid_drop_p (extension used by ResidualConnection): returns resid + dropout(attention output),
            fused into the epilogue of the output projection.
        '''
        if Q.dim() != 3:
            raise RuntimeError("MultiheadedAttention expects (B, S, D) inputs")
        s_core, s_out = self._call_salts()
        if self.memory_kv is not None:
            if self.training:
                raise RuntimeError("MultiheadedAttention.memory_kv is an inference-only cache")
            k, v = self.memory_kv
            q = ops.linear(Q, self.linear_Q2d.weight, self.linear_Q2d.bias)
            o = ops.mha_core(q, k, v, mask, self.H, causal=causal, drop_p=0.0, salt=s_core)
            return ops.linear(o, self.linear_d2Q.weight, self.linear_d2Q.bias, resid=resid, drop_p=0.0, salt=s_out)
        if ops.BF16 and K is V and resid is None:
            from .. import ops16
            kv = None if K is Q else K
            if ops16.attention_block_supported(self, Q, kv):
                return ops16.attention_block(Q, kv, mask, self, None, resid=False, p_res=0.0, causal=causal)
        # the three projections are independent: concurrent branches (and so are their backward passes)
        q, k, v = ops.parallel(Q.device,
                               lambda: ops.linear(Q, self.linear_Q2d.weight, self.linear_Q2d.bias),
                               lambda: ops.linear(K, self.linear_K2d.weight, self.linear_K2d.bias),
                               lambda: ops.linear(V, self.linear_V2d.weight, self.linear_V2d.bias))
        p = self.dropout.p if self.training else 0.0
        o = ops.mha_core(q, k, v, mask, self.H, causal=causal, drop_p=p, salt=s_core)
        return ops.linear(o, self.linear_d2Q.weight, self.linear_d2Q.bias, resid=resid, drop_p=resid_drop_p,
                          salt=s_out)
