"""model/blocks.py of the reference on sm_100a kernels (same class names, constructor arguments, state_dict keys)."""
from copy import deepcopy
import math

import torch
import torch.nn as nn

from .. import ops


class LayerStack(nn.Module):
    """model/blocks.py:10-21."""

    def __init__(self, layer, N):
        super(LayerStack, self).__init__()
        self.layers = clone(layer, N)

    # optional callable(i, x) -> x applied to the output of layer i < N-1 (the data-parallel trainer marks the layer
    # boundaries so that gradient buckets of finished layers can be reduced while backward continues below them)
    layer_boundary_hook = None

    def forward(self, x, masks):
        hook = self.layer_boundary_hook
        for i, layer in enumerate(self.layers):
            x = layer(x, masks)
            if hook is not None and i + 1 < len(self.layers):
                x = hook(i, x)
        return x


def clone(module, N):
    return nn.ModuleList([deepcopy(module) for _ in range(N)])


class Identity(nn.Module):

    def __init__(self):
        super(Identity, self).__init__()

    def forward(self, x):
        return x


class VocabularyEmbedder(nn.Module):
    """model/blocks.py:35-69: Embedding(tok) * sqrt(emb_dim)."""

    def __init__(self, voc_size, emb_dim):
        super(VocabularyEmbedder, self).__init__()
        self.voc_size = voc_size
        self.emb_dim = emb_dim
        self.embedder = nn.Embedding(voc_size, emb_dim)

    def forward(self, x):  # x - tokens (B, seq_len)
        if isinstance(self.embedder, nn.Sequential):  # pretrained vectors of another width (:61-67)
            e = ops.embed(x, self.embedder[0].weight, 1.0)
            e = ops.linear(e, self.embedder[1].weight, self.embedder[1].bias, act=1)
            return e * math.sqrt(self.emb_dim)
        return ops.embed(x, self.embedder.weight, math.sqrt(self.emb_dim))

    def init_word_embeddings(self, weight_matrix, emb_weights_req_grad=True):
        if weight_matrix is None:
            return
        pretrained_voc_size, pretrained_emb_dim = weight_matrix.shape
        if self.emb_dim == pretrained_emb_dim:
            self.embedder = self.embedder.from_pretrained(weight_matrix)
            self.embedder.weight.requires_grad = emb_weights_req_grad
        else:
            self.embedder = nn.Sequential(
                nn.Embedding(self.voc_size, pretrained_emb_dim).from_pretrained(weight_matrix),
                nn.Linear(pretrained_emb_dim, self.emb_dim),
                nn.ReLU()
            )
            self.embedder[0].weight.requires_grad = emb_weights_req_grad


class FeatureEmbedder(nn.Module):
    """model/blocks.py:72-85: relu(Linear(x) * sqrt(d_model)) == sqrt(d_model) * relu(Linear(x))."""

    def __init__(self, d_feat, d_model):
        super(FeatureEmbedder, self).__init__()
        self.d_model = d_model
        self.embedder = nn.Linear(d_feat, d_model)
        self.activation = nn.ReLU()

    def forward(self, x):
        return ops.linear(x, self.embedder.weight, self.embedder.bias, act=1) * math.sqrt(self.d_model)


class PositionalEncoder(nn.Module):
    """model/blocks.py:88-112.  The reference keeps an fp64 table on the CPU and copies a slice to the device on
    every call; here the fp32 cast of the same table is cached on the device and the add (+ dropout) is one kernel."""

    def __init__(self, d_model, dout_p, seq_len=3660):
        super(PositionalEncoder, self).__init__()
        self.d_model = d_model
        self.seq_len = seq_len
        self.dropout = nn.Dropout(dout_p)
        self._salt = None

    @property
    def pos_enc_mat(self):
        """(1, seq_len, d_model) fp64 table, as the reference attribute of the same name."""
        from ..ops import positional_table_f64
        return positional_table_f64(self.seq_len, self.d_model).unsqueeze(0)

    def forward(self, x, x2=None):
        """x2 (extension): a second tensor added before the encoding (fuses V = rgb + flow)."""
        if len(x.shape) == 3:
            if self._salt is None:
                self._salt = ops.next_salt()
            p = self.dropout.p if self.training else 0.0
            x = ops.add_posenc(x, x2, drop_p=p, salt=self._salt, max_len=self.seq_len)
        return x


class Transpose(nn.Module):

    def __init__(self):
        super(Transpose, self).__init__()

    def forward(self, x):
        return x.permute(0, 2, 1)


class ResidualConnection(nn.Module):
    """model/blocks.py:128-144: x + dropout(sublayer(LayerNorm(x))).

    A sublayer that exposes `fused_residual(normed, x, p)` gets the residual add and the dropout folded into the
    epilogue of its last GEMM; any other callable goes through the generic dropout+add kernel."""

    def __init__(self, size, dout_p):
        super(ResidualConnection, self).__init__()
        self.norm = nn.LayerNorm(size)
        self.dropout = nn.Dropout(dout_p)
        self._salt = None
        # bf16 math mode: also emit the bf16 twin of the block's output (set where that output is the key / value
        # source of a later cross-attention, so that no separate cast kernel runs)
        self.emit_twin = False

    def forward(self, x, sublayer):
        p = self.dropout.p if self.training else 0.0
        if ops.BF16 and x.dim() == 3:
            # one autograd node for the whole block: LayerNorm -> sublayer -> dropout -> + x, bf16 interior tensors
            block = getattr(sublayer, "block16", None)
            if block is not None:
                out = block(x, self.norm, p, self.emit_twin)
                if out is not None:
                    return out
        res = ops.layer_norm(x, self.norm.weight, self.norm.bias, self.norm.eps)
        fused = getattr(sublayer, "fused_residual", None)
        if fused is not None:
            return fused(res, x, p)
        res = sublayer(res)
        if self._salt is None:
            self._salt = ops.next_salt()
        return ops.dropout_add(x, res, p, self._salt)


class BridgeConnection(nn.Module):
    """model/blocks.py:147-161: relu(dropout(Linear(LayerNorm(x))))."""

    def __init__(self, in_dim, out_dim, dout_p):
        super(BridgeConnection, self).__init__()
        self.norm = nn.LayerNorm(in_dim)
        self.linear = nn.Linear(in_dim, out_dim)
        self.dropout = nn.Dropout(dout_p)
        self.activation = nn.ReLU()
        self._salt = None

    def forward(self, x):
        if self._salt is None:
            self._salt = ops.next_salt()
        p = self.dropout.p if self.training else 0.0
        x = ops.layer_norm(x, self.norm.weight, self.norm.bias, self.norm.eps)
        # relu(dropout(z)) == dropout(relu(z)): the mask and the 1/(1-p) scale commute with relu
        return ops.linear(x, self.linear.weight, self.linear.bias, act=1, drop_p=p, salt=self._salt)


class PositionwiseFeedForward(nn.Module):
    """model/blocks.py:164-187: fc2(dropout(relu(fc1(x))))."""

    def __init__(self, d_model, d_ff, dout_p):
        super(PositionwiseFeedForward, self).__init__()
        self.d_model = d_model
        self.d_ff = d_ff
        self.dout_p = dout_p
        self.fc1 = nn.Linear(d_model, d_ff)
        self.fc2 = nn.Linear(d_ff, d_model)
        self.dropout = nn.Dropout(dout_p)
        self._salts = None

    def forward(self, x, resid=None, resid_drop_p=0.0):
        '''In, Out: (B, S, D)'''
        if self._salts is None:
            self._salts = (ops.next_salt(), ops.next_salt())
        p = self.dropout.p if self.training else 0.0
        h = ops.linear(x, self.fc1.weight, self.fc1.bias, act=1, drop_p=p, salt=self._salts[0])
        return ops.linear(h, self.fc2.weight, self.fc2.bias, resid=resid, drop_p=resid_drop_p, salt=self._salts[1])

    def fused_residual(self, normed, x, p):
        return self.forward(normed, resid=x, resid_drop_p=p)

    def block16(self, x, norm, p, emit_twin=False):
        """bf16 math mode: x + dropout(fc2(dropout(relu(fc1(LayerNorm(x)))))) as one node (ops16._FfnBlock)."""
        from .. import ops16
        if self.d_model % 4 or self.d_ff % 4:
            return None
        return ops16.ffn_block(x, self, norm, resid=True, p_res=p, emit_twin=emit_twin)


class AttentionSublayer:
    """The `lambda x: att(x, K, V, mask)` closures of the reference's layers (model/bm_hrl_agent.py:358-364,89-94)
    as an object that ResidualConnection can fuse with.  K / V = None means self-attention on the normed input."""

    def __init__(self, att, K=None, V=None, mask=None, causal=False):
        self.att, self.K, self.V, self.mask, self.causal = att, K, V, mask, causal

    def __call__(self, q):
        return self.att(q, q if self.K is None else self.K, q if self.V is None else self.V, self.mask, self.causal)

    def fused_residual(self, normed, x, p):
        return self.att(normed, normed if self.K is None else self.K, normed if self.V is None else self.V, self.mask,
                        self.causal, resid=x, resid_drop_p=p)

    def block16(self, x, norm, p, emit_twin=False):
        """bf16 math mode: x + dropout(attention(LayerNorm(x), K, V)) as one node (ops16._AttnBlock); None when the
        shape has no one-tile kernel or keys and values come from different tensors (the caller then takes the
        per-op path)."""
        from .. import ops16
        if self.K is not self.V or self.att.memory_kv is not None:
            return None
        if not ops16.attention_block_supported(self.att, x, self.K):
            return None
        return ops16.attention_block(x, self.K, self.mask, self.att, norm, resid=True, p_res=p, causal=self.causal,
                                     emit_twin=emit_twin)
