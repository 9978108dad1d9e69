"""model/bm_hrl_agent.py of the reference on sm_100a kernels.

Same class names, constructor / forward signatures, attribute names and state_dict keys (so reference checkpoints
load and `teach_*` phase switching works); every forward launches kernels of libbmhrl_b200 through bmhrl_b200.ops.
Deliberate differences from the reference file, all listed in SURVEY.md section 0.1:
  * no debug prints of whole tensors (F8);
  * segment labels keep their (B, L) shape for B == 1 or L == 1 (F6: `.squeeze()` -> per-row kernel output);
  * `Manager.expand_goals` is a device-side scan instead of a host loop over `torch.nonzero` (same results,
    including the tail-zeroing quirks), exploration noise is drawn on the device (F7);
  * never-called parameters (BMFusionLayer.feed_forward, Manager.core -- F11) are still constructed and saved.
The ablation agents (UnimodalAgent / AudioAgent / VideoAgent) are out of scope.
"""
import os

import torch
import torch.nn as nn

from .. import ops
from .blocks import (AttentionSublayer, LayerStack, PositionalEncoder, PositionwiseFeedForward, ResidualConnection,
                     VocabularyEmbedder, clone)
from .multihead_attention import MultiheadedAttention
from ..scripts.device import get_device


class AReLU(nn.Module):
    """model/bm_hrl_agent.py:13-23."""

    def __init__(self, alpha=0.90, beta=2.0):
        super(AReLU, self).__init__()
        self.alpha = nn.Parameter(torch.tensor([alpha]))
        self.beta = nn.Parameter(torch.tensor([beta]))

    def forward(self, input):
        with torch.no_grad():  # only used inside the frozen critic
            return ops.arelu(input, self.alpha, self.beta)


class ModelBase(nn.Module):
    def __init__(self, name) -> None:
        super(ModelBase, self).__init__()
        self.name = name

    def save_model(self, checkpoint_dir):
        torch.save(self.state_dict(), checkpoint_dir + f"/{self.name}.pt")

    def load_model(self, checkpoint_dir):
        self.load_state_dict(torch.load(checkpoint_dir + f"/{self.name}.pt"))


class ModalityProjection(nn.Module):
    """model/bm_hrl_agent.py:40-51: dropout(Linear(LayerNorm(x)))."""

    def __init__(self, d_mod, d_out, p_dout) -> None:
        super(ModalityProjection, self).__init__()
        self.linear = nn.Linear(d_mod, d_out)
        self.norm = nn.LayerNorm(d_mod)
        self.dropout = nn.Dropout(p_dout)
        self._salt = None

    def forward(self, x):
        if self._salt is None:
            self._salt = ops.next_salt()
        x = ops.layer_norm(x, self.norm.weight, self.norm.bias, self.norm.eps)
        return ops.linear(x, self.linear.weight, self.linear.bias, drop_p=self.dropout.p if self.training else 0.0,
                          salt=self._salt)


class BMFusionLayer(nn.Module):
    """model/bm_hrl_agent.py:54-117."""

    def __init__(self, d_model_A, d_model_V, d_model_C, d_model, d_ff_c, dout_p, H) -> None:
        super(BMFusionLayer, self).__init__()
        self.res_layer_self_att = ResidualConnection(d_model_C, dout_p)
        self.self_att = MultiheadedAttention(d_model_C, d_model_C, d_model_C, H, dout_p, d_model)

        self.res_layer_enc_att_A = ResidualConnection(d_model_C, dout_p)
        self.res_layer_enc_att_V = ResidualConnection(d_model_C, dout_p)
        self.enc_att_A = MultiheadedAttention(d_model_C, d_model_A, d_model_A, H, dout_p, d_model)
        self.enc_att_V = MultiheadedAttention(d_model_C, d_model_V, d_model_V, H, dout_p, d_model)

        self.feed_forward = PositionwiseFeedForward(d_model_C, d_ff_c, dout_p)  # constructed, never called (F11)

        self.normCA = nn.LayerNorm(d_model_C)
        self.normCV = nn.LayerNorm(d_model_C)

        self.a_v_constant = nn.Parameter(torch.tensor([0.0]))

    def forward(self, x, masks):
        '''
        x (C, memory): C (B, Sc, Dc); memory (Av (B, Sa, Da), Va (B, Sv, Dv))
        masks: V_mask (B, 1, Sv); A_mask (B, 1, Sa); C_mask (B, Sc, Sc)
        '''
        C, memory = x
        Av, Va = memory

        C = self.res_layer_self_att(C, AttentionSublayer(self.self_att, None, None, masks['C_mask']))
        def branch_a():
            Ca = self.res_layer_enc_att_A(C, AttentionSublayer(self.enc_att_A, Av, Av, masks['A_mask']))
            return ops.layer_norm(Ca, self.normCA.weight, self.normCA.bias, self.normCA.eps)

        def branch_v():
            Cv = self.res_layer_enc_att_V(C, AttentionSublayer(self.enc_att_V, Va, Va, masks['V_mask']))
            return ops.layer_norm(Cv, self.normCV.weight, self.normCV.bias, self.normCV.eps)

        Cv, Ca = ops.parallel(C.device, branch_v, branch_a)

        # sigmoid(clamp(a_v_constant, -2, 2)) * Cv + (1 - .) * Ca
        fused_features = ops.gated_mix(Cv, Ca, self.a_v_constant)
        return fused_features, memory


class BMFusion(nn.Module):
    """model/bm_hrl_agent.py:120-130."""

    def __init__(self, d_model_A, d_model_V, d_model_C, d_model, d_ff_c, dout_p, H, N) -> None:
        super(BMFusion, self).__init__()
        layer = BMFusionLayer(d_model_A, d_model_V, d_model_C, d_model, d_ff_c, dout_p, H)
        self.decoder = LayerStack(layer, N)

    def forward(self, x, masks):
        C, memory = self.decoder(x, masks)
        return C


class SegmentCritic(nn.Module):
    """model/bm_hrl_agent.py:186-215: frozen LSTM x4 -> AReLU -> GRU x2 -> AReLU -> Linear(600 -> 1), no_grad.
    nn.LSTM / nn.GRU are parameter containers only (reference key names); the recurrences run in rnn.cu."""

    def __init__(self, cfg):
        super(SegmentCritic, self).__init__()
        self.name = "SegmentCritic"
        embed_dim = cfg.d_model_caps

        self.lstm = nn.LSTM(embed_dim, 2 * embed_dim, num_layers=4, batch_first=True)
        self.gru = nn.GRU(2 * embed_dim, 2 * embed_dim, num_layers=2, batch_first=True)
        self.lin = nn.Linear(2 * embed_dim, 1)
        self.relu = AReLU()
        self.relu2 = AReLU()

        for name, param in self.named_parameters():
            param.requires_grad = False

        path = getattr(cfg, "rl_critic_path", None)
        if path is not None:  # the reference always loads; None (extension) keeps the random init
            self.load_state_dict(torch.load(path))

    wavefront = True  # all six recurrent layers in one cooperative launch (ops.critic_wavefront); False: layer by layer

    def hidden(self, embedded_indices):
        with torch.no_grad():
            if self.wavefront and os.environ.get("BMHRL_CRITIC_WAVEFRONT", "0") != "0":
                h = ops.critic_wavefront(embedded_indices, self.lstm, self.gru, self.relu.alpha, self.relu.beta)
                if h is not None:
                    return self.relu2(h)
            h = embedded_indices
            for l in range(self.lstm.num_layers):
                h = ops.rnn_layer("lstm", h, getattr(self.lstm, f"weight_ih_l{l}"), getattr(self.lstm, f"weight_hh_l{l}"),
                                  getattr(self.lstm, f"bias_ih_l{l}"), getattr(self.lstm, f"bias_hh_l{l}"))
            h = self.relu(h)
            for l in range(self.gru.num_layers):
                h = ops.rnn_layer("gru", h, getattr(self.gru, f"weight_ih_l{l}"), getattr(self.gru, f"weight_hh_l{l}"),
                                  getattr(self.gru, f"bias_ih_l{l}"), getattr(self.gru, f"bias_hh_l{l}"))
            return self.relu2(h)

    def forward(self, embedded_indices):
        with torch.no_grad():
            return ops.segment_labels(self.hidden(embedded_indices), self.lin.weight, self.lin.bias, 0.0)[0]

    def step_labels(self, state, x_p, p, threshold):
        """Incremental form of `scores_and_labels` (the critic is causal: zero initial state, unidirectional LSTM / GRU,
        position-wise AReLU and Linear): x_p (B,1,E) is the embedding of the token at position p, positions 0..p-1 have
        been fed before.  Writes the label of position p into state.labels[:, p] -- the same bits as running the whole
        prefix (one `rnn_step` per layer instead of a p+1-step recurrence per layer per generated token)."""
        with torch.no_grad():
            h = x_p
            for l, (gx, out, c) in enumerate(state.lstm):
                g = ops.linear(h, getattr(self.lstm, f"weight_ih_l{l}"), getattr(self.lstm, f"bias_ih_l{l}"))
                gx[:, p].copy_(g[:, 0])
                ops.rnn_step("lstm", gx, getattr(self.lstm, f"weight_hh_l{l}"), getattr(self.lstm, f"bias_hh_l{l}"),
                             out, c, p)
                h = out[:, p:p + 1]
            h = self.relu(h)
            for l, (gx, out) in enumerate(state.gru):
                g = ops.linear(h, getattr(self.gru, f"weight_ih_l{l}"), getattr(self.gru, f"bias_ih_l{l}"))
                gx[:, p].copy_(g[:, 0])
                ops.rnn_step("gru", gx, getattr(self.gru, f"weight_hh_l{l}"), getattr(self.gru, f"bias_hh_l{l}"),
                             out, None, p)
                h = out[:, p:p + 1]
            h = self.relu2(h)
            _, labels = ops.segment_labels(h, self.lin.weight, self.lin.bias, threshold)
            state.labels[:, p].copy_(labels[:, 0])

    def scores_and_labels(self, embedded_indices, threshold):
        """(B,L,1) raw scores and int32 (B,L) labels sigmoid(score) > threshold in one pass."""
        with torch.no_grad():
            return ops.segment_labels(self.hidden(embedded_indices), self.lin.weight, self.lin.bias, threshold)


class SegmentCriticDecodeState:
    """Caller-kept buffers of `SegmentCritic.step_labels` (greedy decoding appends one token per step): for every
    recurrent layer the input contributions gx, the hidden states of all positions and the LSTM cell state after each
    step; `labels` (B, Lmax) int32 collects the segment labels of the positions seen so far."""

    def __init__(self, critic, B, max_len, device):
        H = critic.lstm.hidden_size
        f = dict(device=device, dtype=torch.float32)
        self.lstm = [(torch.zeros(B, max_len, 4 * H, **f), torch.zeros(B, max_len, H, **f),
                      torch.zeros(max_len, B, H, **f)) for _ in range(critic.lstm.num_layers)]
        self.gru = [(torch.zeros(B, max_len, 3 * H, **f), torch.zeros(B, max_len, H, **f))
                    for _ in range(critic.gru.num_layers)]
        self.labels = torch.zeros(B, max_len, device=device, dtype=torch.int32)
        self.max_len = max_len


class BMEncoder(nn.Module):
    """model/bm_hrl_agent.py:218-235."""

    def __init__(self, d_model_M1, d_model_M2, d_model, d_ff_M1, d_ff_M2, dout_p, H, N) -> None:
        super(BMEncoder, self).__init__()
        enc_layer = BMEncoderLayer(d_model_M1, d_model_M2, d_model, d_ff_M1, d_ff_M2, dout_p, H)
        self.encoder = LayerStack(enc_layer, N)

    def forward(self, x, masks):
        V, A = x
        Av, Va = self.encoder((V, A), (masks['V_mask'], masks['A_mask']))
        return (Av, Va)


class BMWorkerValueFunction(ModelBase):
    """model/bm_hrl_agent.py:251-269: Linear(300->1)(relu(FFN(w_feat))); the goal of the input tuple is ignored."""

    def __init__(self, cfg) -> None:
        super(BMWorkerValueFunction, self).__init__("bm_worker_value_function")
        d_worker_feat = cfg.d_model_caps
        dout_p = cfg.dout_p
        input_d = d_worker_feat
        self.value_function = PositionwiseFeedForward(input_d, input_d * 2, dout_p)
        self.projection = nn.Linear(input_d, 1)
        self.activation = nn.ReLU()

    def forward(self, x):
        w_feat, _ = x
        return _value_head(self, w_feat)


class BMManagerValueFunction(ModelBase):
    """model/bm_hrl_agent.py:272-286."""

    def __init__(self, cfg) -> None:
        super(BMManagerValueFunction, self).__init__("bm_manager_value_function")
        d_manager_feat = cfg.d_model_caps
        dout_p = cfg.dout_p
        self.value_function = PositionwiseFeedForward(d_manager_feat, d_manager_feat * 2, dout_p)
        self.projection = nn.Linear(d_manager_feat, 1)
        self.activation = nn.ReLU()

    def forward(self, x):
        return _value_head(self, x)


def _value_head(m, feat):
    ff = m.value_function
    if ff._salts is None:
        ff._salts = (ops.next_salt(), ops.next_salt())
    p = ff.dropout.p if ff.training else 0.0
    h = ops.linear(feat, ff.fc1.weight, ff.fc1.bias, act=1, drop_p=p, salt=ff._salts[0])
    h = ops.linear(h, ff.fc2.weight, ff.fc2.bias, act=1)  # the ReLU after the FFN folded into fc2's epilogue
    return ops.linear(h, m.projection.weight, m.projection.bias)


class BMEncoderLayer(nn.Module):
    """model/bm_hrl_agent.py:328-384."""

    def __init__(self, d_model_M1, d_model_M2, d_model, d_ff_M1, d_ff_M2, dout_p, H):
        super(BMEncoderLayer, self).__init__()
        self.self_att_M1 = MultiheadedAttention(d_model_M1, d_model_M1, d_model_M1, H, dout_p, d_model)
        self.self_att_M2 = MultiheadedAttention(d_model_M2, d_model_M2, d_model_M2, H, dout_p, d_model)
        self.bi_modal_att_M1 = MultiheadedAttention(d_model_M1, d_model_M2, d_model_M2, H, dout_p, d_model)
        self.bi_modal_att_M2 = MultiheadedAttention(d_model_M2, d_model_M1, d_model_M1, H, dout_p, d_model)

        self.feed_forward_M1 = PositionwiseFeedForward(d_model_M1, d_ff_M1, dout_p)
        self.feed_forward_M2 = PositionwiseFeedForward(d_model_M2, d_ff_M2, dout_p)

        self.res_layers_M1 = clone(ResidualConnection(d_model_M1, dout_p), 3)
        self.res_layers_M2 = clone(ResidualConnection(d_model_M2, dout_p), 3)
        # outputs that become the key / value source of a later cross-attention (the other modality's bimodal
        # attention; after the last layer: the caption side's eight encoder-memory attentions)
        for r in (self.res_layers_M1[0], self.res_layers_M2[0], self.res_layers_M1[2], self.res_layers_M2[2]):
            r.emit_twin = True

    # optional callable((M1, M2)) -> (M1, M2) applied between the self-attention blocks and the bimodal blocks: both
    # tensors feed BOTH bimodal blocks, so when either one's gradient is complete the bimodal and FFN blocks of this
    # layer have finished their backward (their gradients can be reduced / stepped while the self-attention runs)
    mid_hook = None

    def forward(self, x, masks):
        '''x (M1, M2): (B, Sm, Dm); masks (M1, M2): (B, 1, Sm) -> M1m2 (B, Sm1, Dm1), M2m1 (B, Sm2, Dm2)'''
        M1, M2 = x
        M1_mask, M2_mask = masks

        # 1. self attention
        # (the two modalities are independent inside each stage: they run as concurrent branches)
        dev = M1.device
        M1, M2 = ops.parallel(
            dev, lambda: self.res_layers_M1[0](M1, AttentionSublayer(self.self_att_M1, None, None, M1_mask)),
            lambda: self.res_layers_M2[0](M2, AttentionSublayer(self.self_att_M2, None, None, M2_mask)))
        if self.mid_hook is not None:   # data-parallel trainer: marks the end (in backward) of the bimodal + FFN blocks
            M1, M2 = self.mid_hook((M1, M2))
        # 2. bimodal attention (queries are normed, keys / values are the un-normed other modality), then
        # 3. position-wise feed forward of the same modality
        M1m2, M2m1 = ops.parallel(
            dev, lambda: self.res_layers_M1[2](
                self.res_layers_M1[1](M1, AttentionSublayer(self.bi_modal_att_M1, M2, M2, M2_mask)),
                self.feed_forward_M1),
            lambda: self.res_layers_M2[2](
                self.res_layers_M2[1](M2, AttentionSublayer(self.bi_modal_att_M2, M1, M1, M1_mask)),
                self.feed_forward_M2))
        return M1m2, M2m1


class LinearCore(nn.Module):
    """model/bm_hrl_agent.py:387-396: Linear(dropout(x))."""

    def __init__(self, d_model_caps, d_goal, dout_p) -> None:
        super(LinearCore, self).__init__()
        self.linear = nn.Linear(d_model_caps, d_goal)
        self.dropout = nn.Dropout(dout_p)
        self._salt = None

    def forward(self, x):
        if self._salt is None:
            self._salt = ops.next_salt()
        p = self.dropout.p if self.training else 0.0
        if p > 0:
            x = ops.dropout_add(torch.zeros_like(x), x, p, self._salt)
        return ops.linear(x, self.linear.weight, self.linear.bias)


class Manager(nn.Module):
    """model/bm_hrl_agent.py:399-454."""

    def __init__(self, device, d_model_caps, d_goal, dout_p, core=None, exploration=True, ) -> None:
        super(Manager, self).__init__()
        self.device = device
        if core is not None:
            self.core = core
        else:
            self.core = LinearCore(d_model_caps, d_goal, dout_p)
        self.linear = nn.Linear(d_model_caps, d_goal)
        self.dropout = nn.Dropout(dout_p)
        self.exploration = exploration
        self.d_goal = d_goal

        self.mean_factor = 10
        self.std_factor = 5
        self.noise_eps = None  # tests: explicit N(0,1) draws (d_goal,) instead of the device RNG
        self._salts = None

    def expand_goals(self, x, segment_mask):
        """:415-429 -- every position takes the goal at the next segment boundary of its row (device-side scan)."""
        return ops.expand_goals(x, segment_mask.int())

    def nanstd(self, o):
        return ops.exploration_noise(o, self.d_goal, eps_normal=torch.zeros(self.d_goal, device=o.device),
                                     return_stats=True)[1][1]

    def forward(self, x, critic_mask):
        if self._salts is None:
            self._salts = (ops.next_salt(), ops.next_salt())
        p = self.dropout.p if self.training else 0.0
        if ops.BF16 and x.shape[-1] % 4 == 0 and self.d_goal % 4 == 0:
            from .. import ops16
            x = ops16.linear16(x, self.linear.weight, self.linear.bias, drop_p=p, salt=self._salts[0])
        else:
            x = ops.linear(x, self.linear.weight, self.linear.bias, drop_p=p, salt=self._salts[0])
        noise = None
        if self.exploration:
            if self.mean_factor != 10 or self.std_factor != 5:
                raise RuntimeError("Manager: the exploration kernel is built for mean_factor 10 / std_factor 5")
            noise = ops.exploration_noise(x, self.d_goal, eps_normal=self.noise_eps, salt=self._salts[1])
        return ops.expand_goals(x, critic_mask.int(), noise)


class WorkerCore(nn.Module):
    """model/bm_hrl_agent.py:456-466: LogSoftmax(Linear(364 -> V))."""

    def __init__(self, voc_size, d_in, d_goal) -> None:
        super(WorkerCore, self).__init__()
        self.projection = nn.Linear(in_features=d_in + d_goal, out_features=voc_size)
        self.logsoftmax = nn.LogSoftmax(dim=-1)
        self.softmax = nn.Softmax(dim=-1)

    def forward(self, x):
        if ops.BF16 and x.shape[-1] % 4 == 0 and self.projection.out_features % 4 == 0:
            from .. import ops16
            return ops.log_softmax(ops16.linear16(x, self.projection.weight, self.projection.bias))
        return ops.log_softmax(ops.linear(x, self.projection.weight, self.projection.bias))


class Worker(nn.Module):
    """model/bm_hrl_agent.py:468-487."""

    def __init__(self, voc_size, d_in, d_goal, dout_p, d_model, core=None) -> None:
        super(Worker, self).__init__()
        heads = 2
        if core is not None:
            self.core = core
        else:
            self.core = WorkerCore(voc_size, d_in, d_goal)
        self.goal_attention = MultiheadedAttention(d_goal, d_in, d_in, heads, dout_p, d_model)
        self.logsoftmax = nn.LogSoftmax(dim=-1)
        self.softmax = nn.Softmax(dim=-1)

    # trainer-only switch (bf16 math mode): hand back the 364-d features + projection (ops16.VocabHead) instead of the
    # (B,L,V) log-probabilities, so that the losses are evaluated by the fused vocabulary kernel without the logits
    fused_vocab = False

    def forward(self, x, goal, mask):
        goal_completion = self.goal_attention(goal, x, x, mask)
        if not self.core.__class__ == torch.nn.Identity:
            x = ops.concat_last(x, goal_completion)
            if self.fused_vocab and ops.BF16 and isinstance(self.core, WorkerCore) and x.shape[-1] % 4 == 0:
                from .. import ops16
                return ops16.VocabHead(x, self.core.projection)
            return self.core(x)
        return goal_completion


class BMHrlAgent(nn.Module):
    """model/bm_hrl_agent.py:491-661."""

    def __init__(self, cfg, train_dataset):
        super(BMHrlAgent, self).__init__()
        self.name = "bm_hrl_agent"
        self.d_video = cfg.d_vid
        self.d_audio = cfg.d_aud
        self.d_proj = cfg.rl_projection_d
        self.d_model_caps = cfg.d_model_caps
        self.d_model = cfg.d_model
        self.att_heads = cfg.rl_att_heads
        self.att_layers = cfg.rl_att_layers
        self.dout_p = cfg.dout_p
        self.d_goal = cfg.rl_goal_d
        self.voc_size = train_dataset.trg_voc_size
        self.device = get_device(cfg)

        self.critic_score_threshhold = cfg.rl_critic_score_threshhold

        self.pos_enc_A = PositionalEncoder(cfg.d_model_audio, cfg.dout_p)
        self.pos_enc_V = PositionalEncoder(cfg.d_model_video, cfg.dout_p)
        self.pos_enc_C = PositionalEncoder(cfg.d_model_caps, cfg.dout_p)

        self.critic = SegmentCritic(cfg)

        self.emb_C = VocabularyEmbedder(train_dataset.trg_voc_size, cfg.d_model_caps)
        self.emb_C.init_word_embeddings(train_dataset.train_vocab.vectors, cfg.unfreeze_word_emb)

        self.bm_enc = BMEncoder(d_model_M1=self.d_video, d_model_M2=self.d_audio, d_model=self.d_model,
                                d_ff_M1=cfg.rl_ff_v, d_ff_M2=cfg.rl_ff_a, dout_p=self.dout_p, H=self.att_heads,
                                N=self.att_layers)

        self.bm_worker_fus = BMFusion(
            cfg.d_model_audio, cfg.d_model_video, cfg.d_model_caps, cfg.d_model, cfg.rl_ff_c, self.dout_p,
            self.att_heads, self.att_layers
        )

        self.bm_manager_fus = BMFusion(
            cfg.d_model_audio, cfg.d_model_video, cfg.d_model_caps, cfg.d_model, cfg.rl_ff_c, self.dout_p,
            self.att_heads, self.att_layers
        )
        self.manager_core = LinearCore(cfg.d_model_caps, cfg.rl_goal_d, cfg.dout_p)

        self.manager = Manager(self.device, self.d_model_caps, self.d_goal, self.dout_p, self.manager_core)
        self.worker = Worker(voc_size=self.voc_size, d_in=self.d_model_caps, d_goal=self.d_goal, dout_p=self.dout_p,
                             d_model=self.d_model)

        self.teach_warmstart()
        self.warmstarting = True
        self.teaching_worker = True

        self.sigmoid_epoch_offset = torch.tensor(-1)

        self.worker_modules = [self.bm_enc, self.bm_worker_fus, self.worker]
        self.manager_modules = [self.bm_manager_fus, self.manager]

    def save_model(self, checkpoint_dir):
        torch.save(self.state_dict(), checkpoint_dir + f"/{self.name}.pt")

    def load_model(self, checkpoint_dir):
        self.load_state_dict(torch.load(checkpoint_dir + f"/{self.name}.pt"))

    def _set_worker_grad(self, enabled):
        for name, param in self.worker.named_parameters():
            param.requires_grad = enabled
        for name, param in self.bm_worker_fus.named_parameters():
            param.requires_grad = enabled

    def _set_manager_grad(self, enabled):
        for name, param in self.manager.named_parameters():
            param.requires_grad = enabled
        for name, param in self.bm_manager_fus.named_parameters():
            param.requires_grad = enabled

    def _set_module_grads(self, modules, enable):
        for module in modules:
            for name, param in module.named_parameters():
                param.requires_grad = enable

    def teach_warmstart(self):
        # like the reference (:572-575) this does NOT re-enable bm_enc after a manager epoch (F10)
        self.warmstarting = True
        self._set_worker_grad(True)
        self._set_manager_grad(True)

    def teach_worker(self):
        self.warmstarting = False
        self.teaching_worker = True
        self._set_module_grads(self.worker_modules, True)
        self._set_module_grads(self.manager_modules, False)
        self.manager.exploration = False

    def teach_manager(self):
        self.warmstarting = False
        self.teaching_worker = False
        self._set_module_grads(self.worker_modules, False)
        self._set_module_grads(self.manager_modules, True)
        self.manager.exploration = True

    def set_inference_mode(self, inference):
        self.manager.exploration = not inference

    def warmstart(self, x, trg, mask):
        return self.prediction(x, trg, mask)

    def _encode_features(self, x):
        x_video, x_audio = x
        if isinstance(x_video, (tuple, list)):  # extension: (rgb, flow) -> the add is fused into the encoding
            V = self.pos_enc_V(x_video[0], x_video[1])
        else:
            V = self.pos_enc_V(x_video)
        return V, self.pos_enc_A(x_audio)

    def mixed_prediction(self, x, trgs, mask, mix_factor):
        """:611-625 -- blend of the embeddings of two token tensors."""
        y_trg, yhat_trg = trgs
        C = self.emb_C(y_trg) * (1 - mix_factor) + self.emb_C(yhat_trg) * mix_factor
        V, A = self._encode_features(x)
        return self.predict_with_features(C, V, A, mask)

    def prediction(self, x, trg, mask):
        C = self.emb_C(trg)
        V, A = self._encode_features(x)
        return self.predict_with_features(C, V, A, mask)

    concurrent_branches = True  # critic / worker fusion stack on side streams (same results, fewer idle SMs)

    # data-parallel hook (set by the trainer): called in backward once every gradient of the modules AFTER the encoder
    # has been enqueued, so their all-reduce overlaps the encoder's backward (bmhrl_b200/ddp.py)
    grad_ready_callback = None

    def _mark_encoder_outputs(self, Va, Av):
        cb = self.grad_ready_callback
        if cb is None:
            return Va, Av
        return ops.grad_marker(Va, cb), ops.grad_marker(Av, cb)

    # Trainer hook: segment labels of THIS batch computed ahead of time (the critic is frozen and sees only the caption
    # tokens, so train.StepRunner evaluates it for the NEXT batch underneath the current step's backward); None = run
    # the critic here, like the reference
    precomputed_segment_labels = None

    def segment_labels_of(self, trg):
        """Segment labels (B,L) int32 of a token tensor: emb_C + frozen critic + threshold (:630-640), nothing else."""
        with torch.no_grad():
            return self.critic.scores_and_labels(self.emb_C(trg), self.critic_score_threshhold)[1]

    def predict_with_features(self, C, V, A, mask):
        """:637-651.  The frozen critic (depends only on the caption embedding) and the worker fusion stack
        (independent of the manager one, model/bm_hrl_agent.py:646-647) are forked onto side streams."""
        if self.precomputed_segment_labels is not None:
            segment_labels = self.precomputed_segment_labels
            C = self.pos_enc_C(C)
            Va, Av = self.bm_enc((V, A), mask)
            Va, Av = self._mark_encoder_outputs(Va, Av)
            worker_feat, manager_feat = ops.parallel(C.device, lambda: self.bm_worker_fus((C, (Av, Va)), mask),
                                                     lambda: self.bm_manager_fus((C, (Av, Va)), mask))
            goals = self.manager(manager_feat, segment_labels)
            pred = self.worker(worker_feat, goals, mask["C_mask"])
            return pred, worker_feat, manager_feat, goals, segment_labels
        if not (self.concurrent_branches and ops.CONCURRENT):
            _, segment_labels = self.critic.scores_and_labels(C, self.critic_score_threshhold)
            C = self.pos_enc_C(C)
            Va, Av = self.bm_enc((V, A), mask)
            Va, Av = self._mark_encoder_outputs(Va, Av)
            worker_feat = self.bm_worker_fus((C, (Av, Va)), mask)
            manager_feat = self.bm_manager_fus((C, (Av, Va)), mask)
        else:
            cur = torch.cuda.current_stream(C.device)
            s_critic, s_worker = ops.side_stream(C.device, 8), ops.side_stream(C.device, 9)
            s_critic.wait_stream(cur)
            with torch.cuda.stream(s_critic):
                _, segment_labels = self.critic.scores_and_labels(C, self.critic_score_threshhold)
            C = self.pos_enc_C(C)
            Va, Av = self.bm_enc((V, A), mask)
            Va, Av = self._mark_encoder_outputs(Va, Av)
            s_worker.wait_stream(cur)
            with torch.cuda.stream(s_worker):
                worker_feat = self.bm_worker_fus((C, (Av, Va)), mask)
            manager_feat = self.bm_manager_fus((C, (Av, Va)), mask)
            cur.wait_stream(s_critic)
            cur.wait_stream(s_worker)
            if not torch.cuda.is_current_stream_capturing():  # eager mode: tell the allocator about the cross-stream uses
                for t in (C, Av, Va):
                    t.record_stream(s_worker)
                segment_labels.record_stream(cur)
                worker_feat.record_stream(cur)
        goals = self.manager(manager_feat, segment_labels)
        pred = self.worker(worker_feat, goals, mask["C_mask"])
        return pred, worker_feat, manager_feat, goals, segment_labels

    def inference(self, x, trg, mask):
        return self.prediction(x, trg, mask)[0]

    def forward(self, x, trg, mask, factor=1):
        if type(trg) is tuple:
            return self.mixed_prediction(x, trg, mask, factor)
        return self.prediction(x, trg, mask)
