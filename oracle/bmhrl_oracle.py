"""CPU oracle for the BMHRL hot path -- TEST INFRASTRUCTURE ONLY.

A plain-PyTorch (CPU, fp32 or fp64) *functional* restatement of the reference algorithm for the
path named in BASELINE.json `north_star`.  It takes a reference-layout ``state_dict`` (a dict of
tensors keyed by the reference's parameter names) and plain tensors; it never builds nn.Modules and
never touches CUDA.  Gradients come from torch autograd over these plain ops.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this module, and only as the checker / reported CPU baseline.  The product package
``bmhrl_b200`` never imports it.

Pinning status: the reference ships NO tests or golden vectors for this path (SURVEY.md section 4 and
8(c)).  This restatement is instead pinned against outputs of the reference itself, run in the build
container by ``oracle/pin_oracle.py`` (imports /root/reference through ``oracle/refboot.py``); the
fixtures that script writes are committed under ``tests/golden/`` and re-checked on every run of the
CPU test-suite (tests/test_oracle_golden.py).

Every function cites the reference file:line (paths relative to /root/reference) it follows.
"""
import math

import torch
import torch.nn.functional as F

NEG_FILL = -1e9  # model/multihead_attention.py:21-22 -- masked scores are REPLACED by -1e9


# --------------------------------------------------------------------------------------------------
# blocks (model/blocks.py)
# --------------------------------------------------------------------------------------------------
def positional_table(seq_len, d_model):
    """model/blocks.py:95-103.  fp64 table; column j uses exponent j/d_model itself (not 2*floor(j/2)):
    even j -> sin(pos / 10000**(j/D)), odd j -> cos(pos / 10000**(j/D))."""
    pos = torch.arange(seq_len, dtype=torch.float64)[:, None]
    j = torch.arange(d_model, dtype=torch.float64)[None, :]
    ang = pos / torch.pow(torch.tensor(10000.0, dtype=torch.float64), j / d_model)
    even = (torch.arange(d_model) % 2 == 0)[None, :]
    return torch.where(even, torch.sin(ang), torch.cos(ang))


# Dropout probability at the reference's dropout sites (PositionalEncoder model/blocks.py:112, attention output
# model/multihead_attention.py:29-30, ResidualConnection model/blocks.py:142-144, FFN model/blocks.py:183, Manager
# model/bm_hrl_agent.py:440).  0 for every parity check (the goldens are dropout-free); bench.py's timed CPU arm sets
# it to the benchmark's 0.1 so that both arms do the same work (torch's own CPU RNG: timing only, no parity claim).
DROPOUT_P = 0.0


def _drop(x):
    return F.dropout(x, DROPOUT_P, training=True) if DROPOUT_P > 0 else x


def add_positional(x):
    """model/blocks.py:105-112 (table cast to x's dtype by type_as), dropout p = DROPOUT_P."""
    S, D = x.shape[-2], x.shape[-1]
    return _drop(x + positional_table(S, D).to(x.dtype)[None])


def embed_tokens(sd, tokens, prefix="emb_C.embedder."):
    """model/blocks.py:44-48: Embedding(tok) * sqrt(emb_dim)."""
    w = sd[prefix + "weight"]
    return w[tokens] * math.sqrt(w.shape[1])


def linear(sd, prefix, x):
    return F.linear(x, sd[prefix + "weight"], sd[prefix + "bias"])


def layer_norm(sd, prefix, x, eps=1e-5):
    """nn.LayerNorm (model/blocks.py:132): biased variance, eps 1e-5, affine."""
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * sd[prefix + "weight"] + sd[prefix + "bias"]


def ffn(sd, prefix, x):
    """model/blocks.py:175-187 with dropout p=0: fc2(relu(fc1(x)))."""
    return linear(sd, prefix + "fc2.", _drop(torch.relu(linear(sd, prefix + "fc1.", x))))


# --------------------------------------------------------------------------------------------------
# masks (model/masking.py)
# --------------------------------------------------------------------------------------------------
def make_masks(rgb, audio, captions, pad_idx):
    """model/masking.py:18-25,44-47 for modality 'audio_video': feature masks look at component 0 only
    (of rgb, before flow is added); caption mask = not-pad AND lower-triangular."""
    masks = {"V_mask": (rgb[:, :, 0] != 0)[:, None, :], "A_mask": (audio[:, :, 0] != 0)[:, None, :]}
    if captions is not None:
        L = captions.shape[-1]
        tril = torch.tril(torch.ones(1, L, L, dtype=torch.bool))
        masks["C_mask"] = (captions != pad_idx)[:, None, :] & tril
    return masks


# --------------------------------------------------------------------------------------------------
# attention (model/multihead_attention.py)
# --------------------------------------------------------------------------------------------------
def attention_core(q, k, v, mask, causal=False):
    """model/multihead_attention.py:7-31.  q (B,H,Sq,dk), k/v (B,H,Sk,dk); mask bool (B,1,1,Sk) or
    (B,1,Sq,Sk) or None.  Masked scores are replaced (not added) by -1e9, so a fully masked row becomes
    a uniform distribution over ALL keys, padding included."""
    d_k = q.shape[-1]
    s = q @ k.transpose(-1, -2) / math.sqrt(d_k)
    if mask is not None:
        if causal:  # :18-21, the extra tril is only applied when a mask is given
            Sk = mask.shape[-1]
            tril = torch.tril(torch.ones(Sk, Sk, dtype=torch.bool))
            s = s.masked_fill(~tril[None, None], NEG_FILL)
        s = s.masked_fill(~mask, NEG_FILL)
    return _drop(torch.softmax(s, dim=-1) @ v)


def mha(sd, prefix, Q, K, V, mask, H, causal=False):
    """model/multihead_attention.py:60-92 with dropout p=0."""
    B, Sq, _ = Q.shape
    q = linear(sd, prefix + "linear_Q2d.", Q)
    k = linear(sd, prefix + "linear_K2d.", K)
    v = linear(sd, prefix + "linear_V2d.", V)
    D = q.shape[-1]
    dk = D // H
    q = q.view(B, -1, H, dk).transpose(1, 2)
    k = k.view(B, -1, H, dk).transpose(1, 2)
    v = v.view(B, -1, H, dk).transpose(1, 2)
    m = mask[:, None] if mask is not None else None
    o = attention_core(q, k, v, m, causal)
    o = o.transpose(1, 2).reshape(B, Sq, D)
    return linear(sd, prefix + "linear_d2Q.", o)


# --------------------------------------------------------------------------------------------------
# bimodal encoder / fusion decoder (model/bm_hrl_agent.py)
# --------------------------------------------------------------------------------------------------
def bm_encoder_layer(sd, p, M1, M2, M1_mask, M2_mask, H):
    """model/bm_hrl_agent.py:344-384: pre-LN residual blocks; cross attention uses LN(query) but the
    UN-normed other modality as key/value."""
    def res(i, mod, x, f):  # model/blocks.py:135-144
        return x + _drop(f(layer_norm(sd, f"{p}res_layers_{mod}.{i}.norm.", x)))

    M1 = res(0, "M1", M1, lambda n: mha(sd, p + "self_att_M1.", n, n, n, M1_mask, H))
    M2 = res(0, "M2", M2, lambda n: mha(sd, p + "self_att_M2.", n, n, n, M2_mask, H))
    M1m2 = res(1, "M1", M1, lambda n: mha(sd, p + "bi_modal_att_M1.", n, M2, M2, M2_mask, H))
    M2m1 = res(1, "M2", M2, lambda n: mha(sd, p + "bi_modal_att_M2.", n, M1, M1, M1_mask, H))
    M1m2 = res(2, "M1", M1m2, lambda n: ffn(sd, p + "feed_forward_M1.", n))
    M2m1 = res(2, "M2", M2m1, lambda n: ffn(sd, p + "feed_forward_M2.", n))
    return M1m2, M2m1


def bm_encoder(sd, p, V, A, masks, H, N):
    """model/bm_hrl_agent.py:224-235 (LayerStack, model/blocks.py:10-21).  Returns (video-shaped,
    audio-shaped); the reference names them (Av, Va) at :234 and swaps them again at :644."""
    for i in range(N):
        V, A = bm_encoder_layer(sd, f"{p}encoder.layers.{i}.", V, A, masks["V_mask"], masks["A_mask"], H)
    return V, A


def bm_fusion_layer(sd, p, C, mem_A, mem_V, masks, H):
    """model/bm_hrl_agent.py:73-117.  feed_forward exists but is never called (F11)."""
    n = layer_norm(sd, p + "res_layer_self_att.norm.", C)
    C = C + _drop(mha(sd, p + "self_att.", n, n, n, masks["C_mask"], H))
    Ca = C + _drop(mha(sd, p + "enc_att_A.", layer_norm(sd, p + "res_layer_enc_att_A.norm.", C), mem_A, mem_A,
                       masks["A_mask"], H))
    Cv = C + _drop(mha(sd, p + "enc_att_V.", layer_norm(sd, p + "res_layer_enc_att_V.norm.", C), mem_V, mem_V,
                       masks["V_mask"], H))
    Ca = layer_norm(sd, p + "normCA.", Ca)
    Cv = layer_norm(sd, p + "normCV.", Cv)
    g = torch.sigmoid(torch.clamp(sd[p + "a_v_constant"], min=-2, max=2))
    return g * Cv + (1 - g) * Ca


def bm_fusion(sd, p, C, mem_A, mem_V, masks, H, N):
    """model/bm_hrl_agent.py:120-130."""
    for i in range(N):
        C = bm_fusion_layer(sd, f"{p}decoder.layers.{i}.", C, mem_A, mem_V, masks, H)
    return C


# --------------------------------------------------------------------------------------------------
# frozen segment critic (model/bm_hrl_agent.py:186-215) -- explicit LSTM / GRU recurrences
# --------------------------------------------------------------------------------------------------
def arelu(sd, p, x):
    """model/bm_hrl_agent.py:13-23."""
    alpha = torch.clamp(sd[p + "alpha"], min=0.01, max=0.99)
    beta = 1 + torch.sigmoid(sd[p + "beta"])
    return torch.relu(x) * beta - torch.relu(-x) * alpha


def lstm_stack(sd, p, x, layers):
    """nn.LSTM(batch_first) gate order i,f,g,o; zero initial state."""
    B, L, _ = x.shape
    for l in range(layers):
        w_ih, w_hh = sd[f"{p}weight_ih_l{l}"], sd[f"{p}weight_hh_l{l}"]
        b = sd[f"{p}bias_ih_l{l}"] + sd[f"{p}bias_hh_l{l}"]
        Hd = w_hh.shape[1]
        h = x.new_zeros(B, Hd)
        c = x.new_zeros(B, Hd)
        gx = x @ w_ih.t() + b
        out = []
        for t in range(L):
            g = gx[:, t] + h @ w_hh.t()
            i, f, gg, o = g.chunk(4, dim=-1)
            c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
            h = torch.sigmoid(o) * torch.tanh(c)
            out.append(h)
        x = torch.stack(out, dim=1)
    return x


def gru_stack(sd, p, x, layers):
    """nn.GRU(batch_first) gate order r,z,n; n = tanh(W_in x + b_in + r*(W_hn h + b_hn))."""
    B, L, _ = x.shape
    for l in range(layers):
        w_ih, w_hh = sd[f"{p}weight_ih_l{l}"], sd[f"{p}weight_hh_l{l}"]
        b_ih, b_hh = sd[f"{p}bias_ih_l{l}"], sd[f"{p}bias_hh_l{l}"]
        Hd = w_hh.shape[1]
        h = x.new_zeros(B, Hd)
        gx = x @ w_ih.t() + b_ih
        out = []
        for t in range(L):
            gh = h @ w_hh.t() + b_hh
            xr, xz, xn = gx[:, t].chunk(3, dim=-1)
            hr, hz, hn = gh.chunk(3, dim=-1)
            r = torch.sigmoid(xr + hr)
            z = torch.sigmoid(xz + hz)
            n = torch.tanh(xn + r * hn)
            h = (1 - z) * n + z * h
            out.append(h)
        x = torch.stack(out, dim=1)
    return x


def segment_critic(sd, C_emb, p="critic."):
    """model/bm_hrl_agent.py:204-215 (no_grad): LSTM x4 -> AReLU -> GRU x2 -> AReLU -> Linear(600->1)."""
    with torch.no_grad():
        h = arelu(sd, p + "relu.", lstm_stack(sd, p + "lstm.", C_emb, 4))
        h = arelu(sd, p + "relu2.", gru_stack(sd, p + "gru.", h, 2))
        return linear(sd, p + "lin.", h)


def segment_labels_from_scores(scores, threshold=0.25):
    """model/bm_hrl_agent.py:638-640, with squeeze(-1) (F6) so B=1 / L=1 keep their (B,L) shape."""
    return (torch.sigmoid(scores) > threshold).squeeze(-1).int()


# --------------------------------------------------------------------------------------------------
# manager / worker heads
# --------------------------------------------------------------------------------------------------
def expand_goals(x, seg):
    """model/bm_hrl_agent.py:415-429, restated without the in-place host loop.

    For a position (b,l) with a segment boundary at l' >= l in the same row: out = x[b, min such l'].
    Otherwise (no boundary at or after l in row b):
      * row b has a boundary (l lies after its last one): zero if ANY later row has a boundary, else raw;
      * row b has no boundary: raw -- except row 0, which the loop zeroes wholesale when a later row has
        a boundary (old_b starts at 0, :418).
    """
    B, L, _ = x.shape
    seg = seg.bool()
    idx = torch.arange(L)
    big = L + 1
    cand = torch.where(seg, idx[None, :].expand(B, L), torch.full((B, L), big))
    nxt = torch.flip(torch.cummin(torch.flip(cand, [1]), dim=1).values, [1])  # next boundary at/after l
    has_next = nxt < big
    row_has = seg.any(dim=1)
    later_has = torch.zeros(B, dtype=torch.bool)
    for b in range(B - 2, -1, -1):
        later_has[b] = later_has[b + 1] | row_has[b + 1]
    gathered = torch.gather(x, 1, nxt.clamp(max=L - 1)[:, :, None].expand_as(x))
    zero_tail = (row_has & later_has)[:, None] & ~has_next
    zero_tail = zero_tail | ((~row_has) & later_has & (torch.arange(B) == 0))[:, None].expand(B, L)
    out = torch.where(has_next[:, :, None], gathered, x)
    return torch.where(zero_tail[:, :, None], torch.zeros_like(out), out)


def manager_head(sd, manager_feat, seg, noise=None, p="manager."):
    """model/bm_hrl_agent.py:437-454 with dropout p=0.  `noise`: the explicit 64-d exploration vector
    (already including the -0.5*mean shift), or None when exploration is off."""
    g = _drop(linear(sd, p + "linear.", manager_feat))
    if noise is not None:
        g = g + noise
    return expand_goals(g, seg)


def exploration_noise_params(goals_pre):
    """model/bm_hrl_agent.py:444-451: (mean/10, std/5) of the whole pre-noise goal tensor, detached."""
    mean = torch.nanmean(goals_pre)
    std = torch.sqrt(torch.nanmean(torch.pow(torch.abs(goals_pre - mean), 2)))
    return (mean / 10).detach(), (std / 5).detach()


def worker_head(sd, worker_feat, goals, C_mask, p="worker."):
    """model/bm_hrl_agent.py:480-487, 463-466: goal attention (H=2), concat, Linear(364->V), LogSoftmax."""
    gc = mha(sd, p + "goal_attention.", goals, worker_feat, worker_feat, C_mask, 2)
    x = torch.cat([worker_feat, gc], dim=-1)
    return torch.log_softmax(linear(sd, p + "core.projection.", x), dim=-1)


def value_function(sd, feat, p=""):
    """model/bm_hrl_agent.py:263-269 / :281-286: Linear(300->1)(relu(FFN(feat)))."""
    return linear(sd, p + "projection.", torch.relu(ffn(sd, p + "value_function.", feat)))


def agent_forward(sd, V, A, trg, masks, H=4, N=2, threshold=0.25, noise=None):
    """model/bm_hrl_agent.py:627-651 (prediction -> predict_with_features), dropout p=0.
    V = rgb + flow (caller adds, epoch_loops/captioning_bmrl_loops.py:498).  Returns the reference's
    5-tuple (pred, worker_feat, manager_feat, goals, segment_labels)."""
    C = embed_tokens(sd, trg)
    Vp = add_positional(V)
    Ap = add_positional(A)
    seg = segment_labels_from_scores(segment_critic(sd, C), threshold)
    C = add_positional(C)
    mem_V, mem_A = bm_encoder(sd, "bm_enc.", Vp, Ap, masks, H, N)
    worker_feat = bm_fusion(sd, "bm_worker_fus.", C, mem_A, mem_V, masks, H, N)
    manager_feat = bm_fusion(sd, "bm_manager_fus.", C, mem_A, mem_V, masks, H, N)
    goals = manager_head(sd, manager_feat, seg, noise)
    pred = worker_head(sd, worker_feat, goals, masks["C_mask"])
    return pred, worker_feat, manager_feat, goals, seg


# --------------------------------------------------------------------------------------------------
# DETR-variant layers (model/encoder.py, model/decoder.py) -- API rows a8
# --------------------------------------------------------------------------------------------------
def detr_encoder_layer(sd, p, src, mask, H, pos=add_positional):
    """model/encoder.py:59-68 (post-LN): q = k = pos(src), value = src."""
    qk = pos(src)
    src = layer_norm(sd, p + "norm1.", src + mha(sd, p + "self_attn.", qk, qk, src, mask, H))
    ff = linear(sd, p + "linear2.", torch.relu(linear(sd, p + "linear1.", src)))
    return layer_norm(sd, p + "norm2.", src + ff)


def detr_decoder_layer(sd, p, tgt, memory, memory_mask, query_mask, H, goal=None, goal_mask=None,
                       detected=None, obj_mask=None, pos=add_positional, query_pos=add_positional,
                       goal_pos=add_positional, add_pos=False):
    """model/decoder.py:66-100 (post-LN variant with the LN-before-residual quirk at :77-78)."""
    if not add_pos:
        causal = True
        q = k = query_pos(tgt)
    else:
        causal = False
        q = k = tgt + query_pos
    t2 = mha(sd, p + "self_attn.", q, k, tgt, query_mask, H, causal=causal)
    tgt = layer_norm(sd, p + "norm1.", tgt) + t2
    tgt = layer_norm(sd, p + "norm2.", tgt + mha(sd, p + "multihead_attn.", q, pos(memory), memory, memory_mask, H))
    if goal is not None:
        tgt = layer_norm(sd, p + "norm4.", tgt + mha(sd, p + "goal_attention.", query_pos(tgt), goal_pos(goal),
                                                      goal, goal_mask, H))
    if detected is not None:
        tgt = layer_norm(sd, p + "norm5.", tgt + mha(sd, p + "detected_attention.", q, detected, detected,
                                                      obj_mask, H))
    ff = linear(sd, p + "linear2.", torch.relu(linear(sd, p + "linear1.", tgt)))
    return layer_norm(sd, p + "norm3.", tgt + ff)


# --------------------------------------------------------------------------------------------------
# losses (loss/label_smoothing.py, loss/biased_kl.py)
# --------------------------------------------------------------------------------------------------
def _xlogy_kl(pred, dist):
    """F.kl_div(pred, dist, reduction='none') = xlogy(dist, dist) - dist * pred."""
    return torch.xlogy(dist, dist) - dist * pred


def _keep_rows(target, pad_idx):
    """loss/label_smoothing.py:26-30, loss/biased_kl.py:46-50: rows whose target is pad are zeroed -- but only
    `if mask.sum() > 0 and len(mask) > 0`, where `mask = nonzero(target == pad)` holds the flat ROW INDICES.  When
    the only padded target is flat row 0 the index sum is 0 and the reference zeroes nothing: that row keeps its
    smoothed distribution (minus the pad column) and contributes loss and gradient."""
    is_pad = target == pad_idx
    idx = torch.nonzero(is_pad).reshape(-1)
    if idx.numel() > 0 and int(idx.sum()) > 0:
        return ~is_pad
    return torch.ones_like(is_pad)


def label_smoothing_kl(pred, target, smoothing, pad_idx):
    """loss/label_smoothing.py:12-32.  Returns the (B*L, V) element-wise divergence."""
    V = pred.shape[-1]
    pred = pred.reshape(-1, V)
    target = target.reshape(-1)
    dist = torch.full_like(pred, smoothing / (V - 2))
    dist.scatter_(1, target[:, None], 1 - smoothing)
    dist[:, pad_idx] = 0
    dist = dist * _keep_rows(target, pad_idx)[:, None].to(dist.dtype)
    return _xlogy_kl(pred, dist)


def biased_kl_divergence(pred, trg, biased_trg, biased_offset, smoothing, pad_idx):
    """loss/biased_kl.py:22-53.  biased_offset (B,L) carries gradient into the TARGET distribution;
    the +1e-8 is applied to every entry, including pad rows and the pad column."""
    V = pred.shape[-1]
    tf = 1 - smoothing
    p2 = pred.reshape(-1, V)
    t = trg.reshape(-1)
    trg_ampl = tf * (1 - biased_offset).reshape(-1)
    dist = torch.full_like(p2, smoothing / (V - 2))
    dist = dist.scatter(1, t[:, None], trg_ampl[:, None])
    colmask = torch.ones(V, dtype=p2.dtype)
    colmask[pad_idx] = 0
    dist = dist * colmask[None, :]
    dist = dist + torch.zeros_like(p2).scatter(1, biased_trg.reshape(-1, 1), (biased_offset * tf).reshape(-1, 1))
    dist = dist * _keep_rows(t, pad_idx)[:, None].to(dist.dtype)
    return _xlogy_kl(p2, dist + 1e-8)


def reinforce_loss(pred_probs, action, value, critic_value, eps=1e-5):
    """loss/biased_kl.py:69-81: -mean(adv.detach * log pi(a)) + mean(adv^2)."""
    pi = torch.clamp(pred_probs, eps, 1 - eps)
    pa = torch.gather(pi, -1, action.reshape(pi.shape[:-1] + (1,))).squeeze(-1)
    adv = value - critic_value
    return -torch.mean(adv.detach().squeeze() * torch.log(pa)) + torch.mean(adv ** 2)


# --------------------------------------------------------------------------------------------------
# RL step pieces (epoch_loops/captioning_bmrl_loops.py)
# --------------------------------------------------------------------------------------------------
def sample_tokens(pred, exp_noise=None):
    """epoch_loops/captioning_bmrl_loops.py:273-285.  Worker (exp_noise given): Categorical(exp(pred))
    .sample() == argmax((p / sum p) / q), q ~ Exp(1) of pred's shape (torch.multinomial's one-sample
    path).  Manager (exp_noise None): argmax.  Returns (tokens int64 (B,L), sampled_probs (B,L) w/ grad)."""
    probs = torch.exp(pred)
    if exp_noise is not None:
        p_hat = probs / probs.sum(-1, keepdim=True)
        tok = torch.argmax(p_hat.detach() / exp_noise, dim=-1)
    else:
        tok = torch.argmax(probs, dim=-1)
    return tok, torch.gather(probs, 2, tok[:, :, None]).squeeze(-1)


def segment_prod_and_sum(sampled_probs, expected, seg):
    """epoch_loops/captioning_bmrl_loops.py:299-317 (manager branch host loop).

    segment_prob[b, s..e] = prod(sampled_probs[b, s..e]) for every segment [s..e] ending at a boundary e;
    positions not covered by a closed segment stay 0.  expected[b, s..e] = sum(expected[b, s..e]); the tail
    after a row's last boundary is zeroed iff a later row has a boundary; boundary-free rows are untouched
    except row 0 (zeroed when a later row has a boundary).  `expected` is modified like the reference does
    (it aliases the value-function output that the value loss later reads)."""
    B, L = sampled_probs.shape
    segb = seg.bool()
    seg_prob = torch.zeros_like(sampled_probs)
    exp_out = expected.clone()
    rows = [b for b in range(B) if bool(segb[b].any())]
    for b in range(B):
        later = any(r > b for r in rows)
        if b in rows:
            start = 0
            for l in range(L):
                if segb[b, l]:
                    seg_prob[b, start:l + 1] = torch.prod(sampled_probs[b, start:l + 1])
                    exp_out[b, start:l + 1] = expected[b, start:l + 1].sum()
                    start = l + 1
            if later:
                exp_out[b, start:] = 0
        elif b == 0 and later:
            exp_out[b, :] = 0
    return seg_prob, exp_out


def biased_kl_rl(pred, trg, score, expected, loss_mask, seg, train_worker, stabilize, smoothing, pad_idx,
                 exp_noise=None):
    """epoch_loops/captioning_bmrl_loops.py:271-334 with the scorer factored out: `score` (B,L) is what
    get_score() returned.  Returns (divergence (B*L,V), score_used, tokens, expected_after)."""
    tok, sp = sample_tokens(pred, exp_noise if train_worker else None)
    if train_worker:
        norm = loss_mask.sum(-1).reshape(-1, 1)  # :414-416
    else:
        norm = seg.sum(-1).reshape(-1, 1)
        score = score * seg.to(score.dtype)
        sp, expected = segment_prod_and_sum(sp, expected, seg)
    if stabilize:
        score = (score - expected) * loss_mask.to(score.dtype)
    amplitude = torch.clamp(score.to(pred.dtype) * sp * norm.to(pred.dtype), 0, 1)  # :409-411
    div = biased_kl_divergence(pred, trg, tok, amplitude, smoothing, pad_idx)
    return div, score, tok, expected


def adam_step(params, grads, state, lr=1e-4, betas=(0.9, 0.999), eps=1e-8):
    """torch.optim.Adam defaults as built at scripts/train_rl_captioning_module.py:81-83 (wd 0).
    state: dict name -> (m, v, t)."""
    for name, p in params.items():
        g = grads.get(name)
        if g is None:
            continue
        m, v, t = state.get(name, (torch.zeros_like(p), torch.zeros_like(p), 0))
        t += 1
        m = betas[0] * m + (1 - betas[0]) * g
        v = betas[1] * v + (1 - betas[1]) * g * g
        denom = v.sqrt() / math.sqrt(1 - betas[1] ** t) + eps
        params[name] = p - (lr / (1 - betas[0] ** t)) * m / denom
        state[name] = (m, v, t)
    return params, state


def greedy_decode(sd, rgb, flow, audio, max_len, start_idx, end_idx, pad_idx, H=4, N=2, threshold=0.25):
    """epoch_loops/captioning_bmrl_loops.py:127-152 with the 3-argument inference form (F5): the whole
    model is re-run for every generated token."""
    B = audio.shape[0]
    trg = torch.full((B, 1), start_idx, dtype=torch.long)
    done = torch.zeros(B, 1, dtype=torch.bool)
    last = []
    with torch.no_grad():
        while trg.shape[-1] <= max_len and not bool(done.all()):
            masks = make_masks(rgb, audio, trg, pad_idx)
            pred = agent_forward(sd, rgb + flow, audio, trg, masks, H, N, threshold)[0]
            last.append(pred[:, -1])
            nxt = pred[:, -1].max(dim=-1)[1][:, None]
            trg = torch.cat([trg, nxt], dim=-1)
            done = done | (nxt == end_idx)
    return trg, torch.stack(last, dim=1)


# --------------------------------------------------------------------------------------------------
# input pipeline
# --------------------------------------------------------------------------------------------------
def crop_bounds(S, start, end, duration):
    """captioning_datasets/load_features.py:14-31: [start_idx, end_idx) of a segment in an S-frame stack."""
    start_idx = int(S * (start / duration))
    end_idx = int(S * (end / duration))
    if start_idx == end_idx:
        if start_idx == S:
            start_idx -= 1
        else:
            end_idx += 1
    return start_idx, end_idx


def pad_batch(stacks, pad_value):
    """torch pad_sequence(batch_first=True, padding_value=pad_value) as used at captioning_dataset.py:296-299."""
    S = max(s.shape[0] for s in stacks)
    out = torch.full((len(stacks), S, stacks[0].shape[1]), float(pad_value), dtype=stacks[0].dtype)
    for i, s in enumerate(stacks):
        out[i, : s.shape[0]] = s
    return out


def generate_synonyms(caption_idx, p, vocab_size, u1, u2, rand_words, end_idx=3, pad_idx=1):
    """epoch_loops/captioning_bmrl_loops.py:510-528 with the `random` draws laid out per position: u1[i,j] is the
    `random.random()` compared with p at token (i,j), u2[i,j] the second draw (read only when u1 < p), rand_words[i,j]
    the `random.randint(2, V-1)` result (read only when u2 >= 0.9)."""
    out = caption_idx.clone()
    for i in range(out.shape[0]):
        for j in range(out.shape[1]):
            w = int(caption_idx[i, j])
            if w == end_idx:
                out[i, j] = pad_idx
                break
            if float(u1[i, j]) < p:
                if float(u2[i, j]) < 0.8:
                    out[i, j] = pad_idx
                elif float(u2[i, j]) >= 0.9:
                    out[i, j] = int(rand_words[i, j])
    return out
