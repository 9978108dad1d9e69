"""GPU: every kernel of libbmhrl_b200 (called through the C ABI via bmhrl_b200.ops / the reference-API modules)
against the CPU oracle and the committed golden vectors.  fp32 tolerance 1e-3 relative (north_star); integer
outputs (tokens, labels, maps) bit-exact."""
import math

import pytest
import torch

from oracle import bmhrl_oracle as O
from conftest import PAD, SMOOTH, rel

pytestmark = pytest.mark.gpu
TOL = 1e-3    # north_star fp32 bound
TIGHT = 2e-5  # what the exact-fp32 kernels actually reach on these sizes


def _g(seed):
    return torch.Generator().manual_seed(seed)


@pytest.mark.parametrize("M,K,N", [(37, 300, 64), (320, 364, 1000), (480, 1024, 1024), (5, 64, 1024), (129, 128, 512)])
@pytest.mark.parametrize("act,resid", [(0, False), (1, False), (0, True)])
def test_linear(gpu_lib, M, K, N, act, resid):
    from bmhrl_b200 import ops
    g = _g(M + K + N)
    x, w, b = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g) / math.sqrt(K), torch.randn(N, generator=g)
    r = torch.randn(M, N, generator=g) if resid else None
    dy = torch.randn(M, N, generator=g)
    ref_in = [t.clone().requires_grad_(True) for t in (x, w, b)] + ([r.clone().requires_grad_(True)] if resid else [])
    y = torch.nn.functional.linear(ref_in[0], ref_in[1], ref_in[2])
    if act:
        y = torch.relu(y)
    if resid:
        y = y + ref_in[3]
    y.backward(dy)
    dev_in = [t.clone().cuda().requires_grad_(True) for t in (x, w, b)] + ([r.clone().cuda().requires_grad_(True)] if resid else [])
    yd = ops.linear(dev_in[0], dev_in[1], dev_in[2], resid=dev_in[3] if resid else None, act=act)
    yd.backward(dy.cuda())
    assert rel(yd, y) < TIGHT
    for a, c in zip(dev_in, ref_in):
        assert rel(a.grad, c.grad) < TIGHT


@pytest.mark.parametrize("D", [128, 300, 364, 1024, 20])
def test_layer_norm(gpu_lib, D):
    from bmhrl_b200 import ops
    g = _g(D)
    x = torch.randn(3, 17, D, generator=g) * 2 + 0.5
    w, b = 1 + 0.1 * torch.randn(D, generator=g), 0.1 * torch.randn(D, generator=g)
    dy = torch.randn(3, 17, D, generator=g)
    ri = [t.clone().requires_grad_(True) for t in (x, w, b)]
    O.layer_norm({"weight": ri[1], "bias": ri[2]}, "", ri[0]).backward(dy)
    di = [t.clone().cuda().requires_grad_(True) for t in (x, w, b)]
    y = ops.layer_norm(di[0], di[1], di[2])
    y.backward(dy.cuda())
    assert rel(y, O.layer_norm({"weight": w, "bias": b}, "", x)) < TIGHT
    for a, c in zip(di, ri):
        assert rel(a.grad, c.grad) < 1e-4


def _mha_sd(dq, dk_, dv, D, g):
    sd = {}
    for n, (o, i) in {"linear_Q2d": (D, dq), "linear_K2d": (D, dk_), "linear_V2d": (D, dv), "linear_d2Q": (dq, D)}.items():
        sd[n + ".weight"] = torch.randn(o, i, generator=g) / math.sqrt(i)
        sd[n + ".bias"] = 0.1 * torch.randn(o, generator=g)
    return sd


@pytest.mark.parametrize("case", ["key_pad", "full", "causal", "all_masked_row", "none"])
@pytest.mark.parametrize("dims", [(128, 128, 128, 1024, 4, 9, 9), (300, 1024, 1024, 1024, 4, 7, 12),
                                  (64, 300, 300, 1024, 2, 6, 6), (32, 32, 32, 64, 4, 70, 133)])
def test_multiheaded_attention(gpu_lib, case, dims):
    """MultiheadedAttention (module API) vs oracle mha: all mask kinds incl. the replace-by--1e9 semantics
    (a fully masked row attends uniformly over ALL keys), forward and all gradients."""
    from bmhrl_b200.model.multihead_attention import MultiheadedAttention
    dq, dk_, dv, D, H, Sq, Sk = dims
    if case in ("full", "causal") and Sq != Sk:
        pytest.skip("square masks only")
    g = _g(hash((case, dims)) % 1000)
    B = 3
    sd = _mha_sd(dq, dk_, dv, D, g)
    Q, K = torch.randn(B, Sq, dq, generator=g), torch.randn(B, Sk, dk_, generator=g)
    Vv = K if dv == dk_ else torch.randn(B, Sk, dv, generator=g)
    causal = False
    if case == "none":
        mask = None
    elif case in ("key_pad", "all_masked_row"):
        lens = torch.tensor([Sk, max(1, Sk // 2), 0 if case == "all_masked_row" else max(1, Sk - 1)])
        mask = (torch.arange(Sk)[None, :] < lens[:, None])[:, None, :]
    else:
        lens = torch.tensor([Sk, max(1, Sk // 2), max(1, Sk - 1)])
        mask = (torch.arange(Sk)[None, :] < lens[:, None])[:, None, :] & torch.tril(torch.ones(1, Sq, Sk, dtype=torch.bool))
        if case == "causal":
            mask = (torch.arange(Sk)[None, :] < lens[:, None])[:, None, :]
            causal = True
    p = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    Qr, Kr = Q.clone().requires_grad_(True), K.clone().requires_grad_(True)
    out_ref = O.mha(p, "", Qr, Kr, Kr if Vv is K else Vv, mask, H, causal=causal)
    dy = torch.randn(out_ref.shape, generator=g)
    out_ref.backward(dy)
    m = MultiheadedAttention(dq, dk_, dv, H, 0.0, D)
    m.load_state_dict(sd)
    m = m.cuda()
    Qd, Kd = Q.clone().cuda().requires_grad_(True), K.clone().cuda().requires_grad_(True)
    out = m(Qd, Kd, Kd if Vv is K else Vv.cuda(), None if mask is None else mask.cuda(), causal)
    out.backward(dy.cuda())
    assert rel(out, out_ref) < TIGHT * 5
    assert rel(Qd.grad, Qr.grad) < 1e-4 and rel(Kd.grad, Kr.grad) < 1e-4
    for n, q in m.named_parameters():
        if n == "linear_K2d.bias":  # mathematically zero
            assert float(q.grad.norm()) < 1e-4 * (1 + float(p["linear_Q2d.bias"].grad.norm()))
            continue
        assert rel(q.grad, p[n].grad) < 2e-4, n


def test_attention_function(gpu_lib):
    from bmhrl_b200.model.multihead_attention import attention
    g = _g(2)
    B, H, Sq, Sk, dk = 2, 4, 5, 8, 16
    q, k, v = (torch.randn(B, H, s, dk, generator=g) for s in (Sq, Sk, Sk))
    mask = torch.ones(B, 1, 1, Sk, dtype=torch.bool)
    mask[1, ..., 5:] = False
    ref = O.attention_core(q, k, v, mask)
    out = attention(q.cuda(), k.cuda(), v.cuda(), mask.cuda())
    assert out.shape == ref.shape and rel(out, ref) < TIGHT


def test_kl_losses_golden(gpu_lib, golden):
    """LabelSmoothing / BiasedKL / Reinforce modules vs outputs of the reference itself."""
    from bmhrl_b200.loss.label_smoothing import LabelSmoothing
    from bmhrl_b200.loss.biased_kl import BiasedKL, Reinforce
    from bmhrl_b200 import ops
    gl = golden["losses"]
    i = gl["inputs"]
    for name in ("ls", "bkl"):
        logits = i["logits"].clone().cuda().requires_grad_(True)
        lp = ops.log_softmax(logits)
        lp.retain_grad()
        off = i["off"].clone().cuda().requires_grad_(True)
        if name == "ls":
            out = LabelSmoothing(SMOOTH, PAD)(lp, i["trg"].cuda())
            rs = LabelSmoothing(SMOOTH, PAD).row_sums(lp, i["trg"].cuda())
        else:
            out = BiasedKL(SMOOTH, PAD)(lp, i["trg"].cuda(), i["samp"].cuda(), off)
            rs = BiasedKL(SMOOTH, PAD).row_sums(lp, i["trg"].cuda(), i["samp"].cuda(), off)
        assert out.shape == gl[name]["out"].shape
        (out * gl[name]["w"].cuda()).sum().backward()
        assert rel(out, gl[name]["out"]) < TIGHT
        assert rel(lp.grad, gl[name]["dpred"]) < TIGHT
        assert rel(rs, gl[name]["out"].sum(-1)) < TIGHT
        if name == "bkl":
            assert rel(off.grad, gl[name]["doffset"]) < 1e-4
    r = gl["reinforce"]
    probs = torch.softmax(i["logits"], -1).cuda().requires_grad_(True)
    cv = r["critic"].clone().cuda().requires_grad_(True)
    out = Reinforce()(probs, i["samp"].cuda(), r["value"].cuda(), cv)
    out.backward()
    assert rel(out, r["out"]) < TIGHT and rel(probs.grad, r["dprobs"]) < TIGHT and rel(cv.grad, r["dcritic"]) < TIGHT


def test_kl_losses_only_row0_padded(gpu_lib, golden):
    """Reference quirk (loss/label_smoothing.py:26-30): pad rows are zeroed only when the flat row indices of the pad
    targets sum to > 0 -- a lone pad target in flat row 0 keeps its distribution.  Kernel vs the reference's output."""
    from bmhrl_b200.loss.label_smoothing import LabelSmoothing
    from bmhrl_b200.loss.biased_kl import BiasedKL
    from bmhrl_b200 import ops
    gl = golden["losses_row0"]
    i = gl["inputs"]
    for name in ("ls", "bkl"):
        lp = ops.log_softmax(i["logits"].clone().cuda().requires_grad_(True))
        lp.retain_grad()
        if name == "ls":
            out = LabelSmoothing(SMOOTH, PAD)(lp, i["trg"].cuda())
            rs = LabelSmoothing(SMOOTH, PAD).row_sums(lp, i["trg"].cuda())
        else:
            out = BiasedKL(SMOOTH, PAD)(lp, i["trg"].cuda(), i["samp"].cuda(), i["off"].clone().cuda())
            rs = BiasedKL(SMOOTH, PAD).row_sums(lp, i["trg"].cuda(), i["samp"].cuda(), i["off"].clone().cuda())
        (out * gl[name]["w"].cuda()).sum().backward()
        assert rel(out, gl[name]["out"]) < TIGHT
        assert rel(lp.grad, gl[name]["dpred"]) < TIGHT
        assert rel(rs, gl[name]["out"].sum(-1)) < TIGHT


def test_kl_rowsum_backward_full_vocab(gpu_lib):
    from bmhrl_b200 import ops
    g = _g(8)
    B, L, V = 4, 6, 10172
    logits = torch.randn(B, L, V, generator=g)
    trg = torch.randint(2, V, (B, L), generator=g)
    trg[1, 4:] = PAD
    samp = torch.randint(0, V, (B, L), generator=g)
    off = torch.rand(B, L, generator=g)
    lr = logits.clone().requires_grad_(True)
    offr = off.clone().requires_grad_(True)
    ref = O.biased_kl_divergence(torch.log_softmax(lr, -1), trg, samp, offr, SMOOTH, PAD).sum() / 7.0
    ref.backward()
    ld = logits.clone().cuda().requires_grad_(True)
    offd = off.clone().cuda().requires_grad_(True)
    out = ops.biased_kl(ops.log_softmax(ld), trg.cuda(), samp.cuda(), offd, SMOOTH, PAD, rowsum=True).sum() / 7.0
    out.backward()
    assert rel(out, ref) < TIGHT and rel(ld.grad, lr.grad) < 1e-4 and rel(offd.grad, offr.grad) < 1e-4


def test_expand_goals_golden(gpu_lib, golden):
    from bmhrl_b200 import ops
    for c in golden["expand_goals"]:
        x = c["x"].clone().cuda().requires_grad_(True)
        out = ops.expand_goals(x * 1.0, c["seg"].cuda())
        (out * c["w"].cuda()).sum().backward()
        assert torch.equal(out.detach().cpu(), c["out"])
        assert torch.allclose(x.grad.cpu(), c["dx"])


def test_manager_exploration_golden(gpu_lib, golden):
    from bmhrl_b200.model.bm_hrl_agent import Manager
    e = golden["exploration"]
    m = Manager("cuda", 12, 6, 0.0)
    m.linear.weight.data.copy_(e["w"])
    m.linear.bias.data.copy_(e["b"])
    m = m.cuda()
    m.exploration = True
    # fixture: noise = N(mean/10, std/5) - 0.5*mean/10 with e["mean"], e["std"] already divided -> recover the N(0,1) draws
    z = (e["noise"] - 0.5 * e["mean"]) / e["std"]
    m.noise_eps = z.cuda()
    out = m(e["x"].cuda(), e["seg"].cuda())
    assert rel(out, e["out"]) < 1e-5
    m.noise_eps = None  # device RNG: statistics only
    out2 = m(e["x"].cuda(), e["seg"].cuda())
    assert out2.shape == e["out"].shape and bool(torch.isfinite(out2).all())


def test_sampling_bit_exact(gpu_lib):
    from bmhrl_b200 import ops
    g = _g(4)
    B, L, V = 5, 9, 10172
    pred = torch.log_softmax(torch.randn(B, L, V, generator=g) * 2, -1)
    q = torch.empty(B, L, V).exponential_(1, generator=g)
    for noise in (None, q):
        tok_ref, p_ref = O.sample_tokens(pred, noise)
        pd = pred.clone().cuda().requires_grad_(True)
        tok, pr = ops.sample_tokens(pd, "argmax" if noise is None else "noise", None if noise is None else noise.cuda())
        assert torch.equal(tok.cpu(), tok_ref)
        assert rel(pr, p_ref) < TIGHT
        pr.sum().backward()
        pr_ref = pred.clone().requires_grad_(True)
        O.sample_tokens(pr_ref, noise)[1].sum().backward()
        assert rel(pd.grad, pr_ref.grad) < TIGHT
    tok, pr = ops.sample_tokens(pred.cuda(), "philox")
    assert tok.min() >= 0 and tok.max() < V and rel(pr, torch.exp(pred).gather(2, tok.cpu()[..., None]).squeeze(-1)) < TIGHT


def test_philox_sampling_distribution(gpu_lib):
    from bmhrl_b200 import ops
    V, n = 7, 40000
    p = torch.tensor([0.02, 0.3, 0.08, 0.25, 0.05, 0.2, 0.1])
    pred = torch.log(p)[None, None, :].expand(1, n, V).contiguous().cuda()
    ops.manual_seed(123)
    ops.new_rng_epoch("cuda")
    tok, _ = ops.sample_tokens(pred, "philox", salt=5)
    freq = torch.bincount(tok.flatten().cpu(), minlength=V).float() / n
    assert float((freq - p).abs().max()) < 0.01


def test_segment_scan(gpu_lib):
    from bmhrl_b200 import ops
    g = _g(6)
    B, L = 6, 9
    for t in range(5):
        seg = (torch.rand(B, L, generator=g) < 0.3).int()
        if t == 0:
            seg[0] = 0
        pr = torch.rand(B, L, generator=g) * 0.8 + 0.1
        ex = torch.randn(B, L, generator=g)
        prr = pr.clone().requires_grad_(True)
        sp_ref, ex_ref = O.segment_prod_and_sum(prr, ex, seg)
        w = torch.randn(B, L, generator=g)
        (sp_ref * w).sum().backward()
        prd = pr.clone().cuda().requires_grad_(True)
        sp, exo = ops.segment_scan(prd, ex.cuda(), seg.cuda())
        (sp * w.cuda()).sum().backward()
        assert rel(sp, sp_ref) < TIGHT and rel(exo, ex_ref) < TIGHT and rel(prd.grad, prr.grad) < 1e-4


def test_fused_adam_golden(gpu_lib, golden):
    from bmhrl_b200 import ops
    a = golden["adam"]
    p = torch.nn.Parameter(a["p0"].clone().cuda())
    frozen = torch.nn.Parameter(torch.ones(5).cuda(), requires_grad=False)
    opt = ops.FusedAdam([p, frozen], lr=1e-4)
    for gr in a["grads"]:
        opt.zero_grad()
        p.grad.copy_(gr.cuda())
        opt.step()
    assert rel(p, a["p3"]) < 1e-6
    assert torch.equal(frozen.cpu(), torch.ones(5))


def test_embed_posenc(gpu_lib, golden):
    from bmhrl_b200 import ops
    g = _g(12)
    W = torch.randn(50, 300, generator=g)
    tok = torch.randint(0, 50, (3, 11), generator=g)
    emb = ops.embed(tok.cuda(), W.cuda(), math.sqrt(300))
    assert rel(emb, O.embed_tokens({"emb_C.embedder.weight": W}, tok)) < 1e-6
    for D in (300, 128, 1024, 30):
        x = torch.randn(2, 13, D, generator=g)
        assert rel(ops.add_posenc(x.cuda()), O.add_positional(x)) < 1e-6
    x2 = torch.randn(2, 13, 1024, generator=g)
    assert rel(ops.add_posenc(x.new_ones(2, 13, 1024).cuda(), x2.cuda()), O.add_positional(1 + x2)) < 1e-6
    assert torch.equal(ops.positional_table_f64(40, 300)[::7, ::13], golden["posenc"]["t300"])


def test_segment_critic(gpu_lib, agent_sd):
    """LSTM x4 -> AReLU -> GRU x2 -> AReLU -> Linear -> sigmoid > 0.25: scores to fp32 round-off, labels exact."""
    from types import SimpleNamespace as NS
    from bmhrl_b200.model.bm_hrl_agent import SegmentCritic
    g = _g(20)
    crit = SegmentCritic(NS(d_model_caps=300, rl_critic_path=None))
    crit.load_state_dict({k[len("critic."):]: v for k, v in agent_sd.items() if k.startswith("critic.")})
    crit = crit.cuda()
    for B, L in ((4, 9), (16, 20), (19, 1), (1, 5)):
        x = torch.randn(B, L, 300, generator=g) * math.sqrt(300) * 0.3
        ref = O.segment_critic(agent_sd, x)
        scores, labels = crit.scores_and_labels(x.cuda(), 0.25)
        assert scores.shape == (B, L, 1) and labels.shape == (B, L)
        assert float((scores.cpu() - ref).abs().max()) < 1e-4
        near = (torch.sigmoid(ref) - 0.25).abs().squeeze(-1) < 1e-5
        assert torch.equal(labels.cpu()[~near], O.segment_labels_from_scores(ref)[~near])
        assert rel(crit(x.cuda()), ref) < 1e-4


def test_segment_critic_incremental_steps_equal_whole_prefix(gpu_lib, agent_sd):
    """SegmentCritic.step_labels (one rnn_step per layer per appended position, states kept by the caller) gives the
    labels AND hidden states of the whole-prefix recurrence bit for bit; a repeated step changes nothing (the LSTM cell
    state is kept per step)."""
    from types import SimpleNamespace as NS
    from bmhrl_b200 import ops
    from bmhrl_b200.model.bm_hrl_agent import SegmentCritic, SegmentCriticDecodeState
    g = _g(23)
    crit = SegmentCritic(NS(d_model_caps=300, rl_critic_path=None))
    crit.load_state_dict({k[len("critic."):]: v for k, v in agent_sd.items() if k.startswith("critic.")})
    crit = crit.cuda()
    for B, L, Lmax in ((3, 7, 10), (1, 12, 12), (17, 4, 6)):
        x = (torch.randn(B, L, 300, generator=g) * math.sqrt(300) * 0.3).cuda()
        state = SegmentCriticDecodeState(crit, B, Lmax, "cuda")
        for p in range(L):
            crit.step_labels(state, x[:, p:p + 1], p, 0.25)
            if p % 3 == 1:
                crit.step_labels(state, x[:, p:p + 1], p, 0.25)  # a repeated step (warm-up, then capture) is harmless
            _, labels = crit.scores_and_labels(x[:, :p + 1].contiguous(), 0.25)
            assert torch.equal(state.labels[:, :p + 1], labels), (B, p)
        assert torch.equal(state.gru[-1][1][:, :L], _critic_last_gru(crit, x))
    with pytest.raises(RuntimeError):
        ops.rnn_step("lstm", state.lstm[0][0], crit.lstm.weight_hh_l0, crit.lstm.bias_hh_l0, state.lstm[0][1], None, 0)
    with pytest.raises(RuntimeError):
        ops.rnn_step("gru", state.gru[0][0], crit.gru.weight_hh_l0, crit.gru.bias_hh_l0, state.gru[0][1], None, Lmax)


def _critic_last_gru(crit, x):
    from bmhrl_b200 import ops
    h = x
    for l in range(crit.lstm.num_layers):
        h = ops.rnn_layer("lstm", h, getattr(crit.lstm, f"weight_ih_l{l}"), getattr(crit.lstm, f"weight_hh_l{l}"),
                          getattr(crit.lstm, f"bias_ih_l{l}"), getattr(crit.lstm, f"bias_hh_l{l}"))
    h = crit.relu(h)
    for l in range(crit.gru.num_layers):
        h = ops.rnn_layer("gru", h, getattr(crit.gru, f"weight_ih_l{l}"), getattr(crit.gru, f"weight_hh_l{l}"),
                          getattr(crit.gru, f"bias_ih_l{l}"), getattr(crit.gru, f"bias_hh_l{l}"))
    return h


def test_detr_layers_golden(gpu_lib, golden):
    from bmhrl_b200.model.encoder import TransformerEncoder, TransformerEncoderLayer
    from bmhrl_b200.model.decoder import TransformerDecoder, TransformerDecoderLayer
    from bmhrl_b200.model.blocks import PositionalEncoder
    d = golden["detr"]
    dm = d["dims"]
    enc = TransformerEncoderLayer(dm["d_model"], dm["H"], dim_feedforward=dm["ff_e"], dropout=0.0, embed_size=dm["dC"])
    enc.load_state_dict(d["enc_sd"])
    enc = enc.cuda()
    pe_m, pe_c, pe_g = PositionalEncoder(dm["d_model"], 0.0), PositionalEncoder(dm["dC"], 0.0), PositionalEncoder(dm["dG"], 0.0)
    out = enc(d["src"].cuda(), d["smask"].cuda(), pe_m)
    assert rel(out, d["enc_out"]) < TIGHT * 5
    dec = TransformerDecoderLayer(dm["d_model"], dm["H"], dm["dC"], dm["dG"], dim_feedforward=dm["ff_d"], dropout=0.0)
    dec.load_state_dict(d["dec_sd"])
    dec = dec.cuda()
    o2 = dec(d["tgt"].cuda(), d["enc_out"].cuda(), d["smask"].cuda(), pe_m, pe_c, d["qmask"].cuda(), d["goal"].cuda(),
             d["gmask"].cuda(), pe_g, add_pos=False, detected_objects=d["det"].cuda())
    assert rel(o2, d["dec_out"]) < TIGHT * 5
    stack = TransformerEncoder(enc, 2, norm=None)(d["src"].cuda(), d["smask"].cuda(), pe_m)
    assert stack.shape == (2,) + tuple(d["src"].shape)
    st2 = TransformerDecoder(dec, 2)(d["tgt"].cuda(), d["enc_out"].cuda(), d["smask"].cuda(), pe_m, pe_c,
                                     d["qmask"].cuda(), d["goal"].cuda(), d["gmask"].cuda(), pe_g)
    assert st2.shape == (2,) + tuple(d["tgt"].shape)


def test_dropout_statistics_and_backward_mask(gpu_lib):
    from bmhrl_b200 import ops
    ops.manual_seed(7)
    ops.new_rng_epoch("cuda")
    x = torch.zeros(64, 1024, device="cuda")
    r = torch.ones(64, 1024, device="cuda", requires_grad=True)
    y = ops.dropout_add(x, r, 0.1, salt=3)
    keep = (y != 0).float().mean().item()
    assert abs(keep - 0.9) < 0.01
    assert abs(float(y.detach()[y.detach() != 0].mean()) - 1 / 0.9) < 1e-5
    y.sum().backward()
    assert torch.equal((r.grad != 0), (y != 0))
    y2 = ops.dropout_add(x, r.detach(), 0.1, salt=3)
    assert torch.equal(y2, y.detach())              # same epoch + salt -> same mask
    ops.new_rng_epoch("cuda")
    y3 = ops.dropout_add(x, r.detach(), 0.1, salt=3)
    assert not torch.equal(y3, y.detach())          # next step -> new mask


def test_errors_are_loud(gpu_lib):
    from bmhrl_b200 import ops
    with pytest.raises(RuntimeError):
        ops.linear(torch.randn(4, 8, device="cuda"), torch.randn(3, 7, device="cuda"))
    with pytest.raises(RuntimeError):
        ops.mha_core(torch.randn(2, 3, 10, device="cuda"), torch.randn(2, 3, 10, device="cuda"),
                     torch.randn(2, 3, 10, device="cuda"), None, 4)
    with pytest.raises(RuntimeError):
        ops.layer_norm(torch.randn(4, 8, device="cuda").double(), torch.ones(8, device="cuda"), torch.zeros(8, device="cuda"))


def test_make_masks_fused_equals_reference_masks(gpu_lib):
    """ops.make_masks_fused (one launch) against model/masking.py:28-53 as restated in bmhrl_b200.model.masking and in the
    oracle, including the reference's quirks (component 0 of rgb only; a zero first component masks a real frame), ragged
    padding, non-contiguous feature views, and the non-pad target count of the loss normaliser.  Bit-exact."""
    from bmhrl_b200 import ops
    from bmhrl_b200.model.masking import make_masks
    g = _g(61)
    for B, Sv, Sa, L in ((16, 30, 80, 20), (3, 7, 5, 1), (1, 1, 1, 9)):
        rgb = torch.randn(B, Sv, 1024, generator=g)
        aud = torch.randn(B, Sa, 128, generator=g)
        rgb[0, Sv // 2:, :] = 0
        aud[B - 1, :, 0] = 0            # real frames whose first component is 0 are masked too
        rgb[B // 2, 0, 0] = 0
        cap = torch.randint(4, 50, (B, L + 1), generator=g)
        cap[0, L // 2:] = 1
        cap_in, y = cap[:, :-1].contiguous(), cap[:, 1:].contiguous()
        want = make_masks({"rgb": rgb, "flow": rgb, "audio": aud}, cap_in, "audio_video", 1)
        got, nt = ops.make_masks_fused(rgb.cuda(), aud.cuda(), cap_in.cuda(), 1, targets=y.cuda())
        for k in ("V_mask", "A_mask", "C_mask"):
            assert got[k].dtype == torch.bool and got[k].shape == want[k].shape, k
            assert torch.equal(got[k].cpu(), want[k].bool()), (k, B)
        assert int(nt) == int((y != 1).sum())
        om = O.make_masks(rgb, aud, cap_in, 1)
        assert torch.equal(got["C_mask"].cpu(), om["C_mask"]) and torch.equal(got["V_mask"].cpu(), om["V_mask"])
        # a strided view (every second frame) and no captions / targets
        got2, nt2 = ops.make_masks_fused(rgb.cuda()[:, ::2], aud.cuda()[:, ::2], None, 1)
        assert nt2 is None and "C_mask" not in got2
        assert torch.equal(got2["V_mask"].cpu(), (rgb[:, ::2, 0] != 0).unsqueeze(1))
        assert torch.equal(got2["A_mask"].cpu(), (aud[:, ::2, 0] != 0).unsqueeze(1))


@pytest.mark.parametrize("cell", ["lstm", "gru"])
def test_rnn_layer_variants_bit_identical(gpu_lib, cell):
    """The three forms of one recurrent layer -- tagged-exchange persistent kernel (default), grid-barrier persistent
    kernel, one launch per step -- give the same bits; repeated launches on a recycled workspace (stale tags of the
    previous call in the exchange buffer) included, and batches of more than one 16-row group."""
    from bmhrl_b200 import ops
    from bmhrl_b200._lib import LIB
    g = _g(77)
    NG = 4 if cell == "lstm" else 3
    for H, E in ((600, 300), (64, 40)):
        w_ih = (torch.randn(NG * H, E, generator=g) / math.sqrt(E)).cuda()
        w_hh = (torch.randn(NG * H, H, generator=g) / math.sqrt(H)).cuda()
        b_ih = (torch.randn(NG * H, generator=g) * 0.1).cuda()
        b_hh = (torch.randn(NG * H, generator=g) * 0.1).cuda()
        for B, L in ((16, 20), (3, 7), (35, 4), (1, 1), (16, 2)):
            x = torch.randn(B, L, E, generator=g).cuda()
            outs = {}
            try:
                for name, (persistent, exchange) in {"exchange": (1, 1), "barrier": (1, 0), "steps": (0, 0)}.items():
                    LIB.call("bmhrl_rnn_set_persistent", persistent)
                    LIB.call("bmhrl_rnn_set_exchange", exchange)
                    outs[name] = [ops.rnn_layer(cell, x, w_ih, w_hh, b_ih, b_hh).clone() for _ in range(3)]
            finally:
                LIB.call("bmhrl_rnn_set_persistent", 1)
                LIB.call("bmhrl_rnn_set_exchange", -1)
            torch.cuda.synchronize()
            for name, o in outs.items():
                assert torch.isfinite(o[0]).all()
                assert torch.equal(o[0], o[1]) and torch.equal(o[0], o[2]), (name, H, B, L)
                assert torch.equal(o[0], outs["steps"][0]), (name, H, B, L)
            ref = getattr(torch.nn, cell.upper())(E, H, batch_first=True)   # CPU fp32 (no TF32 in the way)
            with torch.no_grad():
                ref.weight_ih_l0.copy_(w_ih); ref.weight_hh_l0.copy_(w_hh); ref.bias_ih_l0.copy_(b_ih); ref.bias_hh_l0.copy_(b_hh)
                want = ref(x.cpu())[0]
            assert rel(outs["exchange"][0], want) < 1e-4


@pytest.mark.parametrize("B,L", [(3, 7), (16, 20), (17, 5), (1, 1)])
def test_critic_wavefront_equals_layer_by_layer(gpu_lib, agent_sd, B, L):
    """ops.critic_wavefront (all recurrent layers of the SegmentCritic in one cooperative launch, L + 5 grid barriers)
    against the layer-by-layer kernels and the fp32 oracle: hidden states to 1e-5, segment labels exact."""
    from types import SimpleNamespace as NS
    from bmhrl_b200 import ops
    from bmhrl_b200.model.bm_hrl_agent import SegmentCritic
    g = _g(31 + B)
    crit = SegmentCritic(NS(d_model_caps=300, rl_critic_path=None))
    csd = {k[len("critic."):]: v for k, v in agent_sd.items() if k.startswith("critic.")}
    crit.load_state_dict(csd)
    crit = crit.cuda()
    x = (torch.randn(B, L, 300, generator=g) * math.sqrt(300) * 0.3)
    hw = ops.critic_wavefront(x.cuda(), crit.lstm, crit.gru, crit.relu.alpha, crit.relu.beta)
    assert hw is not None
    ref = _critic_last_gru(crit, x.cuda())
    assert rel(hw, ref) < 1e-5
    import os
    os.environ["BMHRL_CRITIC_WAVEFRONT"] = "1"
    try:
        s1, l1 = crit.scores_and_labels(x.cuda(), 0.25)
    finally:
        os.environ["BMHRL_CRITIC_WAVEFRONT"] = "0"
    s0, l0 = crit.scores_and_labels(x.cuda(), 0.25)
    assert torch.equal(l1, l0) and rel(s1, s0) < 1e-5
    so = O.segment_critic({"critic." + k: v for k, v in csd.items()}, x)
    assert torch.equal(l1.cpu(), O.segment_labels_from_scores(so, 0.25).to(l1.dtype).cpu().reshape(l1.shape))
