"""GPU: the bf16 math mode (bmhrl_b200/ops16.py: grouped bf16 tcgen05 GEMMs, fused one-tile attention forward + backward,
block-level autograd nodes) against the fp32 oracle and the reference's golden vectors.  Bound: the north_star's bf16
bound 2e-2 on logits, losses and gradients; integer outputs (segment labels) exact; observed errors are printed."""
import math

import pytest
import torch
import torch.nn as nn

from oracle import bmhrl_oracle as O
from bmhrl_b200 import synth
from conftest import PAD, SMOOTH, VOC, build_agent, rel

pytestmark = pytest.mark.gpu
TOL = 2e-2


@pytest.fixture()
def bf16(gpu_lib):
    from bmhrl_b200 import ops
    ops.set_math_mode("bf16")
    yield ops
    ops.set_math_mode("fp32")


def _g(seed):
    return torch.Generator().manual_seed(seed)


def _sd(mod, prefix):
    return {prefix + k: v.detach().cpu() for k, v in mod.state_dict().items()}


@pytest.mark.parametrize("B,Sq,Sk,dq,dkv,H,kind", [
    (4, 30, 30, 1024, None, 4, "pad"),       # video self-attention
    (4, 80, 80, 128, None, 4, "pad"),        # audio self-attention
    (4, 30, 80, 1024, 128, 4, "pad"),        # video <- audio bimodal attention
    (4, 80, 30, 128, 1024, 4, "pad"),        # audio <- video
    (4, 20, 20, 300, None, 4, "full"),       # caption self-attention, (B,L,L) mask
    (4, 20, 30, 300, 1024, 4, "pad"),        # caption <- video memory
])
def test_attention_residual_block(bf16, B, Sq, Sk, dq, dkv, H, kind):
    from bmhrl_b200.model.blocks import AttentionSublayer, ResidualConnection
    from bmhrl_b200.model.multihead_attention import MultiheadedAttention
    g = _g(B + Sq + Sk + dq)
    att = MultiheadedAttention(dq, dkv or dq, dkv or dq, H, 0.0, 1024).cuda()
    res = ResidualConnection(dq, 0.0).cuda()
    with torch.no_grad():
        res.norm.weight.copy_(1 + 0.1 * torch.randn(dq, generator=g))
        res.norm.bias.copy_(0.1 * torch.randn(dq, generator=g))
    x = torch.randn(B, Sq, dq, generator=g)
    kv = torch.randn(B, Sk, dkv, generator=g) if dkv else None
    if kind == "pad":
        lens = torch.randint(1, Sk + 1, (B,), generator=g)
        lens[0] = Sk
        mask = (torch.arange(Sk)[None, :] < lens[:, None])[:, None, :]
        mask[1] = False
    else:
        mask = torch.tril(torch.ones(Sq, Sk, dtype=torch.bool))[None].repeat(B, 1, 1)
        mask[0, :, Sk // 2:] = False
    dy = torch.randn(B, Sq, dq, generator=g)
    # oracle (fp32, CPU)
    sd = {**_sd(att, "a."), **_sd(res.norm, "n.")}
    xr = x.clone().requires_grad_(True)
    kvr = kv.clone().requires_grad_(True) if dkv else None
    pr = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    n = O.layer_norm(pr, "n.", xr)
    yr = xr + O.mha(pr, "a.", n, n if kvr is None else kvr, n if kvr is None else kvr, mask, H)
    yr.backward(dy)
    # device
    xd = x.cuda().requires_grad_(True)
    kvd = kv.cuda().requires_grad_(True) if dkv else None
    y = res(xd, AttentionSublayer(att, kvd, kvd, mask.cuda()))
    y.backward(dy.cuda())
    errs = {"y": rel(y, yr), "dx": rel(xd.grad, xr.grad)}
    if dkv:
        errs["dkv"] = rel(kvd.grad, kvr.grad)
    for name, p in list(att.named_parameters()) + [("norm." + k, v) for k, v in res.norm.named_parameters()]:
        key = ("a." + name) if not name.startswith("norm.") else ("n." + name[5:])
        if name == "linear_K2d.bias":   # softmax is shift invariant: the true gradient is 0, only noise is left
            continue
        errs[name] = rel(p.grad, pr[key].grad)
    print("attention block errors", {k: f"{v:.1e}" for k, v in errs.items()})
    assert max(errs.values()) < TOL, errs


def _rt(t):
    """bf16 rounding with a straight-through gradient: the oracle evaluated on the operands the tensor cores see."""
    return t + (t.detach().to(torch.bfloat16).float() - t.detach())


def test_ffn_residual_block_and_linear16(bf16):
    """The FFN block.  ReLU makes the gradient discontinuous: where a pre-activation is within the bf16 operand
    rounding of zero its gate can come out on the other side, and ONE flipped gate changes that element's gradient by
    100 %.  So (1) against the oracle on the SAME bf16-rounded operands (only accumulation order differs: essentially no
    flips) every output / gradient must meet the 2e-2 bound, and (2) against the exact fp32 oracle the non-gated
    quantities meet it too, while the gated ones (fc1, norm) are reported and bounded by the near-zero fraction."""
    from bmhrl_b200 import ops16
    from bmhrl_b200.model.blocks import PositionwiseFeedForward, ResidualConnection
    g = _g(5)
    B, S, D, F_ = 4, 30, 1024, 1024
    ff = PositionwiseFeedForward(D, F_, 0.0).cuda()
    res = ResidualConnection(D, 0.0).cuda()
    x = torch.randn(B, S, D, generator=g)
    dy = torch.randn(B, S, D, generator=g)
    sd = {**_sd(ff, "f."), **_sd(res.norm, "n.")}
    refs = {}
    for kind in ("exact", "rounded"):
        pr = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
        xr = x.clone().requires_grad_(True)
        n = O.layer_norm(pr, "n.", xr)
        if kind == "exact":
            yr = xr + O.ffn(pr, "f.", n)
        else:
            h = torch.relu(torch.nn.functional.linear(_rt(n), _rt(pr["f.fc1.weight"]), pr["f.fc1.bias"]))
            yr = xr + torch.nn.functional.linear(_rt(h), _rt(pr["f.fc2.weight"]), pr["f.fc2.bias"])
        yr.backward(dy)
        refs[kind] = (yr.detach(), xr.grad, {k: v.grad for k, v in pr.items()})
    xd = x.cuda().requires_grad_(True)
    y = res(xd, ff)
    y.backward(dy.cuda())
    for kind, (yr, dxr, gr) in refs.items():
        errs = {"y": rel(y, yr), "dx": rel(xd.grad, dxr)}
        for name, p in ff.named_parameters():
            errs[name] = rel(p.grad, gr["f." + name])
        for name, p in res.norm.named_parameters():
            errs["norm." + name] = rel(p.grad, gr["n." + name])
        print(f"ffn block errors vs the {kind} oracle", {k: f"{v:.1e}" for k, v in errs.items()})
        if kind == "rounded":
            assert max(errs.values()) < TOL, errs
        else:
            gated = ("fc1.weight", "fc1.bias", "norm.weight", "norm.bias", "dx")
            assert max(v for k, v in errs.items() if k not in gated) < TOL, errs
            assert max(errs[k] for k in gated) < 5e-2, errs   # gate flips of near-zero pre-activations (randn data)
    # plain linear (Manager.linear 300 -> 64, the vocabulary projection 364 -> V)
    for (M, K, N) in [(320, 300, 64), (320, 364, 10172)]:
        lin = nn.Linear(K, N).cuda()
        x = torch.randn(M, K, generator=g)
        dy = torch.randn(M, N, generator=g)
        xr = x.clone().requires_grad_(True)
        w, b = lin.weight.detach().cpu().requires_grad_(True), lin.bias.detach().cpu().requires_grad_(True)
        (torch.nn.functional.linear(xr, w, b)).backward(dy)
        xd = x.cuda().requires_grad_(True)
        ops16.linear16(xd, lin.weight, lin.bias).backward(dy.cuda())
        # (the bias gradient is the fixed-order column sum of the bf16 dZ: ~2^-9 per element)
        assert rel(xd.grad, xr.grad) < TOL and rel(lin.weight.grad, w.grad) < TOL and rel(lin.bias.grad, b.grad) < 5e-3


@pytest.mark.parametrize("tag", ["small", "c16"])
def test_agent_bf16_vs_reference_golden(bf16, golden, agent_sd, tag):
    """Whole agent forward / backward in bf16 math mode vs the reference's fp32 outputs."""
    from bmhrl_b200.model.masking import make_masks
    from bmhrl_b200.loss.label_smoothing import LabelSmoothing
    g = golden[f"agent_{tag}"]
    c = g["cfg"]
    model, _ = build_agent(agent_sd)
    model.train()
    model.teach_warmstart()
    model.manager.exploration = False
    b = {k: v.cuda() for k, v in synth.synthetic_batch(B=c["B"], Sv=c["Sv"], Sa=c["Sa"], L=c["L"], V=VOC, seed=c["seed"],
                                                        ragged=c["ragged"]).items()}
    cin, cy = b["caption"][:, :-1], b["caption"][:, 1:]
    out = model((b["rgb"] + b["flow"], b["audio"]), cin, make_masks(b, cin, "audio_video", PAD))
    loss = torch.sum(LabelSmoothing(SMOOTH, PAD)(out[0], cy)) / (cy != PAD).sum()
    loss.backward()
    bf16.join_side_streams(loss.device)
    assert torch.equal(out[4].cpu(), g["segment_labels"])
    errs = {"pred": rel(out[0][:, :, ::101], g["pred_slice"]), "worker_feat": rel(out[1], g["worker_feat"]),
            "manager_feat": rel(out[2], g["manager_feat"]), "goals": rel(out[3], g["goals"]), "loss": rel(loss, g["loss"])}
    params = dict(model.named_parameters())
    per = {}
    for n, gn in g["grad_norm"].items():
        if n.endswith("linear_K2d.bias"):
            continue
        sl = params[n].grad.flatten()[:: max(1, params[n].grad.numel() // 64)][:64].cpu()
        per[n] = max(float((sl - g["grad_slice"][n]).norm()) / (gn + 1e-12), abs(float(params[n].grad.norm()) - gn) / (gn + 1e-12))
    top = sorted(per.items(), key=lambda kv: -kv[1])[:8]
    # a_v_constant is ONE scalar per fusion layer: its gradient sum_i g_i (Cv_i - Ca_i) is a heavily cancelling sum over
    # all B*L*300 elements (|sum| << sum|.|), so rounding anywhere upstream is amplified by that condition number; it is
    # reported and held to 1e-1, every tensor-valued gradient to the 2e-2 bound
    scalars = {n: e for n, e in per.items() if n.endswith("a_v_constant")}
    errs["grads_worst"] = max(e for n, e in per.items() if n not in scalars)
    print("bf16 errors", tag, {k: f"{v:.2e}" for k, v in errs.items()}, "top:", [(n, f"{e:.2e}") for n, e in top])
    assert max(errs.values()) < TOL, (errs, top)
    assert max(scalars.values()) < 1e-1, scalars


def _trainer(agent_sd, **kw):
    from bmhrl_b200.model.bm_hrl_agent import BMManagerValueFunction, BMWorkerValueFunction
    from bmhrl_b200.train import BMHRLTrainer
    model, cfg = build_agent(agent_sd, **kw)
    tr = BMHRLTrainer(model, BMWorkerValueFunction(cfg).cuda(), BMManagerValueFunction(cfg).cuda(), pad_idx=PAD)
    tr.teach_warmstart()
    model.manager.exploration = False
    return model, tr


def test_step_bf16_adam_keeps_shadows_fresh_and_is_deterministic(bf16, agent_sd):
    """(1) after FusedAdam steps every bf16 weight shadow equals bf16(parameter) bit for bit (the Adam kernel writes
    them); (2) the step has no floating-point atomics: two runs from the same state give bit-identical parameters;
    (3) the CUDA-graph replay of the step equals the eager step."""
    from bmhrl_b200 import ops16
    from bmhrl_b200.train import StepRunner
    b = synth.synthetic_batch(B=4, Sv=10, Sa=14, L=7, V=VOC, seed=2, ragged=True)
    finals = []
    for use_graph in (False, False, True):
        bf16.manual_seed(7)
        bf16.RNG.salt = 5000    # the same dropout call-site salts in every run
        model, tr = _trainer(agent_sd, dout_p=0.1)
        run = StepRunner(tr, "warmstart", b, use_graph=use_graph, warmup=2 if use_graph else 0)
        n0 = 2 if use_graph else 0
        losses = [float(run.step(b)[0]) for _ in range(4 - n0)]
        torch.cuda.synchronize()
        finals.append((losses, {n: p.detach().clone() for n, p in model.named_parameters()}))
        if not use_graph:
            checked = 0
            for n, p in model.named_parameters():
                e = ops16.SHADOWS.entry(p)
                if e is None or e[3] != 0:
                    continue
                for key, grp in ops16.SHADOWS.groups.items():
                    if id(p) in key:
                        r0 = sum(q.shape[0] for q in [pp for pp in model.parameters() if id(pp) in key[:key.index(id(p))]])
                        sh = grp["w16"][r0:r0 + p.shape[0]]
                        assert torch.equal(sh, p.detach().to(torch.bfloat16)), n
                        checked += 1
            assert checked > 40
    assert finals[0][0] == finals[1][0]
    for n in finals[0][1]:
        assert torch.equal(finals[0][1][n], finals[1][1][n]), f"run-to-run difference in {n}"
    worst = max(rel(finals[2][1][n], finals[0][1][n]) for n in finals[0][1])
    assert worst < 1e-6, worst
    assert finals[0][0][-1] < finals[0][0][0]


def test_backward_gradient_twin_equals_separate_cast(bf16, agent_sd, monkeypatch):
    """A block's LayerNorm backward emits bf16(dropout-mask o dx) for the block that produced its input
    (ops16.hint_backward_cast / grad16_of): same parameters after two steps, bit for bit, as with the separate mask + cast
    launch at the head of every block's backward -- and the emitted copy is really used where a block's output has a single
    consumer (6 of the 25 block heads of this model per step: the FFN <- bimodal-attention and the layer-to-layer hand-overs
    of the encoder; everywhere else the gradient is a sum over several consumers and the cast stays)."""
    from bmhrl_b200 import ops16
    from bmhrl_b200.train import StepRunner
    b = synth.synthetic_batch(B=4, Sv=10, Sa=14, L=7, V=VOC, seed=3, ragged=True)
    finals, taken = [], []
    real = ops16.grad16_of
    for fused in (True, False):
        hits = [0, 0]

        def spy(dy, rows, cols, drop_p, salt, fused=fused, hits=hits):
            t = real(dy, rows, cols, drop_p, salt) if fused else None
            hits[0 if t is not None else 1] += 1
            return t
        monkeypatch.setattr(ops16, "grad16_of", spy)
        bf16.manual_seed(11)
        bf16.RNG.salt = 7000
        model, tr = _trainer(agent_sd, dout_p=0.1)
        run = StepRunner(tr, "warmstart", b, use_graph=False)
        losses = [float(run.step(b)[0]) for _ in range(2)]
        torch.cuda.synchronize()
        finals.append((losses, {n: p.detach().clone() for n, p in model.named_parameters()}))
        taken.append(hits)
    monkeypatch.setattr(ops16, "grad16_of", real)
    print("blocks that started from an emitted gradient copy / from a cast:", taken)
    assert taken[0][0] >= 10 and taken[1][0] == 0
    assert finals[0][0] == finals[1][0]
    for n in finals[0][1]:
        assert torch.equal(finals[0][1][n], finals[1][1][n]), n


def test_warmstart_step_bf16_matches_oracle_adam(bf16, agent_sd):
    """Two warm-start steps in bf16 math mode vs oracle autograd + oracle Adam: losses within 2e-2; parameters move by
    lr-sized steps, so they are compared through the size of the update."""
    model, tr = _trainer(agent_sd)
    b = synth.synthetic_batch(B=3, Sv=8, Sa=10, L=6, V=VOC, seed=31, ragged=True)
    cin, cy = b["caption"][:, :-1], b["caption"][:, 1:]
    params = {k: v.clone() for k, v in agent_sd.items()}
    state, losses = {}, []
    for _ in range(2):
        p = {k: (v.clone().requires_grad_(True) if v.dtype.is_floating_point else v) for k, v in params.items()}
        pred = O.agent_forward(p, b["rgb"] + b["flow"], b["audio"], cin, O.make_masks(b["rgb"], b["audio"], cin, PAD))[0]
        loss = O.label_smoothing_kl(pred, cy, SMOOTH, PAD).sum() / (cy != PAD).sum()
        loss.backward()
        grads = {k: v.grad for k, v in p.items() if torch.is_tensor(v) and v.grad is not None and not k.startswith("manager.core.")}
        params, state = O.adam_step({k: v.detach() for k, v in p.items()}, grads, state)
        losses.append(float(loss.detach()))
    dl = [float(tr.warmstart_step({k: v.cuda() for k, v in b.items()})) for _ in range(2)]
    assert abs(dl[0] - losses[0]) < TOL * abs(losses[0]) and abs(dl[1] - losses[1]) < TOL * abs(losses[1])
    moved = 0
    for n, q in model.named_parameters():
        if q.requires_grad and n in params and not n.endswith("linear_K2d.bias"):
            step = (params[n] - agent_sd[n]).norm()
            if float(step) > 0:
                # Adam normalises the gradient: a 2e-2 gradient error moves a weight by at most ~2e-2 of its step
                assert float((q.detach().cpu() - params[n]).norm()) < 0.25 * float(step) + 1e-7, n
                moved += 1
    assert moved > 50


def test_fused_vocab_stats_and_kl_vs_oracle(bf16):
    """K4 fused: vocabulary projection + log-softmax + KL losses without the (M,V) logits.  Statistics and tokens against
    fp32 torch on the same bf16-rounded operands; LabelSmoothing / BiasedKL row sums and ALL gradients (features,
    projection weight / bias, biased offset) against the oracle's element-wise divergence on log_softmax of those logits;
    covers pad rows, sampled == target, sampled == pad and the reference's flat-row-0 quirk."""
    from bmhrl_b200 import ops16
    g = _g(13)
    M, K, V = 41, 364, 10172
    lin = nn.Linear(K, V).cuda()
    x = torch.randn(M, K, generator=g)
    xb = x.to(torch.bfloat16).float()
    wb = lin.weight.detach().cpu().to(torch.bfloat16).float()
    bias = lin.bias.detach().cpu()
    for case in ("plain", "row0_quirk"):
        trg = torch.randint(2, V, (M,), generator=g)
        if case == "plain":
            trg[5] = PAD; trg[17] = PAD; trg[0] = PAD
        else:
            trg[0] = PAD                      # the ONLY pad target is flat row 0: the reference zeroes nothing
        logits = xb @ wb.t() + bias
        # ---- statistics + arg-max / noisy arg-max tokens
        st, tok = ops16.vocab_stats(x.cuda(), lin.weight, lin.bias, trg.cuda(), PAD, "argmax")
        ref = torch.stack([torch.logsumexp(logits, -1), logits.sum(-1), logits.gather(1, trg[:, None])[:, 0], logits[:, PAD],
                           logits.max(-1).values], -1)
        assert rel(st, ref) < 1e-5
        assert torch.equal(tok.cpu(), logits.argmax(-1))
        q = torch.empty(M, V).exponential_(1, generator=g)
        st2, tok2 = ops16.vocab_stats(x.cuda(), lin.weight, lin.bias, trg.cuda(), PAD, "noise", q.cuda())
        key = logits - q.log()
        assert float((tok2.cpu() == key.argmax(-1)).float().mean()) > 0.95
        flipped = tok2.cpu() != key.argmax(-1)
        if flipped.any():   # only exact near-ties may differ
            a = key.gather(1, tok2.cpu()[:, None])[:, 0][flipped]
            assert float((key.max(-1).values[flipped] - a).abs().max()) < 1e-3
        # ---- losses: label smoothing and biased KL with awkward sampled tokens
        samp = torch.randint(0, V, (M,), generator=g)
        samp[3] = trg[3]
        samp[4] = PAD
        samp[0] = 7
        off = torch.rand(M, generator=g)
        w = torch.randn(M, generator=g)
        for mode in ("ls", "bkl"):
            xr = xb.clone().requires_grad_(True)
            wr = wb.clone().requires_grad_(True)
            br = bias.clone().requires_grad_(True)
            offr = off.clone().requires_grad_(True)
            lp = torch.log_softmax(xr @ wr.t() + br, -1)
            if mode == "ls":
                rows_r = O.label_smoothing_kl(lp[None], trg[None], SMOOTH, PAD).sum(-1)
            else:
                rows_r = O.biased_kl_divergence(lp[None], trg[None], samp[None], offr[None], SMOOTH, PAD).sum(-1)
            (rows_r * w).sum().backward()
            xd = x.cuda().requires_grad_(True)
            offd = off.cuda().requires_grad_(True)
            lin.zero_grad()
            if mode == "ls":
                std, _ = ops16.vocab_stats(xd, lin.weight, lin.bias, trg.cuda(), PAD, "argmax")
                rows = ops16.kl_rowsum_from_stats(std, trg.cuda(), SMOOTH, PAD, V)
            else:
                # the token the statistics carry must be the sampled one: feed it through the noise channel
                qq = torch.ones(M, V)
                qq[torch.arange(M), samp] = 1e-30
                std, tk = ops16.vocab_stats(xd, lin.weight, lin.bias, trg.cuda(), PAD, "noise", qq.cuda())
                assert torch.equal(tk.cpu(), samp)
                rows = ops16.kl_rowsum_from_stats(std, trg.cuda(), SMOOTH, PAD, V, samp.cuda(), offd)
            (rows * w.cuda()).sum().backward()
            bf16.join_side_streams("cuda")
            errs = {"rows": rel(rows, rows_r), "dx": rel(xd.grad, xr.grad), "dW": rel(lin.weight.grad, wr.grad),
                    "db": rel(lin.bias.grad, br.grad)}
            if mode == "bkl":
                errs["doffset"] = rel(offd.grad, offr.grad)
            print("fused vocab", case, mode, {k: f"{v:.1e}" for k, v in errs.items()})
            assert errs["rows"] < 1e-4 and max(errs.values()) < TOL, (case, mode, errs)


@pytest.mark.parametrize("kind", ["warmstart", "worker", "manager"])
def test_trainer_fused_vocab_equals_materialised(bf16, agent_sd, golden, kind):
    """The trainer's fused vocabulary path (no (B*L,V) tensor) against its materialising path in the same math mode:
    same losses, same tokens, same gradients."""
    res = []
    b = synth.synthetic_batch(B=4, Sv=10, Sa=14, L=7, V=VOC, seed=2, ragged=True)
    b = {k: v.cuda() for k, v in b.items()}
    for fuse in (False, True):
        bf16.manual_seed(3)
        bf16.RNG.salt = 7000
        model, tr = _trainer(agent_sd)
        tr.fuse_vocab = fuse
        if kind == "warmstart":
            tr.cap_opt.zero_grad()
            loss = None
            x, cap_in, y, masks = tr._inputs(b)
            pred = tr._forward(x, cap_in, masks)[0]
            from bmhrl_b200 import ops16
            if fuse:
                assert isinstance(pred, ops16.VocabHead)
                st, _ = pred.stats(y, PAD)
                loss = ops16.kl_rowsum_from_stats(st, y, SMOOTH, PAD, VOC).sum() / (y != PAD).sum()
            else:
                loss = tr.ls.row_sums(pred, y).sum() / (y != PAD).sum()
            loss.backward()
            tok = None
        else:
            (tr.teach_worker if kind == "worker" else tr.teach_manager)()
            model.manager.exploration = False
            tr.cap_opt.zero_grad()
            torch.manual_seed(5)
            q = torch.empty(4, 7, VOC).exponential_(1).cuda() if kind == "worker" else None
            loss, vloss, tok, _, _ = tr.rl_losses(b, kind == "worker", False, q)
            loss.backward()
        bf16.join_side_streams("cuda")
        torch.cuda.synchronize()
        res.append((float(loss), tok, {n: p.grad.detach().clone() for n, p in model.named_parameters() if p.grad is not None}))
    assert abs(res[0][0] - res[1][0]) < 2e-3 * abs(res[0][0]), (res[0][0], res[1][0])
    if res[0][1] is not None:
        assert float((res[0][1] == res[1][1]).float().mean()) > 0.9
    if res[0][1] is None or torch.equal(res[0][1], res[1][1]):
        worst = max(rel(res[1][2][n], res[0][2][n]) for n in res[0][2] if not n.endswith("linear_K2d.bias")
                    and not n.endswith("a_v_constant"))
        assert worst < 1e-2, worst
