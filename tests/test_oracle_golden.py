"""CPU: the oracle (oracle/bmhrl_oracle.py) re-checked against the committed golden vectors, which are outputs of
the UNMODIFIED reference (written by oracle/pin_oracle.py in the build container)."""
import torch

from oracle import bmhrl_oracle as O
from bmhrl_b200 import synth
from conftest import PAD, SMOOTH, VOC, rel

TOL = 2e-5


def _params(sd):
    return {k: (v.clone().requires_grad_(True) if v.dtype.is_floating_point else v) for k, v in sd.items()}


def test_positional_table(golden):
    assert torch.equal(O.positional_table(40, 300)[::7, ::13], golden["posenc"]["t300"])


def test_losses(golden):
    g = golden["losses"]
    i = g["inputs"]
    for name in ("ls", "bkl"):
        lp = torch.log_softmax(i["logits"].clone().requires_grad_(True), -1)
        lp.retain_grad()
        off = i["off"].clone().requires_grad_(True)
        if name == "ls":
            o = O.label_smoothing_kl(lp, i["trg"], SMOOTH, PAD)
        else:
            o = O.biased_kl_divergence(lp, i["trg"], i["samp"], off, SMOOTH, PAD)
        (o * g[name]["w"]).sum().backward()
        assert rel(o, g[name]["out"]) < TOL
        assert rel(lp.grad, g[name]["dpred"]) < TOL
        if name == "bkl":
            assert rel(off.grad, g[name]["doffset"]) < TOL
    r = g["reinforce"]
    probs = torch.softmax(i["logits"], -1).clone().requires_grad_(True)
    cv = r["critic"].clone().requires_grad_(True)
    o = O.reinforce_loss(probs, i["samp"], r["value"], cv)
    o.backward()
    assert rel(o, r["out"]) < TOL and rel(probs.grad, r["dprobs"]) < TOL and rel(cv.grad, r["dcritic"]) < TOL


def test_losses_only_row0_padded(golden):
    """The reference zeroes pad rows only `if mask.sum() > 0` over the flat row INDICES of the pad targets
    (loss/label_smoothing.py:26-30): a lone pad target in flat row 0 keeps its smoothed distribution."""
    g = golden["losses_row0"]
    i = g["inputs"]
    assert int(i["trg"][0, 0]) == PAD and int((i["trg"] == PAD).sum()) == 1
    for name in ("ls", "bkl"):
        lp = torch.log_softmax(i["logits"].clone().requires_grad_(True), -1)
        lp.retain_grad()
        if name == "ls":
            o = O.label_smoothing_kl(lp, i["trg"], SMOOTH, PAD)
        else:
            o = O.biased_kl_divergence(lp, i["trg"], i["samp"], i["off"].clone(), SMOOTH, PAD)
        (o * g[name]["w"]).sum().backward()
        assert float(g[name]["out"][0].abs().sum()) > 0  # row 0 was NOT zeroed by the reference
        assert rel(o, g[name]["out"]) < TOL
        assert rel(lp.grad, g[name]["dpred"]) < TOL


def test_expand_goals(golden):
    for c in golden["expand_goals"]:
        x = c["x"].clone().requires_grad_(True)
        o = O.expand_goals(x * 1.0, c["seg"])
        (o * c["w"]).sum().backward()
        assert torch.equal(o.detach(), c["out"])
        assert torch.allclose(x.grad, c["dx"])


def test_exploration(golden):
    e = golden["exploration"]
    sd = {"manager.linear.weight": e["w"], "manager.linear.bias": e["b"]}
    assert rel(O.manager_head(sd, e["x"], e["seg"], e["noise"]), e["out"]) < TOL
    pre = torch.nn.functional.linear(e["x"], e["w"], e["b"])
    mean, std = O.exploration_noise_params(pre)
    assert rel(mean, e["mean"]) < TOL and rel(std, e["std"]) < TOL


def test_adam(golden):
    a = golden["adam"]
    params, state = {"p": a["p0"].clone()}, {}
    for gr in a["grads"]:
        params, state = O.adam_step(params, {"p": gr}, state)
    assert rel(params["p"], a["p3"]) < 1e-6


def test_detr_layers(golden):
    d = golden["detr"]
    H = d["dims"]["H"]
    o = O.detr_encoder_layer(d["enc_sd"], "", d["src"], d["smask"], H)
    assert rel(o, d["enc_out"]) < TOL
    o2 = O.detr_decoder_layer(d["dec_sd"], "", d["tgt"], o, d["smask"], d["qmask"], H, goal=d["goal"],
                              goal_mask=d["gmask"], detected=d["det"], obj_mask=None)
    assert rel(o2, d["dec_out"]) < TOL


def _agent_case(golden, agent_sd, tag):
    g = golden[f"agent_{tag}"]
    c = g["cfg"]
    b = synth.synthetic_batch(B=c["B"], Sv=c["Sv"], Sa=c["Sa"], L=c["L"], V=VOC, seed=c["seed"], ragged=c["ragged"])
    cin, cy = b["caption"][:, :-1], b["caption"][:, 1:]
    p = _params(agent_sd)
    masks = O.make_masks(b["rgb"], b["audio"], cin, PAD)
    out = O.agent_forward(p, b["rgb"] + b["flow"], b["audio"], cin, masks)
    loss = O.label_smoothing_kl(out[0], cy, SMOOTH, PAD).sum() / (cy != PAD).sum()
    loss.backward()
    assert rel(out[0][:, :, ::101], g["pred_slice"]) < TOL
    assert rel(out[1], g["worker_feat"]) < TOL and rel(out[2], g["manager_feat"]) < TOL
    assert rel(out[3], g["goals"]) < TOL
    assert torch.equal(out[4], g["segment_labels"])
    assert rel(loss, g["loss"]) < TOL
    for n, gs in g["grad_slice"].items():
        og = p[n].grad
        assert og is not None, n
        if n.endswith("linear_K2d.bias"):  # mathematically zero: both sides hold round-off only
            continue
        assert abs(float(og.norm()) - g["grad_norm"][n]) <= 5e-4 * g["grad_norm"][n] + 1e-9, n
        sl = og.flatten()[:: max(1, og.numel() // 64)][:64]
        assert float((sl - gs).norm()) <= 5e-4 * g["grad_norm"][n] + 1e-9, n
    for n in g["never_grad"]:
        assert p[n].grad is None, n


def test_agent_small(golden, agent_sd):
    _agent_case(golden, agent_sd, "small")


def test_rl_steps(golden, agent_sd):
    for tag in ("worker_raw", "worker_stab", "manager_raw", "manager_stab"):
        g = golden[f"rl_{tag}"]
        c = g["cfg"]
        train_worker, stabilize = tag.startswith("worker"), tag.endswith("stab")
        b = synth.synthetic_batch(B=c["B"], Sv=c["Sv"], Sa=c["Sa"], L=c["L"], V=VOC, seed=c["seed"], ragged=True)
        cin, cy = b["caption"][:, :-1], b["caption"][:, 1:]
        p = _params(agent_sd)
        pv = _params(synth.fill_state_dict(golden["value_fn_shapes"], seed=c["vseed"]))
        masks = O.make_masks(b["rgb"], b["audio"], cin, PAD)
        pred, wf, mf, gf, seg = O.agent_forward(p, b["rgb"] + b["flow"], b["audio"], cin, masks)
        ev = O.value_function(pv, (wf if train_worker else mf).detach()).squeeze(-1)
        torch.manual_seed(c["noise_seed"])
        q = torch.empty(c["B"] * c["L"], VOC).exponential_(1).view(c["B"], c["L"], VOC) if train_worker else None
        loss_mask = cy != PAD
        div, score, tok, oexp = O.biased_kl_rl(pred, cy, b["score"] * c["score_scale"], ev.detach(), loss_mask, seg,
                                               train_worker, stabilize, SMOOTH, PAD, q)
        cap = div.sum() / (loss_mask.sum() * 0.2)
        cap.backward()
        assert torch.equal(tok, g["tokens"])
        assert torch.equal(seg, g["segment_labels"])
        assert rel(score, g["score"]) < TOL and rel(div.sum(-1), g["div_rowsum"]) < TOL
        assert rel(cap, g["cap_loss"]) < TOL
        ev_used = ev + (oexp - ev).detach()
        lm = loss_mask if train_worker else seg.float()
        vl = (((ev_used - score.float()) ** 2) * lm.float()).mean()
        vl.backward()
        assert rel(vl, g["value_loss"]) < TOL
        for n, gn in g["grad_norm"].items():
            if n.endswith("linear_K2d.bias"):
                continue
            assert abs(float(p[n].grad.norm()) - gn) <= 2e-4 * gn + 1e-9, n
        for n, gn in g["vgrad_norm"].items():
            assert abs(float(pv[n].grad.norm()) - gn) <= 2e-4 * gn + 1e-9, n


def test_greedy_fp16_inputs(golden, agent_sd):
    g = golden["greedy"]
    rgb, flow, aud = g["rgb"].float(), g["flow"].float(), g["audio"].float()
    assert int((aud[0, :, 0] != 0).sum()) == g["audio_mask_true"]
    tok, last = O.greedy_decode(agent_sd, rgb, flow, aud, 4, 2, 3, PAD)
    assert torch.equal(tok, g["tokens_fp16in"][:, :5])
    assert rel(last[:, :, ::101], g["last_slice_fp16in"][:, :4]) < TOL
