"""Input pipeline: crop / pad / rgb+flow / masks / generate_synonyms.
CPU: the host logic (bmhrl_b200.data) and the oracle against goldens written from the UNMODIFIED reference
(oracle/pin_pipeline.py).  GPU: the two kernels through the C ABI against the same goldens and the oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import bmhrl_oracle as O
from bmhrl_b200 import data
from conftest import ROOT


@pytest.fixture(scope="module")
def pgold():
    return torch.load(os.path.join(ROOT, "tests", "golden", "pipeline_golden.pt"), weights_only=False)


def test_crop_bounds_match_reference(pgold):
    for S, a, b, dur, s, e, n in pgold["crop"]:
        assert data.crop_bounds(S, a, b, dur) == (s, e) == O.crop_bounds(S, a, b, dur)
        feat = np.arange(S, dtype=np.float32)[:, None]
        r = data.crop_a_segment(feat, a, b, dur)
        assert (0 if r is None else len(r)) == n


def test_generate_synonyms_oracle_vs_reference(pgold):
    for c in pgold["synonyms"]:
        assert torch.equal(O.generate_synonyms(c["cap"], c["p"], c["V"], c["u1"], c["u2"], c["rw"]), c["out"])


def test_feature_loader_and_staging(tmp_path):
    """np.load + crop + missing-file substitution, then the packed pinned layout (host side only, no GPU)."""
    rng = np.random.RandomState(0)
    vdir, adir = tmp_path / "v", tmp_path / "a"
    vdir.mkdir(); adir.mkdir()
    for vid, S, Sa in (("v1", 20, 50), ("v2", 7, 19)):
        np.save(vdir / f"{vid}_rgb.npy", rng.rand(S, 16).astype(np.float32))
        np.save(vdir / f"{vid}_flow.npy", rng.rand(S, 16).astype(np.float32))
        np.save(adir / f"{vid}.npy", rng.rand(Sa, 8).astype(np.float32))
    ld = data.FeatureLoader(str(vdir), str(adir), d_video=16, d_audio=8, workers=2)
    clips = ld.submit([("v1", 10.0, 40.0, 100.0), ("v2", 0.0, 60.0, 60.0), ("missing", 0.0, 1.0, 2.0)]).result()
    s, e = O.crop_bounds(20, 10.0, 40.0, 100.0)
    assert clips[0][0].shape == (e - s, 16) and np.array_equal(clips[0][0], np.load(vdir / "v1_rgb.npy")[s:e])
    assert clips[1][2].shape == (19, 8)
    assert clips[2][0].shape == (1, 16) and not clips[2][0].any() and clips[2][2].shape == (1, 8)
    if not torch.cuda.is_available():
        return
    rb = data.RaggedBatch(3, 32, 64, 5, d_video=16, d_audio=8)
    cap = torch.arange(15).view(3, 5)
    Sv, Sa = rb.stage(clips, cap)
    assert (Sv, Sa) == (max(c[0].shape[0] for c in clips), max(c[2].shape[0] for c in clips))


@pytest.mark.gpu
def test_generate_synonyms_kernel_vs_reference(gpu_lib, pgold):
    from bmhrl_b200 import ops
    for c in pgold["synonyms"]:
        out = data.generate_synonyms(c["cap"].cuda(), c["p"], c["V"], draws=(c["u1"], c["u2"], c["rw"]))
        assert torch.equal(out.cpu(), c["out"]), c["p"]
    # Philox mode (the training path): replacement statistics of the reference's rules; rows stop at the end token
    ops.manual_seed(11)
    ops.new_rng_epoch("cuda")
    B, L, V = 4096, 20, 10172
    cap = torch.randint(4, V, (B, L))
    cap[:, 10] = 3
    cap[:, 11:] = 1
    out = data.generate_synonyms(cap.cuda(), 0.3, V).cpu()
    assert bool((out[:, 10] == 1).all()) and torch.equal(out[:, 11:], cap[:, 11:])
    body, ref = out[:, :10], cap[:, :10]
    frac_pad = float((body == 1).float().mean())
    frac_rand = float(((body != ref) & (body != 1)).float().mean())
    assert abs(frac_pad - 0.3 * 0.8) < 0.01 and abs(frac_rand - 0.3 * 0.1) < 0.005
    assert int(body.min()) >= 1 and int(body.max()) <= V - 1


@pytest.mark.gpu
def test_ragged_batch_to_device(gpu_lib):
    """ONE staged H2D copy + pad / rgb+flow / mask kernels == pad_sequence + add + make_masks of the reference path, for
    both padding conventions (reference: rgb / audio padded with the vocabulary pad index 1 -> padding stays visible)."""
    g = torch.Generator().manual_seed(4)
    B, L, dv, da = 5, 9, 1024, 128
    lens_v, lens_a = [7, 30, 1, 12, 30], [19, 80, 1, 33, 64]
    clips = []
    for i in range(B):
        rgb = torch.relu(torch.randn(lens_v[i], dv, generator=g))
        rgb[:, 0] = torch.rand(lens_v[i], generator=g) + 0.05
        if i == 3:
            rgb[2, 0] = 0.0   # a real frame whose first component is 0 is masked (model/masking.py:18-25)
        clips.append((rgb.numpy(), torch.randn(lens_v[i], dv, generator=g).numpy(),
                      torch.relu(torch.randn(lens_a[i], da, generator=g)).numpy()))
    cap = torch.randint(4, 100, (B, L), generator=g)
    for pad in (1.0, 0.0):
        rb = data.RaggedBatch(B, 40, 96, L, device="cuda", feature_pad_value=pad)
        Sv, Sa = rb.stage(clips, cap)
        out = rb.to_device()
        torch.cuda.synchronize()
        rgb_p = O.pad_batch([torch.from_numpy(c[0]) for c in clips], pad)
        flow_p = O.pad_batch([torch.from_numpy(c[1]) for c in clips], 0)
        aud_p = O.pad_batch([torch.from_numpy(c[2]) for c in clips], pad)
        masks = O.make_masks(rgb_p, aud_p, None, 1)
        assert torch.equal(out["V"].cpu(), rgb_p + flow_p)
        assert torch.equal(out["audio"].cpu(), aud_p)
        assert torch.equal(out["V_mask"].cpu(), masks["V_mask"]) and torch.equal(out["A_mask"].cpu(), masks["A_mask"])
        assert torch.equal(out["caption"].cpu(), cap)
        assert 0 <= rb.h2d_bytes - (4 * (2 * sum(lens_v) * dv + sum(lens_a) * da) + 8 * (B + 1) + 8 * B * L) < 6 * 16
        if pad == 0.0:
            assert out["V_mask"].sum(-1).flatten().tolist() == [7, 30, 1, 11, 30]


@pytest.mark.gpu
def test_step_fed_by_ragged_pipeline_equals_padded_batch(gpu_lib, agent_sd):
    """StepRunner.step_ragged (one staged H2D + pad / rgb+flow / mask kernels into the static buffers of the captured
    step) trains exactly like StepRunner.step on the pre-padded batch."""
    from bmhrl_b200 import synth
    from bmhrl_b200.model.bm_hrl_agent import BMManagerValueFunction, BMWorkerValueFunction
    from bmhrl_b200.train import BMHRLTrainer, StepRunner
    from conftest import PAD, VOC, build_agent
    b = synth.synthetic_batch(B=4, Sv=10, Sa=14, L=7, V=VOC, seed=2, ragged=True)
    lens_v = (b["rgb"][:, :, 0] != 0).sum(1).tolist()
    lens_a = (b["audio"][:, :, 0] != 0).sum(1).tolist()
    clips = [(b["rgb"][i, :lens_v[i]].numpy(), b["flow"][i, :lens_v[i]].numpy(), b["audio"][i, :lens_a[i]].numpy())
             for i in range(4)]
    res = []
    for ragged in (False, True):
        model, cfg = build_agent(agent_sd)
        tr = BMHRLTrainer(model, BMWorkerValueFunction(cfg).cuda(), BMManagerValueFunction(cfg).cuda(), pad_idx=PAD)
        tr.teach_warmstart()
        model.manager.exploration = False
        if ragged:
            rb = data.RaggedBatch(4, 16, 16, b["caption"].shape[1], device="cuda", feature_pad_value=0.0)
            rb.stage(clips, b["caption"])
            ex = rb.to_device()
            ex = {k: ex[k] for k in ("V", "audio", "V_mask", "A_mask", "caption")}
            run = StepRunner(tr, "warmstart", ex, use_graph=True, warmup=2)
            losses = []
            for _ in range(3):
                rb.stage(clips, b["caption"])
                losses.append(float(run.step_ragged(rb)[0]))
        else:
            run = StepRunner(tr, "warmstart", b, use_graph=True, warmup=2)
            losses = [float(run.step(b)[0]) for _ in range(3)]
        res.append(losses)
    for a, c in zip(*res):
        assert abs(a - c) < 1e-5 * abs(c), res
