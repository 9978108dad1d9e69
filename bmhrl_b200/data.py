"""Input pipeline of the training step.

The reference loads every sample synchronously inside `Dataset.__getitem__` (np.load x3, crop, torch pad_sequence, one
`.to(device)` per modality; captioning_datasets/load_features.py:14-99, captioning_dataset.py:246-316) and corrupts the
caption tokens in a Python loop over every token (epoch_loops/captioning_bmrl_loops.py:510-528).  Here:

  host    `FeatureLoader` reads + crops the .npy stacks on worker threads (prefetching the next batches), `RaggedBatch`
          packs the whole ragged batch -- rgb | flow | audio frames back to back, offsets, captions -- into ONE pinned
          staging buffer (double buffered);
  device  one host->device copy, then `bmhrl_pad_features_f32` (pad + rgb + flow + masks) twice and
          `bmhrl_generate_synonyms` -- all graph-capturable, on the caller's stream.

Dataset / vocabulary handling (torchtext, spaCy, GloVe) stays out of scope: captions arrive as token-id tensors.
"""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

from ._lib import LIB


def crop_bounds(S, start, end, duration):
    """Frame range of a segment, captioning_datasets/load_features.py:14-37 (`crop_a_segment`): quantile -> index with
    int() truncation, an empty range is widened by one frame (to the left at the end of the clip)."""
    start_idx = int(S * (start / duration))
    end_idx = int(S * (end / duration))
    if start_idx == end_idx:
        if start_idx == S:
            start_idx -= 1
        else:
            end_idx += 1
    return start_idx, end_idx


def crop_a_segment(feature, start, end, duration):
    """(S, D) array -> the rows of the segment, or None when nothing is left (load_features.py:33-37)."""
    s, e = crop_bounds(feature.shape[0], start, end, duration)
    out = feature[s:e]
    return None if len(out) == 0 else out


class FeatureLoader:
    """load_features_from_npy (load_features.py:49-99) + the missing-file / empty-crop substitution of
    captioning_dataset.py:276-284 (a single all-zero frame), on a thread pool with look-ahead."""

    def __init__(self, video_features_path, audio_features_path, d_video=1024, d_audio=128, workers=4):
        self.vpath, self.apath = video_features_path, audio_features_path
        self.d_video, self.d_audio = d_video, d_audio
        self.pool = ThreadPoolExecutor(max_workers=workers)

    def load_clip(self, video_id, start, end, duration):
        try:
            aud = np.load(os.path.join(self.apath, f"{video_id}.npy")).astype(np.float32, copy=False)
            aud = crop_a_segment(aud, start, end, duration)
        except FileNotFoundError:
            aud = None
        try:
            rgb = np.load(os.path.join(self.vpath, f"{video_id}_rgb.npy")).astype(np.float32, copy=False)
            flow = np.load(os.path.join(self.vpath, f"{video_id}_flow.npy")).astype(np.float32, copy=False)
            if rgb.shape != flow.shape:
                raise RuntimeError(f"{video_id}: rgb {rgb.shape} and flow {flow.shape} differ")
            rgb, flow = crop_a_segment(rgb, start, end, duration), crop_a_segment(flow, start, end, duration)
        except FileNotFoundError:
            rgb = flow = None
        if rgb is None or flow is None:
            rgb = flow = np.zeros((1, self.d_video), np.float32)
        if aud is None:
            aud = np.zeros((1, self.d_audio), np.float32)
        return rgb, flow, aud

    def submit(self, samples):
        """samples: iterable of (video_id, start, end, duration) -> future of the list of (rgb, flow, audio) arrays."""
        samples = list(samples)
        return self.pool.submit(lambda: [self.load_clip(*s) for s in samples])


class RaggedBatch:
    """Pinned staging of one ragged batch and its device-side expansion.

    Layout of the staging buffer (bytes, every section 16-byte aligned, packed to the batch's fill level):
    [rgb frames | flow frames | audio frames | v_off (B+1) i32 | a_off (B+1) i32 | captions (B, L) i64].
    `stage()` fills it on the host; `to_device()` issues ONE cudaMemcpyAsync of the used bytes and the expansion kernels
    on the current stream and returns (or fills) the padded tensors."""

    def __init__(self, B, max_video_frames, max_audio_frames, L, d_video=1024, d_audio=128, device="cuda",
                 feature_pad_value=1.0, n_buffers=2):
        self.B, self.L, self.dv, self.da = B, L, d_video, d_audio
        self.max_v, self.max_a = max_video_frames, max_audio_frames
        self.pad = float(feature_pad_value)   # the reference pads rgb / audio with the vocabulary pad index (1), flow with 0
        self.nbytes = 4 * B * (2 * max_video_frames * d_video + max_audio_frames * d_audio) + 8 * (B + 1) + 8 * B * L + 6 * 16
        self.device = torch.device(device)
        self.host = [torch.empty(self.nbytes, dtype=torch.uint8).pin_memory() for _ in range(n_buffers)]
        self.dev = torch.empty(self.nbytes, dtype=torch.uint8, device=self.device)
        self.cur = 0
        self.layout = None
        self.h2d_bytes = 0

    def stage(self, clips, captions):
        """clips: list of B (rgb, flow, audio) float32 arrays; captions (B, L) int64 tensor / array.  Returns (Sv, Sa)."""
        if len(clips) != self.B:
            raise ValueError(f"RaggedBatch was built for B={self.B}")
        self.cur = (self.cur + 1) % len(self.host)
        buf = self.host[self.cur].numpy()
        voff, aoff = [0], [0]
        for rgb, flow, aud in clips:
            voff.append(voff[-1] + rgb.shape[0])
            aoff.append(aoff[-1] + aud.shape[0])
        nv, na = voff[-1], aoff[-1]
        Sv = max(voff[i + 1] - voff[i] for i in range(self.B))
        Sa = max(aoff[i + 1] - aoff[i] for i in range(self.B))
        if Sv > self.max_v or Sa > self.max_a:
            raise ValueError(f"clip longer than the staging buffer: Sv={Sv} (max {self.max_v}), Sa={Sa} (max {self.max_a})")
        a16 = lambda n: (n + 15) & ~15
        o_rgb = 0
        o_flow = a16(o_rgb + 4 * nv * self.dv)
        o_aud = a16(o_flow + 4 * nv * self.dv)
        o_voff = a16(o_aud + 4 * na * self.da)
        o_aoff = a16(o_voff + 4 * (self.B + 1))
        o_cap = a16(o_aoff + 4 * (self.B + 1))
        used = a16(o_cap + 8 * self.B * self.L)
        f = lambda off, n: buf[off:off + 4 * n].view(np.float32)
        rgb_all = f(o_rgb, nv * self.dv).reshape(nv, self.dv)
        flow_all = f(o_flow, nv * self.dv).reshape(nv, self.dv)
        aud_all = f(o_aud, na * self.da).reshape(na, self.da)
        for i, (rgb, flow, aud) in enumerate(clips):
            rgb_all[voff[i]:voff[i + 1]] = rgb
            flow_all[voff[i]:voff[i + 1]] = flow
            aud_all[aoff[i]:aoff[i + 1]] = aud
        buf[o_voff:o_voff + 4 * (self.B + 1)].view(np.int32)[:] = voff
        buf[o_aoff:o_aoff + 4 * (self.B + 1)].view(np.int32)[:] = aoff
        cap = captions.numpy() if torch.is_tensor(captions) else np.asarray(captions)
        buf[o_cap:o_cap + 8 * self.B * self.L].view(np.int64).reshape(self.B, self.L)[:] = cap
        self.layout = (Sv, Sa, nv, na, o_rgb, o_flow, o_aud, o_voff, o_aoff, o_cap, used)
        self.h2d_bytes = used   # exactly what to_device() copies
        return Sv, Sa

    def to_device(self, out=None, synonyms_p=0.0, vocab_size=None, end_idx=3, pad_idx=1, draws=None, salt=0, rng=None):
        """ONE H2D copy + the expansion kernels.  Returns dict(V, audio, V_mask, A_mask, caption[, caption_in]); `out`
        (same keys, e.g. the static input buffers of a captured step) is filled in place when given."""
        Sv, Sa, nv, na, o_rgb, o_flow, o_aud, o_voff, o_aoff, o_cap, used = self.layout
        st = torch.cuda.current_stream(self.device).cuda_stream
        self.dev[:used].copy_(self.host[self.cur][:used], non_blocking=True)   # ONE cudaMemcpyAsync of the staged batch
        B, L = self.B, self.L
        f32 = lambda off, n: self.dev[off:off + 4 * n].view(torch.float32)
        rgb, flow, aud = f32(o_rgb, nv * self.dv), f32(o_flow, nv * self.dv), f32(o_aud, na * self.da)
        voff = self.dev[o_voff:o_voff + 4 * (B + 1)].view(torch.int32)
        aoff = self.dev[o_aoff:o_aoff + 4 * (B + 1)].view(torch.int32)
        cap = self.dev[o_cap:o_cap + 8 * B * L].view(torch.int64).view(B, L)
        if out is not None:
            V, A, vm, am = out["V"], out["audio"], out["V_mask"], out["A_mask"]
            if V.shape != (B, Sv, self.dv) or A.shape != (B, Sa, self.da) or not (V.is_contiguous() and A.is_contiguous()):
                raise ValueError(f"RaggedBatch.to_device: output buffers are {tuple(V.shape)} / {tuple(A.shape)}, the staged "
                                 f"batch needs ({B},{Sv},{self.dv}) / ({B},{Sa},{self.da})")
        else:
            V = torch.empty(B, Sv, self.dv, device=self.device)
            A = torch.empty(B, Sa, self.da, device=self.device)
            vm = torch.empty(B, 1, Sv, device=self.device, dtype=torch.bool)
            am = torch.empty(B, 1, Sa, device=self.device, dtype=torch.bool)
        LIB.call("bmhrl_pad_features_f32", rgb.data_ptr(), flow.data_ptr(), voff.data_ptr(), V.data_ptr(), vm.data_ptr(), B, Sv,
                 self.dv, self.pad, 0.0, st)
        LIB.call("bmhrl_pad_features_f32", aud.data_ptr(), None, aoff.data_ptr(), A.data_ptr(), am.data_ptr(), B, Sa, self.da,
                 self.pad, 0.0, st)
        if out is not None:
            out["caption"].copy_(cap)
            res = out
        else:
            res = {"V": V, "audio": A, "V_mask": vm, "A_mask": am, "caption": cap.clone()}
        if synonyms_p > 0:
            res["caption_in"] = generate_synonyms(res["caption"][:, :-1].contiguous(), synonyms_p, vocab_size, end_idx, pad_idx,
                                                  draws, salt, rng)
        return res


def generate_synonyms(caption_idx, p, vocab_size, end_idx=3, pad_idx=1, draws=None, salt=0, rng=None):
    """epoch_loops/captioning_bmrl_loops.py:510-528 on the device.  draws: None (Philox, the training path) or a tuple
    (u1, u2, rand_words) of (B, L) tensors -- the reference's `random.random()` / `random.randint()` results laid out per
    position -- which reproduces the reference's output exactly."""
    if not caption_idx.is_cuda or caption_idx.dtype != torch.int64:
        raise RuntimeError("generate_synonyms: caption_idx must be a CUDA int64 tensor (there is no CPU fallback)")
    from . import ops
    cap = caption_idx.contiguous()
    B, L = cap.shape
    out = torch.empty_like(cap)
    st = torch.cuda.current_stream(cap.device).cuda_stream
    if draws is not None:
        u1, u2, rw = draws
        u1 = u1.to(cap.device, torch.float32).contiguous()
        u2 = u2.to(cap.device, torch.float32).contiguous()
        rw = rw.to(cap.device, torch.int64).contiguous()
        LIB.call("bmhrl_generate_synonyms", cap.data_ptr(), out.data_ptr(), B, L, float(p), int(vocab_size), int(end_idx),
                 int(pad_idx), u1.data_ptr(), u2.data_ptr(), rw.data_ptr(), None, 0, st)
    else:
        if rng is None:
            rng = ops.RNG.current(cap.device)
        LIB.call("bmhrl_generate_synonyms", cap.data_ptr(), out.data_ptr(), B, L, float(p), int(vocab_size), int(end_idx),
                 int(pad_idx), None, None, None, rng.data_ptr(), int(salt), st)
    return out
