"""metrics/util.py:56-88 `discontinue_reward` and metrics/batched_meteor.py:19-36 `segment_reward` with the
reference's signatures; the loops run in libbmhrl_reward.so (fp32, the reference's accumulation order)."""
import torch

from ._lib import RLIB


def _f32_cpu(t):
    return torch.as_tensor(t).detach().to("cpu", torch.float32).contiguous()


def _sections_u8(sections):
    return (torch.as_tensor(sections).detach().to("cpu") != 0).to(torch.uint8).contiguous()


def discontinue_reward(cider_diff, gamma, n_step=100, segments=None):
    """Without `segments`: out[b, l] = sum_{i < n_step} gamma^i * cider_diff[b, l + i] as a new CPU fp32 tensor
    (:56-66, :90).  With `segments`: the reference's segment-end discount, written IN PLACE into `cider_diff`
    (with all its quirks, see include/bmhrl_reward.h) and returned as a clone (:67-88)."""
    x = _f32_cpu(cider_diff)
    if x.dim() != 2:
        raise ValueError("discontinue_reward expects a (B, L) tensor")
    B, L = x.shape
    if segments is None:
        out = torch.empty_like(x)
        RLIB.call("bmhrl_discount_rewards", x.data_ptr(), B, L, float(gamma), int(n_step), out.data_ptr())
        return out
    if not torch.is_tensor(cider_diff):
        raise TypeError("discontinue_reward(..., segments=...) edits cider_diff in place: pass a tensor")
    sec = _sections_u8(segments)
    if tuple(sec.shape) != (B, L):
        raise ValueError("segments must have the shape of cider_diff")
    if x.data_ptr() == cider_diff.data_ptr():
        work = x
    else:
        work = x.clone()
    RLIB.call("bmhrl_discount_segments", work.data_ptr(), sec.data_ptr(), B, L, float(gamma))
    if work.data_ptr() != cider_diff.data_ptr():
        cider_diff.copy_(work.to(cider_diff.dtype))
    return cider_diff.clone()


def segment_reward(reward, sections):
    """Returns (segment_reward (B, L) fp32 on CPU, torch.nonzero(sections)) like batched_meteor.py:19-36."""
    x = _f32_cpu(reward)
    sec = _sections_u8(sections)
    if x.dim() != 2 or tuple(sec.shape) != tuple(x.shape):
        raise ValueError("segment_reward expects reward and sections of the same (B, L) shape")
    out = torch.empty_like(x)
    RLIB.call("bmhrl_segment_reward", x.data_ptr(), sec.data_ptr(), x.shape[0], x.shape[1], out.data_ptr())
    return out, torch.nonzero(torch.as_tensor(sections))
