"""model/masking.py of the reference, kept on the host side: boolean/integer glue whose outputs are INPUTS of the
attention kernels.  Semantics are the reference's, including its quirks: feature masks
look at component 0 only (of rgb, before flow is added), so a real frame whose first component is 0 is masked."""
import torch


def subsequent_mask(size, device=None):
    """model/masking.py:3-11 -> (1, size, size) uint8 lower-triangular (built on `device`: no H2D copy, so the
    mask construction is CUDA-graph capturable)."""
    return torch.tril(torch.ones(1, size, size, device=device), 0).byte()


def c_mask(trg, pad_idx):
    """model/masking.py:13-15."""
    m = (trg != pad_idx).unsqueeze(-2)
    return m & subsequent_mask(trg.size(-1), trg.device).type_as(m)


def mask(src, trg, pad_idx, data_pad=0):
    """model/masking.py:18-25: src (B,S') -> (B,1,S'); trg -> (B,L,L) pad AND causal."""
    src_mask = (src != data_pad).unsqueeze(1)
    if trg is not None:
        trg_mask = (trg != pad_idx).unsqueeze(-2) & subsequent_mask(trg.size(-1), trg.device).type_as(src_mask.data)
        return src_mask, trg_mask
    return src_mask


def make_masks(feature_stacks, captions, modality, pad_idx):
    """model/masking.py:28-53."""
    masks = {}
    if modality == 'video':
        if captions is None:
            masks['V_mask'] = mask(feature_stacks['rgb'][:, :, 0], None, pad_idx)
        else:
            masks['V_mask'], masks['C_mask'] = mask(feature_stacks['rgb'][:, :, 0], captions, pad_idx)
    elif modality == 'audio':
        assert len(feature_stacks['audio'].shape) == 3
        if captions is None:
            masks['A_mask'] = mask(feature_stacks['audio'][:, :, 0], None, pad_idx)
        else:
            masks['A_mask'], masks['C_mask'] = mask(feature_stacks['audio'][:, :, 0], captions, pad_idx)
    elif modality == 'audio_video':
        assert len(feature_stacks['audio'].shape) == 3
        if captions is None:
            masks['A_mask'] = mask(feature_stacks['audio'][:, :, 0], None, pad_idx)
            masks['V_mask'] = mask(feature_stacks['rgb'][:, :, 0], None, pad_idx)
        else:
            masks['V_mask'], masks['C_mask'] = mask(feature_stacks['rgb'][:, :, 0], captions, pad_idx)
            masks['A_mask'] = mask(feature_stacks['audio'][:, :, 0], None, pad_idx)
    elif modality == 'subs_audio_video':
        assert len(feature_stacks['audio'].shape) == 3
        masks['V_mask'], masks['C_mask'] = mask(feature_stacks['rgb'][:, :, 0], captions, pad_idx)
        masks['A_mask'] = mask(feature_stacks['audio'][:, :, 0], None, pad_idx)
        masks['S_mask'] = mask(feature_stacks['subs'], None, pad_idx)
    return masks
