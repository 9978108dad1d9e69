"""Greedy caption decoding, config 1 (epoch_loops/captioning_bmrl_loops.py:127-152 with the 3-argument
`model.inference` form).  Like the reference the whole model is re-run for every generated token."""
import torch

from .model.masking import make_masks


def greedy_decoder(model, feature_stacks, max_len, start_idx, end_idx, pad_idx, modality, return_last=False):
    with torch.no_grad():
        B = feature_stacks['audio'].shape[0] if 'audio' in feature_stacks else feature_stacks['rgb'].shape[0]
        device = feature_stacks['rgb'].device
        trg = (torch.ones(B, 1) * start_idx).long().to(device)
        completeness_mask = torch.zeros(B, 1).byte().to(device)
        V = feature_stacks['rgb'] + feature_stacks['flow']
        last = []
        while (trg.size(-1) <= max_len) and (not completeness_mask.all()):
            masks = make_masks(feature_stacks, trg, modality, pad_idx)
            preds = model.inference((V, feature_stacks['audio']), trg, masks)
            if return_last:
                last.append(preds[:, -1])
            next_word = preds[:, -1].max(dim=-1)[1].unsqueeze(1)
            trg = torch.cat([trg, next_word], dim=-1)
            completeness_mask = completeness_mask | torch.eq(next_word, end_idx).byte()
        if return_last:
            return trg, torch.stack(last, dim=1)
        return trg


bmhrl_greedy_decoder = greedy_decoder  # the reference's name for the BMHRL variant (captioning_bmrl_loops.py:127)
