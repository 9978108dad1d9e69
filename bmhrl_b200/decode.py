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


def greedy_decoder_cached(model, feature_stacks, max_len, start_idx, end_idx, pad_idx, modality, return_last=False):
    """Same tokens and last-position log-probabilities as `greedy_decoder` (first
    stage): the bimodal encoder sees only the features and their padding masks (model/bm_hrl_agent.py:637-645), so it
    runs ONCE per clip batch instead of once per generated token, and the 364 -> V vocabulary projection + log-softmax
    is evaluated for the last caption position only (the only one the decoder reads, captioning_bmrl_loops.py:146).
    The caption side (critic, both fusion stacks, manager, goal attention) is still re-run over the prefix."""
    from . import ops
    if modality != 'audio_video':
        raise NotImplementedError("greedy_decoder_cached: the BMHRL agent is bimodal (modality 'audio_video')")
    worker = model.worker
    if worker.core.__class__ == torch.nn.Identity:
        raise NotImplementedError("greedy_decoder_cached needs the worker's vocabulary projection")
    with torch.no_grad():
        B = feature_stacks['audio'].shape[0]
        device = feature_stacks['rgb'].device
        trg = (torch.ones(B, 1) * start_idx).long().to(device)
        completeness_mask = torch.zeros(B, 1).byte().to(device)
        masks = make_masks(feature_stacks, trg, modality, pad_idx)
        V, A = model._encode_features((feature_stacks['rgb'] + feature_stacks['flow'], feature_stacks['audio']))
        Va, Av = model.bm_enc((V, A), masks)
        last = []
        while (trg.size(-1) <= max_len) and (not completeness_mask.all()):
            masks = make_masks(feature_stacks, trg, modality, pad_idx)
            C = model.emb_C(trg)
            _, segment_labels = model.critic.scores_and_labels(C, model.critic_score_threshhold)
            C = model.pos_enc_C(C)
            worker_feat = model.bm_worker_fus((C, (Av, Va)), masks)
            manager_feat = model.bm_manager_fus((C, (Av, Va)), masks)
            goals = model.manager(manager_feat, segment_labels)
            goal_completion = worker.goal_attention(goals, worker_feat, worker_feat, masks["C_mask"])
            x_last = ops.concat_last(worker_feat, goal_completion)[:, -1:].contiguous()
            logp = worker.core(x_last)[:, 0]
            if return_last:
                last.append(logp)
            next_word = logp.max(dim=-1)[1].unsqueeze(1)
            trg = torch.cat([trg, next_word], dim=-1)
            completeness_mask = completeness_mask | torch.eq(next_word, end_idx).byte()
        if return_last:
            return trg, torch.stack(last, dim=1)
        return trg
