"""Greedy caption decoding, config 1 (epoch_loops/captioning_bmrl_loops.py:127-152 with the 3-argument
`model.inference` form).

  greedy_decoder           the reference's algorithm: the whole model is re-run for every generated token
  greedy_decoder_cached    encoder once per clip batch, vocabulary projection for the last position only
  GraphedGreedyDecoder     for a stream of clips: static buffers, encoder + encoder-memory K/V once per clip, one CUDA
                           graph replay and one recurrent step of the critic per token

All three generate the same tokens (tests/test_gpu_agent.py); none has a CPU path."""
import torch

from .model.masking import c_mask, make_masks


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


def reference_stop_length(trg, end_idx):
    """Number of columns the reference's greedy loop (captioning_bmrl_loops.py:137-150) would have produced, given a
    token matrix (B, n) that starts with the start token and was generated WITHOUT the completeness test: 1 + the first
    column by which every row has emitted `end_idx`; None if some row has not emitted it yet."""
    ends = (trg[:, 1:] == end_idx)
    if ends.shape[1] == 0 or not bool(ends.any(dim=1).all()):
        return None
    first = ends.int().argmax(dim=1)  # first end position of each row (0-based within columns 1..)
    return int(first.max()) + 2


class GraphedGreedyDecoder:
    """Greedy captioning for a stream of clips (second stage): the same tokens and
    last-position log-probabilities as `greedy_decoder_cached`, with

      * the clip zero-padded into static (B, Sv, .) / (B, Sa, .) feature buffers -- `make_masks` masks the padding
        like any other all-zero frame (model/masking.py:18-25), so one set of CUDA graphs serves every clip;
      * one graph for the encoder, which also projects the encoder memory to the K / V of every caption->audio /
        caption->video attention of both fusion stacks (model/bm_hrl_agent.py:89-94) ONCE per clip instead of once
        per token (`MultiheadedAttention.memory_kv`);
      * one graph per prefix length t (captured on first use, shared memory pool): ONE step of the critic's six
        recurrent layers for the new position (the critic is causal; its states of the earlier positions are kept),
        both fusion stacks, manager, goal attention, last-position vocabulary projection, argmax written into column
        t of the static token buffer -- a token is one graph replay instead of ~350 launches from Python;
      * the completeness test of the reference loop (captioning_bmrl_loops.py:137,150: stop once every row has
        produced `end_idx`) read back every `check_every` tokens; tokens generated past the stopping point are cut
        off on the host, so the result has exactly the reference's length.

    `decoder(feature_stacks)` -> trg (B, n) int64 [, last (B, n-1, V)] for a clip batch of exactly B clips with
    Sv' <= Sv, Sa' <= Sa (the batch size is not padded: `Manager.expand_goals` couples the rows of a batch,
    model/bm_hrl_agent.py:415-429, so extra rows would change the captions of the real ones).  The model must be in eval mode with `set_inference_mode(True)`; graphs are tied to the math mode that
    was active at construction."""

    def __init__(self, model, B, Sv, Sa, max_len, start_idx, end_idx, pad_idx, modality='audio_video',
                 use_graph=True, check_every=4):
        from . import ops
        if modality != 'audio_video':
            raise NotImplementedError("GraphedGreedyDecoder: the BMHRL agent is bimodal (modality 'audio_video')")
        if model.worker.core.__class__ == torch.nn.Identity:
            raise NotImplementedError("GraphedGreedyDecoder needs the worker's vocabulary projection")
        if model.training or model.manager.exploration:
            raise RuntimeError("GraphedGreedyDecoder: call model.eval() and model.set_inference_mode(True) first")
        dev = next(model.parameters()).device
        if dev.type != 'cuda':
            raise RuntimeError("GraphedGreedyDecoder needs the model on a CUDA device (no CPU fallback)")
        self.model, self.B, self.Sv, self.Sa, self.max_len = model, B, Sv, Sa, max_len
        self.start_idx, self.end_idx, self.pad_idx, self.modality = start_idx, end_idx, pad_idx, modality
        self.use_graph, self.check_every = use_graph, max(1, int(check_every))
        self.device = dev
        self.rgb = torch.zeros(B, Sv, model.d_video, device=dev)
        self.flow = torch.zeros(B, Sv, model.d_video, device=dev)
        self.audio = torch.zeros(B, Sa, model.d_audio, device=dev)
        self.trg = torch.full((B, max_len + 1), pad_idx, dtype=torch.long, device=dev)
        self.last = torch.zeros(B, max_len, model.voc_size, device=dev)
        from .model.bm_hrl_agent import SegmentCriticDecodeState
        self.critic_state = SegmentCriticDecodeState(model.critic, B, max_len, dev)
        self.mem = None
        self.graphs = {}
        self.pool = torch.cuda.graph_pool_handle() if use_graph else None
        self.mode = ops.get_math_mode()
        self.replays = 0
        self.captures = 0

    # ---- bodies (eager or captured) ----
    def _encode_body(self):
        m = self.model
        fs = {'rgb': self.rgb, 'flow': self.flow, 'audio': self.audio}
        masks = make_masks(fs, None, self.modality, self.pad_idx)
        V, A = m._encode_features((self.rgb + self.flow, self.audio))
        Va, Av = m.bm_enc((V, A), masks)
        kv = []
        for fus in (m.bm_worker_fus, m.bm_manager_fus):
            for layer in fus.decoder.layers:
                kv.append((layer.enc_att_A, layer.enc_att_A.project_memory(Av, Av)))
                kv.append((layer.enc_att_V, layer.enc_att_V.project_memory(Va, Va)))
        self.mem = (masks, Av, Va, kv)

    def _token_body(self, t):
        from . import ops
        m = self.model
        mem_masks, Av, Va, kv = self.mem
        trg = self.trg[:, :t].contiguous()
        masks = dict(mem_masks)
        masks['C_mask'] = c_mask(trg, self.pad_idx)
        for att, pair in kv:
            att.memory_kv = pair
        try:
            C = m.emb_C(trg)
            # the critic is causal: only the new position runs through its six recurrent layers
            m.critic.step_labels(self.critic_state, C[:, t - 1:t], t - 1, m.critic_score_threshhold)
            segment_labels = self.critic_state.labels[:, :t].contiguous()
            C = m.pos_enc_C(C)
            worker_feat = m.bm_worker_fus((C, (Av, Va)), masks)
            manager_feat = m.bm_manager_fus((C, (Av, Va)), masks)
        finally:
            for att, _ in kv:
                att.memory_kv = None
        goals = m.manager(manager_feat, segment_labels)
        goal_completion = m.worker.goal_attention(goals, worker_feat, worker_feat, masks['C_mask'])
        x_last = ops.concat_last(worker_feat, goal_completion)[:, -1:].contiguous()
        logp = m.worker.core(x_last)[:, 0]
        self.last[:, t - 1].copy_(logp)
        self.trg[:, t].copy_(logp.max(dim=-1)[1])

    def _run(self, key, body):
        if not self.use_graph:
            body()
            return
        g = self.graphs.get(key)
        if g is None:
            from . import ops as _ops
            cur = torch.cuda.current_stream(self.device)
            if key != 'enc' or self.mem is None:
                # one eager pass first: lazily created salts / positional tables must exist before the capture
                side = _ops.warmup_stream(self.device)
                side.wait_stream(cur)
                with torch.cuda.stream(side):
                    if self.mem is None:
                        self._encode_body()
                    if key != 'enc':
                        body()
                cur.wait_stream(side)
                torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, pool=self.pool):
                body()
            self.graphs[key] = g
            self.captures += 1
        g.replay()
        self.replays += 1

    def __call__(self, feature_stacks, return_last=False):
        from . import ops
        if ops.get_math_mode() != self.mode:
            raise RuntimeError(f"GraphedGreedyDecoder was built in math mode {self.mode!r}")
        rgb, flow, audio = feature_stacks['rgb'], feature_stacks['flow'], feature_stacks['audio']
        Bc, Sv, Sa = rgb.shape[0], rgb.shape[1], audio.shape[1]
        if Bc != self.B or Sv > self.Sv or Sa > self.Sa or audio.shape[0] != Bc or flow.shape != rgb.shape:
            raise ValueError(f"clip batch (B={Bc}, Sv={Sv}, Sa={Sa}) does not fit the decoder's static buffers "
                             f"(B={self.B}, Sv={self.Sv}, Sa={self.Sa})")
        with torch.no_grad():
            for buf, x in ((self.rgb, rgb), (self.flow, flow), (self.audio, audio)):
                if x.shape != buf.shape:
                    buf.zero_()
                buf[:x.shape[0], :x.shape[1]].copy_(x, non_blocking=True)
            self.trg.fill_(self.pad_idx)
            self.trg[:, 0] = self.start_idx
            if self.use_graph and 'enc' not in self.graphs:
                self.mem = None
            self._run('enc', self._encode_body)
            n = 1  # columns of trg that are final
            t = 1
            while t <= self.max_len:
                self._run(t, lambda: self._token_body(t))
                t += 1
                if (t - 1) % self.check_every == 0 or t > self.max_len:
                    stop = reference_stop_length(self.trg[:, :t].cpu(), self.end_idx)
                    if stop is not None:
                        n = stop
                        break
                n = t
            trg = self.trg[:, :n].clone()
            if return_last:
                return trg, self.last[:, :n - 1].clone()
            return trg
