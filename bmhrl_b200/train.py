"""Training steps of the BMHRL hot path on one B200 (per data-parallel rank).

Restates the bodies of the reference's epoch loops as fixed-shape device programs:
  * warm-start step  epoch_loops/captioning_bmrl_loops.py:1148-1160  (fwd + LabelSmoothing + bwd + Adam)
  * RL step          epoch_loops/captioning_bmrl_loops.py:837-877    (fwd + value fn + biased_kl() + BiasedKL +
                     bwd + Adam + value-function MSE bwd + Adam), worker and manager phases (F4-free).
The timed step takes the (B, L) score tensor as an input (synthetic rewards in the timed
runs).  With `scorer=` (bmhrl_b200.metrics.CiderScorer, the host C++ restatement of metrics/cider.py) and
batch["trg_caption"] (list of strings) the eager step scores the sampled tokens on the host between sampling and the
loss like the reference does; that variant has a host round trip and is not graph-captured.

Every step can be captured into ONE CUDA graph (`StepRunner.capture`): the library's entry points never
synchronise or allocate, dropout/sampling counters and Adam step counts advance on the device.  No CPU fallback.
"""
import os

import torch

from . import ops, ops16
from .loss.biased_kl import BiasedKL
from .loss.label_smoothing import LabelSmoothing
from .model.masking import make_masks

# parameters the reference constructs but never calls: they never receive a gradient
_NEVER_USED = (".feed_forward.fc", "manager_core.", "manager.core.")


def used_parameters(model):
    """model.parameters() minus the never-called ones, in registration order (named_parameters de-duplicates the
    module that is registered twice, manager_core == manager.core)."""
    return [p for n, p in model.named_parameters()
            if not (("_fus." in n and ".feed_forward.fc" in n) or n.startswith("manager_core.")
                    or n.startswith("manager.core."))]


def get_score(train_worker, scorer, predicted_tokens, caption, mask, segments):
    """captioning_bmrl_loops.py:219-230 for the CIDEr and BLEU scorers (METEOR needs nltk + WordNet: not built)."""
    if scorer.type == "CIDER":
        if train_worker:
            return scorer.delta_cider_worker(predicted_tokens, caption)
        return scorer.delta_cider_manager(predicted_tokens, caption, mask, segments)
    if scorer.type == "BLEU":
        if train_worker:
            return scorer.delta_bleu_worker(predicted_tokens, caption)
        return scorer.delta_bleu_manager(predicted_tokens, caption, mask, segments)
    raise NotImplementedError(f"scorer type {scorer.type!r}: the CIDEr and BLEU reward scorers are implemented")


class BMHRLTrainer:
    """Holds the agent, the two value functions, their fused Adam optimisers (torch.optim.Adam defaults,
    scripts/train_rl_captioning_module.py:81-83) and the criteria (:66-67, :132)."""

    def __init__(self, cap_model, wv_model, mv_model, pad_idx=1, smoothing=0.7, lr=1e-4, modality="audio_video",
                 grad_sync=None, scorer=None, random_synonyms=0.0, end_idx=3, peer=None, owner_only_master=False):
        # feature_getter(..., random_synonyms=0.3) (captioning_bmrl_loops.py:487-528): the reference corrupts the INPUT
        # caption tokens of every training step; 0 (default) switches it off, which is what the parity goldens assume
        self.random_synonyms, self.end_idx = float(random_synonyms), int(end_idx)
        self.scorer = scorer  # bmhrl_b200.metrics.CiderScorer; None = the step takes batch["score"] (synthetic rewards)
        self.cap_model, self.wv_model, self.mv_model = cap_model, wv_model, mv_model
        self.pad_idx, self.modality = pad_idx, modality
        self.ls = LabelSmoothing(smoothing, pad_idx)
        self.bkl = BiasedKL(smoothing, pad_idx)
        # peer: a peer.PeerGroup -- the agent's gradients are then summed over the ranks, Adam-stepped and the parameters
        # broadcast by ONE kernel over NVLink peer memory (ops.FusedAdam.step_peer) instead of NCCL all-reduce + a
        # replicated Adam; grad_sync still carries the scalar loss normaliser and the (tiny) value functions
        self.peer = peer
        self.cap_opt = ops.FusedAdam(used_parameters(cap_model), lr=lr, peer=peer)
        # owner_only_master (peer mode, bf16 math mode only): fp32 masters of GEMM weights stay on the rank that owns their
        # Adam state; the replicas compute with the broadcast bf16 shadows.  Call `cap_opt.gather_params()` before reading
        # fp32 parameters on other ranks (state_dict / checkpoints / leaving bf16 mode).
        self.cap_opt.owner_only_master = bool(owner_only_master) and peer is not None
        if peer is not None:
            n_el = self.cap_opt.flat_grad.numel()
            ops16.SHADOWS.arena = peer.arena(peer.SHADOW, int(2 * n_el * 1.15) + (32 << 20))
            peer.share()
        self.wv_opt = ops.FusedAdam(wv_model.parameters(), lr=lr)
        self.mv_opt = ops.FusedAdam(mv_model.parameters(), lr=lr)
        self.grad_sync = grad_sync  # data-parallel hook: callable(optimizer) all-reducing optimizer.flat_grad
        self.loss_factor = 4.0 / 20.0  # captioning_bmrl_loops.py:831-833
        # bf16 math mode: vocabulary projection + log-softmax + KL loss as the fused kernel pair (no (B*L,V) tensors)
        self.fuse_vocab = True
        # Overlap of the gradient all-reduce (and of Adam) with backward.  Autograd runs ready nodes in reverse creation
        # order, so when a marker placed on the output of a module fires in backward, the backward kernels of everything
        # created after it have been enqueued.  Stage 0 = everything after the encoder (fusion stacks, heads), fired by
        # the marker on the encoder outputs; stage k = encoder layer N-k, fired by the marker on the output of layer
        # N-k-1.  A fired stage is reduced and Adam-stepped on a communication stream while backward continues; what is
        # left (encoder layer 0) follows the final join.
        ids = {id(p): n for n, p in cap_model.named_parameters()}
        pnames = [ids[id(p)] for p in self.cap_opt.params]
        n_enc = len(cap_model.bm_enc.encoder.layers) if hasattr(cap_model, "bm_enc") else 0
        # stage 0: everything behind the encoder.  Then, per encoder layer from the last to the first, TWO stages: the
        # bimodal-attention + FFN blocks (complete when the marker between them and the self-attention blocks fires) and
        # the self-attention blocks (complete at the layer boundary).  What is left for the tail after the final join is
        # the self-attention of encoder layer 0 only (4.7 M of the 47 M parameters).
        self._stages = [frozenset(i for i, n in enumerate(pnames) if not n.startswith("bm_enc."))]
        self._stage_of_mid, self._stage_of_boundary = {}, {}
        is_self = lambda n: (".self_att_" in n) or (".res_layers_M1.0." in n) or (".res_layers_M2.0." in n)
        for k in range(n_enc - 1, -1, -1):
            pre = f"bm_enc.encoder.layers.{k}."
            self._stage_of_mid[k] = len(self._stages)
            self._stages.append(frozenset(i for i, n in enumerate(pnames) if n.startswith(pre) and not is_self(n)))
            if k > 0:
                self._stage_of_boundary[k] = len(self._stages)
                self._stages.append(frozenset(i for i, n in enumerate(pnames) if n.startswith(pre) and is_self(n)))
        self._fired = []          # stages already handed to the communication stream in this backward
        self._comm_stream = None
        self._adam_stream = None
        self.early_adam_ctas = int(os.environ.get("BMHRL_ADAM_CTAS", "296"))   # 0 = unlimited (one CTA per chunk)
        self.overlap_grad_sync = True
        # also on ONE GPU: the Adam update of a finished stage (1.3 GB of optimiser traffic per step in total) then runs
        # under the rest of backward instead of after it
        if n_enc > 0:
            cap_model.grad_ready_callback = lambda: self._early_sync(0)
            cap_model.bm_enc.encoder.layer_boundary_hook = self._mark_layer_boundary
            for k, layer in enumerate(cap_model.bm_enc.encoder.layers):
                layer.mid_hook = (lambda x, k=k: self._mark_mid(k, x))

    # ---- phase switching (captioning_bmrl_loops.py:809-820, :1144) ------------------------------------------------
    def teach_warmstart(self):
        self.cap_model.train()
        self.cap_model.teach_warmstart()
        self.cap_opt.refresh_active()

    def teach_worker(self):
        self.cap_model.train()
        self.wv_model.train()
        self.cap_model.teach_worker()
        self.cap_opt.refresh_active()

    def teach_manager(self):
        self.cap_model.train()
        self.mv_model.train()
        self.cap_model.teach_manager()
        self.cap_opt.refresh_active()

    # ---- pieces ------------------------------------------------------------------------------------------------------
    def _inputs(self, batch):
        """feature_getter (captioning_bmrl_loops.py:488-500): input / target captions, `generate_synonyms` token
        corruption of the input (device kernel, Philox), V = rgb + flow (fused into the positional-encoding kernel), masks.
        A batch expanded by data.RaggedBatch.to_device already carries V and the feature masks."""
        from .model.masking import c_mask
        cap = batch["caption"]
        cap_in, y = cap[:, :-1], cap[:, 1:]
        if "caption_in" in batch:
            cap_in = batch["caption_in"]
        elif self.random_synonyms > 0:
            from . import data
            cap_in = data.generate_synonyms(cap_in.contiguous(), self.random_synonyms, self.cap_model.voc_size, self.end_idx,
                                            self.pad_idx, salt=self._syn_salt())
        if "V" in batch:
            masks = {"V_mask": batch["V_mask"], "A_mask": batch["A_mask"], "C_mask": c_mask(cap_in, self.pad_idx)}
            return (batch["V"], batch["audio"]), cap_in, y.contiguous(), masks
        y = y.contiguous()
        if self.modality == "audio_video" and batch["rgb"].is_cuda and batch["rgb"].dtype == torch.float32:
            # the three masks and the token count of the loss normaliser in one launch (ops.make_masks_fused)
            masks, n_local = ops.make_masks_fused(batch["rgb"], batch["audio"], cap_in, self.pad_idx, targets=y)
            self._n_tokens_local = (y, n_local)
        else:
            masks = make_masks({"rgb": batch["rgb"], "flow": batch["flow"], "audio": batch["audio"]}, cap_in,
                               self.modality, self.pad_idx)
        return ((batch["rgb"], batch["flow"]), batch["audio"]), cap_in, y, masks

    def _syn_salt(self):
        if not hasattr(self, "_synsalt"):
            self._synsalt = ops.next_salt()
        return self._synsalt

    def _forward(self, x, cap_in, masks):
        """cap_model(x, cap_in, masks) with the worker handing back its vocabulary head un-evaluated (bf16 math mode)."""
        w = getattr(self.cap_model, "worker", None)
        fuse = self.fuse_vocab and ops.BF16 and w is not None
        if fuse:
            w.fused_vocab = True
        try:
            return self.cap_model(x, cap_in, masks)
        finally:
            if fuse:
                w.fused_vocab = False

    def _mark_layer_boundary(self, i, x):
        """LayerStack hook of the encoder: output (M1, M2) of layer i -> marked copies; the marker fires in backward
        when layer i+1 (its self-attention blocks last) is done."""
        stage = self._stage_of_boundary[i + 1]
        cb = lambda: self._early_sync(stage)
        return tuple(ops.grad_marker(t, cb) for t in x)

    def _mark_mid(self, k, x):
        """BMEncoderLayer.mid_hook of layer k: fires in backward when the bimodal + FFN blocks of layer k are done."""
        stage = self._stage_of_mid[k]
        cb = lambda: self._early_sync(stage)
        return tuple(ops.grad_marker(t, cb) for t in x)

    def _early_sync(self, stage):
        """Backward-time callback (ops.grad_marker): all-reduce + Adam of one finished stage on the communication
        stream, behind everything enqueued so far on this stream and on the side streams."""
        if not self.overlap_grad_sync or stage in self._fired or stage >= len(self._stages):
            return
        if not (int(os.environ.get("BMHRL_EARLY_STAGES", "255")) >> stage) & 1:   # experiment switch
            return
        if not any(self.cap_opt.params[i].requires_grad for i in self._stages[stage]):
            return   # nothing of this stage is trained in the current phase
        self._fired.append(stage)
        dev = self.cap_opt.flat_grad.device
        if self._comm_stream is None:
            self._comm_stream = torch.cuda.Stream(device=dev)
        comm = self._comm_stream
        comm.wait_stream(torch.cuda.current_stream(dev))
        ops.wait_dirty_side_streams(comm)
        # an update that runs under the rest of backward gets a bounded grid (ops.adam_grid_limit): see there
        lim = ops.adam_grid_limit(self.early_adam_ctas)
        if self.peer is not None:
            with torch.cuda.stream(comm), lim:
                self.cap_opt.step_peer(stage, only=self._stages[stage])
            return
        if self.grad_sync is None:
            with torch.cuda.stream(comm), lim:
                self.cap_opt.step(only=self._stages[stage])
            return
        # data parallel: the all-reduces of successive stages queue on the communication stream, the Adam updates on a
        # stream of their own behind an event each -- the all-reduce of stage k + 1 (NVLink-bound) then runs beside the
        # Adam update of stage k (HBM-bound) instead of behind it
        if self._adam_stream is None:
            self._adam_stream = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(comm):
            self.grad_sync(self.cap_opt, only=self._stages[stage])
        self._adam_stream.wait_stream(comm)
        with torch.cuda.stream(self._adam_stream), lim:
            self.cap_opt.step(only=self._stages[stage])

    def _sync_and_step(self, opt, mean=False):
        """All-reduce (data parallel) + Adam of whatever has not been handled from a backward marker."""
        if opt is self.cap_opt and self.peer is not None:
            # peer-memory data parallel: what no backward marker has handled yet, then the end-of-step barrier (nobody
            # zeroes gradients or reads parameters of the next step while a peer still reads / writes this step's)
            cur = torch.cuda.current_stream(opt.flat_grad.device)
            rest = None
            if self._fired:
                done = frozenset().union(*(self._stages[k] for k in self._fired))
                rest = frozenset(range(len(opt.params))) - done
                self._fired = []
            opt.step_peer(len(self._stages), only=rest)
            if self._comm_stream is not None:
                cur.wait_stream(self._comm_stream)
            self.peer.barrier(self.peer.n_slots - 1)
            return
        if opt is self.cap_opt and self._fired:
            done = frozenset().union(*(self._stages[k] for k in self._fired))
            rest = frozenset(range(len(opt.params))) - done
            self._fired = []
            if self.grad_sync is not None:
                self.grad_sync(opt, mean=mean, only=rest)
            opt.step(only=rest)
            torch.cuda.current_stream(opt.flat_grad.device).wait_stream(self._comm_stream)
            if self._adam_stream is not None:
                torch.cuda.current_stream(opt.flat_grad.device).wait_stream(self._adam_stream)
            return
        if self.grad_sync is not None:
            self.grad_sync(opt, mean=mean)
        opt.step()

    def _global_tokens(self, n_tokens):
        """The reference computes the loss once over the gathered global batch (DataParallel), so the normaliser is
        the GLOBAL token count; ranks contribute sum/global_n and gradients are summed."""
        if self.grad_sync is not None and hasattr(self.grad_sync, "all_reduce_scalar"):
            return self.grad_sync.all_reduce_scalar(n_tokens)
        return n_tokens

    def _global_tokens_early(self, y):
        """The token count depends on the captions only: with more than one rank its all-reduce is issued on the
        communication stream BEFORE the forward pass (the ranks have just left the previous step's gradient
        all-reduce together), instead of as a rendezvous in the middle of the step.  Returns a thunk that makes the
        current stream wait for it and yields the global count."""
        pre = getattr(self, "_n_tokens_local", None)
        self._n_tokens_local = None
        local = pre[1] if (pre is not None and pre[0] is y) else (y != self.pad_idx).sum()
        world = getattr(self.grad_sync, "world", 1) if self.grad_sync is not None else 1
        if world <= 1 or not self.overlap_grad_sync or not local.is_cuda:
            return lambda: self._global_tokens(local)
        dev = local.device
        if self._comm_stream is None:
            self._comm_stream = torch.cuda.Stream(device=dev)
        comm = self._comm_stream
        comm.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(comm):
            total = self.grad_sync.all_reduce_scalar(local)
        done = torch.cuda.Event()
        done.record(comm)

        def get():
            torch.cuda.current_stream(dev).wait_event(done)
            return total
        return get

    # ---- warm-start step -------------------------------------------------------------------------------------------
    def warmstart_step(self, batch):
        """captioning_bmrl_loops.py:1148-1160.  Returns the scalar loss tensor (device)."""
        ops.new_rng_epoch(batch["caption"].device)
        zero_done = self._zero_grads_beside_forward([self.cap_opt])
        x, cap_in, y, masks = self._inputs(batch)
        n_tokens = self._global_tokens_early(y)
        pred = self._forward(x, cap_in, masks)[0]
        n_tokens = n_tokens()
        zero_done()
        if isinstance(pred, ops16.VocabHead):
            stats, _ = pred.stats(y, self.pad_idx)
            rows = ops16.kl_rowsum_from_stats(stats, y, self.ls.smoothing, self.pad_idx, pred.shape[-1])
        else:
            rows = self.ls.row_sums(pred, y)
        loss = rows.sum() / n_tokens
        loss.backward()
        ops.join_side_streams(loss.device)
        self._sync_and_step(self.cap_opt)
        return loss.detach()

    # ---- RL step -----------------------------------------------------------------------------------------------------
    def rl_forward(self, batch, train_worker, exp_noise=None):
        """First half of the RL step, up to the sampled tokens (:843-858, biased_kl() :271-285): everything the host
        reward scorer needs.  Returns the state `rl_tail` continues from."""
        x, cap_in, y, masks = self._inputs(batch)
        n_tokens = self._global_tokens_early(y)
        pred, worker_feat, manager_feat, goal_feat, seg = self._forward(x, cap_in, masks)
        loss_mask = y != self.pad_idx
        n_tokens = n_tokens()
        if train_worker:
            ev = self.wv_model((worker_feat.detach(), goal_feat.detach())).squeeze(-1)
            mode = "noise" if exp_noise is not None else "philox"
        else:
            ev = self.mv_model(manager_feat.detach()).squeeze(-1)
            mode = "argmax"
        stats = None
        if isinstance(pred, ops16.VocabHead):  # sampling inside the fused vocabulary kernel, p(token) from the statistics
            stats, tok = pred.stats(y, self.pad_idx, mode, exp_noise, salt=self._sample_salt())
            sp = torch.exp(stats[:, 4] - stats[:, 0]).view(tok.shape)
        else:
            tok, sp = ops.sample_tokens(pred, mode, exp_noise, salt=self._sample_salt())
        return {"pred": pred, "y": y, "loss_mask": loss_mask, "n_tokens": n_tokens, "ev": ev, "seg": seg, "tok": tok,
                "sp": sp, "stats": stats}

    def rl_tail(self, st, score, seg, train_worker, stabilize=False):
        """Second half (biased_kl() :287-334, value function :873-875) given the (B, L) scores and the segment labels
        the scorer left behind.  Returns (cap_loss, value_loss, tokens, score, divergence row sums)."""
        pred, y, loss_mask, n_tokens, ev, tok, sp = (st[k] for k in ("pred", "y", "loss_mask", "n_tokens", "ev", "tok",
                                                                      "sp"))
        fdt = torch.float32
        score = score.to(fdt)
        lm = loss_mask.to(fdt)
        expected = ev.detach()
        if train_worker:
            norm = lm.sum(-1, keepdim=True)
        else:
            segf = seg.to(fdt)
            norm = segf.sum(-1, keepdim=True)
            score = score * segf
            sp, expected = ops.segment_scan(sp, expected, seg)
        if stabilize:
            score = (score - expected) * lm
        amplitude = torch.clamp(score * sp * norm, 0, 1)
        if st.get("stats") is not None:
            div = ops16.kl_rowsum_from_stats(st["stats"], y, self.bkl.ls, self.pad_idx, pred.shape[-1], tok, amplitude)
        else:
            div = self.bkl.row_sums(pred, y, tok, amplitude)
        cap_loss = div.sum() / (n_tokens * self.loss_factor)
        # value function :873-875; the reference's host loop wrote the segment sums INTO the value output (aliasing)
        ev_used = ev if train_worker else ev + (expected - ev).detach()
        vmask = lm if train_worker else seg.to(fdt)
        value_loss = (((ev_used - score.detach()) ** 2) * vmask).mean()
        return cap_loss, value_loss, tok, score.detach(), div.detach()

    def host_score(self, tok, seg, loss_mask, trg_caption, train_worker):
        """get_score() :219-230, :287 on the host (libbmhrl_reward.so): D2H of the (B, L) tokens (and of the segment
        labels for the manager), H2D of the scores.  The reference's manager scorer closes `segments` in place at the
        reference caption length (metrics/cider.py:74-77) and every later use sees that edit; it is done on a copy
        here because the forward pass has saved `seg` for backward.  Returns (score, seg) on tok's device."""
        seg_h = seg.detach().to("cpu")
        score = get_score(train_worker, self.scorer, tok.detach().to("cpu"), trg_caption, loss_mask, seg_h)[0]
        return score.to(tok.device), seg_h.to(seg.device)

    def rl_losses(self, batch, train_worker, stabilize=False, exp_noise=None):
        """Forward part of train_bimodal_bl (:843-862, :873-875): returns (cap_loss, value_loss, tokens, score, div).
        With a scorer and batch["trg_caption"] the sampled tokens are scored on the host (not graph-capturable as one
        piece: `ScoredStepRunner` splits the step around the host call)."""
        st = self.rl_forward(batch, train_worker, exp_noise)
        if self.scorer is not None and "trg_caption" in batch:
            score, seg = self.host_score(st["tok"], st["seg"], st["loss_mask"], batch["trg_caption"], train_worker)
        else:
            score, seg = batch["score"], st["seg"]
        return self.rl_tail(st, score, seg, train_worker, stabilize)

    def _sample_salt(self):
        if not hasattr(self, "_ssalt"):
            self._ssalt = ops.next_salt()
        return self._ssalt

    def rl_step(self, batch, train_worker, stabilize=False, exp_noise=None):
        """captioning_bmrl_loops.py:837-877.  Returns (cap_loss, value_loss) device scalars."""
        self.rl_begin(batch, train_worker)
        cap_loss, value_loss, _, _, _ = self.rl_losses(batch, train_worker, stabilize, exp_noise)
        return self.rl_finish(cap_loss, value_loss, train_worker)

    def _zero_grads_beside_forward(self, opts):
        """The flat gradient buffers (188 MB for the agent) are first touched by backward: their memset runs on a side
        stream next to the forward pass instead of in front of it.  Returns the join to call before backward."""
        dev = opts[0].flat_grad.device
        if not ops.CONCURRENT:
            for o in opts:
                o.zero_grad()
            return lambda: None
        cur = torch.cuda.current_stream(dev)
        zs = ops.side_stream(dev, 13)
        zs.wait_stream(cur)
        with torch.cuda.stream(zs):
            for o in opts:
                o.zero_grad()
        return lambda: torch.cuda.current_stream(dev).wait_stream(zs)

    def rl_begin(self, batch, train_worker):
        ops.new_rng_epoch(batch["caption"].device)
        self._zero_join = self._zero_grads_beside_forward([self.cap_opt, self.wv_opt if train_worker else self.mv_opt])

    def rl_finish(self, cap_loss, value_loss, train_worker):
        vopt = self.wv_opt if train_worker else self.mv_opt
        if getattr(self, "_zero_join", None) is not None:
            self._zero_join()
            self._zero_join = None
        cap_loss.backward()
        ops.join_side_streams(cap_loss.device)
        self._sync_and_step(self.cap_opt)
        value_loss.backward()
        ops.join_side_streams(value_loss.device)
        self._sync_and_step(vopt, mean=True)  # the value loss is a .mean() over the global (B,L) batch (:874-875)
        return cap_loss.detach(), value_loss.detach()


class PendingLoss:
    """A step's loss on its way to the host (StepRunner.step_async): at most four may be outstanding."""

    def __init__(self, buf, event):
        self.buf, self.event = buf, event

    def value(self):
        self.event.synchronize()
        return float(self.buf)


class StepRunner:
    """One training step as a CUDA graph over static input buffers.

    `step(host_batch)` copies the pinned host batch into the static device buffers (H2D), replays the graph and
    returns the device loss tensors; `replay()` re-runs the step on whatever the buffers hold."""

    def __init__(self, trainer, kind, example_batch, stabilize=False, use_graph=True, warmup=3, pipeline_critic=False):
        """pipeline_critic: the frozen SegmentCritic depends on nothing but the caption tokens, and at B = 16 its six
        recurrent layers are the longest dependency chain of the forward pass.  With this switch a step evaluates the
        critic (and the `generate_synonyms` corruption it sees) for the NEXT batch on a side stream, underneath its own
        forward / backward, and uses the labels the previous step left behind: same arithmetic, off the critical path.
        `step(batch, next_batch)` then needs the next batch's captions (a data loader has them: it prefetches)."""
        self.trainer, self.kind, self.stabilize = trainer, kind, stabilize
        dev = next(trainer.cap_model.parameters()).device
        self.static = {k: torch.empty_like(v, device=dev).copy_(v) for k, v in example_batch.items()
                       if k in ("rgb", "flow", "audio", "caption", "score", "V", "V_mask", "A_mask")}
        self.graph = None
        self.out = None
        self.pipeline_critic = bool(pipeline_critic) and hasattr(trainer.cap_model, "segment_labels_of")
        self.critic_units_per_cta = int(os.environ.get("BMHRL_CRITIC_UPB", "16"))
        if self.pipeline_critic:
            self.static["caption_next"] = self.static["caption"].clone()
            with torch.no_grad():
                self.cap_in_cur = self._corrupt(self.static["caption"]).clone()
                self.labels_cur = trainer.cap_model.segment_labels_of(self.cap_in_cur).clone()
            self.cap_in_next = self.cap_in_cur.clone()
            self.labels_next = self.labels_cur.clone()
            torch.cuda.synchronize(dev)
        self._need_prime = False
        self.prefetch_inputs = True   # step(batch, next_batch): copy the next batch's inputs under this step (see load)
        self._staged = None
        if use_graph:
            side = ops.warmup_stream(dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                for _ in range(warmup):  # lazily-created salts / workspaces / cuBLAS-free paths settle here
                    self._eager()
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            from ._lib import launch_count, reset_launch_count, LIB
            # node_priority (experiment, BMHRL_NODE_PRIORITY=1): the chain kernels of the step are captured with launch
            # priority -1, the weight / bias / affine gradient launches (nothing downstream waits for them) with 0, and the
            # graph is instantiated with cudaGraphInstantiateFlagUseNodePriority through cuda.bindings
            self.node_priority = os.environ.get("BMHRL_NODE_PRIORITY", "0") == "1"
            self.graph = torch.cuda.CUDAGraph(keep_graph=True) if self.node_priority else torch.cuda.CUDAGraph()
            n0 = launch_count()
            if self.node_priority:
                ops.NODE_PRIORITY = True
                ops._PRIO[0] = -1
                LIB.call("bmhrl_set_launch_priority", 1, -1)
            try:
                with torch.cuda.graph(self.graph):
                    self.out = self._eager()
            finally:
                if self.node_priority:
                    ops.NODE_PRIORITY = False
                    LIB.call("bmhrl_set_launch_priority", 0, 0)
            self.captured_launches = launch_count() - n0  # kernels of libbmhrl_b200 inside one graph replay
            self._exec = None
            if self.node_priority:
                from cuda.bindings import runtime as rt
                raw = self.graph.raw_cuda_graph()
                err, ex = rt.cudaGraphInstantiateWithFlags(rt.cudaGraph_t(int(raw)),
                                                           int(rt.cudaGraphInstantiateFlags.cudaGraphInstantiateFlagUseNodePriority))
                if int(err) != 0:
                    raise RuntimeError(f"cudaGraphInstantiateWithFlags(UseNodePriority) failed: {err}")
                self._exec, self._rt = ex, rt
        self._need_prime = self.pipeline_critic

    def _corrupt(self, caption, rng=None):
        """Input tokens the model (and the critic) see: caption[:, :-1] after the reference's generate_synonyms."""
        t = self.trainer
        cap_in = caption[:, :-1].contiguous()
        if t.random_synonyms > 0:
            from . import data
            cap_in = data.generate_synonyms(cap_in, t.random_synonyms, t.cap_model.voc_size, t.end_idx, t.pad_idx,
                                            salt=t._syn_salt(), rng=rng)
        return cap_in

    def _eager(self):
        t = self.trainer
        if not self.pipeline_critic:
            if self.kind == "warmstart":
                return (t.warmstart_step(self.static),)
            return t.rl_step(self.static, self.kind == "worker", self.stabilize)
        dev = self.static["caption"].device
        cur = torch.cuda.current_stream(dev)
        # the Philox state the next batch's token corruption draws from: a copy of the LIVE device state, taken (inside
        # the graph, on this stream) before the step advances it.  Reading ops.RNG.current() on the side branch instead
        # would capture the address of a snapshot made outside the graph -- the same draws in every replay.
        if getattr(self, "rng_next", None) is None:
            self.rng_next = ops.RNG.get_state(dev).clone()
        self.rng_next.copy_(ops.RNG.get_state(dev))
        side = ops.side_stream(dev, 12)
        side.wait_stream(cur)
        with torch.cuda.stream(side), torch.no_grad(), ops.rnn_units_per_cta(self.critic_units_per_cta):
            # the NEXT batch's corruption + critic, beside this step, confined to a quarter of the SMs
            nxt = self._corrupt(self.static["caption_next"], rng=self.rng_next)
            self.cap_in_next.copy_(nxt)
            self.labels_next.copy_(t.cap_model.segment_labels_of(self.cap_in_next))
        batch = {k: v for k, v in self.static.items() if k != "caption_next"}
        batch["caption_in"] = self.cap_in_cur
        t.cap_model.precomputed_segment_labels = self.labels_cur
        try:
            if self.kind == "warmstart":
                out = (t.warmstart_step(batch),)
            else:
                out = t.rl_step(batch, self.kind == "worker", self.stabilize)
        finally:
            t.cap_model.precomputed_segment_labels = None
        cur.wait_stream(side)
        self.cap_in_cur.copy_(self.cap_in_next)
        self.labels_cur.copy_(self.labels_next)
        return out

    def load(self, host_batch, non_blocking=True, next_host_batch=None):
        """Inputs of this step into the static buffers.  When the step is given its successor (`next_host_batch`), that
        batch's host->device copy is issued on a copy stream into a staging set NOW, so it runs under this step's kernels;
        the next `load` of that same batch object is then a device-to-device copy (a few microseconds) instead of 4.6 MB
        over the host link in front of the graph.  The caller must not overwrite the pinned buffers of a batch it has
        handed over as `next_host_batch` before the step that consumes it has been issued (the usual non_blocking rule)."""
        dev = self.static["caption"].device
        cur = torch.cuda.current_stream(dev)
        st = getattr(self, "_staged", None)
        if st is not None and st["src"] is host_batch:
            cur.wait_event(st["ready"])
            for k, buf in self.static.items():
                if k != "caption_next":
                    buf.copy_(st["bufs"][k], non_blocking=True)
            st["free"] = torch.cuda.Event()
            st["free"].record(cur)
        else:
            for k, buf in self.static.items():
                if k != "caption_next":
                    buf.copy_(host_batch[k], non_blocking=non_blocking)
        if "caption_next" in self.static:
            src = (next_host_batch if next_host_batch is not None else host_batch)["caption"]
            self.static["caption_next"].copy_(src, non_blocking=non_blocking)
        if next_host_batch is not None and non_blocking and self.prefetch_inputs:
            if st is None:
                st = self._staged = {"bufs": {k: torch.empty_like(v) for k, v in self.static.items() if k != "caption_next"},
                                     "stream": torch.cuda.Stream(device=dev), "free": None}
            cs = st["stream"]
            if st["free"] is not None:
                cs.wait_event(st["free"])      # the previous device-to-device copy out of the staging set is done
            with torch.cuda.stream(cs):
                for k, buf in st["bufs"].items():
                    buf.copy_(next_host_batch[k], non_blocking=True)
                st["ready"] = torch.cuda.Event()
                st["ready"].record(cs)
            st["src"] = next_host_batch
        elif st is not None:
            st["src"] = None

    def replay(self):
        if self.graph is not None:
            if getattr(self, "_exec", None) is not None:
                err, = self._rt.cudaGraphLaunch(self._exec, torch.cuda.current_stream().cuda_stream)
                if int(err) != 0:
                    raise RuntimeError(f"cudaGraphLaunch failed: {err}")
            else:
                self.graph.replay()
        else:
            self.out = self._eager()
        return self.out

    def step(self, host_batch, next_host_batch=None):
        """One step on `host_batch`.  With pipeline_critic the captions of the batch that FOLLOWS must come along
        (`next_host_batch`; None = the same batch again): this step labels them for the next one."""
        self.load(host_batch, next_host_batch=next_host_batch)
        if self.pipeline_critic and (self._need_prime or next_host_batch is None):
            # first step of a stream, or no look-ahead given: label THIS batch now (the step then labels whatever
            # caption_next holds -- this batch again -- for a possible repeat)
            self.prime()
        return self.replay()

    def step_async(self, host_batch, next_host_batch=None):
        """`step`, plus the device-to-host copy of the step's first result (the loss) into pinned memory, stream-ordered
        behind the step: returns a `PendingLoss` whose `.value()` waits for exactly that copy.  A training loop reads the
        loss of step i AFTER it has issued step i + 1, so the host never stalls the GPU between steps (with `float(loss)`
        right after `step` the GPU idles while the host prepares and launches the next replay: ~0.1 ms per step here)."""
        out = self.step(host_batch, next_host_batch)
        ring = getattr(self, "_loss_ring", None)
        if ring is None:
            ring = self._loss_ring = [torch.empty((), dtype=out[0].dtype).pin_memory() for _ in range(4)]
            self._loss_i = 0
        buf = ring[self._loss_i % len(ring)]
        self._loss_i += 1
        buf.copy_(out[0].detach(), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(out[0].device))
        return PendingLoss(buf, ev)

    def prime(self):
        """(Re)labels the batch that sits in the static buffers: needed once, before the first step of a stream of
        batches (afterwards every step labels its successor)."""
        with torch.no_grad():
            self.cap_in_cur.copy_(self._corrupt(self.static["caption"]))
            self.labels_cur.copy_(self.trainer.cap_model.segment_labels_of(self.cap_in_cur))
        self._need_prime = False

    def step_ragged(self, ragged, score=None, next_caption=None):
        """One step fed by the input pipeline: `ragged` is a data.RaggedBatch whose stage() has been called -- ONE
        host->device copy of the packed clips + captions, the pad / rgb+flow / mask kernels straight into this runner's
        static buffers (the runner must have been built from a batch with V / V_mask / A_mask), then the step."""
        ragged.to_device(out=self.static)
        if score is not None:
            self.static["score"].copy_(score, non_blocking=True)
        if self.pipeline_critic:
            if next_caption is not None:
                self.static["caption_next"].copy_(next_caption, non_blocking=True)
                if self._need_prime:
                    self.prime()
            else:   # no look-ahead available: label this batch now (correct, just not pipelined)
                self.static["caption_next"].copy_(self.static["caption"])
                self.prime()
        return self.replay()


class ScoredStepRunner:
    """The RL step with the REAL reward scorer in the loop (captioning_bmrl_loops.py:837-877 as the reference runs it):
    two CUDA graphs around one host call.

        graph A   H2D'd batch -> forward, value function, token sampling           (static outputs: tokens, segments)
        host      D2H tokens (+ segments) -> libbmhrl_reward.so CIDEr / BLEU rewards -> H2D scores (+ closed segments)
        graph B   biased-KL loss, backward, all-reduce, Adam, value-function step  (reads the static score buffers)

    The autograd graph built while capturing A is consumed while capturing B; both captures share one memory pool, so
    the saved activations of A stay at fixed addresses.  Replays must alternate A, B like the captures did."""

    def __init__(self, trainer, kind, example_batch, stabilize=False, use_graph=True, warmup=3, pipeline_critic=False):
        """pipeline_critic (graphs only): the frozen critic -- the longest chain of graph A, which the host has to wait
        for before it can score -- leaves graph A.  A third graph C labels the NEXT batch's (corrupted) captions on a side
        stream while the host scores this batch and graph B runs; graph A reads the labels the previous step left behind.
        `step(batch, next_batch)` then takes the captions of the batch that follows (None: this batch is labelled inline
        first, correct but not pipelined)."""
        if trainer.scorer is None:
            raise RuntimeError("ScoredStepRunner needs BMHRLTrainer(scorer=...)")
        if kind not in ("worker", "manager"):
            raise ValueError("ScoredStepRunner runs the worker or the manager RL step")
        self.trainer, self.kind, self.stabilize = trainer, kind, stabilize
        self.train_worker = kind == "worker"
        dev = next(trainer.cap_model.parameters()).device
        self.static = {k: torch.empty_like(v, device=dev).copy_(v) for k, v in example_batch.items()
                       if k in ("rgb", "flow", "audio", "caption")}
        self.captions = list(example_batch["trg_caption"])
        self.use_graph = use_graph
        self.graph_a = self.graph_b = self.graph_c = None
        self.score_buf = self.seg_buf = None
        self.scorer_seconds = 0.0
        self.pipeline_critic = bool(pipeline_critic) and use_graph and hasattr(trainer.cap_model, "segment_labels_of")
        self.critic_units_per_cta = int(os.environ.get("BMHRL_CRITIC_UPB", "16"))
        if not use_graph:
            return
        if self.pipeline_critic:
            self.static["caption_next"] = self.static["caption"].clone()
            with torch.no_grad():
                self.cap_in_cur = self._corrupt(self.static["caption"]).clone()
                self.labels_cur = trainer.cap_model.segment_labels_of(self.cap_in_cur).clone()
            self.cap_in_next = self.cap_in_cur.clone()
            self.labels_next = self.labels_cur.clone()
            self.critic_stream = ops.new_stream(dev.index if dev.index is not None else torch.cuda.current_device())
            self.rng_next = ops.RNG.get_state(dev).clone()   # Philox state graph C draws from (see replay)
            self._need_prime = True
            torch.cuda.synchronize(dev)
        side = ops.warmup_stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self._eager()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        from ._lib import launch_count
        n0 = launch_count()
        self.graph_a = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph_a):
            batch_a = {k: v for k, v in self.static.items() if k != "caption_next"}
            if self.pipeline_critic:
                batch_a["caption_in"] = self.cap_in_cur
                trainer.cap_model.precomputed_segment_labels = self.labels_cur
            try:
                trainer.rl_begin(batch_a, self.train_worker)
                self.st = trainer.rl_forward(batch_a, self.train_worker)
            finally:
                trainer.cap_model.precomputed_segment_labels = None
            if getattr(trainer, "_zero_join", None) is not None:   # the gradient memset belongs to THIS graph
                trainer._zero_join()
                trainer._zero_join = None
            ops.join_side_streams(dev)
            self.tok_out = self.st["tok"]
            self.seg_out = self.st["seg"]
        torch.cuda.synchronize(dev)
        B, L = self.tok_out.shape
        self.score_buf = torch.zeros(B, L, device=dev, dtype=torch.float32)
        self.seg_buf = torch.zeros_like(self.seg_out)   # filled by the host part of every replay
        self.graph_b = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph_b, pool=self.graph_a.pool()):
            losses = trainer.rl_tail(self.st, self.score_buf, self.seg_buf, self.train_worker, self.stabilize)
            self.out = trainer.rl_finish(losses[0], losses[1], self.train_worker)
        if self.pipeline_critic:
            # graph C (a pool of its own: it runs BESIDE A / the host part / B): corruption + critic of the next batch
            self.graph_c = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph_c), torch.no_grad(), ops.rnn_units_per_cta(self.critic_units_per_cta):
                self.cap_in_next.copy_(self._corrupt(self.static["caption_next"], rng=self.rng_next))
                self.labels_next.copy_(trainer.cap_model.segment_labels_of(self.cap_in_next))
        self.captured_launches = launch_count() - n0
        del self.st  # the autograd graph has been consumed by the capture of B

    def _corrupt(self, caption, rng=None):
        """caption[:, :-1] after the reference's generate_synonyms (same as StepRunner._corrupt)."""
        t = self.trainer
        cap_in = caption[:, :-1].contiguous()
        if t.random_synonyms > 0:
            from . import data
            cap_in = data.generate_synonyms(cap_in, t.random_synonyms, t.cap_model.voc_size, t.end_idx, t.pad_idx,
                                            salt=t._syn_salt(), rng=rng)
        return cap_in

    def prime(self):
        with torch.no_grad():
            self.cap_in_cur.copy_(self._corrupt(self.static["caption"]))
            self.labels_cur.copy_(self.trainer.cap_model.segment_labels_of(self.cap_in_cur))
        self._need_prime = False

    def _host_part(self):
        import time
        t = self.trainer
        tok_h = self.tok_out  # .to("cpu") inside host_score synchronises with graph A
        t0 = time.perf_counter()
        score, seg = t.host_score(tok_h, self.seg_out, None, self.captions, self.train_worker)
        self.scorer_seconds = time.perf_counter() - t0
        self.score_buf.copy_(score, non_blocking=True)
        self.seg_buf.copy_(seg, non_blocking=True)

    def _eager(self):
        t = self.trainer
        b = {k: v for k, v in self.static.items() if k != "caption_next"}
        b["trg_caption"] = self.captions
        return t.rl_step(b, self.train_worker, self.stabilize)

    def load(self, host_batch, non_blocking=True, next_host_batch=None):
        for k, buf in self.static.items():
            if k == "caption_next":
                src = (next_host_batch if next_host_batch is not None else host_batch)["caption"]
                buf.copy_(src, non_blocking=non_blocking)
            else:
                buf.copy_(host_batch[k], non_blocking=non_blocking)
        if "trg_caption" in host_batch:
            self.captions = list(host_batch["trg_caption"])

    def replay(self):
        if not self.use_graph:
            self.out = self._eager()
            return self.out
        if not self.pipeline_critic:
            self.graph_a.replay()
            self._host_part()
            self.graph_b.replay()
            return self.out
        dev = self.static["caption"].device
        cur = torch.cuda.current_stream(dev)
        self.rng_next.copy_(ops.RNG.get_state(dev))  # before graph A advances the device state: race-free, new per step
        self.critic_stream.wait_stream(cur)          # caption_next is loaded; the previous step's hand-over is done
        self.graph_a.replay()
        with torch.cuda.stream(self.critic_stream):  # beside A, the host part and B
            self.graph_c.replay()
        self._host_part()
        self.graph_b.replay()
        cur.wait_stream(self.critic_stream)
        self.cap_in_cur.copy_(self.cap_in_next)
        self.labels_cur.copy_(self.labels_next)
        return self.out

    def step(self, host_batch, next_host_batch=None):
        self.load(host_batch, next_host_batch=next_host_batch)
        if self.pipeline_critic and (self._need_prime or next_host_batch is None):
            self.prime()    # first step of a stream, or no look-ahead: label THIS batch now
        return self.replay()
