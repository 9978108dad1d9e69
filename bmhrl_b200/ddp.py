"""Data-parallel gradient synchronisation: one process per GPU, torch.distributed (NCCL over NVLink / NVSwitch on
the B200 box, gloo in the CPU tests).

The reference wraps the agent in single-process nn.DataParallel (scripts/train_rl_captioning_module.py:97-99):
outputs of all replicas are gathered and the loss is computed ONCE over the global batch, so
    loss = sum over ALL ranks' rows / n_tokens_global      (captioning_bmrl_loops.py:1156-1158, :846-862)
and the gradient is the SUM of the per-rank gradients of (local row sum / global token count).  GradSync
reproduces exactly that: `all_reduce_scalar` makes the normaliser global, `__call__` sum-reduces the flat gradient
buffer of a FusedAdam in fixed-size buckets (no averaging).  Parameters and Adam state are replicated and never
broadcast per step (DataParallel re-broadcasts 281 MB every step).
"""
import torch
import torch.distributed as dist


class GradSync:
    def __init__(self, group=None, bucket_bytes=128 << 20, value_fn_mean=True):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.bucket_elems = max(1, bucket_bytes // 4)
        self.value_fn_mean = value_fn_mean
        self.calls = 0

    def ranges(self, opt, only=None):
        """Contiguous [start, end) element ranges of opt.flat_grad that hold currently-trainable tensors (of the
        parameter indices in `only`, when given), merged and then split into buckets of at most bucket_elems."""
        spans = []
        for i in opt.active:
            if only is not None and i not in only:
                continue
            s = opt.offsets[i]
            e = s + opt.params[i].numel()
            # FusedAdam starts every slice on a 16-byte boundary: up to 3 zero padding elements between neighbours
            if spans and 0 <= s - spans[-1][1] < 4 and (s == spans[-1][1] or s % 4 == 0):
                spans[-1][1] = e
            else:
                spans.append([s, e])
        out = []
        for s, e in spans:
            while s < e:
                n = min(self.bucket_elems, e - s)
                out.append((s, s + n))
                s += n
        return out

    def __call__(self, opt, mean=False, only=None):
        if self.world == 1:
            return
        for s, e in self.ranges(opt, only):
            dist.all_reduce(opt.flat_grad[s:e], op=dist.ReduceOp.SUM, group=self.group)
            self.calls += 1
        if mean:
            opt.flat_grad.div_(self.world)

    def all_reduce_scalar(self, t):
        if self.world == 1:
            return t
        t = t.clone()
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t
