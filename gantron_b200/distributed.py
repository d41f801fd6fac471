"""Drop-in replacement for the reference's `distributed.apply_gradient_allreduce` (distributed.py:126-173).

The reference flattens ALL gradients into one buffer per dtype after backward has finished, all-reduces it,
divides and copies back (no overlap, 3x the gradient bytes of extra HBM traffic, one broadcast per state_dict
tensor at start-up, and a forward-hook dependency that breaks for the Discriminator).

Here:
  * start-up: ONE flat broadcast of parameters and ONE of floating-point buffers (BatchNorm running stats).
  * gradients live in a few flat buckets (~32 MB, reverse parameter order = the order backward produces them:
    Discriminator -> Postnet -> Decoder -> Encoder).  A per-parameter post-accumulate hook scales the gradient by
    1/world into its bucket slot (so `.grad` becomes a view of the bucket: no unflatten copy later) and, when a
    bucket is complete, launches `all_reduce(async_op=True)`: NCCL runs it on its own stream over NVLink 5 /
    NVSwitch while the BPTT kernel and the deferred weight-gradient GEMMs keep the SMs busy.
  * an end-of-backward engine callback waits for the outstanding reductions (and reduces any partially filled
    bucket), so by the time `backward()` returns every `.grad` is the world average, like the reference.
It keeps the signature `apply_gradient_allreduce(module) -> module` and does not depend on a forward hook.
Works with any torch.distributed backend (NCCL on the B200 box; gloo in the CPU tests).
"""
from __future__ import annotations

from typing import List

import torch
import torch.distributed as dist
from torch.autograd import Variable

BUCKET_BYTES = 32 << 20


class _Bucket:
    def __init__(self, params: List[torch.nn.Parameter]):
        self.params = params
        p0 = params[0]
        self.numel = sum(p.numel() for p in params)
        self.flat = torch.zeros(self.numel, device=p0.device, dtype=p0.dtype)
        self.views = []
        off = 0
        for p in params:
            self.views.append(self.flat[off:off + p.numel()].view_as(p))
            off += p.numel()
        self.ready = [False] * len(params)
        self.n_ready = 0
        self.work = None

    def reset(self):
        self.ready = [False] * len(self.params)
        self.n_ready = 0
        self.work = None


class GradientAllReducer:
    """Holds the buckets and hooks of one module (returned module keeps a reference as `_gtn_reducer`)."""

    def __init__(self, module: torch.nn.Module, bucket_bytes: int = None, group=None):
        if not dist.is_initialized():
            raise RuntimeError("apply_gradient_allreduce: torch.distributed is not initialised")
        self.module, self.group = module, group
        bucket_bytes = BUCKET_BYTES if bucket_bytes is None else bucket_bytes
        self.world = dist.get_world_size(group)
        self.inv_world = 1.0 / self.world
        self._broadcast_state()
        params = [p for p in module.parameters() if p.requires_grad or True]
        params = list(reversed(params))
        self.buckets: List[_Bucket] = []
        cur, cur_bytes, cur_key = [], 0, None
        for p in params:
            key = (p.device, p.dtype)
            nbytes = p.numel() * p.element_size()
            if cur and (key != cur_key or cur_bytes + nbytes > bucket_bytes):
                self.buckets.append(_Bucket(cur))
                cur, cur_bytes = [], 0
            cur.append(p)
            cur_bytes += nbytes
            cur_key = key
        if cur:
            self.buckets.append(_Bucket(cur))
        self._callback_queued = False
        self._task_id = None             # autograd graph-task id of the backward pass the bucket state belongs to
        self.launched_early = 0          # buckets whose all-reduce started before backward finished (overlap evidence)
        self._hooks = []
        for b in self.buckets:
            for i, p in enumerate(b.params):
                self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(b, i)))

    # ---- start-up ------------------------------------------------------------------------------
    @torch.no_grad()
    def _broadcast_state(self):
        # copy INTO the Parameters / buffers themselves (not through `.data`): the in-place copy bumps their version
        # counters, which is what invalidates the kernels' bf16 weight packs (decoder_ops.PackedDecoder.stale)
        tensors = list(self.module.parameters())
        tensors += [b for b in self.module.buffers() if b.is_floating_point()]
        by_key = {}
        for t in tensors:
            by_key.setdefault((t.device, t.dtype), []).append(t)
        for ts in by_key.values():
            flat = torch.cat([t.detach().reshape(-1) for t in ts])
            dist.broadcast(flat, 0, group=self.group)
            off = 0
            for t in ts:
                t.copy_(flat[off:off + t.numel()].view_as(t))
                off += t.numel()

    # ---- per-gradient hook ---------------------------------------------------------------------
    def _make_hook(self, bucket: _Bucket, idx: int):
        def hook(param):
            g = param.grad
            if g is None:
                return
            tid = torch._C._current_graph_task_id()
            if tid != self._task_id:
                # first gradient of a NEW backward pass.  If the previous pass raised before the engine ran its final
                # callbacks, its bucket state (ready flags, launched reductions, the queued flag) is still around: drop it,
                # or no later pass would ever queue _finalize again and the ranks would silently diverge.
                if self._callback_queued or any(b.n_ready or b.work is not None for b in self.buckets):
                    self._abandon()
                self._task_id = tid
            view = bucket.views[idx]
            if g.data_ptr() != view.data_ptr():
                torch.mul(g, self.inv_world, out=view)     # scale while moving into the bucket slot
                param.grad = view
            else:
                view.mul_(self.inv_world)
            if not bucket.ready[idx]:
                bucket.ready[idx] = True
                bucket.n_ready += 1
            if not self._callback_queued:
                Variable._execution_engine.queue_callback(self._finalize)
                self._callback_queued = True
            if bucket.n_ready == len(bucket.params) and bucket.work is None:
                bucket.work = dist.all_reduce(bucket.flat, group=self.group, async_op=True)
                self.launched_early += 1
        return hook

    # ---- end of backward -----------------------------------------------------------------------
    def _finalize(self):
        self._callback_queued = False
        self._task_id = None
        for b in self.buckets:
            if b.n_ready == 0:
                continue                                   # e.g. the frozen Discriminator during a generator step
            if b.work is None:
                # partially filled bucket (some parameters unused this step): slots without a gradient must not
                # contribute stale values
                for i, r in enumerate(b.ready):
                    if not r:
                        b.views[i].zero_()
                b.work = dist.all_reduce(b.flat, group=self.group, async_op=True)
        for b in self.buckets:
            if b.work is not None:
                b.work.wait()
            b.reset()

    def _abandon(self):
        """Forget a backward pass that never reached _finalize (it raised): wait for what was launched, reset the rest."""
        for b in self.buckets:
            if b.work is not None:
                try:
                    b.work.wait()
                except Exception:
                    pass
            b.reset()
        self._callback_queued = False

    def remove(self):
        for h in self._hooks:
            h.remove()
        self._hooks = []


def apply_gradient_allreduce(module: torch.nn.Module) -> torch.nn.Module:
    """Same contract as the reference (distributed.py:126): returns the module, whose `.grad`s are world-averaged
    after every backward.  Calling it twice on the same module (train.py:121-123 and :250-252 both do) is a no-op."""
    if getattr(module, "_gtn_reducer", None) is not None:
        return module
    object.__setattr__(module, "_gtn_reducer", GradientAllReducer(module))
    return module


def reduce_tensor(tensor: torch.Tensor, n_gpus: int) -> torch.Tensor:
    """Scalar averaging helper with the semantics of train.py:72-76."""
    rt = tensor.clone()
    dist.all_reduce(rt, op=dist.ReduceOp.SUM)
    return rt / n_gpus
