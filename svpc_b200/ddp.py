"""Data-parallel training over the videos of a batch: one process per GPU, parameters replicated, each
rank runs the step on its own videos and the fp32 gradients are SUMMED (the reference's loss is a sum over videos,
model.py:1110-1115, 1188, so summing reproduces a single-process run on the concatenated batch) with bucketed NCCL
all-reduces launched from autograd hooks while the rest of the backward is still running.

Parameters that receive no gradient in the current model mode (e.g. ``memory_intermediate.*``) keep ``grad is None``
on every rank, so the reference BertAdam's skip of such parameters (optimization.py:286-287) is preserved.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


class GradAllReducer:
    def __init__(self, model, process_group=None, bucket_bytes: int = 0):
        if bucket_bytes <= 0:
            bucket_bytes = int(float(os.environ.get("SVPC_DDP_BUCKET_MB", "32")) * (1 << 20))
        self.group = process_group
        self.params = [p for p in model.parameters() if p.requires_grad]
        # gradients become ready roughly in reverse registration order
        order = list(reversed(self.params))
        self.buckets, cur, size = [], [], 0
        for p in order:
            cur.append(p)
            size += p.numel() * 4
            if size >= bucket_bytes:
                self.buckets.append(cur)
                cur, size = [], 0
        if cur:
            self.buckets.append(cur)
        self.bucket_of = {}
        for bi, b in enumerate(self.buckets):
            for p in b:
                self.bucket_of[p] = bi
        self._ready = [set() for _ in self.buckets]
        self._launched = [False] * len(self.buckets)
        self._work = []
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]
        self.enabled = dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1

    def _on_grad(self, p):
        if not self.enabled:
            return
        bi = self.bucket_of[p]
        self._ready[bi].add(p)
        if not self._launched[bi] and len(self._ready[bi]) == len(self.buckets[bi]):
            self._launch(bi, self.buckets[bi])

    def _launch(self, bi, params):
        self._launched[bi] = True
        if not params:
            return
        from . import functional as Fn
        Fn.WGRAD.join()          # weight gradients are produced on a side stream: the flatten below reads them on this one
        # one flat bucket, every tensor's slot padded to 16 bytes so that the views handed back as .grad keep the alignment
        # the vectorised optimiser kernels want; the pad elements are never read
        offs, total = [], 0
        for p in params:
            offs.append(total)
            total += (p.numel() + 3) // 4 * 4
        flat = torch.empty(total, dtype=params[0].grad.dtype, device=params[0].grad.device)
        torch._foreach_copy_([flat[o:o + p.numel()].view_as(p) for o, p in zip(offs, params)], [p.grad for p in params])
        h = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        self._work.append((h, flat, params, offs))

    def finish(self):
        """Call after ``loss.backward()``: reduces the buckets whose parameters did not all get a gradient (same set on
        every rank for a given model mode), waits, and writes the summed gradients back."""
        if not self.enabled:
            return
        for bi, b in enumerate(self.buckets):
            if not self._launched[bi]:
                self._launch(bi, [p for p in b if p in self._ready[bi]])
        for h, flat, params, offs in self._work:
            h.wait()
            # the summed gradients stay where the collective left them: every .grad becomes a view of its bucket (no
            # copy-back pass over the 170 MB; inside a captured step the bucket is a static buffer of the graph's pool)
            for off, p in zip(offs, params):
                p.grad = flat[off:off + p.numel()].view_as(p)
        self._work = []
        self._ready = [set() for _ in self.buckets]
        self._launched = [False] * len(self.buckets)

    def remove(self):
        for h in self._hooks:
            h.remove()
