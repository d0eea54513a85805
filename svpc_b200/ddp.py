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
    """Buckets are FLAT PERSISTENT buffers; every parameter owns a 16-byte aligned slot of its bucket.

    * weight-gradient GEMMs write straight into the slots (``LinearFn.backward`` asks ``grad_slot`` for its output buffer;
      the weights of a fused QKV / KV projection get adjacent slots so that the one GEMM that produces all of them writes
      one contiguous block) - only the small tensors (biases, LayerNorm, embeddings) are copied in by the flatten pass;
    * buckets follow the order in which gradients become ready (reverse registration order) and SHRINK towards the end
      of the backward pass: the all-reduce of the last bucket cannot overlap anything, so it is kept small;
    * after the all-reduce every ``.grad`` is a view of its slot (no copy-back)."""

    def __init__(self, model, process_group=None, bucket_bytes: int = 0):
        if bucket_bytes <= 0:
            bucket_bytes = int(float(os.environ.get("SVPC_DDP_BUCKET_MB", "32")) * (1 << 20))
        self.group = process_group
        named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
        self.params = [p for _, p in named]
        # gradients become ready roughly in reverse registration order; the weights of one attention module (query / key /
        # value) stay together, in forward order, because one fused GEMM produces their gradients as one block
        by_name = dict(named)
        order, seen = [], set()
        for n, p in reversed(named):
            if id(p) in seen:
                continue
            grp = [p]
            for tag in (".query.weight", ".key.weight", ".value.weight"):
                if n.endswith(tag):
                    base = n[:-len(tag)]
                    grp = [by_name[base + t] for t in (".query.weight", ".key.weight", ".value.weight") if base + t in by_name]
                    break
            for q in grp:
                if id(q) not in seen:
                    seen.add(id(q))
                    order.append(q)
        # bucket sizes: `bucket_bytes` at the start of the backward pass, halving over the last buckets (the tail)
        total = sum(p.numel() * 4 for p in order)
        tail = []
        sz = max(bucket_bytes // 8, 1 << 20)
        while sz < bucket_bytes and sum(tail) + sz < total // 2:
            tail.append(sz)
            sz *= 2
        head_total = total - sum(tail)
        limits = [bucket_bytes] * max(1, -(-head_total // bucket_bytes)) + list(reversed(tail))
        self.buckets, cur, size, li = [], [], 0, 0
        for p in order:
            cur.append(p)
            size += p.numel() * 4
            if size >= limits[min(li, len(limits) - 1)]:
                self.buckets.append(cur)
                cur, size, li = [], 0, li + 1
        if cur:
            self.buckets.append(cur)
        self.bucket_of, self.offset_of = {}, {}
        self.bucket_numel = []
        for bi, b in enumerate(self.buckets):
            off = 0
            for p in b:
                self.bucket_of[p] = bi
                self.offset_of[p] = off
                off += (p.numel() + 3) // 4 * 4          # 16-byte aligned slots
            self.bucket_numel.append(off)
        self.flat = [None] * len(self.buckets)           # allocated on first use (on the parameters' device)
        self._ready = [set() for _ in self.buckets]
        self._launched = [False] * len(self.buckets)
        self._work = []
        self._handed = set()                             # slots given to a weight-gradient GEMM in the current backward pass
        self._comm = {}                                  # device -> communication stream
        # parameters that get no gradient in the current model mode (dead reference parameters, modules a mode does not
        # run): learned from the first backward pass (agreed between the ranks) - a bucket that holds one is then launched
        # as soon as its LIVE parameters are ready instead of at finish(), where its all-reduce overlaps nothing
        self._absent = None
        self._need = [len(b) for b in self.buckets]
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]
        self.enabled = dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1
        if self.enabled:
            from . import functional as Fn
            Fn.GRAD_SLOTS.register(self)

    # ---- slots
    def _flat(self, bi, like):
        if self.flat[bi] is None or self.flat[bi].device != like.device:
            self.flat[bi] = torch.zeros(self.bucket_numel[bi], dtype=torch.float32, device=like.device)
        return self.flat[bi]

    def slot(self, p):
        bi = self.bucket_of[p]
        off = self.offset_of[p]
        return self._flat(bi, p)[off:off + p.numel()].view_as(p)

    def grad_slot(self, params):
        """One contiguous fp32 buffer (rows of all `params` stacked) inside a bucket, or None when the parameters are not
        adjacent slots of one bucket (or the reducer is idle): the output of a (fused) weight-gradient GEMM."""
        if not self.enabled or any(p not in self.bucket_of or p in self._handed for p in params):
            return None                                   # (a weight used twice: the second gradient must not overwrite the first)
        bi = self.bucket_of[params[0]]
        off = self.offset_of[params[0]]
        end = off
        for p in params:
            if self.bucket_of[p] != bi or self.offset_of[p] != end or p.dim() != 2 or p.shape[1] != params[0].shape[1] or \
                    p.numel() % 4 != 0:
                return None
            end += p.numel()
        rows = sum(p.shape[0] for p in params)
        self._handed.update(params)
        return self._flat(bi, params[0])[off:end].view(rows, params[0].shape[1])

    # ---- hooks
    def _on_grad(self, p):
        if not self.enabled:
            return
        bi = self.bucket_of[p]
        if self._absent is not None and p in self._absent:
            raise RuntimeError("GradAllReducer: a parameter that had no gradient in the first backward pass got one now "
                               "(the set of live parameters changed, e.g. another model mode): call reset_absent() first")
        self._ready[bi].add(p)
        if not self._launched[bi] and len(self._ready[bi]) == self._need[bi]:
            self._launch(bi, [q for q in self.buckets[bi] if q in self._ready[bi]])

    def _launch(self, bi, params):
        self._launched[bi] = True
        if not params:
            return
        from . import functional as Fn
        flat = self._flat(bi, params[0])
        ctx = None
        if flat.is_cuda:
            # the flatten pass and the collective run on a communication stream that waits for the compute stream AND for
            # the weight-gradient stream; the compute stream itself does not wait for either (it goes on with the input-
            # gradient chain) until finish()
            dev = flat.device
            cs = self._comm.get(dev)
            if cs is None:
                cs = self._comm[dev] = torch.cuda.Stream(device=dev)
            cs.wait_stream(torch.cuda.current_stream(dev))
            if dev in Fn.WGRAD.active:
                cs.wait_stream(Fn.WGRAD.streams[dev])
            ctx = torch.cuda.stream(cs)
            ctx.__enter__()
        try:
            # gradients that were not written into their slot by the GEMM (biases, LayerNorm, embeddings, ...) are copied in
            stray = [p for p in params if p.grad.data_ptr() != flat.data_ptr() + 4 * self.offset_of[p]]
            if stray:
                torch._foreach_copy_([self.slot(p) for p in stray], [p.grad for p in stray])
            missing = [p for p in self.buckets[bi] if p not in self._ready[bi]]
            if missing:                                   # parameters without a gradient in this mode: their slots carry zeros
                torch._foreach_zero_([self.slot(p) for p in missing])
            h = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        finally:
            if ctx is not None:
                ctx.__exit__(None, None, None)
        self._work.append((h, params))

    def finish(self):
        """Call after ``loss.backward()``: reduces the buckets whose parameters did not all get a gradient (same set on
        every rank for a given model mode), waits, and leaves every ``.grad`` as a view of its bucket slot."""
        if not self.enabled:
            return
        for bi, b in enumerate(self.buckets):
            if not self._launched[bi]:
                self._launch(bi, [p for p in b if p in self._ready[bi]])
        if self._absent is None:
            self._learn_absent()
        for dev, cs in self._comm.items():
            torch.cuda.current_stream(dev).wait_stream(cs)
        for h, params in self._work:
            h.wait()
            for p in params:
                p.grad = self.slot(p)
        self._work = []
        self._handed = set()
        self._ready = [set() for _ in self.buckets]
        self._launched = [False] * len(self.buckets)

    def _learn_absent(self):
        """After the first backward pass: the parameters without a gradient ON EVERY RANK (one small MAX all-reduce)."""
        dev = self.params[0].device
        has = torch.tensor([1 if p in self._ready[self.bucket_of[p]] else 0 for p in self.params], dtype=torch.int32, device=dev)
        dist.all_reduce(has, op=dist.ReduceOp.MAX, group=self.group)
        has = has.cpu().tolist()
        self._absent = {p for p, h in zip(self.params, has) if not h}
        self._need = [sum(1 for p in b if p not in self._absent) for b in self.buckets]

    def reset_absent(self):
        """Forget which parameters get no gradient (call when the model mode or the set of trained modules changes)."""
        self._absent = None
        self._need = [len(b) for b in self.buckets]

    def remove(self):
        for h in self._hooks:
            h.remove()
        from . import functional as Fn
        Fn.GRAD_SLOTS.unregister(self)
