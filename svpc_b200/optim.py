"""Training-iteration tail of the reference as three multi-tensor CUDA launches.

Mirrors the reference's interface (``src/rtransformer/optimization.py``) so ``src/train.py`` can swap the import:

  * ``BertAdam(params_or_groups, lr, warmup, t_total, schedule, b1, b2, e, weight_decay, max_grad_norm)`` - same
    constructor, same ``ValueError``s (:230-238), same per-parameter state (``step`` / ``next_m`` / ``next_v``, :296-303),
    ``get_lr()`` (:260-270) and ``step()`` (:272-331);
  * ``EMA(decay)`` with ``register`` / ``__call__(model, step)`` / ``assign`` / ``resume`` (:188-216).

``BertAdam.step()`` alone is the reference's ``optimizer.step()``.  ``BertAdam.step(grad_clip=..., ema=..., niter=...)``
additionally folds ``nn.utils.clip_grad_norm_(model.parameters(), grad_clip)`` (train.py:141-142) and ``ema(model, niter)``
(train.py:146-147) into the same pass: one read of every gradient for the norms, one read-modify-write of
(g, p, m, v, shadow).  No device synchronisation: the per-step scalars (scheduled learning rates, EMA decay) live in device
memory and are refreshed by one few-hundred-byte H2D copy in stream order before the launches.

There is no CPU path: parameters must be CUDA fp32 tensors.
"""
from __future__ import annotations

import ctypes as C
import math

import torch
from torch.optim import Optimizer

from . import backend
from .cuda_backend import CudaBackend, SvpcError

CHUNK = 16384                # elements per CTA (64 KB of each stream)
ADAM, EMA_FLAG, WRITE_GRADS = 1, 2, 4


# ---------------------------------------------------------------------------------------------------- schedules
class _LRSchedule:
    """optimization.py:35-78."""

    def __init__(self, warmup=0.002, t_total=-1, **kw):
        if not 0.0 <= warmup < 1.0 and not warmup == -1:
            raise ValueError("Invalid warmup: {} - should be in [0.0, 1.0[ or -1".format(warmup))
        warmup = max(warmup, 0.)
        self.warmup, self.t_total = float(warmup), float(t_total)

    def get_lr(self, step, nowarn=False):
        if self.t_total < 0:
            return 1.
        return self.get_lr_(float(step) / self.t_total)

    def get_lr_(self, progress):
        return 1.


class ConstantLR(_LRSchedule):
    pass


class WarmupConstantSchedule(_LRSchedule):
    def get_lr_(self, progress):                                  # :154-157
        if progress < self.warmup:
            return progress / self.warmup
        return 1.


class WarmupLinearSchedule(_LRSchedule):
    def get_lr_(self, progress):                                  # :165-168
        if progress < self.warmup:
            return progress / self.warmup
        return max((progress - 1.) / (self.warmup - 1.), 0.)


class WarmupCosineSchedule(_LRSchedule):
    def __init__(self, warmup=0.002, t_total=-1, cycles=.5, **kw):
        super().__init__(warmup=warmup, t_total=t_total, **kw)
        self.cycles = cycles

    def get_lr_(self, progress):                                  # :97-102
        if progress < self.warmup:
            return progress / self.warmup
        progress = (progress - self.warmup) / (1 - self.warmup)
        return 0.5 * (1. + math.cos(math.pi * self.cycles * 2 * progress))


SCHEDULES = {None: ConstantLR, "none": ConstantLR, "warmup_cosine": WarmupCosineSchedule,
             "warmup_constant": WarmupConstantSchedule, "warmup_linear": WarmupLinearSchedule}


# ---------------------------------------------------------------------------------------------------- tables
class _TensorTable:
    """Device-side description of a fixed list of fp32 CUDA tensors: chunk table, element counts, pointer tables."""

    def __init__(self, tensors):
        assert tensors, "empty parameter list"
        dev = tensors[0].device
        for t in tensors:
            if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
                raise SvpcError("svpc_b200.optim needs contiguous fp32 CUDA parameters (no CPU path)")
        self.device, self.n = dev, len(tensors)
        ct, cs, first = [], [], [0]
        for i, t in enumerate(tensors):
            for s in range(0, max(t.numel(), 1), CHUNK):
                ct.append(i)
                cs.append(s)
            first.append(len(ct))
        self.n_chunks = len(ct)
        self.chunk_tensor = torch.tensor(ct, dtype=torch.int32, device=dev)
        self.chunk_start = torch.tensor(cs, dtype=torch.int64, device=dev)
        self.tensor_first = torch.tensor(first, dtype=torch.int32, device=dev)
        self.numel = torch.tensor([t.numel() for t in tensors], dtype=torch.int64, device=dev)
        self.chunk_sq = torch.zeros(self.n_chunks, dtype=torch.float32, device=dev)
        self.coef = torch.ones(self.n, dtype=torch.float32, device=dev)
        self.norms = torch.zeros(2, dtype=torch.float32, device=dev)

    def ptr_table(self, tensors, cache=None):
        """int64 device array of data pointers (0 for None); re-uploaded only when a pointer changed."""
        ptrs = [0 if t is None else t.data_ptr() for t in tensors]
        if cache is not None and cache.get("ptrs") == ptrs:
            return cache["dev"]
        dev = torch.tensor(ptrs, dtype=torch.int64).to(self.device, non_blocking=True)
        if cache is not None:
            cache["ptrs"], cache["dev"] = ptrs, dev
        return dev


_BE = None


def _be():
    """The CUDA operator library (never the plain-torch checker the CPU tests may have installed in ``backend``)."""
    global _BE
    be = backend._impl
    if isinstance(be, CudaBackend):
        return be
    if _BE is None:
        _BE = CudaBackend()
    return _BE


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


# ---------------------------------------------------------------------------------------------------- EMA
class EMA(object):
    """optimization.py:183-216; the shadow update runs as one multi-tensor launch (or inside BertAdam.step)."""

    def __init__(self, decay):
        self.decay = decay
        self.shadow = {}
        self.original = {}
        self._ptr_name = {}
        self._table = None

    def register(self, name, val):
        self.shadow[name] = val.clone()
        self._ptr_name[val.data_ptr()] = name       # train.py:348 registers p.data: remembers which parameter this is
        self._table = None

    def shadow_of(self, param):
        """Shadow tensor of a parameter registered through ``register(name, p.data)`` (None if it was not)."""
        name = self._ptr_name.get(param.data_ptr())
        return None if name is None else self.shadow[name]

    def decay_at(self, step):
        return min(self.decay, (1 + step) / (10.0 + step))          # :197

    def _bind(self, model):
        named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
        for n, _ in named:
            assert n in self.shadow
        if self._table is None or self._table["names"] != [n for n, _ in named]:
            params = [p.data for _, p in named]
            tab = _TensorTable(params)
            self._table = {"names": [n for n, _ in named], "tab": tab, "pc": {}, "sc": {},
                           "decay": torch.zeros(2, dtype=torch.float32, device=tab.device)}
        return named, self._table

    def __call__(self, model, step):
        named, T = self._bind(model)
        tab = T["tab"]
        d = self.decay_at(step)
        T["decay"].copy_(torch.tensor([d, 1.0 - d], dtype=torch.float32), non_blocking=True)
        pt = tab.ptr_table([p.data for _, p in named], T["pc"])
        st = tab.ptr_table([self.shadow[n] for n, _ in named], T["sc"])
        _be()._call("svpc_optim_apply", tab.n_chunks, _p(tab.chunk_tensor), _p(tab.chunk_start), CHUNK,
                            None, _p(pt), None, None, _p(st), _p(tab.numel), None, None, None,
                            C.c_double(0.0), C.c_double(0.0), C.c_double(0.0), _p(T["decay"]), EMA_FLAG)

    def _copy_all(self, T, key, srcs, dsts):
        """dst[i] <- src[i] for all registered tensors in ONE launch (multi-tensor fp32 copy through svpc_cast_multi)."""
        sig = ([t.data_ptr() for t in srcs], [t.data_ptr() for t in dsts])
        hit = T.get(key)
        if hit is None or hit[0] != sig:
            n = [t.numel() for t in srcs]
            hit = T[key] = (sig, _be().cast_multi_tables(sig[0], sig[1], n, n, n, CHUNK, T["tab"].device))
        _be().cast_multi(hit[1], torch.float32, CHUNK)

    def assign(self, model):                                          # :205-210
        """Parameters take the shadow values, the originals are kept for ``resume``.  The reference re-points
        ``param.data`` at the shadow tensors (after one clone per parameter); here the VALUES move - two multi-tensor
        launches - and every parameter keeps its address, so captured CUDA graphs, optimiser pointer tables and the
        registered-pointer map stay valid across the evaluation window (train.py:393-405)."""
        named, T = self._bind(model)
        params = [p.data for _, p in named]
        for (n, _), p in zip(named, params):
            o = self.original.get(n)
            if o is None or o.shape != p.shape or o.device != p.device:
                self.original[n] = torch.empty_like(p)
        self._copy_all(T, "save", params, [self.original[n] for n, _ in named])
        self._copy_all(T, "load", [self.shadow[n] for n, _ in named], params)
        torch.autograd.graph.increment_version([p for _, p in named])

    def resume(self, model):                                          # :212-216
        named, T = self._bind(model)
        for n, _ in named:
            assert n in self.original, "resume() without assign()"
        self._copy_all(T, "restore", [self.original[n] for n, _ in named], [p.data for _, p in named])
        torch.autograd.graph.increment_version([p for _, p in named])


# ---------------------------------------------------------------------------------------------------- BertAdam
class BertAdam(Optimizer):
    def __init__(self, params, lr=None, warmup=-1, t_total=-1, schedule="warmup_linear", b1=0.9, b2=0.999, e=1e-6,
                 weight_decay=0.01, max_grad_norm=1.0, **kwargs):
        if lr is None:
            raise ValueError("BertAdam requires a learning rate")
        if lr < 0.0:
            raise ValueError("Invalid learning rate: {} - should be >= 0.0".format(lr))
        if not isinstance(schedule, _LRSchedule) and schedule not in SCHEDULES:
            raise ValueError("Invalid schedule parameter: {}".format(schedule))
        if not 0.0 <= b1 < 1.0:
            raise ValueError("Invalid b1 parameter: {} - should be in [0.0, 1.0[".format(b1))
        if not 0.0 <= b2 < 1.0:
            raise ValueError("Invalid b2 parameter: {} - should be in [0.0, 1.0[".format(b2))
        if not e >= 0.0:
            raise ValueError("Invalid epsilon value: {} - should be >= 0.0".format(e))
        if not isinstance(schedule, _LRSchedule):
            schedule = SCHEDULES[schedule](warmup=warmup, t_total=t_total)
        defaults = dict(lr=lr, schedule=schedule, b1=b1, b2=b2, e=e, weight_decay=weight_decay, max_grad_norm=max_grad_norm)
        super().__init__(params, defaults)
        g0 = self.param_groups[0]
        for g in self.param_groups:
            if (g["b1"], g["b2"], g["e"], g["max_grad_norm"]) != (g0["b1"], g0["b2"], g0["e"], g0["max_grad_norm"]):
                raise SvpcError("the fused step needs b1 / b2 / e / max_grad_norm shared by all groups "
                                        "(learning rate, schedule and weight decay may differ per group)")
        self._tab = None
        self._last_norm = None

    # ------------------------------------------------------------------ reference surface
    def get_lr(self):                                                 # :260-270
        lr = []
        for group in self.param_groups:
            for p in group["params"]:
                state = self.state[p]
                if len(state) == 0:
                    return [0]
                lr.append(group["lr"] * group["schedule"].get_lr(state["step"]))
        return lr

    def total_norm(self):
        """Device scalar: gradient norm before the global clip of the last fused step (clip_grad_norm_'s return value)."""
        return None if self._last_norm is None else self._last_norm[0]

    # ------------------------------------------------------------------ fused step
    def _bind(self):
        if self._tab is None:
            ps = [p for g in self.param_groups for p in g["params"]]
            tab = _TensorTable([p.data for p in ps])
            self._tab = {"tab": tab, "params": ps, "gc": {}, "pc": {}, "mc": {}, "vc": {}, "sc": {},
                         "lr": torch.zeros(tab.n, dtype=torch.float32, device=tab.device),
                         "wd": torch.tensor([g["weight_decay"] for g in self.param_groups for _ in g["params"]],
                                            dtype=torch.float32, device=tab.device),
                         "decay": torch.zeros(2, dtype=torch.float32, device=tab.device)}
        return self._tab

    @torch.no_grad()
    def step(self, closure=None, grad_clip=None, ema=None, niter=None, write_grads=False):
        """``step()``: BertAdam.step of the reference.  ``grad_clip`` (not None / -1): the global clip of train.py:141-142
        happens first, inside the same pass.  ``ema`` + ``niter``: EMA.__call__(model, niter) of train.py:146-147 on the
        updated parameters, inside the same pass (parameters are matched to the tensors given to ``ema.register``).  ``write_grads``: also store the clipped gradients into ``p.grad`` as the reference does in place."""
        loss = closure() if closure is not None else None
        T = self._bind()
        tab, ps = T["tab"], T["params"]
        be = _be()
        g0 = self.param_groups[0]
        lrs, grads, ms, vs = [], [], [], []
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:                                    # :286-287
                    lrs.append(0.0)
                    grads.append(None)
                    st = self.state[p]
                    ms.append(st.get("next_m"))
                    vs.append(st.get("next_v"))
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError("Adam does not support sparse gradients, please consider SparseAdam instead")
                if not (p.grad.is_cuda and p.grad.dtype == torch.float32 and p.grad.is_contiguous()):
                    raise SvpcError("svpc_b200.optim needs contiguous fp32 CUDA gradients (no CPU path)")
                st = self.state[p]
                if len(st) == 0:                                      # :296-303
                    st["step"] = 0
                    st["next_m"] = torch.zeros_like(p.data)
                    st["next_v"] = torch.zeros_like(p.data)
                lrs.append(group["lr"] * group["schedule"].get_lr(st["step"]))      # :325-326
                st["step"] += 1                                       # :331
                grads.append(p.grad)
                ms.append(st["next_m"])
                vs.append(st["next_v"])
        # pageable source: the driver stages these few hundred bytes before returning, so the next step's values cannot
        # overtake a copy that has not run yet (a reused pinned buffer could); no device synchronisation either way
        T["lr"].copy_(torch.tensor(lrs, dtype=torch.float32), non_blocking=True)
        flags = ADAM | (WRITE_GRADS if write_grads else 0)
        st_tab = None
        if ema is not None:
            assert niter is not None, "ema needs the iteration number (train.py:147)"
            d = ema.decay_at(niter)
            T["decay"].copy_(torch.tensor([d, 1.0 - d], dtype=torch.float32), non_blocking=True)
            st_tab = tab.ptr_table([ema.shadow_of(p) for p in ps], T["sc"])
            flags |= EMA_FLAG
        gt = tab.ptr_table(grads, T["gc"])
        pt = tab.ptr_table([p.data for p in ps], T["pc"])
        mt = tab.ptr_table(ms, T["mc"])
        vt = tab.ptr_table(vs, T["vc"])
        gmax = -1.0 if grad_clip is None else float(grad_clip)
        be._call("svpc_optim_sqnorm", tab.n_chunks, _p(tab.chunk_tensor), _p(tab.chunk_start), CHUNK, _p(gt), _p(tab.numel),
                 _p(tab.chunk_sq))
        be._call("svpc_optim_coef", tab.n, _p(tab.tensor_first), _p(tab.chunk_sq), C.c_float(gmax),
                 C.c_float(float(g0["max_grad_norm"])), _p(tab.coef), _p(tab.norms))
        be._call("svpc_optim_apply", tab.n_chunks, _p(tab.chunk_tensor), _p(tab.chunk_start), CHUNK, _p(gt), _p(pt), _p(mt),
                 _p(vt), _p(st_tab), _p(tab.numel), _p(tab.coef), _p(T["lr"]), _p(T["wd"]), C.c_double(g0["b1"]),
                 C.c_double(g0["b2"]), C.c_double(g0["e"]), _p(T["decay"]) if ema is not None else None, flags)
        # the kernels wrote the parameters through raw pointers: bump the version counters so every cache keyed on them
        # (functional.WeightCache's bf16 GEMM-layout copies) sees the update, exactly as after an in-place torch op
        torch.autograd.graph.increment_version([p for p, g in zip(ps, grads) if g is not None])
        self._last_norm = tab.norms
        return loss
