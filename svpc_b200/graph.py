"""CUDA-graph replay of the SVPC training step.

The reference is launch-bound by construction (~430k library kernels per step).  The batched step of
``svpc_b200.model`` still enqueues a few hundred kernels from Python; on a B200 those kernels finish faster than the
interpreter can issue them.  ``GraphedTrainStep`` therefore captures ONE forward + backward (+ the bf16 weight refresh
and, under data parallelism, the bucketed gradient all-reduces) into a CUDA graph whose shapes are fixed by the batch
signature (clips per video, ingredients per video, OOV widths), and replays it on static input buffers:

    step = GraphedTrainStep(model, example_args, gumbel_noise=noise)     # warm-up + capture
    step.load(args, noise)        # copies a new batch into the static buffers (H2D from pinned memory or D2D)
    loss = step.run()             # graph replay; gradients land in p.grad exactly as after loss.backward()

A batch with a different signature needs a new capture (``step.matches(args)`` tells); the data layer can bucket batches
by signature, or pad to the configuration maxima.
"""
from __future__ import annotations

import gc
import os

import torch

from . import functional as Fn


def _is_tensor_list(x):
    return isinstance(x, (list, tuple)) and len(x) > 0 and isinstance(x[0], torch.Tensor)


class InputArena:
    """One contiguous buffer for every tensor input of a step, three times: pinned host memory (the data loader collates
    straight into ``host_views``), a device staging copy (written by ONE H2D memcpy on the copy stream) and the static
    device copy a captured graph reads (refreshed by ONE D2D memcpy).  Replaces ~90 per-tensor copies per step
    (reference train.py:91-112 `prepare_batch_inputs` issues one .cuda() per tensor)."""

    ALIGN = 256

    def __init__(self, tensors, device, used_rows=None):
        """``used_rows``: {index in ``tensors``: r} - of that (B, L, F) tensor only rows [0, r) of every (L, F) block are
        read by the step (video features: the first max_v_len of 125 rows, model.py:889); the upload skips the rest."""
        self.device = device
        if os.environ.get("SVPC_ARENA_PITCHED", "1") == "0":     # diagnostic: upload whole blocks with 1-D copies
            used_rows = None
        self.spec, self.plan, self.nbytes = self.layout(tensors, used_rows)
        self.h2d_nbytes = sum(w * h for _, _, w, h in self.plan)
        self.host = torch.empty(self.nbytes, dtype=torch.uint8).pin_memory()
        self.stage = torch.zeros(self.nbytes, dtype=torch.uint8, device=device)
        self.static = torch.zeros(self.nbytes, dtype=torch.uint8, device=device)
        # zero-copy feature blocks (enable_zero_copy): the step reads the clip features through a device table of
        # pointers, so the pitched blocks are uploaded alternately into `static` and `stage` and the commit only rewrites
        # the table (128 bytes) instead of moving ~120 MB device to device
        self.zc_table, self.zc_tabs, self.zc_entries = None, None, ()
        self.cur, self.up = 0, None                    # buffer (0 static / 1 stage) the table points to / of the pending upload
        self.read_done = [None, None]                  # event after the last replay that read the features of buffer 0 / 1

    @classmethod
    def layout(cls, tensors, used_rows=None):
        """(spec, upload plan, arena bytes) - pure host arithmetic."""
        spec, off = [], 0
        segs = []            # upload plan: (offset, pitch, width, height) in bytes; 1-D copies have height 1
        for i, t in enumerate(tensors):
            n = t.numel() * t.element_size()
            spec.append((off, tuple(t.shape), t.dtype))
            r = (used_rows or {}).get(i)
            if r is not None and t.dim() == 3 and r < t.shape[1]:
                pitch = t.shape[1] * t.shape[2] * t.element_size()
                segs.append([off, pitch, r * t.shape[2] * t.element_size(), t.shape[0]])
            else:
                segs.append([off, 0, n, 1])
            off += (n + cls.ALIGN - 1) // cls.ALIGN * cls.ALIGN
        # merge neighbours: consecutive 1-D pieces into one, consecutive pitched pieces with the same geometry into one
        plan = []
        for sg in segs:
            if plan:
                o, pitch, w, h = plan[-1]
                if pitch == 0 and sg[1] == 0:
                    plan[-1][2] = sg[0] + sg[2] - o
                    continue
                if pitch and sg[1] == pitch and sg[2] == w and sg[0] == o + pitch * h:
                    plan[-1][3] += sg[3]
                    continue
            plan.append(list(sg))
        return spec, [tuple(x) for x in plan], off

    def enable_zero_copy(self, table, feat_idx):
        """``table``: the device int64 array of feature-block addresses the captured kernels read (one per tensor of
        ``feat_idx``, indices into the arena's tensors, currently pointing into ``static``).  Enabled only when every one
        of those tensors lies inside a pitched entry of the upload plan and no other tensor does."""
        pitched = [k for k, (_, pitch, _, _) in enumerate(self.plan) if pitch]
        covered = set()
        for k in pitched:
            o, pitch, _, h = self.plan[k]
            covered |= {i for i, (off, _, _) in enumerate(self.spec) if o <= off < o + pitch * h}
        if not pitched or covered != set(feat_idx) or table.numel() != len(feat_idx):
            return False
        offs = [self.spec[i][0] for i in feat_idx]
        self.zc_tabs = [torch.tensor([b.data_ptr() + o for o in offs], dtype=torch.int64).to(self.device)
                        for b in (self.static, self.stage)]
        assert torch.equal(self.zc_tabs[0], table), "the feature table does not point into the static arena"
        self.zc_table, self.zc_entries = table, tuple(pitched)
        return True

    def reset_zero_copy(self):
        """Point the table back at the static arena (a synchronous ``load`` writes the static views); current stream."""
        if self.zc_table is not None and self.cur != 0:
            self.zc_table.copy_(self.zc_tabs[0], non_blocking=True)
            self.cur = 0

    def mark_read(self):
        """Call after enqueueing a replay: the buffer the table points to may be overwritten only after this event."""
        if self.zc_table is not None:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self.read_done[self.cur] = ev

    def upload(self):
        """pinned host arena -> device on the CURRENT stream: one copy per plan entry (ids/masks, the pitched feature
        block, labels).  The feature block goes straight into the buffer the step is NOT reading (zero-copy), the rest into
        the staging arena."""
        be = Fn.be()
        hp, dp = self.host.data_ptr(), self.stage.data_ptr()
        tgt = (1 - self.cur) if self.zc_table is not None else None
        for k, (o, pitch, w, h) in enumerate(self.plan):
            base = (self.static, self.stage)[tgt].data_ptr() if (tgt is not None and k in self.zc_entries) else dp
            be.memcpy2d_h2d(base + o, pitch or w, hp + o, pitch or w, w, h)
        self.up = tgt

    def commit(self):
        """Make the uploaded batch the one the captured step reads, on the CURRENT stream: staging -> static along the upload
        plan for the small pieces; the feature blocks only switch the pointer table (or, without zero-copy, move too)."""
        be = Fn.be()
        sp, dp = self.stage.data_ptr(), self.static.data_ptr()
        for k, (o, pitch, w, h) in enumerate(self.plan):
            if self.up is not None and k in self.zc_entries:
                continue
            be.memcpy2d_d2d(dp + o, pitch or w, sp + o, pitch or w, w, h)
        if self.up is not None:
            self.zc_table.copy_(self.zc_tabs[self.up], non_blocking=True)
            self.cur, self.up = self.up, None

    def _views(self, buf):
        out = []
        for off, shape, dtype in self.spec:
            n = 1
            for d in shape:
                n *= d
            nb = n * torch.empty((), dtype=dtype).element_size()
            out.append(buf[off:off + nb].view(dtype).view(shape))
        return out

    def host_views(self):
        return self._views(self.host)

    def static_views(self):
        return self._views(self.static)


def _flatten(args, noise):
    """(tensors in a fixed order, rebuild(list of tensors) -> (args, noise))."""
    flat, slots = [], []
    for i, a in enumerate(args):
        if isinstance(a, torch.Tensor):
            slots.append(("t", i, len(flat)))
            flat.append(a)
        elif _is_tensor_list(a):
            slots.append(("l", i, len(flat), len(a)))
            flat += list(a)
    n0 = len(flat)
    if noise is not None:
        flat += list(noise)

    def rebuild(ts):
        out = list(args)
        for sl in slots:
            if sl[0] == "t":
                out[sl[1]] = ts[sl[2]]
            else:
                out[sl[1]] = list(ts[sl[2]:sl[2] + sl[3]])
        return out, (None if noise is None else list(ts[n0:]))
    return flat, rebuild


class GraphedTrainStep:
    N_POS = 13      # positional arguments of StateAwareRecursiveTransformer.forward (reference model.py:1027-1030)

    def __init__(self, model, example_args, gumbel_noise=None, reducer=None, warmup: int = 3, feature_dtype=None):
        """``feature_dtype=torch.bfloat16``: the video features live in the input arena (pinned host, staging, static) as
        bf16 - half the bytes of every upload.  The first kernel that touches them (LN(3072), statistics in fp32) reads
        either type; the loader writes bf16 directly (``FeatureReader`` takes bf16 views), ``load()`` converts."""
        assert len(example_args) == self.N_POS
        self.model, self.reducer = model, reducer
        dev = next(model.parameters()).device
        assert dev.type == "cuda", "GraphedTrainStep needs a CUDA model"
        self.device = dev
        flat, rebuild = _flatten(list(example_args), gumbel_noise)
        feats = {id(t) for t in example_args[1]} if _is_tensor_list(example_args[1]) else set()
        Lv = model.config.max_v_len
        self.feature_dtype = feature_dtype or (example_args[1][0].dtype if feats else torch.float32)
        spec = [torch.empty(t.shape, dtype=self.feature_dtype, device="meta") if id(t) in feats else t for t in flat]
        self.arena = InputArena(spec, dev, used_rows={i: Lv for i, t in enumerate(flat) if id(t) in feats})
        self.arena.feature_dtype = {torch.float32: "fp32", torch.bfloat16: "bf16"}.get(self.feature_dtype, str(self.feature_dtype))
        sv = self.arena.static_views()
        for d, t in zip(sv, flat):
            d.copy_(t)
        self.static, self.noise = rebuild(sv)
        self._rebuild = rebuild
        a = self.static
        self.plan = model.make_plan(a[8], a[9], a[10], a[11], dev)
        self.sig = self.plan.signature()
        # pre-build every cached index array outside the capture
        for C in (1, 2, 3):
            self.plan.cross_lay(C, dev, model.config.hidden_size)
        self.graph = torch.cuda.CUDAGraph()
        self.loss = None
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._eager()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.outputs = None         # drop the warm-up autograd graphs (their AccumulateGrad nodes belong to `side`)
        gc.collect()
        Fn.be().clear_weight_caches()
        Fn.WEIGHTS.invalidate()     # the refresh of the compute-dtype weight copies (one multi-tensor cast) is captured too
        Fn.WEIGHTS.prepare()        # ... its pointer tables are uploaded before the capture starts
        model.zero_grad(set_to_none=True)
        feat_table = Fn.be()._feat_ptrs(self.static[1])[0] if feats else None     # built by the warm-up steps, read by the graph
        with torch.cuda.graph(self.graph):
            self.loss = self._eager(zero=False)
        torch.cuda.synchronize(dev)
        # the graph reads (GEMM operands) and rewrites (the captured weight refresh) the cached weight copies by address
        self._weight_buffers = Fn.WEIGHTS.buffers() + list(Fn.be()._w_cache.values())
        self._feat_table = None
        if feats and os.environ.get("SVPC_ARENA_ZEROCOPY", "1") != "0":
            # the table of feature-block addresses the captured LN(3072) kernels read (kept alive here: the graph reads it)
            self._feat_table = feat_table
            if Fn.be()._feat_ptrs(self.static[1])[0] is feat_table:          # the very tensor the captured kernels were given
                self.arena.enable_zero_copy(feat_table, [i for i, t in enumerate(flat) if id(t) in feats])

    def _clone(self, a):
        if isinstance(a, torch.Tensor):
            return a.to(self.device).clone()
        if _is_tensor_list(a):
            return [t.to(self.device).clone() for t in a]
        return a

    def _eager(self, zero=True):
        if zero:
            self.model.zero_grad(set_to_none=True)
        out = self.model(*self.static, gumbel_noise=self.noise, plan=self.plan)
        loss = out[0]
        loss.backward()
        if self.reducer is not None:
            self.reducer.finish()
        self.outputs = out
        return loss

    # ------------------------------------------------------------------------------------------
    def matches(self, args):
        return self.model.plan_signature(args[8], args[9], args[10]) == self.sig

    def load(self, args, gumbel_noise=None):
        """Copy a batch (host-pinned or device tensors) into the static buffers; asynchronous on the current stream."""
        self.arena.reset_zero_copy()
        for dst, src in zip(self.static, args):
            if isinstance(dst, torch.Tensor):
                dst.copy_(src, non_blocking=True)
            elif _is_tensor_list(dst):
                for d, s in zip(dst, src):
                    d.copy_(s, non_blocking=True)
        if gumbel_noise is not None and self.noise is not None:
            for d, s in zip(self.noise, gumbel_noise):
                d.copy_(s, non_blocking=True)
        if args[9] is not self.static[9] and self.plan.mode != "video":
            if not self.plan.update_tokens(args[9]):
                raise ValueError("ingr_id_dict has a different number of copy tokens than the captured batch: re-capture")
            self.static[9] = args[9]

    def run(self):
        self.graph.replay()
        return self.loss

    # ------------------------------------------------------------------------------------------
    # host -> device input pipeline: the upload of batch i+1 overlaps the replay of batch i
    def host_inputs(self):
        """(args, noise) whose tensors are views of the arena's pinned host buffer: a loader that collates into them
        makes ``prefetch`` a single H2D memcpy."""
        if getattr(self, "_host_args", None) is None:
            self._host_args = self._rebuild(self.arena.host_views())
            self._host_args[0][9] = self.static[9]
        return self._host_args

    def prefetch(self, args, gumbel_noise=None):
        """Start the H2D copy of a batch from pinned host memory into a staging set on a dedicated copy stream.  The
        next ``run_prefetched()`` moves it into the static buffers (one D2D pass) and replays the step, so the upload of
        the following batch can already be in flight while this one computes."""
        self._pf = getattr(self, "_pf", None) or _Prefetcher(self.device)
        ha = getattr(self, "_host_args", None)
        if ha is not None and args is ha[0] and (gumbel_noise is None or gumbel_noise is ha[1]):
            # the batch already sits in the pinned arena: a few large H2D copies (dead feature rows skipped), one D2D
            self._pf.stage_arena(self.arena)
            self._pf_dict = args[9]
            return
        pairs = []
        for dst, src in zip(self.static, args):
            if isinstance(dst, torch.Tensor):
                pairs.append((dst, src))
            elif _is_tensor_list(dst):
                pairs += list(zip(dst, src))
        if gumbel_noise is not None and self.noise is not None:
            pairs += list(zip(self.noise, gumbel_noise))
        self._pf.stage(pairs)
        self._pf_dict = args[9]

    def loss_to_host(self):
        """Asynchronous D2H read of the step that was just enqueued: copies the loss into the next slot of a small pinned
        ring and returns a handle whose ``get()`` waits for that copy only.  Lets a training loop enqueue step i+1 before
        it looks at the loss of step i (the static loss buffer is overwritten by the next replay), so the device never
        waits for the host between steps."""
        ring = getattr(self, "_loss_ring", None)
        if ring is None:
            ring = self._loss_ring = {"buf": torch.zeros(4, dtype=torch.float32).pin_memory(), "i": 0,
                                      "ev": [torch.cuda.Event() for _ in range(4)],
                                      "flags": torch.zeros(4, 2, 4, dtype=torch.int32).pin_memory()}
        i = ring["i"]
        ring["i"] = (i + 1) % 4
        slot, ev = ring["buf"][i:i + 1], ring["ev"][i]
        slot.copy_(self.loss.detach().reshape(1), non_blocking=True)
        flags = None
        if self.model.config.model_mode == "full":
            # the persistent BiLSTM sweeps' barrier time-out flag travels with the loss: a time-out means wrong states
            flags = ring["flags"][i]
            flags.copy_(Fn.be().lstm_sync(self.device), non_blocking=True)
        ev.record(torch.cuda.current_stream(self.device))
        return _HostScalar(slot, ev, flags)

    def run_prefetched(self):
        if not isinstance(self._pf.pairs, InputArena):
            self.arena.reset_zero_copy()        # per-tensor staging lands in the static views
        self._pf.commit()
        d = self._pf_dict
        if d is not None and d is not self.static[9] and self.plan.mode != "video":
            if not self.plan.update_tokens(d):
                raise ValueError("ingr_id_dict has a different number of copy tokens than the captured batch: re-capture")
            self.static[9] = d
        loss = self.run()
        self.arena.mark_read()
        return loss


class _HostScalar:
    def __init__(self, slot, event, lstm_flags=None):
        self.slot, self.event, self.lstm_flags = slot, event, lstm_flags

    def get(self) -> float:
        self.event.synchronize()
        if self.lstm_flags is not None and int(self.lstm_flags[:, 2].sum()) != 0:
            Fn.be().check_lstm_timeout(self.lstm_flags)
        return float(self.slot[0])


class _Prefetcher:
    """Double buffer between pinned host memory and the static inputs of a captured graph.

    ``stage``  (copy stream): waits until the previous staging contents were consumed, then H2D into the staging set.
    ``commit`` (compute stream): waits for the staged copy, D2D staging -> static (HBM-speed, ~0.1 ms for 300 MB)."""

    def __init__(self, device):
        self.device = device
        self.stream = torch.cuda.Stream(device=device)
        self.staged = torch.cuda.Event()
        self.consumed = None
        self.bufs = {}
        self.pairs = None

    def stage(self, pairs, bufs=None):
        cur = torch.cuda.current_stream(self.device)
        if self.consumed is not None:
            self.stream.wait_event(self.consumed)
        else:
            self.stream.wait_stream(cur)
        staged = []
        with torch.cuda.stream(self.stream):
            for i, (dst, src) in enumerate(pairs):
                b = bufs[i] if bufs is not None else self.bufs.get(dst.data_ptr())
                if b is None:
                    b = self.bufs[dst.data_ptr()] = torch.empty_like(dst)
                b.copy_(src, non_blocking=True)
                staged.append((dst, b))
            self.staged.record(self.stream)
        self.pairs = staged

    def stage_arena(self, arena):
        cur = torch.cuda.current_stream(self.device)
        if self.consumed is not None:
            self.stream.wait_event(self.consumed)
        else:
            self.stream.wait_stream(cur)
        if arena.zc_table is not None and arena.read_done[1 - arena.cur] is not None:
            self.stream.wait_event(arena.read_done[1 - arena.cur])      # the last replay that read the target feature buffer
        with torch.cuda.stream(self.stream):
            arena.upload()
            self.staged.record(self.stream)
        self.pairs = arena

    def commit(self):
        assert self.pairs is not None, "commit() without a staged batch"
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self.staged)
        if isinstance(self.pairs, InputArena):
            self.pairs.commit()
        elif len(self.pairs) == 1:
            self.pairs[0][0].copy_(self.pairs[0][1], non_blocking=True)
        else:
            torch._foreach_copy_([d for d, _ in self.pairs], [b for _, b in self.pairs])
        self.consumed = torch.cuda.Event()
        self.consumed.record(cur)
        self.pairs = None


class GraphedDecoder:
    """CUDA-graph replay of ``svpc_b200.decode.greedy_decode`` for one batch signature (clips / ingredients / OOV words
    per video): the ~700 launches of a 25-position decode become one graph launch.

        dec = GraphedDecoder(model, example_inputs)      # warm-up + capture
        dec.load(inputs)                                 # copy a batch (pinned host or device tensors) into the static buffers
        seqs = dec.run()                                 # list of (S_b, max_t_len) int64 extended ids (views of a static buffer)
    """

    TENSOR_LISTS = (0, 1, 2, 3)      # input_ids_list, video_features_list, input_masks_list, token_type_ids_list
    TENSORS = (4, 5, 6)              # ingr_input_ids, ingr_masks, ingr_sep_masks (lists of lists in translate.py:57-70)

    def __init__(self, model, example_inputs, warmup: int = 2, feature_dtype=None):
        """``feature_dtype=torch.bfloat16``: the clip features live in the input arena as bf16 (half the upload; see
        GraphedTrainStep).  The static inputs are views of ONE device arena whose pinned host twin (``host_inputs()``) is
        what a loader fills: ``prefetch`` then uploads it in a few large copies that skip the 25 text rows of every
        (125, 3072) clip block."""
        from .decode import greedy_decode
        self._decode = greedy_decode
        self.model = model
        dev = next(model.parameters()).device
        assert dev.type == "cuda", "GraphedDecoder needs a CUDA model"
        self.device = dev
        st = list(example_inputs)
        for i in self.TENSORS:
            st[i] = torch.as_tensor(st[i]).to(dtype=torch.long)
        flat, slots = [], []
        for i in self.TENSOR_LISTS:
            slots.append((i, len(flat), len(st[i])))
            flat += list(st[i])
        for i in self.TENSORS:
            slots.append((i, len(flat), 0))
            flat.append(st[i])
        self._slots = slots
        feats = {id(t) for t in st[1]}
        self.feature_dtype = feature_dtype or st[1][0].dtype
        spec = [torch.empty(t.shape, dtype=self.feature_dtype, device="meta") if id(t) in feats else t for t in flat]
        Lv = model.config.max_v_len
        self.arena = InputArena(spec, dev, used_rows={k: Lv for k, t in enumerate(flat) if id(t) in feats})
        self.arena.feature_dtype = {torch.float32: "fp32", torch.bfloat16: "bf16"}.get(self.feature_dtype, str(self.feature_dtype))
        sv = self.arena.static_views()
        for d, t in zip(sv, flat):
            d.copy_(t)
        self._place(st, sv)
        self.static = st
        self._host = None
        mode = model.config.model_mode
        steps = [int(x) for x in st[11]]
        B = len(steps)
        if mode == "video":
            ni, n_oov, dicts = [0] * B, [0] * B, [{} for _ in range(B)]
        else:
            ni, n_oov, dicts = model.ingredient_counts(example_inputs[6], st[7]), [len(d) for d in st[8]], st[7]
        self.plan = model.make_plan_raw(steps, ni, n_oov, dicts, dev)
        self.sig = (tuple(steps), tuple(ni), tuple(n_oov))
        self.graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(max(1, warmup)):
                self._decode(model, tuple(self.static), plan=self.plan)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.out = self._decode(model, tuple(self.static), plan=self.plan)
        torch.cuda.synchronize(dev)

    def _place(self, dst, views):
        for i, o, n in self._slots:
            dst[i] = list(views[o:o + n]) if n else views[o]

    def host_inputs(self):
        """The 12-tuple whose tensors are views of the arena's pinned host buffer (lists / dicts shared with the static
        inputs): fill them in place, then ``prefetch(host_inputs())`` is a handful of large H2D copies."""
        if self._host is None:
            h = list(self.static)
            self._place(h, self.arena.host_views())
            self._host = h
        return self._host

    def matches(self, inputs):
        mode = self.model.config.model_mode
        steps = tuple(int(x) for x in inputs[11])
        if mode == "video":
            return steps == self.sig[0]
        return (steps, tuple(len(d) for d in inputs[7]), tuple(len(d) for d in inputs[8])) == self.sig

    def load(self, inputs):
        for i in self.TENSOR_LISTS:
            for d, s in zip(self.static[i], inputs[i]):
                d.copy_(s, non_blocking=True)
        for i in self.TENSORS:
            self.static[i].copy_(torch.as_tensor(inputs[i]), non_blocking=True)
        if inputs[7] is not self.static[7] and self.plan.mode != "video" and self.model.config.model_mode != "video":
            if not self.plan.update_tokens(inputs[7]):
                raise ValueError("ingr_id_dict has a different number of copy tokens than the captured batch: re-capture")
            self.static[7] = inputs[7]

    def run(self):
        self.graph.replay()
        return self.out

    def prefetch(self, inputs):
        """H2D of a batch on the copy stream into a staging set (see GraphedTrainStep.prefetch)."""
        self._pf = getattr(self, "_pf", None) or _Prefetcher(self.device)
        if self._host is not None and inputs is self._host:
            self._pf.stage_arena(self.arena)          # the batch already sits in the pinned arena
            self._pf_dict = inputs[7]
            return
        pairs = []
        for i in self.TENSOR_LISTS:
            pairs += list(zip(self.static[i], inputs[i]))
        for i in self.TENSORS:
            pairs.append((self.static[i], torch.as_tensor(inputs[i])))
        self._pf.stage(pairs)
        self._pf_dict = inputs[7]

    def run_prefetched(self):
        self._pf.commit()
        d = self._pf_dict
        if d is not None and d is not self.static[7] and self.plan.mode != "video" and self.model.config.model_mode != "video":
            if not self.plan.update_tokens(d):
                raise ValueError("ingr_id_dict has a different number of copy tokens than the captured batch: re-capture")
            self.static[7] = d
        return self.run()


class GraphedStepCache:
    """Captured training steps for RAGGED data: a small LRU of ``GraphedTrainStep`` keyed by the batch signature (clips,
    ingredients, OOV width and copy tokens per video - everything the launch shapes depend on).

    A collate function that draws batches from length buckets produces a handful of signatures; each of them is captured
    the ``capture_after``-th time it is seen and replayed from then on, every other batch runs the eager step (same
    kernels, launched from Python).  ``stats()`` gives the hit rate.

        cache = GraphedStepCache(model)
        loss = cache.step(args, gumbel_noise)     # gradients in p.grad exactly as after loss.backward()
    """

    def __init__(self, model, max_graphs: int = 8, capture_after: int = 2, reducer=None, feature_dtype=None):
        self.model, self.reducer, self.feature_dtype = model, reducer, feature_dtype
        self.max_graphs, self.capture_after = int(max_graphs), int(capture_after)
        self.graphs = {}            # signature -> GraphedTrainStep (insertion order = LRU order)
        self.seen = {}
        self.hits = self.misses = self.captures = 0

    def _eager(self, args, gumbel_noise):
        self.model.zero_grad(set_to_none=True)
        loss = self.model(*args, gumbel_noise=gumbel_noise)[0]
        loss.backward()
        if self.reducer is not None:
            self.reducer.finish()
        # detached: a caller holding the loss must not keep this step's autograd graph (and its AccumulateGrad nodes, which
        # are tied to this stream) alive into a later capture
        return loss.detach()

    def step(self, args, gumbel_noise=None):
        sig = self.model.plan_signature(args[8], args[9], args[10])
        g = self.graphs.get(sig)
        if g is not None:
            self.graphs[sig] = self.graphs.pop(sig)           # most recently used last
            g.load(args, gumbel_noise)
            self.hits += 1
            return g.run()
        self.misses += 1
        n = self.seen[sig] = self.seen.get(sig, 0) + 1
        if n >= self.capture_after:
            if len(self.graphs) >= self.max_graphs:
                self.graphs.pop(next(iter(self.graphs)))      # least recently used
            dev = next(self.model.parameters()).device
            to = lambda x: x.to(dev) if isinstance(x, torch.Tensor) else ([t.to(dev) for t in x] if _is_tensor_list(x) else x)
            g = GraphedTrainStep(self.model, [to(a) for a in args], gumbel_noise=None if gumbel_noise is None else [t.to(dev) for t in gumbel_noise],
                                 reducer=self.reducer, warmup=1, feature_dtype=self.feature_dtype)
            self.graphs[sig] = g
            self.captures += 1
            return g.run()
        return self._eager(args, gumbel_noise)

    def stats(self):
        tot = self.hits + self.misses
        return {"steps": tot, "graph_hits": self.hits, "hit_rate": round(self.hits / tot, 4) if tot else 0.0,
                "captures": self.captures, "signatures_seen": len(self.seen), "graphs_resident": len(self.graphs)}
