"""Autograd wiring of the SVPC operators.

Each ``torch.autograd.Function`` here pairs one forward operator of ``svpc_b200.backend`` with its hand-written
backward operator; torch is only the tape (and the owner of device memory).  Nothing in this file computes on the
host: every tensor operation on activations is a backend (CUDA) call.
"""
from __future__ import annotations

import weakref
from contextlib import nullcontext as _nullcontext

import torch
from torch.autograd import Function

from . import backend as B
from .backend import ACT_NONE, ACT_RELU, ACT_GELU, ACT_SIGMOID


def be():
    return B.get()


# ------------------------------------------------------------------------------------------------
# weights in GEMM layout / dtype, refreshed when the optimiser has touched the parameters
# ------------------------------------------------------------------------------------------------
class WeightCache:
    """Concatenated (along the output dim) copies of weight parameters in the compute dtype.

    Keyed by the identity of the parameters, invalidated by their autograd version counters (an in-place optimiser
    update bumps them), so the casts run once per optimiser step, not once per use.  ``refresh_all`` brings every stale
    entry up to date with ONE multi-tensor launch that writes into the entries' existing buffers (stable addresses: a
    captured CUDA graph keeps reading them); a key seen for the first time is built by ``get`` with single casts."""

    CHUNK = 16384

    def __init__(self):
        self.store = {}          # key -> [version tuple or None, tensor, weak refs to the params, dtype]
        self._tables = {}        # (dtype, keys of the stale set) -> device tables of the multi-tensor launch

    @staticmethod
    def _version(ws):
        return tuple(w._version for w in ws) + tuple(w.data_ptr() for w in ws)

    @staticmethod
    def _params(e):
        """The entry's parameters, or None when one of them no longer exists."""
        ws = [r() for r in e[2]]
        return None if any(w is None for w in ws) else ws

    def _purge(self):
        dead = [k for k, e in self.store.items() if self._params(e) is None]
        for k in dead:
            del self.store[k]
        if dead:
            self._tables.clear()

    def get(self, ws, dtype):
        if len(ws) == 1 and ws[0].dtype == dtype and ws[0].is_contiguous():
            return ws[0].detach()
        key = tuple(id(w) for w in ws) + (dtype,)
        ver = self._version(ws)
        hit = self.store.get(key)
        if hit is not None and hit[0] == ver and self._params(hit) is not None:
            return hit[1]
        with torch.no_grad():
            parts = [be().cast(w.detach().contiguous(), dtype) for w in ws]
            t = parts[0] if len(parts) == 1 else torch.cat(parts, dim=0)
        if hit is not None and self._params(hit) is not None and hit[1].shape == t.shape and hit[1].device == t.device \
                and hit[1].stride() == t.stride():
            # stale entry of the same parameters (optimiser step, EMA swap, a decode between two replays): rewrite the
            # EXISTING buffer - a captured CUDA graph holds its address (cast_multi of refresh_all, every GEMM operand)
            with torch.no_grad():
                hit[1].copy_(t)
            hit[0] = ver
            return hit[1]
        self._purge()            # a first sight / reshaped / moved key: also forget entries of parameters that are gone
        self.store[key] = [ver, t, [weakref.ref(w) for w in ws], dtype]
        return t

    def get_split(self, ws):
        """The (concatenated) fp32 weight as split-bf16 planes (``Split2``): operand of the fp32 path's tensor-core route.
        Cached and refreshed in place like the compute-dtype copies."""
        from .cuda_backend import Split2
        src = self.get(ws, torch.float32)
        key = tuple(id(w) for w in ws) + ("split2",)
        ver = self._version(ws)
        hit = self.store.get(key)
        N, K = src.shape
        ld = (K + 7) // 8 * 8
        if hit is None or self._params(hit) is None or hit[1].shape != (2, N, ld) or hit[1].device != src.device:
            self._purge()
            hit = self.store[key] = [None, torch.zeros(2, N, ld, dtype=torch.bfloat16, device=src.device),
                                     [weakref.ref(w) for w in ws], "split2"]
        if hit[0] != ver:
            with torch.no_grad():
                be().split2_into(src, hit[1])
            hit[0] = ver
        return Split2(hit[1][0, :, :K], hit[1][1, :, :K], src)

    def invalidate(self):
        """Mark every entry stale but keep its buffer (next ``refresh_all`` rewrites them in place)."""
        for e in self.store.values():
            e[0] = None

    def prepare(self):
        """Build the device tables of the next ``refresh_all`` without launching (a CUDA-graph capture cannot upload them)."""
        self.refresh_all(launch=False)

    def refresh_all(self, launch=True):
        """One launch per compute dtype for all entries whose parameters changed since they were cast."""
        stale, splits = [], []
        for k, e in self.store.items():
            ws = self._params(e)
            if ws is None or e[0] == self._version(ws):
                continue
            if e[3] == "split2":
                splits.append((e, ws))
                continue
            if all(w.is_cuda and w.dtype == torch.float32 and w.dim() <= 2 and w.is_contiguous() for w in ws) and \
                    e[1].device == ws[0].device:
                stale.append((k, e, ws))
        if not hasattr(be(), "cast_multi"):
            return
        for dtype in {e[3] for _, e, _ in stale}:
            group = [(k, e, ws) for k, e, ws in stale if e[3] == dtype]
            tkey = (dtype,) + tuple(k for k, _, _ in group) + tuple(w.data_ptr() for _, _, ws in group for w in ws)
            tab = self._tables.get(tkey)
            if tab is None:
                if len(self._tables) > 16:
                    self._tables.clear()
                srcs, dsts, numel, cols, lds = [], [], [], [], []
                for _, e, ws in group:
                    t, off = e[1], 0                      # off: element offset of the next part inside the entry's buffer
                    esz = t.element_size()
                    for w in ws:
                        c = w.shape[1] if (w.dim() == 2 and t.dim() == 2) else w.numel()
                        ld = t.stride(0) if (w.dim() == 2 and t.dim() == 2) else c
                        srcs.append(w.data_ptr())
                        dsts.append(t.data_ptr() + off * esz)
                        numel.append(w.numel())
                        cols.append(c)
                        lds.append(ld)
                        off += w.shape[0] * t.stride(0) if (w.dim() == 2 and t.dim() == 2) else w.numel()
                tab = self._tables[tkey] = be().cast_multi_tables(srcs, dsts, numel, cols, lds, self.CHUNK,
                                                                   group[0][2][0].device)
            if launch:
                be().cast_multi(tab, dtype, self.CHUNK)
                for _, e, ws in group:
                    e[0] = self._version(ws)
        if launch:
            # split-bf16 planes of the fp32 path: one split launch per entry, from the (just refreshed) fp32 source
            for e, ws in splits:
                if all(w.is_cuda for w in ws):
                    with torch.no_grad():
                        be().split2_into(self.get(ws, torch.float32), e[1])
                    e[0] = self._version(ws)

    def buffers(self):
        """Every cached tensor and every device table of the multi-tensor refresh (a captured graph keeps them alive: it
        reads and rewrites their addresses, and its captured refresh launch reads the pointer tables)."""
        out = [e[1] for e in self.store.values()]
        for tab in self._tables.values():
            out += [v for v in tab.values() if isinstance(v, torch.Tensor)]
        return out

    def clear(self):
        self.store.clear()
        self._tables.clear()


WEIGHTS = WeightCache()
# fused backward hand-offs (``Link``); SVPC_LINKS=0 falls back to one element-wise pass per activation / fan-out
import os as _os
LINKS = _os.environ.get("SVPC_LINKS", "1") != "0"
# fusing the activation / dropout backward into ln_bwd (grad_link=) saves one read of the gradient but makes the LayerNorm
# backward kernel carry a third accumulator set: measured slower (34 us against 11.5 + 14 us at 19200 x 768), so off by default
LN_FUSE = _os.environ.get("SVPC_LN_FUSE", "0") != "0"


class _WgradStream:
    """Weight-gradient GEMMs on a second CUDA stream.

    In the backward pass of a Linear the input gradient is on the critical path (the next node needs it), the weight
    gradient is a leaf: nothing consumes it before the optimiser / the gradient all-reduce.  ``LinearFn.backward`` therefore
    enqueues it on a side stream that waits for the operands (``fork``); the main stream goes on with the dgrad chain, so the
    two overlap - which matters where the chain is made of kernels that leave most SMs idle (the 192-row step-level modules:
    24-72 CTAs per GEMM, the entity-state recurrence: 16 CTAs).  The main stream re-joins once, when the backward pass ends
    (an autograd engine callback), and before every gradient bucket is flattened for its all-reduce (``join``).  Operands are
    kept referenced until the join so that the caching allocator cannot hand their memory to a main-stream kernel while the
    side stream still reads it.  Works unchanged under CUDA-graph capture (the fork / join become graph edges).
    SVPC_WGRAD_STREAM=0 keeps everything on one stream."""

    def __init__(self):
        self.enabled = _os.environ.get("SVPC_WGRAD_STREAM", "1") != "0"
        self.eager_too = _os.environ.get("SVPC_WGRAD_STREAM", "1") == "2"
        self.streams = {}
        self.keep = []
        self.active = set()
        self.cb_queued = False

    def fork(self, device):
        # outside a CUDA-graph capture the step is bound by the interpreter issuing launches, not by the GPU: the stream
        # switches would only add host time there (SVPC_WGRAD_STREAM=2 forces the side stream in eager mode too)
        if self.eager_too is False and not torch.cuda.is_current_stream_capturing():
            return None
        main = torch.cuda.current_stream(device)
        side = self.streams.get(device)
        if side is None:
            side = self.streams[device] = torch.cuda.Stream(device=device)
        if side == main:
            return None
        side.wait_stream(main)
        self.active.add(device)
        if not self.cb_queued:
            torch.autograd.Variable._execution_engine.queue_callback(self.join)
            self.cb_queued = True
        return side

    def join(self):
        """The current stream of every device with outstanding side-stream work waits for it."""
        for device in list(self.active):
            torch.cuda.current_stream(device).wait_stream(self.streams[device])
        self.active.clear()
        self.keep.clear()
        self.cb_queued = False


class Link:
    """Hand-off between two nodes that are neighbours in the BACKWARD pass, so that a gradient is produced in its final form
    by the kernel that computes it instead of by extra element-wise passes (autograd's fan-out adds, activation backward):

    * skip connections.  x feeds a Linear and, further on, is added back as a residual (reference model.py:232, 288, 645,
      653).  The node that owns the residual add (``res_link=``: LinearFn with ``residual=x``, LayerNormFn with ``res=x``)
      leaves the residual's gradient in ``res_grad`` and returns None for it; the Linear that consumes x (``in_link=``) adds
      it in the epilogue of its input-gradient GEMM.  The producer always runs first: it is downstream of the consumer.
    * activation / dropout backward.  A Linear with ``out_link=`` records its activation, reference tensor and dropout site;
      whoever computes the gradient of its output - the LayerNorm behind it (``grad_link=``, fused into ``ln_bwd``) or the
      next Linear (``in_link=``, fused into that GEMM's epilogue) - writes dz = dy * dropscale * act'(ref) and the bias
      gradient here, and LinearFn.backward skips its own ``act_bwd`` pass."""
    __slots__ = ("act", "ref", "drop", "has_bias", "dz", "dbias", "res_grad")

    def __init__(self):
        self.act, self.ref, self.drop, self.has_bias = ACT_NONE, None, None, False
        self.dz = self.dbias = self.res_grad = None

    def take_res_grad(self):
        g, self.res_grad = self.res_grad, None
        return g


WGRAD = _WgradStream()


class _GradSlots:
    """Where weight-gradient GEMMs write: a data-parallel reducer (svpc_b200.ddp.GradAllReducer) registers itself here and
    hands out slots of its flat all-reduce buckets, so the summed gradients never need a flatten copy."""

    def __init__(self):
        self.reducers = []

    def register(self, r):
        if r not in self.reducers:
            self.reducers.append(r)

    def unregister(self, r):
        if r in self.reducers:
            self.reducers.remove(r)

    def get(self, params):
        for r in self.reducers:
            if r.enabled and params and params[0] in r.bucket_of:
                return r.grad_slot(params)
        return None


GRAD_SLOTS = _GradSlots()


class _LinMeta:
    __slots__ = ("n_x", "n_w", "has_bias", "act", "out_dtype", "cd", "w_is_kn", "has_res", "cache", "drop", "in_link",
                 "out_link", "res_link")

    def __init__(self, **kw):
        for k, v in kw.items():
            setattr(self, k, v)


class LinearFn(Function):
    """y = act([x_0 | x_1 | ...] @ [w_0 ; w_1 ; ...]^T + [b_0 ; ...]) (+ residual).

    Inputs are concatenated along K, weights along N (fused QKV / KV projections; ``W2`` of the reasoner takes the
    two-piece input [h ; ac], reference model.py:803).  ``w_is_kn``: the single weight is stored (K, N) and used
    untransposed (``x @ W``; the action-embedding table, model.py:798)."""

    @staticmethod
    def forward(ctx, meta, residual, *args):
        xs = args[:meta.n_x]
        ws = args[meta.n_x:meta.n_x + meta.n_w]
        bs = args[meta.n_x + meta.n_w:]
        cd = meta.cd
        split = cd == torch.float32 and getattr(be(), "tc3", False) and not meta.w_is_kn
        if split and meta.cache:
            W = WEIGHTS.get_split(ws)                 # fp32 path on the tensor cores: hi / lo bf16 planes, cached per version
        else:
            W = WEIGHTS.get(ws, cd) if meta.cache else be().cast(ws[0].detach().contiguous(), cd)
        bias = None
        if meta.has_bias:
            bias = bs[0].detach() if len(bs) == 1 else WEIGHTS.get(bs, torch.float32)
        xs_c = [x if x.dtype == cd else be().cast(x, cd) for x in xs]
        if split:
            xs_c = [be().maybe_split(x) for x in xs_c]   # split once, reused by the weight-gradient GEMM of the backward
        want_pre = meta.act == ACT_GELU
        if meta.w_is_kn:
            assert meta.n_x == 1 and meta.n_w == 1
            out = be().gemm(xs_c[0], W.t(), meta.out_dtype, bias, meta.act, residual, want_pre=want_pre, drop=meta.drop)
        elif meta.n_x == 1:
            out = be().gemm(xs_c[0], W, meta.out_dtype, bias, meta.act, residual, want_pre=want_pre, drop=meta.drop)
        else:
            assert meta.act == ACT_NONE and residual is None and meta.drop is None
            k0 = 0
            out = None
            for j, x in enumerate(xs_c):
                k1 = k0 + x.shape[1]
                if j == 0:
                    out = be().gemm(x, W[:, k0:k1], meta.out_dtype, bias)
                else:
                    be().gemm(x, W[:, k0:k1], meta.out_dtype, None, out=out, accumulate=True)
                k0 = k1
        pre = None
        if want_pre:
            out, pre = out
        ctx.meta = meta
        ctx.shapes = [tuple(w.shape) for w in ws]
        ctx.x_dtypes = [x.dtype for x in xs]
        ref = pre if meta.act == ACT_GELU else (out if meta.act in (ACT_RELU, ACT_SIGMOID) else None)
        ctx.save_for_backward(ref)
        ctx.ops = (W, xs_c)          # tensors or Split2 operands (not autograd leaves: plain references)
        ctx.split = split
        ctx.w_leaf = all(w.is_leaf for w in ws)
        ctx.w_params = list(ws) if ctx.w_leaf else []
        if meta.out_link is not None:
            L = meta.out_link
            L.act, L.ref, L.drop, L.has_bias = meta.act, ref, meta.drop, meta.has_bias
        return out

    @staticmethod
    def backward(ctx, dy):
        meta = ctx.meta
        (ref,) = ctx.saved_tensors
        W, xs_c = ctx.ops
        cd = meta.cd
        if not (dy.dim() == 2 and dy.stride(1) == 1 and meta.act == ACT_NONE and meta.drop is None):
            dy = dy.contiguous()                      # (a row-pitched gradient - the LM head's - is consumed as it lies)
        need = ctx.needs_input_grad
        d_res = dy if (meta.has_res and need[1]) else None
        L = meta.out_link
        if L is not None and L.dz is not None:
            dz, dbias = L.dz, L.dbias                                         # produced by ln_bwd / the next Linear's dgrad
            L.dz = L.dbias = L.ref = None
        elif meta.act != ACT_NONE or meta.drop is not None:
            dz, dbias = be().act_bwd(dy, ref, meta.act, drop=meta.drop)     # dy * dropout scale * act'
        else:
            dz = dy
            dbias = be().colsum(dy) if meta.has_bias else None
        dzc = dz if dz.dtype == cd else be().cast(dz, cd)
        if ctx.split:
            dzc = be().maybe_split(dzc)              # one split serves the input- and the weight-gradient GEMM
        n_x, n_w = meta.n_x, meta.n_w
        # ---- skip-connection gradient of x: added in the epilogue of the input-gradient GEMM instead of by autograd
        extra = None
        if meta.res_link is not None and d_res is not None:
            if meta.res_link is meta.in_link:
                extra = d_res                        # residual = this Linear's own input (decoder output block, model.py:662)
            else:
                assert meta.res_link.res_grad is None
                meta.res_link.res_grad = d_res
            d_res = None
        Lin = meta.in_link
        grad_act = None
        if Lin is not None:
            if extra is None:
                extra = Lin.take_res_grad()
            if Lin.act != ACT_NONE and Lin.drop is None and Lin.ref is not None and n_x == 1 and need[2] and \
                    Lin.ref.dtype == ctx.x_dtypes[0]:
                grad_act = (Lin.act, Lin.ref)        # x = act(pre): write d(pre) directly
        # ---- input gradients
        dxs = [None] * n_x
        k0 = 0
        for j, x in enumerate(xs_c):
            k1 = k0 + x.shape[1]
            if need[2 + j]:
                res_j = extra if j == 0 else None
                if res_j is not None and res_j.dtype != ctx.x_dtypes[j]:
                    res_j = res_j.to(ctx.x_dtypes[j])
                if meta.w_is_kn:
                    dxs[j] = be().gemm(dzc, W, ctx.x_dtypes[j], residual=res_j, grad_act=grad_act)                  # dz (M,N) @ W(K,N)^T
                else:
                    dxs[j] = be().gemm(dzc, W[:, k0:k1].t(), ctx.x_dtypes[j], residual=res_j, grad_act=grad_act)   # dz (M,N) @ W(N,K)
            k0 = k1
        if grad_act is not None:
            Lin.dz = dxs[0]
            Lin.dbias = be().colsum(dxs[0]) if Lin.has_bias else None
        # ---- weight gradients (fp32, written straight into one buffer in the parameters' layout) - on the side stream
        dws = [None] * n_w
        if any(need[2 + n_x:2 + n_x + n_w]):
            dW = GRAD_SLOTS.get(ctx.w_params) if (GRAD_SLOTS.reducers and ctx.w_leaf) else None   # a reducer's bucket slot
            if dW is not None and meta.w_is_kn and tuple(dW.shape) != (xs_c[0].shape[1], dzc.shape[1]):
                dW = None
            if dW is None:
                if meta.w_is_kn:
                    dW = torch.empty(xs_c[0].shape[1], dzc.shape[1], dtype=torch.float32, device=dy.device)
                else:
                    dW = torch.empty(W.shape[0], W.shape[1], dtype=torch.float32, device=dy.device)
            # side stream only when nothing can read dW before the join: the weights are autograd leaves without an
            # accumulated .grad (AccumulateGrad then just adopts the tensor; a non-leaf weight - the LN-folded video
            # projection in eval mode - hands dW to further autograd nodes at once)
            side = WGRAD.fork(dy.device) if (WGRAD.enabled and dy.is_cuda and ctx.w_leaf and
                                             all(w.grad is None for w in ctx.w_params)) else None
            with (torch.cuda.stream(side) if side is not None else _nullcontext()):
                if meta.w_is_kn:
                    be().gemm(xs_c[0].t(), dzc.t(), torch.float32, out=dW)             # (K,N) = x^T dz
                    dws[0] = dW
                else:
                    k0 = 0
                    for x in xs_c:
                        k1 = k0 + x.shape[1]
                        be().gemm(dzc.t(), x.t(), torch.float32, out=dW[:, k0:k1])
                        k0 = k1
                    n0 = 0
                    for i, shp in enumerate(ctx.shapes):
                        dws[i] = dW[n0:n0 + shp[0]]
                        n0 += shp[0]
            if side is not None:
                WGRAD.keep.append((dzc, xs_c))       # (not dW: an extra reference would make AccumulateGrad clone it)
        dbs = []
        if meta.has_bias:
            n0 = 0
            for i, shp in enumerate(ctx.shapes):
                dbs.append(dbias[n0:n0 + shp[0]])
                n0 += shp[0]
        return (None, d_res, *dxs, *dws, *dbs)


def linear(xs, ws, bs, *, act=ACT_NONE, out_dtype=None, cd=torch.float32, residual=None, w_is_kn=False, cache=True,
           drop=None, in_link=None, out_link=None, res_link=None):
    """drop = (seed tensor, site, p): nn.Dropout applied after the activation and before the residual add.
    in_link / out_link / res_link: see ``Link``."""
    xs = list(xs) if isinstance(xs, (list, tuple)) else [xs]
    ws = list(ws) if isinstance(ws, (list, tuple)) else [ws]
    bs = [] if bs is None else (list(bs) if isinstance(bs, (list, tuple)) else [bs])
    if not LINKS:
        in_link = out_link = res_link = None
    meta = _LinMeta(n_x=len(xs), n_w=len(ws), has_bias=len(bs) > 0, act=act, out_dtype=out_dtype or cd, cd=cd,
                    w_is_kn=w_is_kn, has_res=residual is not None, cache=cache or len(ws) > 1, drop=drop,
                    in_link=in_link, out_link=out_link, res_link=res_link if residual is not None else None)
    return LinearFn.apply(meta, residual, *xs, *ws, *bs)


# ------------------------------------------------------------------------------------------------
class LayerNormFn(Function):
    """y = LN(x (+ res)) * gamma + beta (+ pe[row % period]) (+ tt[tt_ids[row]])  (reference model.py:152-156).
    grad_link: ``Link`` of the Linear that produced x (its activation / dropout backward is fused into ln_bwd);
    res_link: ``Link`` that takes the gradient of ``res`` (see ``Link``)."""

    @staticmethod
    def forward(ctx, x, res, gamma, beta, eps, out_dtype, pe, pe_period, tt, tt_ids, grad_link=None, res_link=None):
        y, pre, mean, rstd = be().ln_fwd(x, res, gamma.detach(), beta.detach(), eps, out_dtype, pe, pe_period,
                                         None if tt is None else tt.detach(), tt_ids)
        ctx.save_for_backward(pre, mean, rstd, gamma.detach(), tt_ids)
        ctx.has_res = res is not None
        ctx.has_tt = tt is not None
        ctx.n_tt = 0 if tt is None else tt.shape[0]
        ctx.grad_link, ctx.res_link = grad_link, (res_link if res is not None else None)
        ctx.x_plain = res is None          # the LN input is the Linear's output itself
        return y

    @staticmethod
    def backward(ctx, dy):
        pre, mean, rstd, gamma, tt_ids = ctx.saved_tensors
        dy = dy.contiguous()
        L = ctx.grad_link
        if L is not None and (L.act != ACT_NONE or L.drop is not None) and (L.ref is None or L.ref.dtype == pre.dtype):
            ref = L.ref if not (L.act == ACT_RELU and ctx.x_plain) else None
            if L.act == ACT_NONE:
                ref = None
            dpre, dgamma, dbeta, L.dz, L.dbias = be().ln_bwd(dy, pre, mean, rstd, gamma, fuse=(L.act, ref, L.drop, L.has_bias))
        else:
            dpre, dgamma, dbeta = be().ln_bwd(dy, pre, mean, rstd, gamma)
        dtt = be().index_colsum(dy, tt_ids, ctx.n_tt) if ctx.has_tt else None
        dres = dpre if ctx.has_res else None
        if ctx.res_link is not None and ctx.needs_input_grad[1]:
            assert ctx.res_link.res_grad is None
            ctx.res_link.res_grad = dpre
            dres = None
        return dpre, dres, dgamma, dbeta, None, None, None, None, dtt, None, None, None


def layer_norm(x, res, gamma, beta, eps, out_dtype=None, pe=None, pe_period=0, tt=None, tt_ids=None, grad_link=None,
               res_link=None):
    if not LINKS:
        grad_link = res_link = None
    if not LN_FUSE:
        grad_link = None
    return LayerNormFn.apply(x, res, gamma, beta, eps, out_dtype or x.dtype, pe, pe_period, tt, tt_ids, grad_link, res_link)


class EmbedLNFn(Function):
    """LN_W(table[ids])  (reference model.py:492-494 / 519-521); row ``padding_idx`` of the table gets no gradient."""

    @staticmethod
    def forward(ctx, table, ids, gamma, beta, eps, out_dtype, padding_idx, drop=None):
        y, mean, rstd = be().embed_ln_fwd(table.detach(), ids, gamma.detach(), beta.detach(), eps, out_dtype, drop=drop)
        ctx.save_for_backward(table.detach(), ids, mean, rstd, gamma.detach())
        ctx.eps, ctx.padding_idx, ctx.drop = eps, padding_idx, drop
        return y

    @staticmethod
    def backward(ctx, dy):
        table, ids, mean, rstd, gamma = ctx.saved_tensors
        dtable, dgamma, dbeta = be().embed_ln_bwd(dy.contiguous(), table, ids, mean, rstd, gamma, ctx.eps, ctx.padding_idx,
                                                  drop=ctx.drop)
        return dtable, None, dgamma, dbeta, None, None, None, None


class AttentionFn(Function):
    """softmax(QK^T/sqrt(dh) + (1-mask)*-1e4) V over ragged short sequences (reference model.py:181-220).

    q / k / v are column blocks (width D) of projection outputs: ``qsrc[:, q_col:q_col+D]`` and, from ``kvsrc``
    (``qsrc`` itself when None, i.e. a fused QKV projection), the blocks at ``k_col`` / ``v_col``.  The gradient
    comes back in the same packing, so the fused projection's backward sees one contiguous tensor."""

    @staticmethod
    def forward(ctx, qsrc, kvsrc, cols, lay, key_mask, causal, n_heads, out_dtype, drop=None):
        D = lay["D"]
        src2 = qsrc if kvsrc is None else kvsrc
        q = qsrc[:, cols[0]:cols[0] + D]
        k = src2[:, cols[1]:cols[1] + D]
        v = src2[:, cols[2]:cols[2] + D]
        out = be().attn_fwd(q, k, v, lay["q_off"], lay["q_len"], lay["k_off"], lay["k_len"], key_mask, causal, n_heads,
                            lay["max_lq"], lay["max_lk"], out_dtype, drop=drop)
        ctx.save_for_backward(qsrc, kvsrc, key_mask)
        ctx.lay, ctx.causal, ctx.n_heads, ctx.cols, ctx.drop = lay, causal, n_heads, cols, drop
        return out

    @staticmethod
    def backward(ctx, dctx):
        qsrc, kvsrc, key_mask = ctx.saved_tensors
        lay, cols, D = ctx.lay, ctx.cols, ctx.lay["D"]
        src2 = qsrc if kvsrc is None else kvsrc
        dqsrc = torch.empty_like(qsrc)
        dkvsrc = dqsrc if kvsrc is None else torch.empty_like(kvsrc)
        sl = lambda t, c: t[:, c:c + D]
        be().attn_bwd(dctx.contiguous(), sl(qsrc, cols[0]), sl(src2, cols[1]), sl(src2, cols[2]), lay["q_off"], lay["q_len"],
                      lay["k_off"], lay["k_len"], key_mask, ctx.causal, ctx.n_heads, lay["max_lq"], lay["max_lk"],
                      sl(dqsrc, cols[0]), sl(dkvsrc, cols[1]), sl(dkvsrc, cols[2]), drop=ctx.drop)
        return dqsrc, (None if kvsrc is None else dkvsrc), None, None, None, None, None, None, None


def attention(qsrc, kvsrc, cols, lay, key_mask, causal, n_heads, out_dtype, drop=None):
    return AttentionFn.apply(qsrc, kvsrc, cols, lay, key_mask, causal, n_heads, out_dtype, drop)


class VideoLNDropFn(Function):
    """Train-mode staging of the video features: drop(LN_F(x) * gamma + beta) in the compute dtype (reference
    model.py:549-550).  The features are inputs, so the backward only produces dgamma / dbeta."""

    @staticmethod
    def forward(ctx, gamma, beta, feats_list, clip_t, clip_b, Lv, eps, out_dtype, drop):
        y, mean, rstd = be().video_ln_pack(feats_list, clip_t, clip_b, Lv, eps, out_dtype, gamma=gamma.detach(),
                                           beta=beta.detach(), drop=drop)
        ctx.save_for_backward(clip_t, clip_b, mean, rstd)
        ctx.feats, ctx.Lv, ctx.drop = feats_list, Lv, drop
        return y

    @staticmethod
    def backward(ctx, dy):
        clip_t, clip_b, mean, rstd = ctx.saved_tensors
        dgamma, dbeta = be().video_ln_affine_bwd(dy.contiguous(), ctx.feats, clip_t, clip_b, ctx.Lv, mean, rstd, drop=ctx.drop)
        return dgamma, dbeta, None, None, None, None, None, None, None


class SegMeanFn(Function):
    """Ingredient vectors: mean of the token vectors between [SEP]s, + pe[j]  (reference model.py:116-140)."""

    @staticmethod
    def forward(ctx, x, sep_mask, cu_ingr, pe, NI):
        ctx.save_for_backward(sep_mask, cu_ingr)
        ctx.Li, ctx.dt = sep_mask.shape[1], x.dtype
        return be().segmean_fwd(x, sep_mask, cu_ingr, pe, NI)

    @staticmethod
    def backward(ctx, dE):
        sep_mask, cu_ingr = ctx.saved_tensors
        return be().segmean_bwd(dE.contiguous(), sep_mask, cu_ingr, ctx.Li, ctx.dt), None, None, None, None


class RaggedMeanFn(Function):
    @staticmethod
    def forward(ctx, x, cu):
        ctx.save_for_backward(cu)
        ctx.NI = x.shape[0]
        return be().ragged_mean_fwd(x, cu)

    @staticmethod
    def backward(ctx, dy):
        return be().ragged_mean_bwd(dy.contiguous(), ctx.saved_tensors[0], ctx.NI), None


class RowNormalizeFn(Function):
    """x / sum(x, -1)  (reference model.py:798)."""

    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(x)
        return be().row_normalize_fwd(x)

    @staticmethod
    def backward(ctx, dy):
        return be().row_normalize_bwd(dy.contiguous(), ctx.saved_tensors[0])


class ReasonerFn(Function):
    """The sequential entity-state recurrence of EntitiyReasoningNetwork.forward (reference model.py:792-820)."""

    @staticmethod
    def forward(ctx, E0, w2v, clog, w4s, lay):
        ent, ebar, Es, saved = be().reasoner_fwd(E0, lay, w2v, clog, w4s)
        ctx.lay = lay
        ctx.n_saved = len(saved)
        ctx.save_for_backward(E0, w2v, clog, w4s, ent, ebar, Es, *saved)
        return ent, ebar, Es

    @staticmethod
    def backward(ctx, d_ent, d_ebar, d_Es):
        E0, w2v, clog, w4s, ent, ebar, Es, *saved = ctx.saved_tensors
        lay = ctx.lay
        z = lambda g, like: torch.zeros_like(like) if g is None else g.contiguous()
        dE0, dw2v, dclog, dw4s = be().reasoner_bwd(z(d_ent, ent), z(d_ebar, ebar), z(d_Es, Es), E0, lay, w2v, clog, w4s,
                                                   ent, ebar, Es, tuple(saved))
        return dE0, dw2v, dclog, dw4s, None


class PointerFn(Function):
    """pointer_generator_network + LabelSmoothingLoss (reference model.py:896-923, 37-55) -> (P flat, caption loss)."""

    @staticmethod
    def forward(ctx, dec, G, Esrc, wpgen, bpgen, logits, lay, dec_link=None):
        P, rows, saved = be().pointer_fwd(dec, G, Esrc, wpgen.detach(), bpgen.detach(), logits, lay)
        ctx.lay = lay
        ctx.dec_link = dec_link if LINKS else None
        ctx.save_for_backward(dec, G, Esrc, wpgen.detach(), bpgen.detach(), logits, P, *saved)
        return P, be().sum_all(rows)

    @staticmethod
    def backward(ctx, dP, dloss):
        dec, G, Esrc, wpgen, bpgen, logits, P, *saved = ctx.saved_tensors
        if dloss is None:
            dloss = torch.zeros((), dtype=torch.float32, device=P.device)
        ddec, dG, dE, dw, db, dlog = be().pointer_bwd(None if dP is None else dP.contiguous(), dloss, dec, G, Esrc,
                                                      wpgen, bpgen, logits, ctx.lay, P, tuple(saved))
        if ctx.dec_link is not None and ctx.needs_input_grad[0]:
            # dec also feeds the LM head: its transform Linear adds this piece in the epilogue of its input-gradient GEMM
            assert ctx.dec_link.res_grad is None
            ctx.dec_link.res_grad = ddec
            ddec = None
        return ddec, dG, dE, dw, db, dlog, None, None


class SoftmaxLSFn(Function):
    """softmax over the vocabulary + LabelSmoothingLoss on probabilities ('video' mode, reference model.py:1011-1014)."""

    @staticmethod
    def forward(ctx, logits, labels, V, smoothing):
        P, rows = be().softmax_ls_fwd(logits, labels, V, smoothing)
        ctx.save_for_backward(logits, labels, P)
        ctx.V, ctx.smoothing = V, smoothing
        return P, be().sum_all(rows)

    @staticmethod
    def backward(ctx, dP, dloss):
        logits, labels, P = ctx.saved_tensors
        if dloss is None:
            dloss = torch.zeros((), dtype=torch.float32, device=P.device)
        return be().softmax_ls_bwd(None if dP is None else dP.contiguous(), dloss, logits, labels, ctx.V, ctx.smoothing, P), None, None, None


class GumbelBowFn(Function):
    """Hard Gumbel-softmax word samples (injected noise) -> masked mean of their word vectors
    (reference model.py:1018-1021)."""

    @staticmethod
    def forward(ctx, P, g, emb, lay, text_mask):
        x, idx, yval = be().gumbel_fwd(P, g, emb.detach(), lay, text_mask)
        ctx.lay = lay
        ctx.save_for_backward(P, g, emb.detach(), text_mask, idx, yval)
        return x

    @staticmethod
    def backward(ctx, dx):
        P, g, emb, text_mask, idx, yval = ctx.saved_tensors
        dP, demb = be().gumbel_bwd(dx.contiguous(), P, g, emb, ctx.lay, text_mask, idx, yval)
        return dP, None, demb, None, None


class BiLSTMFn(Function):
    """nn.LSTM(W->D, bidirectional), batch of ragged sequences, directions summed (reference model.py:865, 1022-1024)."""

    @staticmethod
    def forward(ctx, x, lay, keys, *params):
        p = {k: v.detach() for k, v in zip(keys, params)}
        out, saved = be().bilstm_fwd(x, lay["cu_steps"], lay["steps"], p, cd=lay.get("cd", torch.float32))
        ctx.lay, ctx.keys, ctx.n_p = lay, keys, len(params)
        ctx.save_for_backward(x, *[v.detach() for v in params], *saved)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, *rest = ctx.saved_tensors
        params, saved = rest[:ctx.n_p], rest[ctx.n_p:]
        p = dict(zip(ctx.keys, params))
        dx, grads = be().bilstm_bwd(dout.contiguous(), x, ctx.lay["cu_steps"], ctx.lay["steps"], p, tuple(saved),
                                    cd=ctx.lay.get("cd", torch.float32))
        return (dx, None, None, *[grads[k] for k in ctx.keys])


class BCEFn(Function):
    """nn.BCELoss(reduction='sum') on probabilities (reference model.py:871)."""

    @staticmethod
    def forward(ctx, p, y):
        ctx.save_for_backward(p, y)
        return be().bce_fwd(p, y)

    @staticmethod
    def backward(ctx, g):
        p, y = ctx.saved_tensors
        return be().bce_bwd(p, y, g.contiguous()), None


class ASLFn(Function):
    """AsymmetricLoss on probabilities over the rows with >=1 positive (reference libs/ASL/.../losses.py:15-50,
    model.py:1103-1104)."""

    @staticmethod
    def forward(ctx, p, y):
        ctx.save_for_backward(p, y)
        return be().asl_fwd(p, y)

    @staticmethod
    def backward(ctx, g):
        p, y = ctx.saved_tensors
        return be().asl_bwd(p, y, g.contiguous()), None
