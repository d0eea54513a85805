"""Autograd wiring of the SVPC operators.

Each ``torch.autograd.Function`` here pairs one forward operator of ``svpc_b200.backend`` with its hand-written
backward operator; torch is only the tape (and the owner of device memory).  Nothing in this file computes on the
host: every tensor operation on activations is a backend (CUDA) call.
"""
from __future__ import annotations

import weakref

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

    def invalidate(self):
        """Mark every entry stale but keep its buffer (next ``refresh_all`` rewrites them in place)."""
        for e in self.store.values():
            e[0] = None

    def prepare(self):
        """Build the device tables of the next ``refresh_all`` without launching (a CUDA-graph capture cannot upload them)."""
        self.refresh_all(launch=False)

    def refresh_all(self, launch=True):
        """One launch per compute dtype for all entries whose parameters changed since they were cast."""
        stale = []
        for k, e in self.store.items():
            ws = self._params(e)
            if ws is None or e[0] == self._version(ws):
                continue
            if all(w.is_cuda and w.dtype == torch.float32 and w.dim() <= 2 and w.is_contiguous() for w in ws) and \
                    e[1].device == ws[0].device:
                stale.append((k, e, ws))
        if not stale or not hasattr(be(), "cast_multi"):
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

    def buffers(self):
        """Every cached tensor (a captured graph keeps them alive: it reads and rewrites their addresses)."""
        return [e[1] for e in self.store.values()]

    def clear(self):
        self.store.clear()
        self._tables.clear()


WEIGHTS = WeightCache()


class _LinMeta:
    __slots__ = ("n_x", "n_w", "has_bias", "act", "out_dtype", "cd", "w_is_kn", "has_res", "cache", "drop")

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
        W = WEIGHTS.get(ws, cd) if meta.cache else be().cast(ws[0].detach().contiguous(), cd)
        bias = None
        if meta.has_bias:
            bias = bs[0].detach() if len(bs) == 1 else WEIGHTS.get(bs, torch.float32)
        xs_c = [x if x.dtype == cd else be().cast(x, cd) for x in xs]
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
        ctx.save_for_backward(W, ref, *xs_c)
        return out

    @staticmethod
    def backward(ctx, dy):
        meta = ctx.meta
        W, ref, *xs_c = ctx.saved_tensors
        cd = meta.cd
        dy = dy.contiguous()
        need = ctx.needs_input_grad
        d_res = dy if (meta.has_res and need[1]) else None
        if meta.act != ACT_NONE or meta.drop is not None:
            dz, dbias = be().act_bwd(dy, ref, meta.act, drop=meta.drop)     # dy * dropout scale * act'
        else:
            dz = dy
            dbias = be().colsum(dy) if meta.has_bias else None
        dzc = dz if dz.dtype == cd else be().cast(dz, cd)
        n_x, n_w = meta.n_x, meta.n_w
        # ---- input gradients
        dxs = [None] * n_x
        k0 = 0
        for j, x in enumerate(xs_c):
            k1 = k0 + x.shape[1]
            if need[2 + j]:
                if meta.w_is_kn:
                    dxs[j] = be().gemm(dzc, W, ctx.x_dtypes[j])                     # dz (M,N) @ W(K,N)^T
                else:
                    dxs[j] = be().gemm(dzc, W[:, k0:k1].t(), ctx.x_dtypes[j])      # dz (M,N) @ W(N,K)
            k0 = k1
        # ---- weight gradients (fp32, written straight into one buffer in the parameters' layout)
        dws = [None] * n_w
        if any(need[2 + n_x:2 + n_x + n_w]):
            if meta.w_is_kn:
                dws[0] = be().gemm(xs_c[0].t(), dzc.t(), torch.float32)            # (K,N) = x^T dz
            else:
                Ntot = W.shape[0]
                Ktot = W.shape[1]
                dW = torch.empty(Ntot, Ktot, dtype=torch.float32, device=dy.device)
                k0 = 0
                for x in xs_c:
                    k1 = k0 + x.shape[1]
                    be().gemm(dzc.t(), x.t(), torch.float32, out=dW[:, k0:k1])
                    k0 = k1
                n0 = 0
                for i, shp in enumerate(ctx.shapes):
                    dws[i] = dW[n0:n0 + shp[0]]
                    n0 += shp[0]
        dbs = []
        if meta.has_bias:
            n0 = 0
            for i, shp in enumerate(ctx.shapes):
                dbs.append(dbias[n0:n0 + shp[0]])
                n0 += shp[0]
        return (None, d_res, *dxs, *dws, *dbs)


def linear(xs, ws, bs, *, act=ACT_NONE, out_dtype=None, cd=torch.float32, residual=None, w_is_kn=False, cache=True,
           drop=None):
    """drop = (seed tensor, site, p): nn.Dropout applied after the activation and before the residual add."""
    xs = list(xs) if isinstance(xs, (list, tuple)) else [xs]
    ws = list(ws) if isinstance(ws, (list, tuple)) else [ws]
    bs = [] if bs is None else (list(bs) if isinstance(bs, (list, tuple)) else [bs])
    meta = _LinMeta(n_x=len(xs), n_w=len(ws), has_bias=len(bs) > 0, act=act, out_dtype=out_dtype or cd, cd=cd,
                    w_is_kn=w_is_kn, has_res=residual is not None, cache=cache or len(ws) > 1, drop=drop)
    return LinearFn.apply(meta, residual, *xs, *ws, *bs)


# ------------------------------------------------------------------------------------------------
class LayerNormFn(Function):
    """y = LN(x (+ res)) * gamma + beta (+ pe[row % period]) (+ tt[tt_ids[row]])  (reference model.py:152-156)."""

    @staticmethod
    def forward(ctx, x, res, gamma, beta, eps, out_dtype, pe, pe_period, tt, tt_ids):
        y, pre, mean, rstd = be().ln_fwd(x, res, gamma.detach(), beta.detach(), eps, out_dtype, pe, pe_period,
                                         None if tt is None else tt.detach(), tt_ids)
        ctx.save_for_backward(pre, mean, rstd, gamma.detach(), tt_ids)
        ctx.has_res = res is not None
        ctx.has_tt = tt is not None
        ctx.n_tt = 0 if tt is None else tt.shape[0]
        return y

    @staticmethod
    def backward(ctx, dy):
        pre, mean, rstd, gamma, tt_ids = ctx.saved_tensors
        dy = dy.contiguous()
        dpre, dgamma, dbeta = be().ln_bwd(dy, pre, mean, rstd, gamma)
        dtt = be().index_colsum(dy, tt_ids, ctx.n_tt) if ctx.has_tt else None
        return dpre, (dpre if ctx.has_res else None), dgamma, dbeta, None, None, None, None, dtt, None


def layer_norm(x, res, gamma, beta, eps, out_dtype=None, pe=None, pe_period=0, tt=None, tt_ids=None):
    return LayerNormFn.apply(x, res, gamma, beta, eps, out_dtype or x.dtype, pe, pe_period, tt, tt_ids)


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
    def forward(ctx, dec, G, Esrc, wpgen, bpgen, logits, lay):
        P, rows, saved = be().pointer_fwd(dec, G, Esrc, wpgen.detach(), bpgen.detach(), logits, lay)
        ctx.lay = lay
        ctx.save_for_backward(dec, G, Esrc, wpgen.detach(), bpgen.detach(), logits, P, *saved)
        return P, be().sum_all(rows)

    @staticmethod
    def backward(ctx, dP, dloss):
        dec, G, Esrc, wpgen, bpgen, logits, P, *saved = ctx.saved_tensors
        if dloss is None:
            dloss = torch.zeros((), dtype=torch.float32, device=P.device)
        ddec, dG, dE, dw, db, dlog = be().pointer_bwd(None if dP is None else dP.contiguous(), dloss, dec, G, Esrc,
                                                      wpgen, bpgen, logits, ctx.lay, P, tuple(saved))
        return ddec, dG, dE, dw, db, dlog, None


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
