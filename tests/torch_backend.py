"""TEST INFRASTRUCTURE ONLY - plain-torch statement of every operator contract in ``svpc_b200.backend``.

Never imported by the product.  Two uses:
  * CPU tests install it (``backend.set_backend(TorchBackend())``) so the host-side logic of
    ``svpc_b200.model`` (batching of clips, ragged layouts, CSR building, autograd wiring) can be checked
    against the reference golden fixtures without a GPU;
  * GPU tests call it on the same inputs as the CUDA operators as the per-operator checker.
Backward operators are obtained by differentiating the forward statement with torch autograd, so they are
independent of the hand-derived backward kernels they check.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

ACT_NONE, ACT_RELU, ACT_GELU, ACT_SIGMOID = 0, 1, 2, 3


def _act(x, act):
    if act == ACT_RELU:
        return torch.relu(x)
    if act == ACT_GELU:
        return x * 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0)))
    if act == ACT_SIGMOID:
        return torch.sigmoid(x)
    return x


def _vjp(fn, inputs, grads):
    """inputs: list of tensors (float ones get requires_grad); grads: list matching fn outputs (None allowed)."""
    leaves = [t.detach().clone().requires_grad_(True) if (t is not None and t.is_floating_point()) else t for t in inputs]
    with torch.enable_grad():
        outs = fn(*leaves)
    if not isinstance(outs, (tuple, list)):
        outs = (outs,)
    pairs = [(o, g) for o, g in zip(outs, grads) if g is not None and o.requires_grad]
    req = [t for t in leaves if t is not None and t.is_floating_point()]
    res = torch.autograd.grad([o for o, _ in pairs], req, [g.to(o.dtype) for o, g in pairs], allow_unused=True)
    it = iter(res)
    out = []
    for t in leaves:
        if t is not None and t.is_floating_point():
            g = next(it)
            out.append(torch.zeros_like(t) if g is None else g)
        else:
            out.append(None)
    return out


class TorchBackend:
    name = "torch-test"
    math_dtype = torch.float32

    def __init__(self, math_dtype=torch.float32):
        self.math_dtype = math_dtype

    def f(self, x):
        return None if x is None else x.to(self.math_dtype)

    # ---------------------------------------------------------------- dense
    def gemm(self, a, bt, out_dtype, bias=None, act=0, residual=None, out=None, accumulate=False, want_pre=False,
             route=None, drop=None, grad_act=None):
        assert drop is None, "the plain-torch statements are eval-mode (no dropout)"
        pre = self.f(a) @ self.f(bt).t()
        if bias is not None:
            pre = pre + self.f(bias)
        y = _act(pre, act)
        if residual is not None:
            y = y + self.f(residual)
        if grad_act is not None:
            assert out is None and not accumulate
            y = self.f(self.act_bwd(y.to(out_dtype), grad_act[1], grad_act[0])[0])
        if out is not None:
            if accumulate:
                out.add_(y.to(out.dtype))
            else:
                out.copy_(y.to(out.dtype))
            y = out
        else:
            assert not accumulate
            y = y.to(out_dtype)
        if want_pre:
            return y, pre.to(out_dtype)
        return y

    def act_bwd(self, dy, ref, act, drop=None):
        dy32 = self.f(dy)
        if act == ACT_RELU:
            dz = dy32 * (self.f(ref) > 0)
        elif act == ACT_GELU:
            x = self.f(ref)
            cdf = 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0)))
            pdf = torch.exp(-0.5 * x * x) / math.sqrt(2.0 * math.pi)
            dz = dy32 * (cdf + x * pdf)
        elif act == ACT_SIGMOID:
            y = self.f(ref)
            dz = dy32 * y * (1 - y)
        else:
            dz = dy32
        return dz.to(dy.dtype), dz.sum(0).float()

    def colsum(self, x):
        return self.f(x).sum(0).float()

    def cast(self, x, dtype):
        return x.to(dtype)

    # ---------------------------------------------------------------- layer norm / embeddings
    def _ln(self, pre, gamma, beta, eps):
        u = pre.mean(-1, keepdim=True)
        s = (pre - u).pow(2).mean(-1, keepdim=True)
        rstd = 1.0 / torch.sqrt(s + eps)
        return (pre - u) * rstd * gamma + beta, u.squeeze(-1), rstd.squeeze(-1)

    def ln_fwd(self, x, res, gamma, beta, eps, out_dtype, pe=None, pe_period=0, tt=None, tt_ids=None):
        pre = self.f(x) if res is None else self.f(x) + self.f(res)
        y, mean, rstd = self._ln(pre, self.f(gamma), self.f(beta), eps)
        if pe is not None:
            rows = torch.arange(y.shape[0], device=y.device) % pe_period
            y = y + self.f(pe)[rows]
        if tt is not None:
            y = y + self.f(tt)[tt_ids.long()]
        pre_out = x if res is None else pre.to(x.dtype)
        return y.to(out_dtype), pre_out, mean.float(), rstd.float()

    def ln_bwd(self, dy, pre, mean, rstd, gamma, fuse=None):
        g = self.f(dy) * self.f(gamma)
        xhat = (self.f(pre) - self.f(mean)[:, None]) * self.f(rstd)[:, None]
        dpre = self.f(rstd)[:, None] * (g - g.mean(-1, keepdim=True) - xhat * (g * xhat).mean(-1, keepdim=True))
        out = (dpre.to(pre.dtype), (self.f(dy) * xhat).sum(0).float(), self.f(dy).sum(0).float())
        if fuse is None:
            return out
        act, ref, drop, want_dbias = fuse
        assert drop is None, "the plain-torch statements are eval-mode (no dropout)"
        dz, dbias = self.act_bwd(out[0], pre if ref is None else ref, act)
        return out + (dz, dbias if want_dbias else None)

    def index_colsum(self, dy, ids, nbins):
        out = torch.zeros(nbins, dy.shape[1], dtype=self.math_dtype, device=dy.device)
        out.index_add_(0, ids.long(), self.f(dy))
        return out.float()

    def embed_ln_fwd(self, table, ids, gamma, beta, eps, out_dtype, drop=None):
        y, mean, rstd = self._ln(self.f(table)[ids.long()], self.f(gamma), self.f(beta), eps)
        return y.to(out_dtype), mean.float(), rstd.float()

    def embed_ln_bwd(self, dy, table, ids, mean, rstd, gamma, eps, padding_idx, drop=None):
        def fn(t, g):
            return self._ln(t[ids.long()], g, torch.zeros_like(g), eps)[0]
        dtable, dgamma = _vjp(fn, [self.f(table), self.f(gamma)], [self.f(dy)])
        if padding_idx is not None and padding_idx >= 0:
            dtable[padding_idx].zero_()
        return dtable.float(), dgamma.float(), self.f(dy).sum(0).float()

    def video_ln_pack(self, feats_list, clip_t, clip_b, Lv, eps, out_dtype, gamma=None, beta=None, drop=None):
        assert gamma is None and drop is None
        x = torch.stack([f for f in feats_list])                      # (S, B, L, F)
        x = self.f(x[clip_t.long(), clip_b.long(), :Lv])              # (Nc, Lv, F)
        x = x.reshape(-1, x.shape[-1])
        u = x.mean(-1, keepdim=True)
        s = (x - u).pow(2).mean(-1, keepdim=True)
        rstd = 1.0 / torch.sqrt(s + eps)
        return ((x - u) * rstd).to(out_dtype), u.squeeze(-1).float(), rstd.squeeze(-1).float()

    # ---------------------------------------------------------------- attention
    def _attn(self, q, k, v, q_off, q_len, k_off, k_len, key_mask, causal, H):
        D = q.shape[1]
        dh = D // H
        outs = []
        ctx = torch.zeros(q.shape[0], D, dtype=q.dtype, device=q.device)
        pieces = []
        for n in range(len(q_off)):
            qo, ql, ko, kl = int(q_off[n]), int(q_len[n]), int(k_off[n]), int(k_len[n])
            qq = q[qo:qo + ql].view(ql, H, dh).permute(1, 0, 2)
            kk = k[ko:ko + kl].view(kl, H, dh).permute(1, 0, 2)
            vv = v[ko:ko + kl].view(kl, H, dh).permute(1, 0, 2)
            m = torch.ones(ql, kl, dtype=q.dtype, device=q.device)
            if key_mask is not None:
                m = m * key_mask[ko:ko + kl].to(q.dtype)[None, :]
            if causal:
                m = m * torch.tril(torch.ones(ql, kl, dtype=q.dtype, device=q.device))
            s = qq @ kk.transpose(-1, -2) / math.sqrt(dh) + (1 - m) * -10000.0
            p = torch.softmax(s, dim=-1)
            pieces.append((qo, ql, (p @ vv).permute(1, 0, 2).reshape(ql, D)))
        idx = torch.cat([torch.arange(qo, qo + ql, device=q.device) for qo, ql, _ in pieces])
        ctx = ctx.index_copy(0, idx, torch.cat([c for _, _, c in pieces]))
        return ctx

    def attn_fwd(self, q, k, v, q_off, q_len, k_off, k_len, key_mask, causal, n_heads, max_lq, max_lk, out_dtype, drop=None):
        return self._attn(self.f(q), self.f(k), self.f(v), q_off.tolist(), q_len.tolist(), k_off.tolist(),
                          k_len.tolist(), key_mask, causal, n_heads).to(out_dtype)

    def attn_bwd(self, dctx, q, k, v, q_off, q_len, k_off, k_len, key_mask, causal, n_heads, max_lq, max_lk, dq, dk, dv,
                 drop=None):
        qo, ql, ko, kl = q_off.tolist(), q_len.tolist(), k_off.tolist(), k_len.tolist()
        g = _vjp(lambda a, b, c: self._attn(a, b, c, qo, ql, ko, kl, key_mask, causal, n_heads),
                 [self.f(q), self.f(k), self.f(v)], [self.f(dctx)])
        dq.copy_(g[0].to(dq.dtype)); dk.copy_(g[1].to(dk.dtype)); dv.copy_(g[2].to(dv.dtype))

    # ---------------------------------------------------------------- ingredients
    def _segments(self, sep_mask):
        segs = []
        for b in range(sep_mask.shape[0]):
            prev = 0
            for s in (sep_mask[b] == 1).nonzero().view(-1).tolist():
                segs.append((b, prev, s))
                prev = s + 1
        return segs

    def _segmean(self, x, sep_mask, pe, cu):
        Li = sep_mask.shape[1]
        out = []
        cu = cu.tolist()
        segs = self._segments(sep_mask)
        for gi, (b, lo, hi) in enumerate(segs):
            j = gi - cu[b]
            out.append(x[b * Li + lo: b * Li + hi].mean(dim=0) + pe[j])
        return torch.stack(out)

    def segmean_fwd(self, x, sep_mask, cu_ingr, pe, NI=None):
        return self._segmean(self.f(x), sep_mask, self.f(pe), cu_ingr).float()

    def segmean_bwd(self, dE0, sep_mask, cu_ingr, Li, out_dtype):
        B = sep_mask.shape[0]
        x = torch.zeros(B * Li, dE0.shape[1], dtype=self.math_dtype, device=dE0.device)
        pe = torch.zeros(Li, dE0.shape[1], dtype=self.math_dtype, device=dE0.device)
        (dx, _) = _vjp(lambda a, p: self._segmean(a, sep_mask, p, cu_ingr), [x, pe], [self.f(dE0)])
        return dx.to(out_dtype)

    def ragged_mean_fwd(self, x, cu):
        c = cu.tolist()
        return torch.stack([self.f(x)[c[b]:c[b + 1]].mean(0) for b in range(len(c) - 1)]).float()

    def ragged_mean_bwd(self, dy, cu, NI=None):
        c = cu.tolist()
        return torch.cat([(self.f(dy)[b] / (c[b + 1] - c[b])).expand(c[b + 1] - c[b], -1) for b in range(len(c) - 1)]).float()

    def row_normalize_fwd(self, x):
        return (self.f(x) / self.f(x).sum(-1, keepdim=True)).float()

    def row_normalize_bwd(self, dy, x):
        return _vjp(lambda a: a / a.sum(-1, keepdim=True), [self.f(x)], [self.f(dy)])[0].float()

    # ---------------------------------------------------------------- visual simulator recurrence
    def _reasoner(self, E0, w2v, clog, w4s, cu_ingr, cu_steps):
        ci, cs = cu_ingr.tolist(), cu_steps.tolist()
        ents, ebars, Es, As = [], [], [], []
        for b in range(len(cs) - 1):
            E = E0[ci[b]:ci[b + 1]]
            e_prev = E.new_zeros(E.shape[0])
            for n in range(cs[b], cs[b + 1]):
                e = torch.sigmoid(E @ w2v[n])
                c = torch.softmax(clog[n], dim=-1)
                a = c[0] * e + c[1] * e_prev
                ebar = (a / a.sum()).unsqueeze(0) @ E
                k = torch.relu(w4s[n] * ebar)
                E = a.unsqueeze(1) @ k + (1 - a).unsqueeze(1) * E
                e_prev = e
                ents.append(e); ebars.append(ebar[0]); Es.append(E); As.append(a)
        return torch.cat(ents), torch.stack(ebars), torch.cat(Es), torch.cat(As)

    def reasoner_fwd(self, E0, lay, w2v, clog, w4s):
        cu_ingr, cu_steps = lay["cu_ingr"], lay["cu_steps"]
        ent, ebar, Es, a = self._reasoner(self.f(E0), self.f(w2v), self.f(clog), self.f(w4s), cu_ingr, cu_steps)
        return ent.float(), ebar.float(), Es.float(), (a.float(),)

    def reasoner_bwd(self, d_ent, d_ebar, d_Es, E0, lay, w2v, clog, w4s, ent, ebar, Es, saved):
        cu_ingr, cu_steps = lay["cu_ingr"], lay["cu_steps"]
        g = _vjp(lambda a, b, c, d: self._reasoner(a, b, c, d, cu_ingr, cu_steps)[:3],
                 [self.f(E0), self.f(w2v), self.f(clog), self.f(w4s)], [self.f(d_ent), self.f(d_ebar), self.f(d_Es)])
        return tuple(x.float() for x in g)

    # ---------------------------------------------------------------- pointer-generator + caption loss
    @staticmethod
    def _ls_rows(P, labels, smoothing):
        """LabelSmoothingLoss per row on probabilities P (R, Vx); rows with label -1 give 0."""
        Vx = P.shape[1]
        valid = labels != -1
        q = P.new_full((P.shape[0], Vx), smoothing / (Vx - 1))
        q[:, -1] = 0
        q.scatter_(1, labels.clamp(min=0).long().unsqueeze(1), 1.0 - smoothing)
        logp = torch.log(P + 1e-12)
        rows = (torch.xlogy(q, q) - q * logp).sum(1)
        return rows * valid

    def _pointer(self, dec, G, Esrc, wpgen, bpgen, logits, lay):
        Lt, V = lay["Lt"], lay["V"]
        src_off, nis, gis, vxs = lay["clip_src_off"].tolist(), lay["clip_ni"].tolist(), lay["clip_gi"].tolist(), lay["clip_vx"].tolist()
        tok_ptr, tok_ids = lay["tok_ptr"].tolist(), lay["tok_ids"].tolist()
        labels = lay["labels"]
        Ps, losses = [], []
        D = dec.shape[1]
        for n in range(len(src_off)):
            d = dec[n * Lt:(n + 1) * Lt]                                   # (Lt, D)
            g = G[src_off[n]:src_off[n] + nis[n]]                          # (Ni, D)
            e = Esrc[src_off[n]:src_off[n] + nis[n]]
            alpha = torch.softmax(g @ d.t(), dim=0)                        # (Ni, Lt), over ingredients
            ctx = alpha.t() @ e                                            # (Lt, D)
            pg = torch.sigmoid(d @ wpgen[:D] + ctx @ wpgen[D:] + bpgen)    # (Lt,)
            P = pg[:, None] * torch.softmax(logits[n * Lt:(n + 1) * Lt, :V], dim=-1)
            al = alpha * (1 - pg)[None, :]
            if vxs[n] > V:
                P = torch.cat([P, P.new_zeros(Lt, vxs[n] - V)], dim=1)
            add = torch.zeros_like(P)
            for i in range(nis[n]):
                lo, hi = tok_ptr[gis[n] + i], tok_ptr[gis[n] + i + 1]
                for tok in tok_ids[lo:hi]:
                    add[:, tok] = add[:, tok] + al[i] / (hi - lo)
            P = P + add
            Ps.append(P.reshape(-1))
            losses.append(self._ls_rows(P, labels[n * Lt:(n + 1) * Lt], lay["smoothing"]))
        return torch.cat(Ps), torch.cat(losses)

    def pointer_fwd(self, dec, G, Esrc, wpgen, bpgen, logits, lay):
        P, rows = self._pointer(self.f(dec), self.f(G), self.f(Esrc), self.f(wpgen), self.f(bpgen), self.f(logits), lay)
        return P.float(), rows.float(), ()

    def pointer_bwd(self, dP, dloss, dec, G, Esrc, wpgen, bpgen, logits, lay, P, saved):
        def fn(a, b, c, d, e, f):
            P_, rows = self._pointer(a, b, c, d, e, f, lay)
            return P_, rows.sum()
        g = _vjp(fn, [self.f(dec), self.f(G), self.f(Esrc), self.f(wpgen), self.f(bpgen), self.f(logits)],
                 [None if dP is None else self.f(dP), self.f(dloss)])
        ddec, dG, dE, dw, db, dlog = g
        return ddec.to(dec.dtype), dG.float(), dE.float(), dw.float(), db.float(), dlog.float()

    def softmax_ls_fwd(self, logits, labels, V, smoothing):
        P = torch.softmax(self.f(logits)[:, :V], dim=-1)
        return P.float(), self._ls_rows(P, labels, smoothing).float()

    def softmax_ls_bwd(self, dP, dloss, logits, labels, V, smoothing, P):
        def fn(l):
            P_ = torch.softmax(l[:, :V], dim=-1)
            return P_, self._ls_rows(P_, labels, smoothing).sum()
        return _vjp(fn, [self.f(logits)], [None if dP is None else self.f(dP), self.f(dloss)])[0].float()

    # ---------------------------------------------------------------- textual re-simulator
    def _gumbel(self, P, g, emb, lay, text_mask):
        Lt, V, tau = lay["Lt"], lay["V"], lay["tau"]
        offs, vxs = lay["clip_P_off"].tolist(), lay["clip_vx"].tolist()
        xs, idxs, yvals = [], [], []
        for n in range(len(offs)):
            p = P[offs[n]:offs[n] + Lt * vxs[n]].view(Lt, vxs[n])
            gg = g[offs[n]:offs[n] + Lt * vxs[n]].view(Lt, vxs[n])
            y_soft = ((torch.log(p + 1e-12) + gg) / tau).softmax(-1)
            index = y_soft.max(-1, keepdim=True)[1]
            y_hard = torch.zeros_like(p).scatter_(-1, index, 1.0)
            y = y_hard - y_soft.detach() + y_soft
            bow = y[:, :V] @ emb
            m = (text_mask[n * Lt:(n + 1) * Lt] == 1).to(bow.dtype)
            xs.append((bow * m[:, None]).sum(0) / m.sum())
            idxs.append(index.view(-1))
            yvals.append(y.gather(1, index).view(-1))
        return torch.stack(xs), torch.cat(idxs).int(), torch.cat(yvals)

    def gumbel_noise(self, n, seed, site):
        return -torch.empty(int(n), dtype=torch.float32, device=seed.device).exponential_().log()

    def gumbel_fwd(self, P, g, emb, lay, text_mask):
        x, idx, yv = self._gumbel(self.f(P), self.f(g), self.f(emb), lay, text_mask)
        return x.float(), idx, yv.detach().float()

    def gumbel_bwd(self, dx, P, g, emb, lay, text_mask, idx, yval):
        dP, demb = _vjp(lambda a, b: self._gumbel(a, self.f(g), b, lay, text_mask)[0], [self.f(P), self.f(emb)], [self.f(dx)])
        return dP.float(), demb.float()

    def _bilstm(self, x, cs, p):
        D = p["weight_hh_l0"].shape[1]
        outs = []
        for b in range(len(cs) - 1):
            xb = x[cs[b]:cs[b + 1]]
            S = xb.shape[0]

            def run(suffix, order):
                h = xb.new_zeros(D); c = xb.new_zeros(D)
                out = [None] * S
                for t in order:
                    gates = F.linear(xb[t], p["weight_ih_l0" + suffix], p["bias_ih_l0" + suffix]) + \
                        F.linear(h, p["weight_hh_l0" + suffix], p["bias_hh_l0" + suffix])
                    i, f, gg, o = gates.chunk(4)
                    c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
                    h = torch.sigmoid(o) * torch.tanh(c)
                    out[t] = h
                return torch.stack(out)
            outs.append(run("", range(S)) + run("_reverse", range(S - 1, -1, -1)))
        return torch.cat(outs)

    LSTM_KEYS = tuple(f"{k}_l0{s}" for s in ("", "_reverse") for k in ("weight_ih", "weight_hh", "bias_ih", "bias_hh"))

    def bilstm_fwd(self, x, cu_steps, steps, params, cd=None):
        p = {k: self.f(v) for k, v in params.items()}
        return self._bilstm(self.f(x), cu_steps.tolist(), p).float(), ()

    def bilstm_bwd(self, dout, x, cu_steps, steps, params, saved, cd=None):
        keys = list(self.LSTM_KEYS)
        g = _vjp(lambda a, *ws: self._bilstm(a, cu_steps.tolist(), dict(zip(keys, ws))),
                 [self.f(x)] + [self.f(params[k]) for k in keys], [self.f(dout)])
        return g[0].float(), {k: v.float() for k, v in zip(keys, g[1:])}

    # ---------------------------------------------------------------- losses
    @staticmethod
    def _bce(p, y):
        return -(y * torch.log(p).clamp(min=-100) + (1 - y) * torch.log(1 - p).clamp(min=-100)).sum()

    def bce_fwd(self, p, y):
        return F.binary_cross_entropy(self.f(p), self.f(y), reduction="sum").float()

    def bce_bwd(self, p, y, g):
        return _vjp(lambda a: F.binary_cross_entropy(a, self.f(y), reduction="sum"), [self.f(p)], [self.f(g)])[0].float()

    @staticmethod
    def _asl(p, y):
        rows = (y == 1).any(dim=1)
        p, y = p[rows], y[rows]
        xs_neg = (1 - p + 0.05).clamp(max=1)
        loss = y * torch.log(p.clamp(min=1e-8)) + (1 - y) * torch.log(xs_neg.clamp(min=1e-8))
        pt = p * y + xs_neg * (1 - y)
        w = torch.pow(1 - pt, 1.0 * y + 4.0 * (1 - y))
        return -(loss * w).sum()

    def asl_fwd(self, p, y):
        return self._asl(self.f(p), self.f(y)).float()

    def asl_bwd(self, p, y, g):
        return _vjp(lambda a: self._asl(a, self.f(y)), [self.f(p)], [self.f(g)])[0].float()

    def sum_all(self, x):
        return self.f(x).sum().float()

    # ---------------------------------------------------------------- greedy decoding
    def greedy_pick(self, scores, row_off, ld, row_width, width, row_noov, unk, next_in, ext_out, pos):
        flat = scores.reshape(-1)
        for r in range(next_in.numel()):
            o = int(row_off[r]) if row_off is not None else r * int(ld)
            w = int(row_width[r]) if row_width is not None else int(width)
            s = flat[o:o + w].clone().float()
            s[unk] = -1e10
            bi = int(torch.argmax(s))
            bi = int((s == s[bi]).nonzero()[0])          # lowest index among ties
            noov = int(row_noov[r]) if row_noov is not None else 0
            ext_out[r, pos] = bi
            next_in[r] = unk if bi >= w - noov else bi
