"""B200-native drop-in for ``StateAwareRecursiveTransformer`` (reference src/rtransformer/model.py:826-1189).

Same constructor argument (the EasyDict-like config), same ``forward`` signature and return values, same module
attribute tree and ``state_dict`` layout, so ``src/train.py``, ``src/translator.py`` and
``src/dump_memories.py`` work against it unchanged.  What is different is *how* the step runs:

* the reference loops over clips (model.py:1038), videos (:1059), reasoner steps (:792) and ingredients
  (:125-134, :912-914) in Python and issues ~125k tiny library kernels per step; here every clip of every video goes
  through each stage as ONE batched problem over ragged index arrays built once per call on the host
  (``StepPlan``), and only the entity-state recurrence and the BiLSTM are sequential (one persistent kernel each);
* every device computation is a hand-written sm_100a kernel behind the C-ABI of ``libsvpc_b200.so``
  (``svpc_b200.backend`` / ``svpc_b200.functional``); torch is the tape and the allocator.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import torch
import torch.nn as nn

from . import functional as Fn
from .backend import ACT_NONE, ACT_RELU, ACT_GELU, ACT_SIGMOID
from .config import SvpcConfig, PAD, UNK, IGNORE
from .synth import state_dict_spec, sinusoid_table
from .compat import make_node as _make_node

GUMBEL_SITE = 0x6b6d     # hash site of the Gumbel noise (dropout sites count from 1)
LSTM_KEYS = tuple(f"{k}_l0{s}" for s in ("", "_reverse") for k in ("weight_ih", "weight_hh", "bias_ih", "bias_hh"))


# ------------------------------------------------------------------------------------------------
# host-side plan: ragged layouts of one batch (no device synchronisation)
# ------------------------------------------------------------------------------------------------
class StepPlan:
    """Index arrays describing one collated batch.

    clips are ordered video-major: clip n = (video b, step t) with n = cu_steps[b] + t.
      cu_steps (B+1)        clip offsets per video
      cu_ingr  (B+1)        ingredient offsets per video (NI = cu_ingr[B])
      clip_ent_off (Nc)     row of clip n's (Ni_b) block in the per-step entity arrays (ent, E_stack):
                            ent_off[b] + t * Ni_b, ent_off[b] = sum_{b'<b} S_b' * Ni_b'
      clip_P_off (Nc)       element offset of clip n's (Lt, Vx_b) block in the flat prediction buffer
      tok_ptr / tok_ids     CSR ingredient (global index) -> extended token ids (reference ingr_id_dict)
    """

    def __init__(self, cfg, steps, ni, extra_zeros, ingr_id_dict, device, mode):
        Lv, Lt, V, D = cfg.max_v_len, cfg.max_t_len, cfg.vocab_size, cfg.hidden_size
        B = len(steps)
        self.B, self.steps, self.ni = B, list(steps), list(ni)
        self.Nc = sum(steps)
        cu_steps = [0]
        for s in steps:
            cu_steps.append(cu_steps[-1] + s)
        cu_ingr = [0]
        for n in ni:
            cu_ingr.append(cu_ingr[-1] + n)
        self.NI = cu_ingr[-1]
        self.cu_steps, self.cu_ingr = cu_steps, cu_ingr
        clip_t, clip_b = [], []
        for b in range(B):
            clip_t += list(range(steps[b]))
            clip_b += [b] * steps[b]
        ent_off, P_off = [0], [0]
        vx = [V + int(z) for z in extra_zeros]
        for b in range(B):
            ent_off.append(ent_off[-1] + steps[b] * ni[b])
            P_off.append(P_off[-1] + steps[b] * Lt * vx[b])
        self.ent_off, self.P_off, self.vx = ent_off, P_off, vx
        self.NE, self.P_numel = ent_off[-1], P_off[-1]
        clip_ent_off = [ent_off[b] + t * ni[b] for b, t in zip(clip_b, clip_t)]
        clip_P_off = [P_off[b] + t * Lt * vx[b] for b, t in zip(clip_b, clip_t)]
        clip_ni = [ni[b] for b in clip_b]
        clip_gi = [cu_ingr[b] for b in clip_b]
        clip_vx = [vx[b] for b in clip_b]
        clip_src_copy = [cu_ingr[b] for b in clip_b]
        self.mode = mode
        tok_ptr, tok_ids = self._tok_csr(ingr_id_dict)
        Nc = self.Nc
        ints = OrderedDict(
            cu_steps=cu_steps, cu_ingr=cu_ingr, clip_t=clip_t, clip_b=clip_b, clip_ent_off=clip_ent_off,
            clip_P_off=clip_P_off, clip_ni=clip_ni, clip_gi=clip_gi, clip_vx=clip_vx, clip_src_copy=clip_src_copy,
            tok_ptr=tok_ptr, tok_ids=tok_ids if tok_ids else [0],
            enc_q_off=[n * Lv for n in range(Nc)], enc_len=[Lv] * Nc,
            cls_q_off=list(range(Nc)), ones=[1] * Nc,
            step_len=list(steps), step_off=cu_steps[:-1],
            dec_off=[n * Lt for n in range(Nc)], dec_len=[Lt] * Nc,
        )
        flat, spans = [], {}
        for k, v in ints.items():
            spans[k] = (len(flat), len(v))
            flat += v
        # one host->device copy for all index arrays
        host = torch.tensor(flat, dtype=torch.int32)
        if device.type == "cuda":
            host = host.pin_memory()
        dev = host.to(device, non_blocking=True)
        self._host, self._dev, self._spans = host, dev, spans
        self._ring, self._ring_i = None, 0       # update_tokens: pinned copies of `host`, each guarded by a CUDA event
        self.t = {k: dev[o:o + n] for k, (o, n) in spans.items()}
        t = self.t
        self.max_ni = max(ni) if ni else 0
        self.enc_lay = dict(q_off=t["enc_q_off"], q_len=t["enc_len"], k_off=t["enc_q_off"], k_len=t["enc_len"],
                            max_lq=Lv, max_lk=Lv, D=D)
        self.cls_lay = dict(q_off=t["cls_q_off"], q_len=t["ones"], k_off=t["enc_q_off"], k_len=t["enc_len"],
                            max_lq=1, max_lk=Lv, D=D)
        self.step_lay = dict(q_off=t["step_off"], q_len=t["step_len"], k_off=t["step_off"], k_len=t["step_len"],
                             max_lq=max(steps), max_lk=max(steps), D=D)
        self.dec_lay = dict(q_off=t["dec_off"], q_len=t["dec_len"], k_off=t["dec_off"], k_len=t["dec_len"],
                            max_lq=Lt, max_lk=Lt, D=D)

    def _tok_csr(self, ingr_id_dict):
        """CSR ingredient (global index) -> extended token ids (reference ingr_id_dict, model.py:912-914)."""
        tok_ptr, tok_ids = [0], []
        if self.mode != "video":
            for b in range(self.B):
                d = ingr_id_dict[b]
                bad = [k for k in d if not (0 <= int(k) < self.ni[b])]
                if bad:
                    raise ValueError(f"ingr_id_dict[{b}] has ingredient indices {sorted(bad)} but the video has "
                                     f"{self.ni[b]} [SEP]-terminated ingredients (reference model.py:1067)")
                for i in range(self.ni[b]):
                    ids = list(d.get(i, []))
                    for t in ids:
                        if not (0 <= int(t) < self.vx[b]):
                            raise ValueError(f"ingr_id_dict[{b}][{i}] holds token id {t} outside [0, {self.vx[b]})")
                    tok_ids += [int(t) for t in ids]
                    tok_ptr.append(len(tok_ids))
        return tok_ptr, tok_ids

    def signature(self):
        """What a captured CUDA graph of the step depends on (shapes of every launch)."""
        return (tuple(self.steps), tuple(self.ni), tuple(self.vx), self._spans["tok_ids"][1])

    def update_tokens(self, ingr_id_dict):
        """Refresh the copy CSR in place (same lengths): used when a captured step is replayed on a new batch.  The
        upload is asynchronous and the host may run several steps ahead (prefetch / run_prefetched / loss_to_host), so
        every call writes its own pinned slot of a small ring and waits only for the copy that last used that slot."""
        tok_ptr, tok_ids = self._tok_csr(ingr_id_dict)
        if not tok_ids:
            tok_ids = [0]
        o, n = self._spans["tok_ids"]
        if len(tok_ids) != n:
            return False
        if self._dev.device.type != "cuda":
            src = self._host
        else:
            if self._ring is None:
                self._ring = [(self._host.clone().pin_memory(), torch.cuda.Event()) for _ in range(4)]
            src, ev = self._ring[self._ring_i]
            self._ring_i = (self._ring_i + 1) % len(self._ring)
            ev.synchronize()                       # no-op unless the copy issued 4 updates ago is still pending
        src[o:o + n] = torch.tensor(tok_ids, dtype=torch.int32)
        o2, n2 = self._spans["tok_ptr"]
        src[o2:o2 + n2] = torch.tensor(tok_ptr, dtype=torch.int32)
        self._dev.copy_(src, non_blocking=True)
        if self._dev.device.type == "cuda":
            ev.record(torch.cuda.current_stream(self._dev.device))
        return True

    def cross_lay(self, C, device, D):
        key = ("cross", C)
        if key not in self.t:
            Nc = self.Nc
            self.t[key] = torch.tensor([n * C for n in range(Nc)] + [C] * Nc, dtype=torch.int32).to(device)
        v = self.t[key]
        return dict(q_off=self.t["dec_off"], q_len=self.t["dec_len"], k_off=v[:self.Nc], k_len=v[self.Nc:],
                    max_lq=self.dec_lay["max_lq"], max_lk=C, D=D)


# ------------------------------------------------------------------------------------------------
# module tree with the reference's attribute names
# ------------------------------------------------------------------------------------------------
class StateAwareRecursiveTransformer(nn.Module):
    """Drop-in for reference model.py:826.  ``compute_dtype``: torch.float32 = parity path (fp32 operands, fp32
    accumulate), torch.bfloat16 = fast path (bf16 operands on tcgen05, fp32 accumulate / statistics / losses)."""

    def __init__(self, config, compute_dtype=torch.float32):
        super().__init__()
        self.config = config if hasattr(config, "hidden_size") else SvpcConfig(config)
        cfg = self.config
        if getattr(cfg, "share_wd_cls_weight", False):
            # reference model.py:854 raises AttributeError (self.embeddings does not exist); keep it loud
            raise AttributeError("'StateAwareRecursiveTransformer' object has no attribute 'embeddings'")
        assert cfg.label_smoothing > 0, "label_smoothing == 0 is a crash path in the reference (model.py:870)"
        self.compute_dtype = compute_dtype
        self.eps_prob = 1e-12
        self.gumbel_noise = None      # optional list of (S_b, Lt, Vx_b) tensors consumed by the next forward
        std = float(cfg.initializer_range)
        H = cfg.hidden_size
        for name, (shape, kind) in state_dict_spec(SvpcConfig(cfg)).items():
            parts = name.split(".")
            node = self
            for i, p in enumerate(parts[:-1]):
                if p not in node._modules:
                    node.add_module(p, _make_node(self, ".".join(parts[:i + 1]), p))
                node = node._modules[p]
            leaf = parts[-1]
            if kind.startswith("pe:"):
                node.register_buffer(leaf, sinusoid_table(int(kind[3:]), shape[1]))
                continue
            t = torch.empty(shape)
            if kind in ("lin_w", "emb", "act_emb"):
                t.normal_(0.0, std)                                  # init_bert_weights, model.py:875-885
            elif kind == "zero":
                t.zero_()
            elif kind == "one":
                t.fill_(1.0)
            elif kind == "lstm":
                k = 1.0 / math.sqrt(H)
                t.uniform_(-k, k)
            node.register_parameter(leaf, nn.Parameter(t))
        self._pcache = {}
        self.last_parts = {}
        self.drop_sites = []
        self._seed_t, self._site = None, 0      # train-mode dropout: device step seed (bumped per forward), site counter

    # ---- parameter lookup by reference key
    def p(self, name):
        t = self._pcache.get(name)
        if t is None:
            obj = self
            for part in name.split("."):
                obj = getattr(obj, part)
            self._pcache[name] = t = obj
        return t

    def _apply(self, fn, *a, **k):
        r = super()._apply(fn, *a, **k)
        self._invalidate()
        return r

    def train(self, mode: bool = True):
        """Switching between train and eval marks the cached compute-dtype weight copies stale: an optimiser that updates
        ``param.data`` in place (the reference's BertAdam, optimization.py:327) does not bump the autograd version counters
        the cache is keyed on."""
        Fn.WEIGHTS.invalidate()
        return super().train(mode)

    def _invalidate(self):
        """Parameters were moved / replaced: drop the name -> tensor cache and mark the compute-dtype weight copies stale.
        Their buffers are NOT freed here: another model's captured graph may be reading them (the cache is keyed by
        parameter identity and version + address, so a moved parameter misses by itself; entries of parameters that no
        longer exist are purged on the next miss)."""
        self._pcache = {}
        Fn.WEIGHTS.invalidate()

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        # a reference checkpoint saved without set_pretrained_embedding holds `reasoner.action_embeddings.weight`
        # (nn.Embedding); after it the key is the bare Parameter `reasoner.action_embeddings` (model.py:773-775)
        for r in ("reasoner", "recipe_reasoner"):
            old = f"{prefix}{r}.action_embeddings.weight"
            if old in state_dict:
                state_dict[f"{prefix}{r}.action_embeddings"] = state_dict.pop(old)
        return super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)

    # ------------------------------------------------------------------------------------------
    # building blocks (all batched over clips)
    # ------------------------------------------------------------------------------------------
    def _begin_step(self, device):
        """Train mode: new dropout masks for this forward (the seed lives on the device so that a captured graph draws
        fresh masks on every replay)."""
        self._site = 0
        self.drop_sites = []
        if self.training:
            if self._seed_t is None or self._seed_t.device != device:
                self._seed_t = torch.randint(1, 2 ** 30, (1,), dtype=torch.int32, device=device)
            else:
                self._seed_t += 1

    def _drop(self, p, tag=None):
        """(seed, site, p) for one nn.Dropout site of the reference, or None in eval mode / p == 0.  ``tag`` names the
        reference module the site belongs to; (site, tag, p) of the current forward are kept in ``self.drop_sites`` so that
        a test can rebuild every mask with ``svpc_dropout_mask`` and hand it to the oracle."""
        if not self.training or p <= 0.0 or self._seed_t is None:
            return None
        self._site += 1
        self.drop_sites.append((self._site, tag, float(p)))
        return (self._seed_t, self._site, float(p))

    def _lin(self, x, prefix, **kw):
        return Fn.linear(x, self.p(prefix + ".weight"), self.p(prefix + ".bias"), cd=self.compute_dtype, **kw)

    def _ln(self, x, res, prefix, **kw):
        return Fn.layer_norm(x, res, self.p(prefix + ".weight"), self.p(prefix + ".bias"), self.config.layer_norm_eps, **kw)

    def _qkv(self, x, prefix, names, **kw):
        return Fn.linear(x, [self.p(f"{prefix}.{n}.weight") for n in names], [self.p(f"{prefix}.{n}.bias") for n in names],
                         cd=self.compute_dtype, **kw)

    def _encoder_layer(self, x, prefix, lay, key_mask, cls_lay=None, cls_stride=0):
        """BertLayerNoMemoryUntied (reference model.py:574-591).  With ``cls_lay`` only the first row of every
        ``cls_stride``-row sequence is produced (the only row the reference consumes of the last clip-encoder layer,
        model.py:1064): keys / values still come from all rows."""
        cfg, cd, D, Hh = self.config, self.compute_dtype, self.config.hidden_size, self.config.num_attention_heads
        ap = prefix + ".attention"
        # backward hand-offs (Fn.Link): Lx / La carry the skip-connection gradients of x / a into the input-gradient GEMMs of
        # the QKV / intermediate projections; Lo, Lh, L2 fuse the dropout / GELU backward into ln_bwd and the dgrad epilogue
        Lx, Lo, La, Lh, L2 = Fn.Link(), Fn.Link(), Fn.Link(), Fn.Link(), Fn.Link()
        if cls_lay is None:
            qkv = self._qkv(x, ap + ".self", ("query", "key", "value"), in_link=Lx)
            ctx = Fn.attention(qkv, None, (0, D, 2 * D), lay, key_mask, False, Hh, cd,
                               drop=self._drop(cfg.attention_probs_dropout_prob, prefix + ".attn"))
            res = x
        else:
            kv = self._qkv(x, ap + ".self", ("key", "value"))
            res = x[::cls_stride].contiguous()
            q = self._lin(res, ap + ".self.query")
            ctx = Fn.attention(q, kv, (0, 0, D), cls_lay, key_mask, False, Hh, cd,
                               drop=self._drop(cfg.attention_probs_dropout_prob, prefix + ".attn"))
            Lx = None
        ph = cfg.hidden_dropout_prob
        a = self._ln(self._lin(ctx, ap + ".output.dense", residual=res, drop=self._drop(ph, prefix + ".attn_out"), res_link=Lx,
                               out_link=Lo), None, ap + ".output.LayerNorm", grad_link=Lo)
        h = self._lin(a, prefix + ".hidden_intermediate.dense", act=ACT_GELU, in_link=La, out_link=Lh)
        return self._ln(self._lin(h, prefix + ".output.dense", residual=a, drop=self._drop(ph, prefix + ".out"), in_link=Lh,
                                  res_link=La, out_link=L2), None, prefix + ".output.LayerNorm", grad_link=L2)

    def _video_embed(self, feats_list, plan, vid_ids):
        """BertEmbeddingsVideoUntied + token-type add (reference model.py:548-562, 889-891).  LN(3072)'s affine is
        folded into the projection (x_hat @ (W*gamma)^T + (b + W beta)), so the normalised features are written once,
        in the compute dtype, and the 90 GFLOP input-gradient GEMM is not needed for gamma / beta."""
        cfg = self.config
        pfx = "video_embeddings.video_embeddings"
        drop = self._drop(cfg.hidden_dropout_prob, "video_embeddings")
        if drop is not None:
            # train mode: dropout sits between LN(3072) and the projection (model.py:549-551), so the affine cannot be
            # folded into the weights; the normalised, scaled and dropped features are staged once in the compute dtype
            a = Fn.VideoLNDropFn.apply(self.p(pfx + ".0.weight"), self.p(pfx + ".0.bias"), list(feats_list), plan.t["clip_t"],
                                       plan.t["clip_b"], cfg.max_v_len, cfg.layer_norm_eps, self.compute_dtype, drop)
            Lz = Fn.Link()
            z = self._lin(a, pfx + ".2", act=ACT_RELU, out_link=Lz)
            return self._ln(z, None, pfx + ".4", pe=self.p("video_embeddings.position_embeddings_video.pe"),
                            pe_period=cfg.max_v_len, tt=self.p("token_type_embeddings.weight"), tt_ids=vid_ids, grad_link=Lz)
        xhat, _, _ = Fn.be().video_ln_pack(feats_list, plan.t["clip_t"], plan.t["clip_b"], cfg.max_v_len,
                                          cfg.layer_norm_eps, self.compute_dtype)
        W, g0, b0 = self.p(pfx + ".2.weight"), self.p(pfx + ".0.weight"), self.p(pfx + ".0.bias")
        Wf = W * g0
        bf = self.p(pfx + ".2.bias") + (W * b0).sum(dim=1)
        Lz = Fn.Link()
        z = Fn.linear(xhat, Wf, bf, act=ACT_RELU, cd=self.compute_dtype, cache=False, out_link=Lz)   # Wf is a per-step temporary
        return self._ln(z, None, pfx + ".4", pe=self.p("video_embeddings.position_embeddings_video.pe"),
                        pe_period=cfg.max_v_len, tt=self.p("token_type_embeddings.weight"), tt_ids=vid_ids, grad_link=Lz)

    def _word_fc(self, which, ids, pe=None, pe_period=0):
        """Embedding -> LN(W) -> Linear -> ReLU -> LN(D) (+pe)  (reference model.py:492-499 / 519-526)."""
        cfg = self.config
        pfx = which + ".word_fc"
        e = Fn.EmbedLNFn.apply(self.p(which + ".word_embeddings.weight"), ids, self.p(pfx + ".0.weight"),
                               self.p(pfx + ".0.bias"), cfg.layer_norm_eps, self.compute_dtype, PAD,
                               self._drop(cfg.hidden_dropout_prob, which))
        Lz = Fn.Link()
        z = self._lin(e, pfx + ".2", act=ACT_RELU, out_link=Lz)
        return self._ln(z, None, pfx + ".4", pe=pe, pe_period=pe_period, grad_link=Lz)

    def _reasoner(self, prefix, V, E0, plan):
        """EntitiyReasoningNetwork.forward (reference model.py:777-823): everything that depends only on the step
        vector is batched over all clips; the entity-state recurrence is one sequential kernel per video."""
        f32 = torch.float32
        h_as = self._lin(V, prefix + ".action_selector.0", act=ACT_RELU,
                       drop=self._drop(0.4, prefix + ".action_selector"))                                # nn.Dropout(p=0.4), model.py:757
        ac = self._lin(h_as, prefix + ".action_selector.3", act=ACT_SIGMOID, out_dtype=f32)            # (Nc, A)
        fbar = Fn.linear(Fn.RowNormalizeFn.apply(ac), self.p(prefix + ".action_embeddings"), None, cd=self.compute_dtype,
                         out_dtype=f32, w_is_kn=True)                                                     # (Nc, W)  Eq. 1
        hhat = self._lin(V, prefix + ".W1.0", act=ACT_RELU)                                              # Eq. 2
        w2v = Fn.linear([hhat, ac], self.p(prefix + ".W2.weight"), self.p(prefix + ".W2.bias"), cd=self.compute_dtype,
                        out_dtype=f32)                                                                    # (Nc, D)
        clog = self._lin(hhat, prefix + ".W3", out_dtype=f32)                                            # (Nc, 3)
        w4s = self._lin(fbar, prefix + ".W4", out_dtype=f32).reshape(-1)                                 # (Nc,)
        lay = dict(cu_ingr=plan.t["cu_ingr"], cu_steps=plan.t["cu_steps"], clip_ent_off=plan.t["clip_ent_off"],
                   max_ni=plan.max_ni, NE=plan.NE, B=plan.B, max_steps=max(plan.steps) if plan.steps else 0)
        ent, ebar, Es = Fn.ReasonerFn.apply(E0, w2v, clog, w4s, lay)
        return ent, ac, ebar, Es, fbar

    def _decoder(self, x, text_mask, ctx, plan, C):
        """BertDecoderNoMemoryUntied (reference model.py:630-694)."""
        cfg, cd, D, Hh = self.config, self.compute_dtype, self.config.hidden_size, self.config.num_attention_heads
        cross = plan.cross_lay(C, x.device, D)
        for l in range(cfg.num_hidden_layers):
            pfx = f"decoder.layer.{l}"
            Lx, L1, L2, Lo = Fn.Link(), Fn.Link(), Fn.Link(), Fn.Link()      # skip-connection / dropout hand-offs (Fn.Link)
            qkv = self._qkv(x, pfx + ".self_attention", ("query", "key", "value"), in_link=Lx)
            pa = cfg.attention_probs_dropout_prob
            sa = Fn.attention(qkv, None, (0, D, 2 * D), plan.dec_lay, text_mask, True, Hh, cd,
                              drop=self._drop(pa, pfx + ".self_attn"))
            h1 = self._ln(sa, x, pfx + ".norm1", res_link=Lx)
            q2 = self._lin(h1, pfx + ".dec_enc_attention.query", in_link=L1)
            kv2 = self._qkv(ctx, pfx + ".dec_enc_attention", ("key", "value"))
            ca = Fn.attention(q2, kv2, (0, 0, D), cross, None, False, Hh, cd, drop=self._drop(pa, pfx + ".cross_attn"))
            h2 = self._ln(ca, h1, pfx + ".norm2", res_link=L1)
            x = self._ln(self._lin(h2, pfx + ".output.dense", residual=h2, drop=self._drop(cfg.hidden_dropout_prob, pfx + ".out"),
                                   in_link=L2, res_link=L2, out_link=Lo), None, pfx + ".output.LayerNorm", grad_link=Lo)
        return x

    def _lm_logits(self, dec, dec_link=None):
        """BertLMPredictionHead (reference model.py:704-709, 735-739) -> fp32 logits (R, V).  dec_link: where the pointer
        generator leaves its share of d(dec) (added in this Linear's input-gradient GEMM)."""
        Lt = Fn.Link()
        t = self._lin(dec, "decoder_classifier.transform.dense", act=ACT_GELU, in_link=dec_link, out_link=Lt)
        t = self._ln(t, None, "decoder_classifier.transform.LayerNorm", grad_link=Lt)
        return Fn.linear(t, self.p("decoder_classifier.decoder.weight"), self.p("decoder_classifier.bias"),
                         cd=self.compute_dtype, out_dtype=torch.float32)

    # ------------------------------------------------------------------------------------------
    def _gather_tokens(self, tensors, plan):
        """list of S (B, L) tensors -> (Nc, L) in clip order."""
        return torch.stack(list(tensors))[plan.t["clip_t"].long(), plan.t["clip_b"].long()]

    def _encode_clips(self, input_ids_list, video_features_list, input_masks_list, plan):
        """forward_step for every clip + step-wise encoder -> step vectors (Nc, D)  (reference model.py:1038-1065)."""
        cfg, cd = self.config, self.compute_dtype
        Lv, D = cfg.max_v_len, cfg.hidden_size
        ids = self._gather_tokens(input_ids_list, plan)
        masks = self._gather_tokens(input_masks_list, plan)
        vid_ids = ids[:, :Lv].reshape(-1).to(torch.int32)
        vid_mask = masks[:, :Lv].reshape(-1).to(torch.float32).contiguous()
        x = self._video_embed(video_features_list, plan, vid_ids)
        nl = cfg.num_hidden_layers
        for l in range(nl):
            last = l == nl - 1
            x = self._encoder_layer(x, f"encoder.layer.{l}", plan.enc_lay, vid_mask,
                                    cls_lay=plan.cls_lay if last else None, cls_stride=Lv)
        cls = x if nl > 0 else x[::Lv]
        pe = self.p("step_positional_encoding.pe")
        x = (cls.float() + pe[plan.t["clip_t"].long()]).to(cd)
        for l in range(nl):
            x = self._encoder_layer(x, f"step_wise_encoder.layer.{l}", plan.step_lay, None)
        return x, cls, ids, masks

    def _ingredient_vectors(self, ingr_input_ids, ingr_sep_masks, plan):
        ids = ingr_input_ids.reshape(-1).to(torch.int32)
        x = self._word_fc("ingredient_embeddings", ids)
        return Fn.SegMeanFn.apply(x, ingr_sep_masks.to(torch.int32).contiguous(), plan.t["cu_ingr"],
                                  self.p("ingredient_embeddings.position_embeddings_ingr.pe"), plan.NI)

    def _context(self, mode, V, E0, ebar, fbar, plan):
        cd = self.compute_dtype
        if mode in ("full", "reason_copy"):
            ent = self._lin(ebar, "Went.0", act=ACT_RELU)
            act = self._lin(fbar, "Wac.0", act=ACT_RELU)
            return torch.stack([V, ent, act], dim=1).reshape(-1, V.shape[1]), 3
        if mode == "copy":
            m = Fn.RaggedMeanFn.apply(E0, plan.t["cu_ingr"])[plan.t["clip_b"].long()].to(cd)
            return torch.stack([V, m], dim=1).reshape(-1, V.shape[1]), 2
        return V, 1

    def make_plan(self, batch_step_num, ingr_id_dict, extra_zeros, alignments, device):
        mode = self.config.model_mode
        steps = [int(s) for s in batch_step_num]
        B = len(steps)
        if mode == "video":
            ni, ez, dct = [0] * B, [0] * B, None
        else:
            ni = self.ingredient_counts(getattr(self, "_sep_counts", None), ingr_id_dict)
            ez = [int(z) for z in extra_zeros]
            dct = ingr_id_dict
        return StepPlan(self.config, steps, ni, ez, dct, device, mode)

    @staticmethod
    def ingredient_counts(ingr_sep_masks, ingr_id_dict):
        """Ingredients per video.  The reference counts the [SEP]s of the ingredient row, (ingr_sep_masks[b] == 1).sum()
        (model.py:1067); with a host-resident mask (lists, CPU tensors - what the loader and the translator hold) that
        count is used, so an ingredient that tokenises to nothing (a key missing from ingr_id_dict) still has its row.
        With a device tensor the count would need a device->host sync per step: the dictionary length is used, and
        StepPlan checks that every key fits."""
        if ingr_sep_masks is not None and not (isinstance(ingr_sep_masks, torch.Tensor) and ingr_sep_masks.is_cuda):
            t = torch.as_tensor(ingr_sep_masks)
            return [int(x) for x in (t == 1).sum(dim=-1).tolist()]
        return [len(d) for d in ingr_id_dict]

    def make_plan_raw(self, steps, ni, extra_zeros, ingr_id_dict, device):
        """StepPlan from explicit per-video counts (piecewise callers: svpc_b200.compat, svpc_b200.decode)."""
        return StepPlan(self.config, [int(s) for s in steps], [int(n) for n in ni], [int(z) for z in extra_zeros],
                        ingr_id_dict, device, "copy")

    # ---- piecewise entry points used by the reference translator (src/translator.py:57-159)
    def forward_step(self, input_ids, video_features, input_masks):
        """Reference model.py:887-894: (N, L) ids, (N, L, F) features, (N, L) masks -> (N, max_v_len, D), all rows of
        the last encoder layer (the batched training step only evaluates its [CLS] row)."""
        cfg = self.config
        Lv, N = cfg.max_v_len, input_ids.shape[0]
        plan = self.make_plan_raw([1] * N, [0] * N, [0] * N, [{} for _ in range(N)], video_features.device)
        vid_ids = input_ids[:, :Lv].reshape(-1).to(torch.int32)
        vid_mask = input_masks[:, :Lv].reshape(-1).to(torch.float32).contiguous()
        x = self._video_embed([video_features.contiguous()], plan, vid_ids)
        for l in range(cfg.num_hidden_layers):
            x = self._encoder_layer(x, f"encoder.layer.{l}", plan.enc_lay, vid_mask)
        return x.reshape(N, Lv, -1)

    def pointer_generator_network(self, decoder_outputs, ingredient_vectors, ingr_id_dict, extra_zero):
        """Reference model.py:896-923 for ONE video: (S, Lt, D), (S, Ni, D), {ingredient: [token ids]}, n_oov ->
        (S, Lt, V + n_oov) probabilities."""
        cfg = self.config
        S, Lt, D = decoder_outputs.shape
        Ni = ingredient_vectors.shape[1]
        plan = self.make_plan_raw([S], [Ni], [extra_zero], [ingr_id_dict], decoder_outputs.device)
        dec = decoder_outputs.reshape(S * Lt, D).to(self.compute_dtype)
        Esrc = ingredient_vectors.reshape(S * Ni, D).float().contiguous()
        P = self._pointer(dec, Esrc, plan, plan.t["clip_ent_off"],
                          torch.full((S * Lt,), IGNORE, dtype=torch.int32, device=dec.device), Lt)[0]
        return P.view(S, Lt, plan.vx[0])

    def _pointer(self, dec, Esrc, plan, clip_src_off, labels, Lt, clip_P_off=None, P_numel=None, G=None):
        cfg = self.config
        if G is None:
            G = self._lin(Esrc, "Wing", out_dtype=torch.float32)
        logits = self._lm_logits(dec)
        lay = dict(Lt=Lt, V=cfg.vocab_size, smoothing=float(cfg.label_smoothing), max_ni=plan.max_ni,
                   clip_src_off=clip_src_off, clip_ni=plan.t["clip_ni"], clip_gi=plan.t["clip_gi"],
                   clip_vx=plan.t["clip_vx"], clip_P_off=plan.t["clip_P_off"] if clip_P_off is None else clip_P_off,
                   P_numel=plan.P_numel if P_numel is None else P_numel, tok_ptr=plan.t["tok_ptr"],
                   tok_ids=plan.t["tok_ids"], labels=labels, max_vx=max(plan.vx))
        return Fn.PointerFn.apply(dec, G, Esrc, self.p("pgen_linear.0.weight").reshape(-1), self.p("pgen_linear.0.bias"),
                                  logits, lay)

    def plan_signature(self, batch_step_num, ingr_id_dict, extra_zeros):
        """Host-only: the StepPlan.signature() a batch would have (what a captured graph of the step depends on)."""
        V = self.config.vocab_size
        steps = tuple(int(s) for s in batch_step_num)
        if self.config.model_mode == "video":
            return (steps, (0,) * len(steps), (V,) * len(steps), 1)
        ni = tuple(len(d) for d in ingr_id_dict)
        ntok = sum(len(d.get(i, [])) for d, n in zip(ingr_id_dict, ni) for i in range(n))
        return (steps, ni, tuple(V + int(z) for z in extra_zeros), max(ntok, 1))

    def forward(self, input_ids_list, video_features_list, input_masks_list, token_type_ids_list, input_labels_list,
                ingr_input_ids, ingr_masks, ingr_sep_masks, batch_step_num, ingr_id_dict, extra_zeros, alignments, actions,
                return_memory=False, predict=False, gumbel_noise=None, plan=None):
        """Reference model.py:1027-1189.  ``token_type_ids_list``, ``ingr_masks`` and ``return_memory`` are accepted and
        unused, as in the reference.  ``gumbel_noise`` (or ``self.gumbel_noise``): list of (S_b, Lt, Vx_b) noise tensors
        for the textual re-simulator; drawn on the device when absent."""
        cfg, cd = self.config, self.compute_dtype
        mode, Lv, Lt, V, D = cfg.model_mode, cfg.max_v_len, cfg.max_t_len, cfg.vocab_size, cfg.hidden_size
        f32 = torch.float32
        dev = video_features_list[0].device
        self._begin_step(dev)
        if self.training:
            Fn.WEIGHTS.invalidate()       # a training step follows an optimiser step: `.data` updates are invisible to versions
        Fn.WEIGHTS.refresh_all()          # compute-dtype weight copies touched by the optimiser: one multi-tensor cast
        if plan is None:
            self._sep_counts = ingr_sep_masks if mode != "video" else None
            plan = self.make_plan(batch_step_num, ingr_id_dict, extra_zeros, alignments, dev)
            self._sep_counts = None
        B, Nc = plan.B, plan.Nc

        Vst, cls, ids, masks = self._encode_clips(input_ids_list, video_features_list, input_masks_list, plan)
        labels = self._gather_tokens(input_labels_list, plan)[:, Lv:]
        text_ids = ids[:, Lv:].reshape(-1).to(torch.int32)
        text_mask = masks[:, Lv:].reshape(-1).to(f32).contiguous()

        E0 = ebar = fbar = Es = ent = ac = None
        if mode != "video":
            E0 = self._ingredient_vectors(ingr_input_ids, ingr_sep_masks, plan)
        if mode in ("full", "reason_copy"):
            ent, ac, ebar, Es, fbar = self._reasoner("reasoner", Vst, E0, plan)
        ctx, C = self._context(mode, Vst, E0, ebar, fbar, plan)
        x = self._word_fc("text_embeddings", text_ids, pe=self.p("text_embeddings.position_embeddings_text.pe"), pe_period=Lt)
        dec = self._decoder(x, text_mask, ctx, plan, C)
        Ld = Fn.Link() if mode != "video" else None
        logits = self._lm_logits(dec, dec_link=Ld)

        if mode == "video":
            lab = torch.where(labels >= V, torch.full_like(labels, cfg.unk_id), labels)          # model.py:1013
            P, caption = Fn.SoftmaxLSFn.apply(logits, lab.reshape(-1).to(torch.int32), V, float(cfg.label_smoothing))
        else:
            Esrc = Es if mode != "copy" else E0
            G = self._lin(Esrc, "Wing", out_dtype=f32)
            lay = dict(Lt=Lt, V=V, smoothing=float(cfg.label_smoothing), max_ni=plan.max_ni,
                       clip_src_off=plan.t["clip_ent_off"] if mode != "copy" else plan.t["clip_src_copy"],
                       clip_ni=plan.t["clip_ni"], clip_gi=plan.t["clip_gi"], clip_vx=plan.t["clip_vx"],
                       clip_P_off=plan.t["clip_P_off"], P_numel=plan.P_numel, tok_ptr=plan.t["tok_ptr"],
                       tok_ids=plan.t["tok_ids"], labels=labels.reshape(-1).to(torch.int32), max_vx=max(plan.vx))
            P, caption = Fn.PointerFn.apply(dec, G, Esrc, self.p("pgen_linear.0.weight").reshape(-1),
                                            self.p("pgen_linear.0.bias"), logits, lay, Ld)

        parts = dict(caption=caption)
        total = caption
        re_ent = re_ac = re_Es = None
        if mode in ("full", "reason_copy"):
            y_al = torch.cat([a.reshape(-1) for a in alignments]).to(f32)
            y_ac = torch.cat(list(actions)).to(f32)
            parts["entity"] = Fn.BCEFn.apply(ent, y_al)
            parts["action"] = Fn.ASLFn.apply(ac, y_ac)
            total = total + parts["entity"] + parts["action"]
            if mode == "full":
                noise = gumbel_noise if gumbel_noise is not None else self.gumbel_noise
                self.gumbel_noise = None
                if noise is None and self.training and self._seed_t is not None:
                    g = Fn.be().gumbel_noise(plan.P_numel, self._seed_t, GUMBEL_SITE)     # drawn on the device, per step seed
                elif noise is None:
                    g = -torch.empty(plan.P_numel, dtype=f32, device=dev).exponential_().log()
                else:
                    g = torch.cat([n.reshape(-1) for n in noise]).to(device=dev, dtype=f32)
                glay = dict(Lt=Lt, V=V, tau=float(cfg.temperature), clip_P_off=plan.t["clip_P_off"], clip_vx=plan.t["clip_vx"],
                            max_vx=max(plan.vx))
                xbow = Fn.GumbelBowFn.apply(P, g, self.p("text_embeddings.word_embeddings.weight"), glay, text_mask)
                llay = dict(cu_steps=plan.t["cu_steps"], steps=plan.steps, cd=self.compute_dtype)
                r = Fn.BiLSTMFn.apply(xbow, llay, LSTM_KEYS, *[self.p("recipe_encoder." + k) for k in LSTM_KEYS])
                re_ent, re_ac, _, re_Es, _ = self._reasoner("recipe_reasoner", r, E0, plan)
                parts["re_entity"] = Fn.BCEFn.apply(re_ent, y_al)
                parts["re_action"] = Fn.ASLFn.apply(re_ac, y_ac)
                total = total + float(cfg.lambda_) * (parts["re_entity"] + parts["re_action"])
        self.last_parts = {k: v.detach() for k, v in parts.items()}

        # ---- per-video views of the flat outputs (the reference returns one tensor per video)
        preds, ents, acts, mems = [], [], [], []
        for b in range(B):
            S = plan.steps[b]
            if mode == "video":
                preds.append(P[plan.cu_steps[b] * Lt:(plan.cu_steps[b] + S) * Lt].view(S, Lt, V))
            else:
                preds.append(P[plan.P_off[b]:plan.P_off[b + 1]].view(S, Lt, plan.vx[b]))
            if mode in ("full", "reason_copy"):
                c0 = plan.cu_steps[b]
                e = ent[plan.ent_off[b]:plan.ent_off[b + 1]].view(S, plan.ni[b])
                a = ac[c0:c0 + S]
                ents.append(e); acts.append(a)
                if predict:
                    i0 = plan.cu_ingr[b]
                    d = {"entity_probs": e, "action_probs": a,
                         "entity_vectors": [E0[i0:i0 + plan.ni[b]], Es[plan.ent_off[b]:plan.ent_off[b + 1]].view(S, plan.ni[b], D)]}
                    if mode == "full":
                        d["re_pred_entity_probs"] = re_ent[plan.ent_off[b]:plan.ent_off[b + 1]].view(S, plan.ni[b])
                        d["re_pred_action_probs"] = re_ac[c0:c0 + S]
                        d["re_entity_vectors"] = re_Es[plan.ent_off[b]:plan.ent_off[b + 1]].view(S, plan.ni[b], D)
                    mems.append(d)
        if predict:
            return mems, ents, acts
        return total, preds, ents, acts

