"""TEST INFRASTRUCTURE ONLY - CPU restatement of the SVPC hot path (never imported by the product).

A from-scratch, functional restatement of ``StateAwareRecursiveTransformer`` (reference
src/rtransformer/model.py) and of ``Translator.translate_batch_greedy`` (reference
src/translator.py:45-192) on plain tensors: a state_dict in the reference checkpoint layout, a config
dict and a collated batch go in; loss components, outputs and (through autograd) gradients come out.
It runs in fp32 or fp64 on the CPU (or any torch device) and is what the CUDA path is compared with.

Parity status: the reference has **no tests, golden vectors or known-answer fixtures** (SURVEY.md §4),
so the oracle is pinned against *outputs of the unmodified reference itself*: ``tests/golden/make_golden.py``
imports the real model (``oracle/refshim.py``) in the build container, runs all four model modes on
seeded synthetic batches and commits the outputs under ``tests/golden/``; ``tests/test_oracle.py``
checks this file against those fixtures everywhere and against the live reference when
``/root/reference`` exists.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs
may import this module.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

PAD, CLS, SEP, VID, BOS, EOS, UNK, IGNORE = 0, 1, 2, 3, 4, 5, 6, -1


# ------------------------------------------------------------------------------------------------
# elementwise building blocks
# ------------------------------------------------------------------------------------------------
def layer_norm(x, w, b, eps=1e-12):
    """BertLayerNorm, TF style, eps inside the sqrt (model.py:152-156)."""
    u = x.mean(-1, keepdim=True)
    s = (x - u).pow(2).mean(-1, keepdim=True)
    return w * ((x - u) / torch.sqrt(s + eps)) + b


def gelu(x):
    """erf-GELU (model.py:58-64)."""
    return x * 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0)))


def linear(x, sd, prefix):
    return F.linear(x, sd[prefix + ".weight"], sd.get(prefix + ".bias"))


class DropMasks:
    """Train mode with EXPLICIT dropout masks: one scale tensor (0 or 1/(1-p)) per reference nn.Dropout site
    (model.py:214, 231, 287, 495, 522, 550, 757), in a per-clip / per-video canonical layout, so that a test can hand the
    oracle exactly the masks the CUDA kernels draw (``svpc_dropout_mask``) - or any other masks.  Keys (Nc = clips in
    video-major order, n = cu_steps[b] + t):
      video_embeddings                       (Nc, Lv, F)      model.py:550
      encoder.layer.l.attn                   (Nc, H, Lv, Lv)  :214      .attn_out / .out  (Nc, Lv, D)  :231 / :287
      step_wise_encoder.layer.l.attn         list of (H, S_b, S_b)      .attn_out / .out  (Nc, D)
      ingredient_embeddings                  (B, Li, W)       :522      text_embeddings   (Nc, Lt, W)  :495
      reasoner.action_selector / recipe_reasoner.action_selector        (Nc, D)           :757
      decoder.layer.l.self_attn (Nc, H, Lt, Lt)   .cross_attn (Nc, H, Lt, C)   .out (Nc, Lt, D)
    A missing key means "no dropout at that site"."""

    def __init__(self, masks, cu_steps):
        self.m, self.cu = masks, list(cu_steps)

    def clips(self, tag, b):
        t = self.m.get(tag)
        return None if t is None else t[self.cu[b]:self.cu[b + 1]]

    def video(self, tag, b):
        t = self.m.get(tag)
        return None if t is None else t[b]


def _dm(x, m):
    return x if m is None else x * m.to(x.dtype)


def attention(xq, xkv, mask, sd, prefix, n_heads, pdrop=None):
    """BertSelfAttention.forward (model.py:181-220). mask: (N, Lq or 1, Lk) of 0/1. No output projection.
    ``pdrop``: (N, H, Lq, Lk) dropout scales of the attention probabilities (:214)."""
    N, Lq, D = xq.shape
    Lk = xkv.shape[1]
    dh = D // n_heads
    add = (1 - mask.unsqueeze(1)) * -10000.0
    q = linear(xq, sd, prefix + ".query").view(N, Lq, n_heads, dh).permute(0, 2, 1, 3)
    k = linear(xkv, sd, prefix + ".key").view(N, Lk, n_heads, dh).permute(0, 2, 1, 3)
    v = linear(xkv, sd, prefix + ".value").view(N, Lk, n_heads, dh).permute(0, 2, 1, 3)
    s = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(dh) + add
    p = _dm(torch.softmax(s, dim=-1), pdrop)
    ctx = torch.matmul(p, v).permute(0, 2, 1, 3).reshape(N, Lq, D)
    return ctx


def encoder_layer(x, key_mask, sd, prefix, n_heads, eps, dm=(None, None, None)):
    """BertLayerNoMemoryUntied.forward with diagonal_mask=False (model.py:574-591, 242-253, 262-265, 285-289).
    ``dm``: dropout scales of (attention probabilities, BertSelfOutput.dense output, BertOutput.dense output)."""
    ctx = attention(x, x, key_mask.unsqueeze(1), sd, prefix + ".attention.self", n_heads, dm[0])
    a = layer_norm(_dm(linear(ctx, sd, prefix + ".attention.output.dense"), dm[1]) + x,
                   sd[prefix + ".attention.output.LayerNorm.weight"], sd[prefix + ".attention.output.LayerNorm.bias"], eps)
    h = gelu(linear(a, sd, prefix + ".hidden_intermediate.dense"))
    return layer_norm(_dm(linear(h, sd, prefix + ".output.dense"), dm[2]) + a,
                      sd[prefix + ".output.LayerNorm.weight"], sd[prefix + ".output.LayerNorm.bias"], eps)


def encoder(x, key_mask, sd, prefix, cfg, dms=None):
    for l in range(cfg["num_hidden_layers"]):
        x = encoder_layer(x, key_mask, sd, f"{prefix}.layer.{l}", cfg["num_attention_heads"], cfg["layer_norm_eps"],
                          (None, None, None) if dms is None else dms[l])
    return x


def decoder_layer(x, dec_mask, ctx, sd, prefix, n_heads, eps, dm=(None, None, None)):
    """BertDecoderLayerNoMemoryUntied.forward (model.py:630-663): causal+pad self-attention, cross-attention to
    the context tokens with an all-ones mask (model.py:947), BertOutput with itself as residual.
    ``dm``: dropout scales of (self-attention probabilities, cross-attention probabilities, BertOutput.dense output)."""
    Lt = x.shape[1]
    m = dec_mask.unsqueeze(1) * torch.tril(dec_mask.new_ones(Lt, Lt))
    h1 = layer_norm(attention(x, x, m, sd, prefix + ".self_attention", n_heads, dm[0]) + x,
                    sd[prefix + ".norm1.weight"], sd[prefix + ".norm1.bias"], eps)
    ones = ctx.new_ones(ctx.shape[0], 1, ctx.shape[1])
    h2 = layer_norm(h1 + attention(h1, ctx, ones, sd, prefix + ".dec_enc_attention", n_heads, dm[1]),
                    sd[prefix + ".norm2.weight"], sd[prefix + ".norm2.bias"], eps)
    return layer_norm(_dm(linear(h2, sd, prefix + ".output.dense"), dm[2]) + h2,
                      sd[prefix + ".output.LayerNorm.weight"], sd[prefix + ".output.LayerNorm.bias"], eps)


def decoder(x, dec_mask, ctx, sd, cfg, dms=None):
    for l in range(cfg["num_hidden_layers"]):
        x = decoder_layer(x, dec_mask, ctx, sd, f"decoder.layer.{l}", cfg["num_attention_heads"], cfg["layer_norm_eps"],
                          (None, None, None) if dms is None else dms[l])
    return x


# ------------------------------------------------------------------------------------------------
# embeddings
# ------------------------------------------------------------------------------------------------
def word_fc(ids, sd, prefix, eps, dm=None):
    """Embedding -> LN(W) -> (dropout) -> Linear(W->D) -> ReLU -> LN(D) (model.py:492-499 / 519-526)."""
    e = sd[prefix + ".word_embeddings.weight"][ids]
    e = _dm(layer_norm(e, sd[prefix + ".word_fc.0.weight"], sd[prefix + ".word_fc.0.bias"], eps), dm)
    e = torch.relu(linear(e, sd, prefix + ".word_fc.2"))
    return layer_norm(e, sd[prefix + ".word_fc.4.weight"], sd[prefix + ".word_fc.4.bias"], eps)


def text_embed(ids, sd, cfg, dm=None):
    """BertEmbeddingsTextUntied.forward (model.py:509-513): ids (N, Lt)."""
    x = word_fc(ids, sd, "text_embeddings", cfg["layer_norm_eps"], dm)
    return x + sd["text_embeddings.position_embeddings_text.pe"][: ids.shape[1]]


def ingredient_vectors(ingr_ids, ingr_sep, sd, cfg, dm=None):
    """BertEmbeddingsIngredientsUntied + IngredientPositionEncoding (model.py:534-537, 116-140) for ONE video.

    ingr_ids, ingr_sep: (Li,). Returns (Ni, D): ingredient j = mean of the token vectors strictly between the
    previous [SEP] and its own [SEP] (the SEP itself excluded, model.py:131-133) plus pe[j].
    """
    x = word_fc(ingr_ids, sd, "ingredient_embeddings", cfg["layer_norm_eps"], dm)
    seps = (ingr_sep == 1).nonzero().view(-1).tolist()
    out, prev = [], 0
    for s in seps:
        out.append(x[prev:s].mean(dim=0))
        prev = s + 1
    out = torch.stack(out)
    return out + sd["ingredient_embeddings.position_embeddings_ingr.pe"][: len(seps)]


def video_embed(feats, ids, sd, cfg, dm=None):
    """BertEmbeddingsVideoUntied.forward + token-type add (model.py:558-562, 889-891). feats (N, Lv, F)."""
    eps = cfg["layer_norm_eps"]
    p = "video_embeddings.video_embeddings"
    x = _dm(layer_norm(feats, sd[p + ".0.weight"], sd[p + ".0.bias"], eps), dm)
    x = torch.relu(linear(x, sd, p + ".2"))
    x = layer_norm(x, sd[p + ".4.weight"], sd[p + ".4.bias"], eps)
    x = x + sd["video_embeddings.position_embeddings_video.pe"][: feats.shape[1]]
    return x + sd["token_type_embeddings.weight"][ids]


def forward_step(ids, feats, masks, sd, cfg, dm=None, dms=None):
    """forward_step (model.py:887-894) for a stack of clips: ids/masks (N, L), feats (N, L, F) -> (N, Lv, D)."""
    Lv = cfg["max_v_len"]
    x = video_embed(feats[:, :Lv], ids[:, :Lv], sd, cfg, dm)
    return encoder(x, masks[:, :Lv], sd, "encoder", cfg, dms)


# ------------------------------------------------------------------------------------------------
# visual simulator / textual re-simulator
# ------------------------------------------------------------------------------------------------
def reasoner(step_vecs, E0, sd, prefix, dm=None):
    """EntitiyReasoningNetwork.forward (model.py:777-823). step_vecs (S, D), E0 (Ni, D); ``dm`` (S, D): scales of the
    nn.Dropout(0.4) inside action_selector (:757), absent = eval mode.

    Returns (ent_probs (S,Ni), ac_probs (S,A), ebar (S,D), E_stack (S,Ni,D), fbar (S,W)) - the reference's
    return order (model.py:823)."""
    S = step_vecs.shape[0]
    A_emb = sd[prefix + ".action_embeddings"]
    E = E0
    e_prev = E0.new_zeros(E0.shape[0])
    ent, acs, ebars, Es, fbars = [], [], [], [], []
    for t in range(S):
        v = step_vecs[t]
        h_as = torch.relu(linear(v, sd, prefix + ".action_selector.0"))
        ac = torch.sigmoid(linear(_dm(h_as, None if dm is None else dm[t]), sd, prefix + ".action_selector.3"))
        fbar = (ac / ac.sum()).unsqueeze(0) @ A_emb                                   # Eq. 1  (1, W)
        hhat = torch.relu(linear(v, sd, prefix + ".W1.0"))                            # Eq. 2
        e = torch.sigmoid(E @ linear(torch.cat([hhat, ac]), sd, prefix + ".W2"))      # (Ni,)
        c = torch.softmax(linear(hhat, sd, prefix + ".W3"), dim=-1)                   # Eq. 3
        a = c[0] * e + c[1] * e_prev                                                  # third term multiplies zeros
        ebar = (a / a.sum()).unsqueeze(0) @ E                                         # Eq. 4-5  (1, D)
        k = torch.relu(linear(fbar, sd, prefix + ".W4") @ ebar)                       # Eq. 6   (1,1)@(1,D)
        E = a.unsqueeze(1) @ k + (1 - a).unsqueeze(1) * E                             # Eq. 7
        e_prev = e                                                                    # model.py:813
        ent.append(e); acs.append(ac); ebars.append(ebar[0]); Es.append(E); fbars.append(fbar[0])
    return torch.stack(ent), torch.stack(acs), torch.stack(ebars), torch.stack(Es), torch.stack(fbars)


def lm_head(x, sd, eps):
    """BertLMPredictionHead.forward (model.py:735-739, 704-709)."""
    h = layer_norm(gelu(linear(x, sd, "decoder_classifier.transform.dense")),
                   sd["decoder_classifier.transform.LayerNorm.weight"], sd["decoder_classifier.transform.LayerNorm.bias"], eps)
    return F.linear(h, sd["decoder_classifier.decoder.weight"]) + sd["decoder_classifier.bias"]


def pointer_generator(dec, E_src, ingr_dict, extra_zero, sd, cfg):
    """pointer_generator_network (model.py:896-923). dec (S,Lt,D); E_src (S,Ni,D). Returns probabilities (S,Lt,V+extra)."""
    G = linear(E_src, sd, "Wing")                                         # (S,Ni,D)
    scores = torch.einsum("sic,slc->isl", G, dec)                         # (Ni,S,Lt)
    alpha = torch.softmax(scores, dim=0)                                  # over the ingredient axis (model.py:900)
    ctx = torch.einsum("isl,sic->slc", alpha, E_src)                      # (S,Lt,D)
    p_gen = torch.sigmoid(linear(torch.cat([dec, ctx], dim=2), sd, "pgen_linear.0"))   # (S,Lt,1)
    P = p_gen * torch.softmax(lm_head(dec, sd, cfg["layer_norm_eps"]), dim=-1)
    alpha = alpha * (1 - p_gen).squeeze(2)
    if extra_zero > 0:
        P = torch.cat([P, P.new_zeros(P.shape[0], P.shape[1], extra_zero)], dim=-1)
    add = torch.zeros_like(P)
    for i, ids in ingr_dict.items():
        for tok in ids:
            add[:, :, tok] = add[:, :, tok] + alpha[i] / len(ids)
    return P + add


def label_smoothing_loss(P, labels, smoothing):
    """LabelSmoothingLoss.forward (model.py:37-55) on probabilities P (R, Vx), labels (R,) with -1 ignored."""
    valid = labels != IGNORE
    tgt = labels[valid]
    logp = torch.log(P[valid] + 1e-12)
    Vx = P.shape[1]
    q = P.new_full((Vx,), smoothing / (Vx - 1))
    q[IGNORE] = 0                                            # one_hot[ignore_index]: zeroes the LAST column
    q = q.repeat(tgt.shape[0], 1)
    q.scatter_(1, tgt.unsqueeze(1), 1.0 - smoothing)
    return F.kl_div(logp, q, reduction="sum")


def gumbel_hard(logits, g, tau):
    """F.gumbel_softmax(hard=True) with the Gumbel noise g injected (torch/nn/functional.py gumbel_softmax)."""
    y_soft = ((logits + g) / tau).softmax(-1)
    index = y_soft.max(-1, keepdim=True)[1]
    y_hard = torch.zeros_like(logits).scatter_(-1, index, 1.0)
    return y_hard - y_soft.detach() + y_soft


def bilstm_sum(x, sd):
    """nn.LSTM(W->D, bidirectional) over x (S, W), batch 1, directions summed (model.py:865, 1022-1024)."""
    D = sd["recipe_encoder.weight_hh_l0"].shape[1]

    def run(suffix, order):
        h = x.new_zeros(D)
        c = x.new_zeros(D)
        out = [None] * x.shape[0]
        for t in order:
            gates = F.linear(x[t], sd["recipe_encoder.weight_ih_l0" + suffix], sd["recipe_encoder.bias_ih_l0" + suffix]) + \
                F.linear(h, sd["recipe_encoder.weight_hh_l0" + suffix], sd["recipe_encoder.bias_hh_l0" + suffix])
            i, f, gg, o = gates.chunk(4)                     # PyTorch gate order i, f, g, o
            c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
            h = torch.sigmoid(o) * torch.tanh(c)
            out[t] = h
        return torch.stack(out)

    S = x.shape[0]
    return run("", range(S)) + run("_reverse", range(S - 1, -1, -1))


def reconstruct(P, text_mask, E0, g, sd, cfg, dm=None):
    """reconstruct (model.py:1017-1025): Gumbel-ST word samples -> bag of words per step -> BiLSTM -> recipe_reasoner."""
    V = cfg["vocab_size"]
    y = gumbel_hard(torch.log(P + 1e-12), g, cfg["temperature"])[:, :, :V]
    bow = y @ sd["text_embeddings.word_embeddings.weight"]               # (S,Lt,W)
    m = (text_mask == 1).to(bow.dtype)
    x = (bow * m.unsqueeze(-1)).sum(1) / m.sum(1, keepdim=True)          # masked mean (model.py:1020-1021)
    return reasoner(bilstm_sum(x, sd), E0, sd, "recipe_reasoner", dm)


def bce_sum(p, y):
    """nn.BCELoss(reduction='sum') (model.py:871): log clamped at -100."""
    return -(y * torch.log(p).clamp(min=-100) + (1 - y) * torch.log(1 - p).clamp(min=-100)).sum()


def asymmetric_loss(p, y, gamma_neg=4.0, gamma_pos=1.0, clip=0.05, eps=1e-8):
    """AsymmetricLoss.forward on probabilities (libs/ASL/src/loss_functions/losses.py:15-50)."""
    xs_neg = (1 - p + clip).clamp(max=1)
    loss = y * torch.log(p.clamp(min=eps)) + (1 - y) * torch.log(xs_neg.clamp(min=eps))
    pt = p * y + xs_neg * (1 - y)
    w = torch.pow(1 - pt, gamma_pos * y + gamma_neg * (1 - y))
    return -(loss * w).sum()


# ------------------------------------------------------------------------------------------------
# whole-model forward (model.py:1027-1189) and greedy decode (translator.py:45-192)
# ------------------------------------------------------------------------------------------------
def _stack_video(batch, b, key, S):
    return torch.stack([batch[key][t][b] for t in range(S)])


def _context(mode, step_vecs, E0, ebar, fbar, sd):
    """Decoder context tokens (model.py:939-946 / 986-990 / 1006): (S, C, D)."""
    if mode in ("full", "reason_copy"):
        ent = torch.relu(linear(ebar, sd, "Went.0"))
        act = torch.relu(linear(fbar, sd, "Wac.0"))
        return torch.stack([step_vecs, ent, act], dim=1)
    if mode == "copy":
        return torch.stack([step_vecs, E0.mean(dim=0).unsqueeze(0).expand_as(step_vecs)], dim=1)
    return step_vecs.unsqueeze(1)


def forward(sd, cfg, batch, dtype=torch.float32, return_parts=False, drop=None):
    """StateAwareRecursiveTransformer.forward (model.py:1027-1189); eval mode (dropout off) unless ``drop`` (a
    ``DropMasks``) supplies the scale tensor of every nn.Dropout site.

    ``batch`` is the dict of ``svpc_b200.synth.make_batch`` (the 13 positionals under their reference names plus
    ``gumbel``).  Returns (total_loss, prediction_scores_list, entity_prob_list, action_prob_list) and, with
    ``return_parts``, a dict of the loss components and intermediates used by the per-kernel tests.
    """
    sd = {k: v.to(dtype) if v.is_floating_point() else v for k, v in sd.items()}
    mode, Lv = cfg["model_mode"], cfg["max_v_len"]
    steps = batch["batch_step_num"]
    B = len(steps)
    parts = dict(caption=0., entity=0., action=0., re_entity=0., re_action=0., cls=[], step_vecs=[], ingr=[],
                 dec=[], re_ent=[], re_ac=[], E_stack=[])
    preds, ents, acts = [], [], []
    nl = cfg["num_hidden_layers"]
    dc = (lambda tag, b: None) if drop is None else drop.clips
    dv = (lambda tag, b: None) if drop is None else drop.video
    for b in range(B):
        S = steps[b]
        ids = _stack_video(batch, b, "input_ids_list", S)
        feats = _stack_video(batch, b, "video_features_list", S).to(dtype)
        masks = _stack_video(batch, b, "input_masks_list", S).to(dtype)
        labels = _stack_video(batch, b, "input_labels_list", S)[:, Lv:].clone()
        enc = forward_step(ids, feats, masks, sd, cfg, dc("video_embeddings", b),
                           [tuple(dc(f"encoder.layer.{l}.{k}", b) for k in ("attn", "attn_out", "out")) for l in range(nl)])
        cls = enc[:, 0, :]                                                               # model.py:1064
        x = (cls + sd["step_positional_encoding.pe"][:S]).unsqueeze(0)
        un = lambda t: None if t is None else t.unsqueeze(0)
        step_vecs = encoder(x, x.new_ones(1, S), sd, "step_wise_encoder", cfg,
                            [(un(dv(f"step_wise_encoder.layer.{l}.attn", b)), un(dc(f"step_wise_encoder.layer.{l}.attn_out", b)),
                              un(dc(f"step_wise_encoder.layer.{l}.out", b))) for l in range(nl)])[0]        # (S, D)
        parts["cls"].append(cls); parts["step_vecs"].append(step_vecs)
        E0 = None
        if mode != "video":
            E0 = ingredient_vectors(batch["ingr_input_ids"][b], batch["ingr_sep_masks"][b], sd, cfg, dv("ingredient_embeddings", b))
            parts["ingr"].append(E0)
        ebar = fbar = E_stack = None
        if mode in ("full", "reason_copy"):
            ent_p, ac_p, ebar, E_stack, fbar = reasoner(step_vecs, E0, sd, "reasoner", dc("reasoner.action_selector", b))
            parts["E_stack"].append(E_stack)
        ctx = _context(mode, step_vecs, E0, ebar, fbar, sd)
        text_ids, text_mask = ids[:, Lv:], masks[:, Lv:]
        dec = decoder(text_embed(text_ids, sd, cfg, dc("text_embeddings", b)), text_mask, ctx, sd, cfg,
                      [tuple(dc(f"decoder.layer.{l}.{k}", b) for k in ("self_attn", "cross_attn", "out")) for l in range(nl)])
        parts["dec"].append(dec)                                                         # (S, Lt, D)
        if mode == "video":
            P = torch.softmax(lm_head(dec, sd, cfg["layer_norm_eps"]), dim=-1)
            labels[labels >= cfg["vocab_size"]] = cfg["unk_id"]                          # model.py:1013
        else:
            src = E_stack if mode != "copy" else E0.unsqueeze(0).expand(S, -1, -1)
            P = pointer_generator(dec, src, batch["ingr_id_dict"][b], batch["extra_zeros"][b], sd, cfg)
        parts["caption"] = parts["caption"] + label_smoothing_loss(P.reshape(-1, P.shape[-1]), labels.reshape(-1),
                                                                   cfg["label_smoothing"])
        preds.append(P)
        if mode in ("full", "reason_copy"):
            y_al, y_ac = batch["alignments"][b].to(dtype), batch["actions"][b].to(dtype)
            rows = (y_ac == 1).any(dim=1)
            parts["entity"] = parts["entity"] + bce_sum(ent_p, y_al)
            parts["action"] = parts["action"] + asymmetric_loss(ac_p[rows], y_ac[rows])
            ents.append(ent_p); acts.append(ac_p)
            if mode == "full":
                re_ent, re_ac, _, _, _ = reconstruct(P, text_mask, E0, batch["gumbel"][b].to(dtype), sd, cfg,
                                                     dc("recipe_reasoner.action_selector", b))
                parts["re_entity"] = parts["re_entity"] + bce_sum(re_ent, y_al)
                parts["re_action"] = parts["re_action"] + asymmetric_loss(re_ac[rows], y_ac[rows])
                parts["re_ent"].append(re_ent); parts["re_ac"].append(re_ac)
    total = parts["caption"] + parts["entity"] + parts["action"] + cfg["lambda_"] * (parts["re_entity"] + parts["re_action"])
    if return_parts:
        return total, preds, ents, acts, parts
    return total, preds, ents, acts


def loss_and_grads(sd, cfg, batch, dtype=torch.float32, drop=None):
    """fwd+bwd: returns (loss, parts, {param name: grad or None})."""
    leaf = {k: (v.to(dtype).clone().requires_grad_(True) if v.is_floating_point() and not k.endswith(".pe") else v)
            for k, v in sd.items()}
    total, preds, ents, acts, parts = forward(leaf, cfg, batch, dtype=dtype, return_parts=True, drop=drop)
    total.backward()
    grads = {k: v.grad for k, v in leaf.items() if v.is_floating_point() and not k.endswith(".pe")}
    # padding_idx=0 rows of both word-embedding tables receive no gradient (model.py:492,519)
    for k in ("text_embeddings.word_embeddings.weight", "ingredient_embeddings.word_embeddings.weight"):
        if grads.get(k) is not None:
            grads[k][0].zero_()
    return total.detach(), parts, grads, preds, ents, acts


def greedy_decode(sd, cfg, batch, dtype=torch.float32):
    """Translator.translate_batch_greedy (translator.py:45-192): list of (S_b, Lt) int64 extended ids."""
    sd = {k: v.to(dtype) if v.is_floating_point() else v for k, v in sd.items()}
    mode, Lv, Lt, V = cfg["model_mode"], cfg["max_v_len"], cfg["max_t_len"], cfg["vocab_size"]
    out = []
    with torch.no_grad():
        for b, S in enumerate(batch["batch_step_num"]):
            ids = _stack_video(batch, b, "input_ids_list", S).clone()
            feats = _stack_video(batch, b, "video_features_list", S).to(dtype)
            masks = _stack_video(batch, b, "input_masks_list", S).to(dtype).clone()
            ids[:, Lv:] = PAD                                    # prepare_video_only_inputs (translator.py:205-228)
            masks[:, Lv:] = 0
            cls = forward_step(ids, feats, masks, sd, cfg)[:, 0, :]
            x = (cls + sd["step_positional_encoding.pe"][:S]).unsqueeze(0)
            step_vecs = encoder(x, x.new_ones(1, S), sd, "step_wise_encoder", cfg)[0]
            E0 = ebar = fbar = E_stack = None
            if mode != "video":
                E0 = ingredient_vectors(batch["ingr_input_ids"][b], batch["ingr_sep_masks"][b], sd, cfg)
            if mode in ("full", "reason_copy"):
                _, _, ebar, E_stack, fbar = reasoner(step_vecs, E0, sd, "reasoner")
            ctx = _context(mode, step_vecs, E0, ebar, fbar, sd)
            n_oov = len(batch["oov_word_dict"][b])
            text_ids, ext_ids, text_mask = ids[:, Lv:], ids[:, Lv:].clone(), masks[:, Lv:]
            nxt = torch.full((S,), BOS, dtype=torch.long)
            nxt_ext = nxt.clone()
            for d in range(Lt):
                text_ids[:, d] = nxt
                ext_ids[:, d] = nxt_ext
                text_mask[:, d] = 1
                dec = decoder(text_embed(text_ids, sd, cfg), text_mask, ctx, sd, cfg)
                if mode == "video":
                    scores = lm_head(dec, sd, cfg["layer_norm_eps"])            # raw logits (translator.py:156-161)
                else:
                    src = E_stack if mode != "copy" else E0.unsqueeze(0).expand(S, -1, -1)
                    scores = pointer_generator(dec, src, batch["ingr_id_dict"][b], n_oov, sd, cfg)
                scores[:, :, UNK] = -1e10
                w = scores[:, d].max(1)[1]
                nxt_ext = w.clone()
                w = w.clone()
                if mode != "video":
                    w[w >= scores.shape[-1] - n_oov] = UNK       # copied OOV word -> UNK as next input (translator.py:110-111)
                nxt = w
            out.append(ext_ids if mode != "video" else text_ids)
    return out
