"""Greedy caption decoding, B200-first (reference ``Translator.translate_batch_greedy``, src/translator.py:45-192).

The reference decodes one video at a time and, for each of the ``max_t_len`` positions, re-runs the text embedding, the
whole decoder and the LM head / pointer-generator over ALL positions (no KV cache): ~12 GFLOP per caption
(SURVEY.md 3.3).  Here

* every clip of every video of the batch is encoded once, as in the training step (``model._encode_clips``),
* all clips decode in lock-step as ONE batch of ``Nc`` rows: per position one row per clip goes through the decoder with
  the keys / values of the earlier positions kept in a per-layer cache (the causal mask makes this exact: position d
  never sees later positions, translator.py:96 + model.py:641-645), the cross-attention keys / values of the context
  tokens are projected once,
* the pointer-generator mix and the greedy pick (``[UNK]`` suppressed, copied OOV ids fed back as ``[UNK]``,
  translator.py:106-112) run on the current position only,

which is ~2.4 GFLOP per caption and ~25 launches per position.  Every device operation is a ``libsvpc_b200.so`` kernel.
Output: the reference's return value, a list of ``(S_b, max_t_len)`` int64 tensors of extended token ids.
"""
from __future__ import annotations

import torch

from . import functional as Fn
from .config import BOS, UNK, IGNORE


def _as_long_tensor(x, device):
    if isinstance(x, torch.Tensor):
        return x.to(device=device, dtype=torch.long)
    return torch.tensor(x, dtype=torch.long, device=device)


@torch.no_grad()
def greedy_decode(model, model_inputs, plan=None):
    """``model_inputs``: the 12-tuple ``Translator.translate_batch`` takes (translate.py:57-70):
    (input_ids_list, video_features_list, input_masks_list, token_type_ids_list, ingr_input_ids, ingr_masks,
    ingr_sep_masks, ingr_id_dict, oov_word_dict, alignments, actions, batch_step_num)."""
    (input_ids_list, video_features_list, input_masks_list, _tt, ingr_input_ids, _im, ingr_sep_masks, ingr_id_dict,
     oov_word_dict, _al, _ac, batch_step_num) = model_inputs
    cfg, cd = model.config, model.compute_dtype
    mode, Lv, Lt, V, D = cfg.model_mode, cfg.max_v_len, cfg.max_t_len, cfg.vocab_size, cfg.hidden_size
    H, nl = cfg.num_attention_heads, cfg.num_hidden_layers
    dev = video_features_list[0].device
    be = Fn.be()
    Fn.WEIGHTS.refresh_all()      # weight copies touched by an optimiser step / EMA swap: rewritten in place, one launch
    f32, i32 = torch.float32, torch.int32
    steps = [int(s) for s in batch_step_num]
    B = len(steps)
    if mode == "video":
        ni, n_oov, dicts = [0] * B, [0] * B, [{} for _ in range(B)]
    else:
        ni = model.ingredient_counts(ingr_sep_masks, ingr_id_dict)
        n_oov = [len(d) for d in oov_word_dict]
        dicts = ingr_id_dict
    if plan is None:
        plan = model.make_plan_raw(steps, ni, n_oov, dicts, dev)
    Nc = plan.Nc

    # ---- encode every clip, build the decoder context (same stages as the training step)
    Vst, _, _, _ = model._encode_clips(input_ids_list, video_features_list, input_masks_list, plan)
    E0 = ebar = fbar = Es = None
    if mode != "video":
        E0 = model._ingredient_vectors(_as_long_tensor(ingr_input_ids, dev), _as_long_tensor(ingr_sep_masks, dev), plan)
    if mode in ("full", "reason_copy"):
        _, _, ebar, Es, fbar = model._reasoner("reasoner", Vst, E0, plan)
    ctx, C = model._context(mode, Vst, E0, ebar, fbar, plan)
    ctx = ctx.contiguous()

    # ---- static pieces of the decoding loop
    ar = torch.arange(Nc, dtype=i32, device=dev)
    one = torch.ones(Nc, dtype=i32, device=dev)
    self_lay = dict(q_off=ar, q_len=one, k_off=ar * Lt)
    cross_lay = dict(q_off=ar, q_len=one, k_off=ar * C, k_len=torch.full((Nc,), C, dtype=i32, device=dev))
    k_len_all = torch.arange(1, Lt + 1, dtype=i32, device=dev)[:, None].expand(Lt, Nc).contiguous()
    pe = model.p("text_embeddings.position_embeddings_text.pe")
    W = lambda names: Fn.WEIGHTS.get([model.p(n + ".weight") for n in names], cd)
    Bv = lambda names: Fn.WEIGHTS.get([model.p(n + ".bias") for n in names], f32)
    layers = []
    for l in range(nl):
        pfx = f"decoder.layer.{l}"
        sa, ca = pfx + ".self_attention", pfx + ".dec_enc_attention"
        kv_ctx = be.gemm(ctx, W([ca + ".key", ca + ".value"]), cd, Bv([ca + ".key", ca + ".value"]))     # (Nc*C, 2D), once
        layers.append(dict(
            pfx=pfx, wq=W([sa + ".query"]), bq=Bv([sa + ".query"]), wkv=W([sa + ".key", sa + ".value"]),
            bkv=Bv([sa + ".key", sa + ".value"]), kv_ctx=kv_ctx,
            cache=torch.zeros(Nc, Lt, 2 * D, dtype=cd, device=dev)))
    ext = torch.zeros(Nc, Lt, dtype=torch.long, device=dev)
    nxt = torch.full((Nc,), BOS, dtype=i32, device=dev)
    ext[:, 0] = BOS
    if mode != "video":
        Esrc = Es if mode != "copy" else E0
        src_off = plan.t["clip_ent_off"] if mode != "copy" else plan.t["clip_src_copy"]
        vx_clip = plan.t["clip_vx"]
        P1_off = (torch.cumsum(vx_clip, 0) - vx_clip).to(i32)                       # clip n's (1, Vx) block
        P1_numel = int(sum(plan.vx[b] * steps[b] for b in range(B)))
        noov_clip = (vx_clip - V).to(i32)
        no_labels = torch.full((Nc,), IGNORE, dtype=i32, device=dev)
        G = model._lin(Esrc, "Wing", out_dtype=f32)                                   # Wing(E) does not depend on the position
    eps = cfg.layer_norm_eps

    # ext[:, d] is the (extended) token FED at position d (translator.py:90-94): ext[:, 0] = [BOS] and the pick made at
    # position d lands in column d + 1; the pick of the last position is never used, so Lt - 1 positions are decoded.
    for d in range(Lt - 1):
        # text embedding of the current position (reference model.py:509-513, pe row d)
        x = model._word_fc("text_embeddings", nxt, pe=pe[d:d + 1], pe_period=1)
        for L in layers:
            pfx = L["pfx"]
            q = be.gemm(x, L["wq"], cd, L["bq"])
            be.gemm(x, L["wkv"], cd, L["bkv"], out=L["cache"][:, d, :])                 # append K | V of position d
            kv = L["cache"].view(Nc * Lt, 2 * D)
            sa = be.attn_fwd(q, kv[:, :D], kv[:, D:], self_lay["q_off"], self_lay["q_len"], self_lay["k_off"], k_len_all[d],
                             None, False, H, 1, Lt, cd)
            h1 = Fn.layer_norm(sa, x, model.p(pfx + ".norm1.weight"), model.p(pfx + ".norm1.bias"), eps)
            q2 = model._lin(h1, pfx + ".dec_enc_attention.query")
            ca = be.attn_fwd(q2, L["kv_ctx"][:, :D], L["kv_ctx"][:, D:], cross_lay["q_off"], cross_lay["q_len"],
                             cross_lay["k_off"], cross_lay["k_len"], None, False, H, 1, C, cd)
            h2 = Fn.layer_norm(ca, h1, model.p(pfx + ".norm2.weight"), model.p(pfx + ".norm2.bias"), eps)
            x = model._ln(model._lin(h2, pfx + ".output.dense", residual=h2), None, pfx + ".output.LayerNorm")
        if mode == "video":
            logits = model._lm_logits(x)                                                # raw logits (translator.py:156-161)
            be.greedy_pick(logits, None, logits.stride(0), None, V, None, UNK, nxt, ext, d + 1)
        else:
            P = model._pointer(x, Esrc, plan, src_off, no_labels, 1, clip_P_off=P1_off, P_numel=P1_numel, G=G)[0]
            be.greedy_pick(P, P1_off, 0, vx_clip, 0, noov_clip, UNK, nxt, ext, d + 1)
    return _split(ext, plan, mode, Lt)


def _split(ext, plan, mode, Lt):
    out = []
    for b, S in enumerate(plan.steps):
        out.append(ext[plan.cu_steps[b]:plan.cu_steps[b] + S])
    return out
