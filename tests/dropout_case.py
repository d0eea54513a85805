"""Explicit dropout masks for the train-mode parity tests.

The oracle (``oracle.svpc_oracle.DropMasks``) takes one scale tensor (0 or 1/(1-p)) per reference nn.Dropout site in a
per-clip / per-video canonical layout.  This module produces them two ways:

* ``random_masks``: Bernoulli draws from a seeded torch generator - used to pin the oracle's train mode against the
  unmodified reference (tests/golden/make_dropout_golden.py, tests/test_oracle.py);
* ``product_masks``: the masks the CUDA kernels actually draw (counter-based hash of (seed, site, element index),
  ``svpc_dropout_mask``) rearranged from the product's batched tensors into the oracle's layout - used by the GPU
  train-mode parity tests.
"""
import torch

from oracle.svpc_oracle import DropMasks

CONTEXT_TOKENS = {"full": 3, "reason_copy": 3, "copy": 2, "video": 1}


def site_specs(cfg, batch):
    """tag -> (p, canonical shape or list of per-video shapes); the nn.Dropout sites of one training forward in the
    given model mode (reference model.py:214, 231, 287, 495, 522, 550, 757)."""
    mode = cfg["model_mode"]
    steps = list(batch["batch_step_num"])
    B, Nc = len(steps), sum(steps)
    Lv, Lt, D, F, W, H = (cfg["max_v_len"], cfg["max_t_len"], cfg["hidden_size"], cfg["video_feature_size"],
                          cfg["word_vec_size"], cfg["num_attention_heads"])
    ph, pa, nl = cfg["hidden_dropout_prob"], cfg["attention_probs_dropout_prob"], cfg["num_hidden_layers"]
    C = CONTEXT_TOKENS[mode]
    s = {"video_embeddings": (ph, (Nc, Lv, F))}
    for l in range(nl):
        s[f"encoder.layer.{l}.attn"] = (pa, (Nc, H, Lv, Lv))
        s[f"encoder.layer.{l}.attn_out"] = (ph, (Nc, Lv, D))
        s[f"encoder.layer.{l}.out"] = (ph, (Nc, Lv, D))
        s[f"step_wise_encoder.layer.{l}.attn"] = (pa, [(H, S, S) for S in steps])
        s[f"step_wise_encoder.layer.{l}.attn_out"] = (ph, (Nc, D))
        s[f"step_wise_encoder.layer.{l}.out"] = (ph, (Nc, D))
        s[f"decoder.layer.{l}.self_attn"] = (pa, (Nc, H, Lt, Lt))
        s[f"decoder.layer.{l}.cross_attn"] = (pa, (Nc, H, Lt, C))
        s[f"decoder.layer.{l}.out"] = (ph, (Nc, Lt, D))
    s["text_embeddings"] = (ph, (Nc, Lt, W))
    if mode != "video":
        s["ingredient_embeddings"] = (ph, (B, batch["ingr_input_ids"].shape[1], W))
    if mode in ("full", "reason_copy"):
        s["reasoner.action_selector"] = (0.4, (Nc, D))
    if mode == "full":
        s["recipe_reasoner.action_selector"] = (0.4, (Nc, D))
    return s


def _cu(steps):
    cu = [0]
    for x in steps:
        cu.append(cu[-1] + int(x))
    return cu


def random_masks(cfg, batch, seed):
    g = torch.Generator().manual_seed(seed)
    draw = lambda shape, p: (torch.rand(shape, generator=g) >= p).float() / (1.0 - p)
    masks = {}
    for tag, (p, shape) in site_specs(cfg, batch).items():
        masks[tag] = [draw(s, p) for s in shape] if isinstance(shape, list) else draw(shape, p)
    return DropMasks(masks, _cu(batch["batch_step_num"]))


def product_masks(model, cfg, batch, be, device="cpu"):
    """The masks of the forward that ``model`` has just run (``model.drop_sites``, ``model._seed_t``) in the oracle's
    layout.  Index spaces of the kernels (tests/test_dropout_gpu.py): GEMM epilogues and embedding rows = row-major
    element index of the batched output; attention = ((sequence * H + head) * max_lq + i) * even(max_lk) + j."""
    specs = site_specs(cfg, batch)
    steps = list(batch["batch_step_num"])
    nl, Lv = cfg["num_hidden_layers"], cfg["max_v_len"]
    seed = model._seed_t
    by_tag = {tag: (site, p) for site, tag, p in model.drop_sites}
    assert set(by_tag) == set(specs), (sorted(set(by_tag) ^ set(specs)))
    even = lambda n: (n + 1) // 2 * 2
    masks = {}
    for tag, (p, shape) in specs.items():
        site, p_used = by_tag[tag]
        assert abs(p_used - p) < 1e-12, (tag, p_used, p)
        raw = lambda n: be.dropout_mask(n, (seed, site, p))
        last_enc = tag.startswith(f"encoder.layer.{nl - 1}.")
        if tag.startswith("step_wise_encoder") and tag.endswith(".attn"):
            B, H, mS = len(steps), shape[0][0], max(steps)
            full = raw(B * H * mS * even(mS)).view(B, H, mS, even(mS))
            masks[tag] = [full[b, :, :S, :S].to(device) for b, S in enumerate(steps)]
        elif last_enc and tag.endswith(".attn"):
            # only the [CLS] query row of the last clip-encoder layer is evaluated (model.py:1064 consumes nothing else)
            Nc, H = shape[0], shape[1]
            m = torch.ones(shape, device=seed.device)
            m[:, :, 0, :] = raw(Nc * H * even(Lv)).view(Nc, H, even(Lv))[:, :, :Lv]
            masks[tag] = m.to(device)
        elif last_enc:
            m = torch.ones(shape, device=seed.device)
            m[:, 0, :] = raw(shape[0] * shape[2]).view(shape[0], shape[2])
            masks[tag] = m.to(device)
        elif tag.endswith("attn"):           # (N, H, Lq, Lk) with the key axis padded to an even stride
            N, H, Lq, Lk = shape
            masks[tag] = raw(N * H * Lq * even(Lk)).view(N, H, Lq, even(Lk))[..., :Lk].to(device)
        else:
            n = 1
            for d in shape:
                n *= d
            masks[tag] = raw(n).view(shape).to(device)
    return DropMasks(masks, _cu(steps))
