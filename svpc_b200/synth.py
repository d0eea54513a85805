"""Synthetic YouCook2-shaped batches and deterministic random-init weights.

There is no dataset or checkpoint on the build or benchmark machines, so both the parity tests and
``bench.py`` run on data with the *shape and layout* the reference data layer would have produced
(reference src/rtransformer/recursive_caption_dataset.py:239-332, 389-416, 528-576 and
src/train.py:91-112).  Everything is generated from ``numpy.random.RandomState`` so the same seed
gives bit-identical inputs on every machine and torch version.
"""
from __future__ import annotations

from collections import OrderedDict
import math

import numpy as np
import torch

from .config import PAD, CLS, SEP, VID, BOS, EOS, UNK, IGNORE, SvpcConfig

FIRST_WORD_ID = 7  # ids 0..6 are the special tokens


# --------------------------------------------------------------------------------------------
# weights
# --------------------------------------------------------------------------------------------
def sinusoid_table(max_len: int, d: int) -> torch.Tensor:
    """PositionEncoding buffer, reference src/rtransformer/model.py:87-92 (same torch ops)."""
    pe = torch.zeros(max_len, d)
    position = torch.arange(0, max_len).float().unsqueeze(1)
    div_term = torch.exp(torch.arange(0, d, 2).float() * -(math.log(10000.0) / d))
    pe[:, 0::2] = torch.sin(position * div_term)
    pe[:, 1::2] = torch.cos(position * div_term)
    return pe


def state_dict_spec(cfg: SvpcConfig) -> "OrderedDict[str, tuple]":
    """name -> (shape, kind) for every entry of the reference checkpoint.

    kind: 'lin_w' N(0, initializer_range) | 'zero' | 'one' | 'emb' word-vector table |
          'act_emb' action-embedding Parameter | 'lstm' U(-1/sqrt(H), 1/sqrt(H)) | 'pe:<len>' buffer
    Order follows ``nn.Module`` registration order of the reference constructor
    (reference src/rtransformer/model.py:827-872) so ``named_parameters()`` order matches.
    """
    D, V, W, A, F = cfg.hidden_size, cfg.vocab_size, cfg.word_vec_size, cfg.action_vocab_size, cfg.video_feature_size
    assert cfg.intermediate_size == D and cfg.lstm_hidden_size == D, \
        "the reference only works with hidden == intermediate == lstm_hidden (model.py:281,525,755)"
    P = cfg.max_position_embeddings
    sd: "OrderedDict[str, tuple]" = OrderedDict()

    def lin(prefix, n_out, n_in, bias=True):
        sd[prefix + ".weight"] = ((n_out, n_in), "lin_w")
        if bias:
            sd[prefix + ".bias"] = ((n_out,), "zero")

    def ln(prefix, n):
        sd[prefix + ".weight"] = ((n,), "one")
        sd[prefix + ".bias"] = ((n,), "zero")

    def word_fc(prefix):
        ln(prefix + ".word_fc.0", W)
        lin(prefix + ".word_fc.2", D, W)
        ln(prefix + ".word_fc.4", D)

    def attn(prefix):
        lin(prefix + ".query", D, D)
        lin(prefix + ".key", D, D)
        lin(prefix + ".value", D, D)

    def enc(prefix):
        for l in range(cfg.num_hidden_layers):
            p = f"{prefix}.layer.{l}"
            attn(p + ".attention.self")
            lin(p + ".attention.output.dense", D, D)
            ln(p + ".attention.output.LayerNorm", D)
            lin(p + ".hidden_intermediate.dense", D, D)
            lin(p + ".memory_intermediate.dense", D, D)   # dead parameters, never used
            lin(p + ".output.dense", D, D)
            ln(p + ".output.LayerNorm", D)

    def reasoner(prefix):
        lin(prefix + ".action_selector.0", D, D)
        lin(prefix + ".action_selector.3", A, D)
        sd[prefix + ".action_embeddings"] = ((A, W), "act_emb")
        lin(prefix + ".W1.0", D, D)
        lin(prefix + ".W2", D, D + A)
        lin(prefix + ".W3", 3, D)
        lin(prefix + ".W4", 1, W)

    sd["ingredient_embeddings.word_embeddings.weight"] = ((V, W), "emb")
    word_fc("ingredient_embeddings")
    sd["ingredient_embeddings.position_embeddings_ingr.pe"] = ((P, D), f"pe:{P}")
    ln("video_embeddings.video_embeddings.0", F)
    lin("video_embeddings.video_embeddings.2", D, F)
    ln("video_embeddings.video_embeddings.4", D)
    sd["video_embeddings.position_embeddings_video.pe"] = ((P, D), f"pe:{P}")
    sd["text_embeddings.word_embeddings.weight"] = ((V, W), "emb")
    word_fc("text_embeddings")
    sd["text_embeddings.position_embeddings_text.pe"] = ((P, D), f"pe:{P}")
    sd["token_type_embeddings.weight"] = ((4, D), "lin_w")
    enc("encoder")
    enc("step_wise_encoder")
    sd["step_positional_encoding.pe"] = ((50, D), "pe:50")
    reasoner("reasoner")
    lin("Wac.0", D, W)
    lin("Went.0", D, D)
    for l in range(cfg.num_hidden_layers):
        p = f"decoder.layer.{l}"
        attn(p + ".self_attention")
        ln(p + ".norm1", D)
        attn(p + ".dec_enc_attention")
        ln(p + ".norm2", D)
        lin(p + ".output.dense", D, D)
        ln(p + ".output.LayerNorm", D)
    sd["decoder_classifier.bias"] = ((V,), "zero")
    lin("decoder_classifier.transform.dense", D, D)
    ln("decoder_classifier.transform.LayerNorm", D)
    lin("decoder_classifier.decoder", V, D, bias=False)
    lin("Wing", D, D)
    lin("pgen_linear.0", 1, 2 * D)
    for suf in ("", "_reverse"):
        sd[f"recipe_encoder.weight_ih_l0{suf}"] = ((4 * D, W), "lstm")
        sd[f"recipe_encoder.weight_hh_l0{suf}"] = ((4 * D, D), "lstm")
        sd[f"recipe_encoder.bias_ih_l0{suf}"] = ((4 * D,), "lstm")
        sd[f"recipe_encoder.bias_hh_l0{suf}"] = ((4 * D,), "lstm")
    reasoner("recipe_reasoner")
    return sd


def is_buffer(name: str) -> bool:
    return name.endswith(".pe")


def init_state_dict(cfg: SvpcConfig, seed: int = 2019, *, entity_logit_scale: float = 0.25,
                    randomize_affine: bool = True) -> "OrderedDict[str, torch.Tensor]":
    """Deterministic 'random-init' checkpoint in the reference layout.

    Follows ``init_bert_weights`` (reference model.py:875-885): Linear/Embedding ~ N(0, 0.02), LSTM
    U(-1/sqrt(H), 1/sqrt(H)); word / action embedding tables get GloVe-like scale 0.4*N(0,1)
    (reference src/build_vocab.py:79).  Two deliberate deviations, both to make parity tests sharper
    (SURVEY.md §7.3-2, appendix D):
      * ``W2`` of both reasoners is scaled by ``entity_logit_scale`` so entity logits stay well inside
        +-10 and the reference's BCE does not hit its fp32 sigmoid-saturation discontinuity;
      * with ``randomize_affine`` LayerNorm weights/biases and Linear biases are perturbed away from
        their 1/0 initial values, otherwise gradient paths through them are tested only trivially.
    """
    rng = np.random.RandomState(seed)
    std = float(cfg.initializer_range)
    H = cfg.hidden_size
    out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for name, (shape, kind) in state_dict_spec(cfg).items():
        if kind == "lin_w":
            a = rng.standard_normal(shape) * std
            if name.endswith("W2.weight"):
                a = a * entity_logit_scale
        elif kind == "zero":
            a = rng.standard_normal(shape) * (0.02 if randomize_affine else 0.0)
        elif kind == "one":
            a = 1.0 + rng.standard_normal(shape) * (0.05 if randomize_affine else 0.0)
        elif kind in ("emb", "act_emb"):
            a = rng.standard_normal(shape) * 0.4
        elif kind == "lstm":
            k = 1.0 / math.sqrt(H)
            a = rng.uniform(-k, k, size=shape)
        elif kind.startswith("pe:"):
            out[name] = sinusoid_table(int(kind[3:]), shape[1])
            continue
        else:  # pragma: no cover
            raise ValueError(kind)
        out[name] = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))
    return out


# --------------------------------------------------------------------------------------------
# batches
# --------------------------------------------------------------------------------------------
def make_batch(cfg: SvpcConfig, B: int = 16, S=12, *, seed: int = 2019, ragged: bool = False,
               n_ingr: int = 10, n_oov: int = 0, empty_action_rows: bool = False,
               feature_dtype=np.float32) -> dict:
    """One collated training batch, as ``caption_collate`` + ``train_epoch`` would hand it to the model.

    Returns a dict with the 13 positional arguments of ``forward`` under their reference names plus
    ``gumbel`` (list of (S_b, Lt, V + n_oov) noise tensors), ``oov_word_dict`` (for the translator) and
    ``n_clips``.  Headline shape: everything at max length; ``ragged=True`` draws video
    length, sentence length, clips per video and ingredient count at random.
    """
    rng = np.random.RandomState(seed)
    Lv, Lt, V, A, F = cfg.max_v_len, cfg.max_t_len, cfg.vocab_size, cfg.action_vocab_size, cfg.video_feature_size
    L = Lv + Lt
    Li = int(cfg.get("max_i_len", 100))
    steps = [int(S)] * B if np.isscalar(S) else [int(s) for s in S]
    assert len(steps) == B
    if ragged and np.isscalar(S):
        steps = [int(rng.randint(max(1, S // 3), S + 1)) for _ in range(B)]
        steps[int(rng.randint(B))] = int(S)
    S_max = max(steps)

    # ---- ingredients (clip_ingredient_to_feature, dataset.py:239-281) ----
    ingr_ids = np.zeros((B, Li), np.int64)
    ingr_mask = np.zeros((B, Li), np.int64)
    ingr_sep = np.zeros((B, Li), np.int64)
    ingr_id_dict, oov_word_dict, n_ingr_list = [], [], []
    for b in range(B):
        toks_per = 2
        ni_max = Li // (toks_per + 1)
        ni = min(n_ingr, ni_max)
        if ragged:
            ni = int(rng.randint(max(1, min(3, ni_max)), min(max(ni_max, 1), n_ingr + 5) + 1))
        d, oov, pos = {}, {}, 0
        for i in range(ni):
            nt = int(rng.randint(1, 4)) if ragged else toks_per
            nt = max(1, min(nt, Li - pos - 1 - (ni - i - 1) * 2))
            ids = []
            for _ in range(nt):
                if len(oov) < n_oov and rng.rand() < 0.5:
                    ext = V + len(oov)
                    oov[f"oov{b}_{len(oov)}"] = ext
                    ingr_ids[b, pos] = UNK
                elif oov and rng.rand() < 0.1:      # repeated OOV word reuses its extended id
                    ext = V + int(rng.randint(len(oov)))
                    ingr_ids[b, pos] = UNK
                else:
                    ext = int(rng.randint(FIRST_WORD_ID, V))
                    ingr_ids[b, pos] = ext
                ids.append(ext)
                ingr_mask[b, pos] = 1
                pos += 1
            ingr_ids[b, pos] = SEP
            ingr_mask[b, pos] = 1
            ingr_sep[b, pos] = 1
            pos += 1
            d[i] = ids
        while len(oov) < n_oov:                      # keep extra_zeros == n_oov even if the coin never fell
            oov[f"oov{b}_{len(oov)}"] = V + len(oov)
        ingr_id_dict.append(d)
        oov_word_dict.append(oov)
        n_ingr_list.append(ni)

    # ---- per-step tensors (clip_sentence_to_feature + caption_collate) ----
    input_ids = np.zeros((S_max, B, L), np.int64)
    input_mask = np.zeros((S_max, B, L), np.float32)
    labels = np.full((S_max, B, L), IGNORE, np.int64)
    feats = np.zeros((S_max, B, L, F), feature_dtype)
    tt = np.zeros((S_max, B, L), np.int64)
    tt[:, :, Lv:] = 1
    pad_clip = None
    for b in range(B):
        for t in range(S_max):
            if t >= steps[b]:
                # padding clips are copies of clip 0 of video 0 with all labels ignored (dataset.py:562-567)
                ids0, m0, f0 = pad_clip
                input_ids[t, b], input_mask[t, b], feats[t, b] = ids0, m0, f0
                continue
            n_vid = Lv - 2 if not ragged else int(rng.randint(max(1, Lv // 5), Lv - 1))
            n_w = Lt - 2 if not ragged else int(rng.randint(min(4, Lt - 2), Lt - 1))
            input_ids[t, b, 0] = CLS
            input_ids[t, b, 1:1 + n_vid] = VID
            input_ids[t, b, 1 + n_vid] = SEP
            input_mask[t, b, :n_vid + 2] = 1
            feats[t, b, 1:1 + n_vid] = rng.standard_normal((n_vid, F)).astype(feature_dtype)
            words = rng.randint(FIRST_WORD_ID, V, size=n_w)
            ext = words.copy()
            for j in range(n_w):                    # sprinkle copies of this video's ingredient words
                if rng.rand() < 0.3 and ingr_id_dict[b]:
                    lst = ingr_id_dict[b][int(rng.randint(n_ingr_list[b]))]
                    e = lst[int(rng.randint(len(lst)))]
                    ext[j] = e
                    words[j] = e if e < V else UNK
            seq = np.concatenate([[BOS], words, [EOS]])
            seq_ext = np.concatenate([[BOS], ext, [EOS]])
            input_ids[t, b, Lv:Lv + n_w + 2] = seq
            input_mask[t, b, Lv:Lv + n_w + 2] = 1
            labels[t, b, Lv:Lv + n_w + 1] = seq_ext[1:]
            if pad_clip is None:
                pad_clip = (input_ids[t, b].copy(), input_mask[t, b].copy(), feats[t, b].copy())

    alignments, actions, gumbel = [], [], []
    for b in range(B):
        al = (rng.rand(steps[b], n_ingr_list[b]) < 0.2).astype(np.float32)
        ac = np.zeros((steps[b], A), np.float32)
        for t in range(steps[b]):
            if empty_action_rows and t % 3 == 2:
                continue
            ac[t, rng.choice(A, size=2, replace=False)] = 1.0
        g = -np.log(rng.exponential(size=(steps[b], Lt, V + n_oov)))
        alignments.append(torch.from_numpy(al))
        actions.append(torch.from_numpy(ac))
        gumbel.append(torch.from_numpy(g.astype(np.float32)))

    return dict(
        input_ids_list=[torch.from_numpy(input_ids[t]) for t in range(S_max)],
        video_features_list=[torch.from_numpy(feats[t]) for t in range(S_max)],
        input_masks_list=[torch.from_numpy(input_mask[t]) for t in range(S_max)],
        token_type_ids_list=[torch.from_numpy(tt[t]) for t in range(S_max)],
        input_labels_list=[torch.from_numpy(labels[t]) for t in range(S_max)],
        ingr_input_ids=torch.from_numpy(ingr_ids),
        ingr_masks=torch.from_numpy(ingr_mask),
        ingr_sep_masks=torch.from_numpy(ingr_sep),
        batch_step_num=steps,
        ingr_id_dict=ingr_id_dict,
        extra_zeros=[n_oov] * B,
        alignments=alignments,
        actions=actions,
        gumbel=gumbel,
        oov_word_dict=oov_word_dict,
        n_clips=int(sum(steps)),
    )


FORWARD_ARG_ORDER = ("input_ids_list", "video_features_list", "input_masks_list", "token_type_ids_list",
                     "input_labels_list", "ingr_input_ids", "ingr_masks", "ingr_sep_masks", "batch_step_num",
                     "ingr_id_dict", "extra_zeros", "alignments", "actions")


def forward_args(batch: dict, device=None) -> list:
    """The 13 positionals of ``forward`` (reference model.py:1027-1030), optionally moved to ``device``."""
    def mv(x):
        if device is None:
            return x
        if isinstance(x, torch.Tensor):
            return x.to(device)
        if isinstance(x, list) and x and isinstance(x[0], torch.Tensor):
            return [t.to(device) for t in x]
        return x
    return [mv(batch[k]) for k in FORWARD_ARG_ORDER]
