"""TEST INFRASTRUCTURE ONLY - imports the *unmodified* reference model from /root/reference.

Only usable in the build container (the reference tree does not travel to the GPU box).  Used by
``tests/golden/make_golden.py`` to generate the committed golden vectors and by the ``not gpu`` tests
that pin ``oracle/svpc_oracle.py`` against the real reference when it is present.

Shims, none of which edits a reference file:
  * ``easydict`` / ``nltk`` stub modules (not installed here),
  * ``torch.Tensor.cuda`` -> identity on CPU (the model hard-codes ``.cuda()``),
  * ``F.gumbel_softmax`` inside the model's namespace replaced by a version that consumes an injected
    noise tensor (same formula as torch 2.11 ``functional.gumbel_softmax``).
"""
from __future__ import annotations

import os
import sys
import types

import torch

REFERENCE_ROOT = os.environ.get("SVPC_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "src", "rtransformer", "model.py"))


class _NoiseQueue:
    """FIFO of Gumbel noise tensors consumed by successive ``gumbel_softmax`` calls."""

    def __init__(self):
        self.q = []

    def push_all(self, tensors):
        self.q = list(tensors)

    def pop(self, like):
        g = self.q.pop(0)
        assert g.shape == like.shape, (g.shape, like.shape)
        return g.to(like.dtype).to(like.device)


NOISE = _NoiseQueue()


def injected_gumbel_softmax(logits, tau=1.0, hard=False, eps=1e-10, dim=-1):
    g = NOISE.pop(logits)
    y_soft = ((logits + g) / tau).softmax(dim)
    if not hard:
        return y_soft
    index = y_soft.max(dim, keepdim=True)[1]
    y_hard = torch.zeros_like(logits, memory_format=torch.legacy_contiguous_format).scatter_(dim, index, 1.0)
    return y_hard - y_soft.detach() + y_soft


def load():
    """Returns the reference ``src.rtransformer.model`` module (imported once, shims installed)."""
    if "svpc_reference_model" in sys.modules:
        return sys.modules["svpc_reference_model"]
    assert available(), f"reference tree not found at {REFERENCE_ROOT}"
    if "easydict" not in sys.modules:
        m = types.ModuleType("easydict")

        class EasyDict(dict):
            def __init__(self, d=None, **kw):
                super().__init__()
                for k, v in dict(d or {}, **kw).items():
                    self[k] = v

            def __getattr__(self, k):
                try:
                    return self[k]
                except KeyError as e:
                    raise AttributeError(k) from e

            def __setattr__(self, k, v):
                self[k] = v

        m.EasyDict = EasyDict
        sys.modules["easydict"] = m
    if "nltk" not in sys.modules:
        sys.modules["nltk"] = types.ModuleType("nltk")
    if not torch.cuda.is_available():
        torch.Tensor.cuda = lambda self, *a, **k: self
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import importlib
    mod = importlib.import_module("src.rtransformer.model")
    fshim = types.SimpleNamespace(**{k: getattr(mod.F, k) for k in dir(mod.F) if not k.startswith("__")})
    fshim.gumbel_softmax = injected_gumbel_softmax
    mod.F = fshim
    sys.modules["svpc_reference_model"] = mod
    return mod


def build_model(cfg, state_dict, train: bool = False):
    """Reference ``StateAwareRecursiveTransformer`` carrying ``state_dict`` (reference key layout)."""
    mod = load()
    from easydict import EasyDict
    model = mod.StateAwareRecursiveTransformer(EasyDict(dict(cfg)))
    # train.py:691-707: GloVe tables are installed through set_pretrained_embedding, which also turns
    # reasoner.action_embeddings into a bare Parameter (model.py:773-775).
    model.ingredient_embeddings.set_pretrained_embedding(
        state_dict["ingredient_embeddings.word_embeddings.weight"].clone(), freeze=False)
    model.text_embeddings.set_pretrained_embedding(
        state_dict["text_embeddings.word_embeddings.weight"].clone(), freeze=False)
    model.reasoner.set_pretrained_embedding(state_dict["reasoner.action_embeddings"].clone(), freeze=False)
    model.recipe_reasoner.set_pretrained_embedding(state_dict["recipe_reasoner.action_embeddings"].clone(), freeze=False)
    missing = model.load_state_dict(state_dict, strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    model.train(train)
    return model


def run_forward(model, batch, predict=False):
    from svpc_b200.synth import forward_args
    NOISE.push_all(batch["gumbel"])
    args = forward_args(batch)
    args[4] = [t.clone() for t in args[4]]          # v mode rewrites labels >= V in place (model.py:1013)
    return model(*args, predict=predict)


def load_translator():
    load()
    import importlib
    ds = types.ModuleType("src.rtransformer.recursive_caption_dataset")

    class RecursiveCaptionDataset:      # only the constants the translator reads (translator.py:14,52)
        PAD, CLS, SEP, VID, BOS, EOS, UNK, IGNORE = 0, 1, 2, 3, 4, 5, 6, -1

    ds.RecursiveCaptionDataset = RecursiveCaptionDataset
    sys.modules.setdefault("src.rtransformer.recursive_caption_dataset", ds)
    return importlib.import_module("src.translator")


# ------------------------------------------------------------------------------------------------
# train mode with explicit dropout masks
# ------------------------------------------------------------------------------------------------
_DROP_SITES = (
    # (regex on the nn.Dropout module's qualified name, oracle tag template, how successive calls map onto the masks)
    (r"^ingredient_embeddings\.word_fc\.1$", "ingredient_embeddings", "batch"),
    (r"^video_embeddings\.video_embeddings\.1$", "video_embeddings", "clip_index"),
    (r"^encoder\.layer\.(\d+)\.attention\.self\.dropout$", "encoder.layer.{}.attn", "clip_index"),
    (r"^encoder\.layer\.(\d+)\.attention\.output\.dropout$", "encoder.layer.{}.attn_out", "clip_index"),
    (r"^encoder\.layer\.(\d+)\.output\.dropout$", "encoder.layer.{}.out", "clip_index"),
    (r"^step_wise_encoder\.layer\.(\d+)\.attention\.self\.dropout$", "step_wise_encoder.layer.{}.attn", "video"),
    (r"^step_wise_encoder\.layer\.(\d+)\.attention\.output\.dropout$", "step_wise_encoder.layer.{}.attn_out", "video_rows"),
    (r"^step_wise_encoder\.layer\.(\d+)\.output\.dropout$", "step_wise_encoder.layer.{}.out", "video_rows"),
    (r"^text_embeddings\.word_fc\.1$", "text_embeddings", "video_clips"),
    (r"^decoder\.layer\.(\d+)\.self_attention\.dropout$", "decoder.layer.{}.self_attn", "video_clips"),
    (r"^decoder\.layer\.(\d+)\.dec_enc_attention\.dropout$", "decoder.layer.{}.cross_attn", "video_clips"),
    (r"^decoder\.layer\.(\d+)\.output\.dropout$", "decoder.layer.{}.out", "video_clips"),
    (r"^reasoner\.action_selector\.2$", "reasoner.action_selector", "clip"),
    (r"^recipe_reasoner\.action_selector\.2$", "recipe_reasoner.action_selector", "clip"),
)


def install_dropout_masks(model, drop, steps):
    """Replace the forward of every nn.Dropout INSTANCE of the (unmodified) reference model by a multiplication with
    the matching slice of ``drop`` (an ``oracle.svpc_oracle.DropMasks``).  The reference calls its dropout modules
    per clip index t over the whole batch (forward_step, model.py:1038-1042), per video (step encoder, decoder,
    model.py:1059-) or per (video, step) (action selector, model.py:797); ``how`` says which, the call counter of the
    module says where in that sequence the call is, and the shape of the argument is asserted against the slice.
    Must be re-installed (fresh counters) before every forward."""
    import re
    import torch.nn as nn
    cu = drop.cu
    B = len(steps)
    unused = []
    for name, mod in model.named_modules():
        if not isinstance(mod, nn.Dropout):
            continue
        hit = None
        for rx, tag, how in _DROP_SITES:
            m = re.match(rx, name)
            if m:
                hit = (tag.format(*m.groups()), how)
                break
        if hit is None:
            def dead(x, _n=name):
                raise AssertionError(f"unexpected dropout call at {_n}")
            mod.forward = dead
            unused.append(name)
            continue
        tag, how = hit
        state = {"i": 0}

        def fwd(x, tag=tag, how=how, state=state, name=name):
            i = state["i"]
            state["i"] += 1
            t = drop.m.get(tag)
            if t is None:
                return x
            if how == "batch":
                m = t
            elif how == "clip_index":          # call i = clip index t, rows = videos; videos with S_b <= t are padding
                m = torch.ones_like(x)
                for b in range(B):
                    if i < steps[b]:
                        m[b] = t[cu[b] + i]
            elif how == "video":
                m = t[i].unsqueeze(0)
            elif how == "video_rows":
                m = t[cu[i]:cu[i + 1]].unsqueeze(0)
            elif how == "video_clips":
                m = t[cu[i]:cu[i + 1]]
            else:                               # "clip": one call per (video, step) in video-major order
                m = t[i]
            assert tuple(m.shape) == tuple(x.shape), (name, i, tuple(m.shape), tuple(x.shape))
            return x * m.to(x.dtype)
        mod.forward = fwd
    return unused
