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
