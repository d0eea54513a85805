"""Drop-in boundary for the piecewise callers: the UNMODIFIED reference ``Translator``
(src/translator.py) drives ``svpc_b200.model.StateAwareRecursiveTransformer`` through its sub-module calls
(``forward_step``, ``step_wise_encoder``, ``reasoner``, ``decoder``, ``pointer_generator_network`` ...) and must
produce the golden greedy captions the reference model produced.  Runs where /root/reference is mounted (build
container), on the plain-torch operator statements; tests/test_decode_gpu.py covers the same calls on the CUDA library."""
import copy
import types

import pytest
import torch

from oracle import refshim
from svpc_b200 import backend
from svpc_b200.model import StateAwareRecursiveTransformer
from tests.helpers import MODES, GREEDY_SHARPEN, load_case
from tests.torch_backend import TorchBackend


@pytest.fixture(autouse=True)
def _torch_backend():
    prev = backend._impl
    backend.set_backend(TorchBackend())
    yield
    backend.set_backend(prev)


def translator_inputs(batch):
    b2 = copy.deepcopy(batch)
    return (b2["input_ids_list"], b2["video_features_list"], b2["input_masks_list"], b2["token_type_ids_list"],
            b2["ingr_input_ids"].tolist(), b2["ingr_masks"].tolist(), b2["ingr_sep_masks"].tolist(),
            b2["ingr_id_dict"], b2["oov_word_dict"], b2["alignments"], b2["actions"], b2["batch_step_num"])


@pytest.mark.skipif(not refshim.available(), reason="reference tree not mounted (GPU box)")
@pytest.mark.parametrize("mode", MODES)
def test_reference_translator_drives_the_dropin(mode):
    cfg, sd, batch, gold = load_case("ragged", mode)
    sd = {k: v.clone() for k, v in sd.items()}
    sd["decoder_classifier.decoder.weight"] *= GREEDY_SHARPEN
    tr = refshim.load_translator()
    from easydict import EasyDict
    m = StateAwareRecursiveTransformer(EasyDict(dict(cfg)))
    T = tr.Translator(types.SimpleNamespace(cuda=False), {"model_cfg": EasyDict(dict(cfg)), "model": sd}, model=m)
    seqs, _ = T.translate_batch(translator_inputs(batch))
    for b, s in enumerate(seqs):
        assert torch.equal(s.cpu(), torch.from_numpy(gold[f"greedy.{b}"])), (mode, b)


def test_set_pretrained_embedding_and_checkpoint_keys():
    """train.py:691-707 and the state_dict layout around it (reference model.py:503-507, 530-532, 773-775)."""
    cfg, sd, _, _ = load_case("ragged", "full")
    m = StateAwareRecursiveTransformer(cfg)
    V, W = sd["text_embeddings.word_embeddings.weight"].shape
    glove = torch.randn(V, W)
    m.ingredient_embeddings.set_pretrained_embedding(glove, freeze=True)
    m.text_embeddings.set_pretrained_embedding(glove, freeze=False)
    verbs = torch.randn_like(sd["reasoner.action_embeddings"])
    m.reasoner.set_pretrained_embedding(verbs, freeze=False)
    m.recipe_reasoner.set_pretrained_embedding(verbs, freeze=True)
    assert hasattr(m.reasoner, "action_embeddings")
    assert torch.equal(m.state_dict()["text_embeddings.word_embeddings.weight"], glove)
    assert not m.ingredient_embeddings.word_embeddings.weight.requires_grad
    assert m.text_embeddings.word_embeddings.weight.requires_grad
    assert not m.recipe_reasoner.action_embeddings.requires_grad
    assert sorted(m.state_dict().keys()) == sorted(sd.keys())
    # a checkpoint written before set_pretrained_embedding holds `reasoner.action_embeddings.weight`
    old = {k: v.clone() for k, v in sd.items()}
    for r in ("reasoner", "recipe_reasoner"):
        old[f"{r}.action_embeddings.weight"] = old.pop(f"{r}.action_embeddings")
    res = StateAwareRecursiveTransformer(cfg).load_state_dict(old, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
