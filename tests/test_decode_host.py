"""Host-side logic of the KV-cached batched greedy decoder (svpc_b200.decode) on the plain-torch operator statements:
token-identical to the golden captions the reference Translator produced (tests/golden/*.npz)."""
import copy

import pytest
import torch

from svpc_b200 import backend
from svpc_b200.decode import greedy_decode
from svpc_b200.model import StateAwareRecursiveTransformer
from tests.helpers import MODES, GREEDY_SHARPEN, load_case
from tests.torch_backend import TorchBackend


@pytest.fixture(autouse=True)
def _torch_backend():
    prev = backend._impl
    backend.set_backend(TorchBackend())
    yield
    backend.set_backend(prev)


def decode_inputs(batch, device=None):
    b2 = copy.deepcopy(batch)
    mv = (lambda t: t.to(device)) if device is not None else (lambda t: t)
    return ([mv(t) for t in b2["input_ids_list"]], [mv(t) for t in b2["video_features_list"]],
            [mv(t) for t in b2["input_masks_list"]], [mv(t) for t in b2["token_type_ids_list"]],
            b2["ingr_input_ids"].tolist(), b2["ingr_masks"].tolist(), b2["ingr_sep_masks"].tolist(),
            b2["ingr_id_dict"], b2["oov_word_dict"], b2["alignments"], b2["actions"], b2["batch_step_num"])


@pytest.mark.parametrize("mode", MODES)
def test_batched_kv_cached_greedy_matches_golden(mode):
    cfg, sd, batch, gold = load_case("ragged", mode)
    sd = {k: v.clone() for k, v in sd.items()}
    sd["decoder_classifier.decoder.weight"] *= GREEDY_SHARPEN
    m = StateAwareRecursiveTransformer(cfg)
    m.load_state_dict(sd)
    m.eval()
    seqs = greedy_decode(m, decode_inputs(batch))
    for b, s in enumerate(seqs):
        assert torch.equal(s.cpu(), torch.from_numpy(gold[f"greedy.{b}"])), (mode, b, s, gold[f"greedy.{b}"])
