"""CPU tests of the host-side logic of svpc_b200.model (batching, ragged layouts, CSR, autograd wiring): the model runs
with the plain-torch operator statements of tests/torch_backend.py (test infrastructure, not a product path) and must
reproduce the reference golden fixtures."""
import pytest
import torch

from svpc_b200 import backend
from svpc_b200.model import StateAwareRecursiveTransformer
from svpc_b200.synth import forward_args
from tests.helpers import CASES, MODES, load_case, check_against_golden
from tests.torch_backend import TorchBackend


@pytest.fixture(autouse=True)
def _torch_backend():
    prev = backend._impl
    backend.set_backend(TorchBackend())
    yield
    backend.set_backend(prev)


def build(cfg, sd):
    m = StateAwareRecursiveTransformer(cfg)
    missing = m.load_state_dict(sd, strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    m.eval()
    return m


@pytest.mark.parametrize("case", list(CASES))
@pytest.mark.parametrize("mode", MODES)
def test_model_matches_golden(case, mode):
    cfg, sd, batch, gold = load_case(case, mode)
    m = build(cfg, sd)
    loss, preds, ents, acts = m(*forward_args(batch), gumbel_noise=batch["gumbel"])
    loss.backward()
    grads = {n: p.grad for n, p in m.named_parameters()}
    check_against_golden(gold, loss.detach(), preds, ents, acts, grads, loss_tol=2e-6, out_tol=1e-5, grad_tol=3e-4,
                         what=f"{case}/{mode}")


def test_state_dict_layout():
    cfg, sd, _, _ = load_case("ragged", "full")
    m = StateAwareRecursiveTransformer(cfg)
    assert sorted(m.state_dict().keys()) == sorted(sd.keys())
    for k, v in m.state_dict().items():
        assert v.shape == sd[k].shape and v.dtype == sd[k].dtype, k
