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


def test_input_arena_upload_plan():
    """Upload plan of the step's input arena (svpc_b200.graph.InputArena.layout): neighbouring tensors merge into one
    copy, the (B, 125, F) feature blocks become ONE pitched copy of their first 100 rows."""
    from svpc_b200.graph import InputArena
    B, L, Lv, Fd, S = 4, 128, 100, 64, 3                      # sizes that are multiples of the arena alignment
    ids = [torch.zeros(B, L, dtype=torch.int64) for _ in range(S)]
    feats = [torch.zeros(B, L, Fd) for _ in range(S)]
    masks = [torch.zeros(B, L) for _ in range(S)]
    odd = [torch.zeros(7, dtype=torch.int32)]                     # 28 bytes: padded to the arena alignment
    flat = ids + feats + masks + odd
    spec, plan, nbytes = InputArena.layout(flat, used_rows={i: Lv for i in range(S, 2 * S)})
    assert len(spec) == len(flat) and nbytes % InputArena.ALIGN == 0
    assert len(plan) == 3
    (o0, p0, w0, h0), (o1, p1, w1, h1), (o2, p2, w2, h2) = plan
    assert (o0, p0, h0) == (0, 0, 1) and w0 == S * B * L * 8
    assert o1 == w0 and p1 == L * Fd * 4 and w1 == Lv * Fd * 4 and h1 == S * B
    assert o2 == o1 + S * B * L * Fd * 4 and p2 == 0 and h2 == 1 and o2 + w2 <= nbytes
    uploaded = sum(w * h for _, _, w, h in plan)
    assert uploaded == nbytes - S * B * (L - Lv) * Fd * 4 - (nbytes - (o2 + w2))
    # without the hint everything is one copy
    assert len(InputArena.layout(flat)[1]) == 1
