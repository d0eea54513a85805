"""Shared helpers for the parity tests."""
import os

import numpy as np
import torch

from svpc_b200.config import tiny_config
from svpc_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# must mirror tests/golden/make_golden.py
CASES = {
    "ragged": (1, dict(B=3, S=4, seed=5, ragged=True, n_ingr=4, n_oov=2)),
    "full_len": (2, dict(B=2, S=3, seed=7, ragged=False, n_ingr=3, n_oov=0)),
}
MODES = ("video", "copy", "reason_copy", "full")
GREEDY_SHARPEN = 20.0


def load_case(case, mode):
    wseed, bkw = CASES[case]
    cfg = tiny_config(mode)
    sd = synth.init_state_dict(cfg, seed=wseed)
    batch = synth.make_batch(cfg, **bkw)
    gold = dict(np.load(os.path.join(GOLDEN, f"{case}_{mode}.npz")))
    return cfg, sd, batch, gold


def rel_err(a, b):
    a = torch.as_tensor(a, dtype=torch.float64)
    b = torch.as_tensor(b, dtype=torch.float64)
    return float((a - b).norm() / (b.norm() + 1e-30))


def check_against_golden(gold, loss, preds, ents, acts, grads, *, loss_tol, out_tol, grad_tol, what=""):
    """Compare one fwd+bwd result with a golden fixture. Gradients are compared per tensor against the
    *global* gradient scale: some true gradients are exactly zero (key biases, Wing.bias: softmax shift
    invariance) and hold only rounding noise in both implementations."""
    assert abs(float(loss) - float(gold["loss"])) <= loss_tol * abs(float(gold["loss"])), (what, float(loss), float(gold["loss"]))
    for b, p in enumerate(preds):
        g = torch.from_numpy(gold[f"pred.{b}"])
        assert p.shape == g.shape, (what, b, p.shape, g.shape)
        err = (p.detach().cpu().double() - g.double()).abs().max().item()
        assert err <= out_tol * max(1e-3, g.abs().max().item()), (what, "pred", b, err)
    for b, (e, a) in enumerate(zip(ents, acts)):
        assert (e.detach().cpu() - torch.from_numpy(gold[f"ent.{b}"])).abs().max().item() <= out_tol, (what, "ent", b)
        assert (a.detach().cpu() - torch.from_numpy(gold[f"act.{b}"])).abs().max().item() <= out_tol, (what, "act", b)
    if grads is None:
        return
    gnames = [k[5:] for k in gold if k.startswith("grad.")]
    gscale = max(float(np.abs(gold["grad." + n]).max()) for n in gnames)
    for n in gnames:
        ref = torch.from_numpy(gold["grad." + n]).double()
        got = grads.get(n)
        assert got is not None, (what, "missing grad", n)
        got = got.detach().cpu().double()
        assert got.shape == ref.shape, (what, n)
        err = (got - ref).abs().max().item()
        tol = grad_tol * max(ref.abs().max().item(), 1e-3 * gscale)
        assert err <= tol, (what, "grad", n, err, tol)
    for n, g in grads.items():
        if n not in gnames:
            assert g is None or float(g.abs().max()) == 0.0, (what, "unexpected grad", n)
