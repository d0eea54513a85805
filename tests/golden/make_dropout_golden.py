"""Generates tests/golden/train_*.npz: the UNMODIFIED reference (oracle/refshim.py) in TRAIN mode with explicit dropout
masks (tests/dropout_case.random_masks, seeded) on the ragged tiny case, all four model modes: loss, prediction scores
and every parameter gradient.  Pins the train mode of oracle/svpc_oracle.py (its ``drop`` argument).

    python tests/golden/make_dropout_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import refshim  # noqa: E402
from tests.dropout_case import random_masks  # noqa: E402
from tests.helpers import MODES, load_case  # noqa: E402

MASK_SEED = 77


def run_case(mode):
    cfg, sd, batch, _ = load_case("ragged", mode)
    drop = random_masks(cfg, batch, MASK_SEED)
    model = refshim.build_model(cfg, sd, train=True)
    refshim.install_dropout_masks(model, drop, list(batch["batch_step_num"]))
    loss, preds, ents, acts = refshim.run_forward(model, batch)
    loss.backward()
    out = {"loss": np.float64(loss.item())}
    for b, p in enumerate(preds):
        out[f"pred.{b}"] = p.detach().numpy()
    for b, (e, a) in enumerate(zip(ents, acts)):
        out[f"ent.{b}"] = e.detach().numpy()
        out[f"act.{b}"] = a.detach().numpy()
    for n, p in model.named_parameters():
        if p.grad is not None:
            out["grad." + n] = p.grad.numpy()
    return out


if __name__ == "__main__":
    assert refshim.available(), "needs /root/reference"
    here = os.path.dirname(os.path.abspath(__file__))
    for mode in MODES:
        out = run_case(mode)
        path = os.path.join(here, f"train_ragged_{mode}.npz")
        np.savez_compressed(path, **out)
        print(path, os.path.getsize(path) // 1024, "KiB", "loss", out["loss"])
