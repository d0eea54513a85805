"""CPU tests: the oracle (oracle/svpc_oracle.py) against the committed reference outputs (tests/golden/, produced by
the unmodified reference through tests/golden/make_golden.py) and, when /root/reference is mounted, against the
live reference."""
import copy

import pytest
import torch

from oracle import svpc_oracle as O
from oracle import refshim
from tests.helpers import CASES, MODES, GREEDY_SHARPEN, load_case, check_against_golden


@pytest.mark.parametrize("case", list(CASES))
@pytest.mark.parametrize("mode", MODES)
def test_oracle_matches_golden(case, mode):
    cfg, sd, batch, gold = load_case(case, mode)
    loss, parts, grads, preds, ents, acts = O.loss_and_grads(sd, cfg, batch)
    check_against_golden(gold, loss, preds, ents, acts, grads, loss_tol=1e-6, out_tol=1e-5, grad_tol=2e-4,
                         what=f"{case}/{mode}")


@pytest.mark.parametrize("mode", MODES)
def test_oracle_greedy_matches_golden(mode):
    cfg, sd, batch, gold = load_case("ragged", mode)
    sd = {k: v.clone() for k, v in sd.items()}
    sd["decoder_classifier.decoder.weight"] *= GREEDY_SHARPEN
    seqs = O.greedy_decode(sd, cfg, batch)
    for b, s in enumerate(seqs):
        assert torch.equal(s, torch.from_numpy(gold[f"greedy.{b}"])), (mode, b)


@pytest.mark.parametrize("mode", ("video", "full"))
def test_oracle_fp64_close_to_fp32(mode):
    cfg, sd, batch, gold = load_case("full_len", mode)
    l32 = O.forward(sd, cfg, batch)[0]
    l64 = O.forward(sd, cfg, batch, dtype=torch.float64)[0]
    assert abs(float(l32) - float(l64)) < 1e-4 * abs(float(l64))


@pytest.mark.skipif(not refshim.available(), reason="reference tree not mounted (GPU box)")
def test_oracle_matches_live_reference():
    cfg, sd, batch, _ = load_case("ragged", "full")
    batch = copy.deepcopy(batch)
    model = refshim.build_model(cfg, sd)
    loss_ref, preds_ref, ents_ref, acts_ref = refshim.run_forward(model, batch)
    loss, preds, ents, acts = O.forward(sd, cfg, batch)
    assert abs(float(loss) - float(loss_ref)) <= 1e-6 * abs(float(loss_ref))
    for a, b in zip(preds, preds_ref):
        assert torch.allclose(a, b, atol=1e-6)


def test_label_smoothing_last_column_quirk():
    # model.py:51: one_hot[ignore_index=-1] = 0 removes the smoothing mass of the LAST column
    P = torch.full((2, 5), 0.2)
    l = O.label_smoothing_loss(P, torch.tensor([1, -1]), 0.1)
    q = torch.tensor([0.025, 0.9, 0.025, 0.025, 0.0])
    exp = (q[q > 0] * (q[q > 0].log() - torch.log(torch.tensor(0.2 + 1e-12)))).sum()
    assert abs(float(l) - float(exp)) < 1e-6


# ---- train mode: explicit dropout masks at every reference nn.Dropout site
@pytest.mark.parametrize("mode", MODES)
def test_oracle_train_mode_matches_golden(mode):
    """oracle(drop=masks) against the unmodified reference run in train mode with the same masks injected into its
    nn.Dropout modules (tests/golden/make_dropout_golden.py)."""
    import numpy as np
    import os
    from tests.dropout_case import random_masks
    from tests.golden.make_dropout_golden import MASK_SEED
    from tests.helpers import GOLDEN
    cfg, sd, batch, eval_gold = load_case("ragged", mode)
    gold = dict(np.load(os.path.join(GOLDEN, f"train_ragged_{mode}.npz")))
    assert abs(float(gold["loss"]) - float(eval_gold["loss"])) > 1e-3 * abs(float(eval_gold["loss"]))   # the masks bite
    loss, parts, grads, preds, ents, acts = O.loss_and_grads(sd, cfg, batch, drop=random_masks(cfg, batch, MASK_SEED))
    check_against_golden(gold, loss, preds, ents, acts, grads, loss_tol=1e-6, out_tol=1e-5, grad_tol=2e-4,
                         what=f"train/{mode}")


@pytest.mark.skipif(not refshim.available(), reason="reference tree not mounted (GPU box)")
def test_oracle_train_mode_matches_live_reference():
    from tests.dropout_case import random_masks
    cfg, sd, batch, _ = load_case("full_len", "full")
    batch = copy.deepcopy(batch)
    drop = random_masks(cfg, batch, 3)
    model = refshim.build_model(cfg, sd, train=True)
    unused = refshim.install_dropout_masks(model, drop, list(batch["batch_step_num"]))
    assert unused == []                      # every nn.Dropout of the reference model is a known site
    loss_ref, preds_ref, _, _ = refshim.run_forward(model, batch)
    loss, preds, _, _ = O.forward(sd, cfg, batch, drop=drop)
    assert abs(float(loss) - float(loss_ref)) <= 1e-6 * abs(float(loss_ref))
    for a, b in zip(preds, preds_ref):
        assert torch.allclose(a, b, atol=1e-6)
