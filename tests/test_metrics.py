"""Bookkeeping row.  CPU: the oracle restatement against the fixture produced by the reference's own
cal_performance / calculate_f1 / compute_total_f1 (and against those functions live where /root/reference exists).
GPU: svpc_b200.metrics.StepMetrics (svpc_caption_accuracy / svpc_f1_counts) against the fixture - counts are exact."""
import os

import numpy as np
import pytest
import torch

from oracle import metrics_oracle as MO
from tests import metrics_case as MC

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "metrics_counts.npz")
HAVE_REF = os.path.isfile("/root/reference/src/train.py")


def check(gold, lpw, acc, ent, ac, counts):
    np.testing.assert_array_equal(np.asarray(counts[:8], np.float64), gold["counts"][:8])       # counts: exact
    assert counts[8] == pytest.approx(float(gold["counts"][8]), rel=1e-6)                       # summed fp32 losses
    assert lpw == pytest.approx(float(gold["loss_per_word"]), rel=1e-6)
    assert acc == float(gold["accuracy"])
    for d, g in ((ent, gold["ent"]), (ac, gold["ac"])):
        assert [d["recall"], d["precision"], d["f1"]] == pytest.approx(list(g), rel=1e-12)


def test_oracle_matches_reference_golden():
    lpw, acc, ent, ac, counts = MO.epoch(MC.make_step(i) for i in range(MC.N_STEPS))
    check(dict(np.load(GOLD)), lpw, acc, ent, ac, counts)


@pytest.mark.skipif(not HAVE_REF, reason="reference tree not present on this box")
def test_golden_is_what_the_reference_gives():
    from tests.golden.make_metrics_golden import run_reference
    live, gold = run_reference(), dict(np.load(GOLD))
    for k in gold:
        np.testing.assert_array_equal(np.asarray(live[k]), gold[k], err_msg=k)


def test_f1_edge_cases():
    assert MO.compute_total_f1(0, 0, 0) == {"recall": 0, "precision": 0, "f1": 0}              # train.py:52-65
    assert MO.compute_total_f1(0, 5, 0)["f1"] == 0
    from svpc_b200.metrics import compute_total_f1
    for c in [(0, 0, 0), (0, 5, 0), (3, 4, 6), (2, 2, 2)]:
        assert compute_total_f1(*c) == MO.compute_total_f1(*c)


@pytest.mark.gpu
def test_device_metrics_match_reference_golden():
    from svpc_b200.metrics import StepMetrics
    dev = torch.device("cuda:0")
    m = StepMetrics(dev)
    for i in range(MC.N_STEPS):
        loss, preds, labels, ent, al, ac, acts = MC.make_step(i)
        m.update(loss.to(dev), [p.to(dev) for p in preds], labels, [e.to(dev) for e in ent], al, [a.to(dev) for a in ac], acts)
    lpw, acc, entd, acd = m.result()
    check(dict(np.load(GOLD)), lpw, acc, entd, acd, m.counts())
    m.reset()
    assert m.counts() == [0.0] * 9


@pytest.mark.gpu
def test_device_metrics_on_model_outputs():
    """Counts on the real ragged model outputs (per-video views of the flat score buffer, OOV-extended widths) equal the
    oracle's on the same tensors; v-mode steps carry no entity / action lists."""
    from svpc_b200 import synth
    from svpc_b200.config import tiny_config
    from svpc_b200.metrics import StepMetrics
    from svpc_b200.model import StateAwareRecursiveTransformer
    dev = torch.device("cuda:0")
    for mode in ("full", "video"):
        cfg = tiny_config(mode)
        model = StateAwareRecursiveTransformer(cfg)
        model.load_state_dict(synth.init_state_dict(cfg, seed=1))
        model.to(dev).eval()
        batch = synth.make_batch(cfg, B=3, S=4, seed=5, ragged=True, n_ingr=4, n_oov=2)
        with torch.no_grad():
            loss, preds, ents, acts = model(*synth.forward_args(batch, dev), gumbel_noise=batch["gumbel"])
        labels = []
        for b, S in enumerate(batch["batch_step_num"]):                                       # train.py:133-138
            labels.append(torch.stack([x[b][cfg.max_v_len:] for x in batch["input_labels_list"][:S]]))
        m = StepMetrics(dev)
        m.update(loss, preds, labels, ents, batch["alignments"] if ents else (), acts, batch["actions"] if acts else ())
        want = MO.epoch([(loss.cpu(), [p.cpu() for p in preds], labels, [e.cpu() for e in ents],
                          batch["alignments"] if ents else [], [a.cpu() for a in acts], batch["actions"] if acts else [])])
        got = m.counts()
        assert got[:8] == [float(v) for v in want[4][:8]], (mode, got, want[4])
        assert got[1] > 0
