"""Generates tests/golden/metrics_counts.npz by executing the UNMODIFIED function definitions cal_performance /
calculate_f1 / compute_total_f1 taken from /root/reference/src/train.py (the module cannot be imported here: tensorboardX,
h5py and spacy are not installed) on the seeded case of tests/metrics_case.py, in the order of train.py:153-180.

    python tests/golden/make_metrics_golden.py
"""
import ast
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import metrics_case as MC  # noqa: E402

REF = os.environ.get("SVPC_REFERENCE_ROOT", "/root/reference")


def load_reference_functions():
    src = open(os.path.join(REF, "src", "train.py")).read()
    tree = ast.parse(src)
    want = {"cal_performance", "calculate_f1", "compute_total_f1"}
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in want]
    assert {n.name for n in body} == want
    ns = {"torch": torch, "RCDataset": types.SimpleNamespace(IGNORE=-1)}
    exec(compile(ast.Module(body=body, type_ignores=[]), os.path.join(REF, "src", "train.py"), "exec"), ns)
    return ns


def run_reference():
    ns = load_reference_functions()
    tot = nwt = nct = 0
    e, a = [0, 0, 0], [0, 0, 0]
    for i in range(MC.N_STEPS):
        loss, preds, labels, ent, al, ac, acts = MC.make_step(i)
        for pred, gold in zip(preds, labels):                       # train.py:153-157
            nct += ns["cal_performance"](pred, gold)
            nwt += gold.ne(-1).sum().item()
        for acc, (probs, golds) in ((e, (ent, al)), (a, (ac, acts))):   # train.py:162-171
            c, r, q = ns["calculate_f1"](probs, golds)
            acc[0] += c; acc[1] += r; acc[2] += q
        tot += loss.item()
    ent_d, ac_d = ns["compute_total_f1"](*e), ns["compute_total_f1"](*a)
    return {"counts": np.asarray([nct, nwt] + e + a + [tot], np.float64),
            "loss_per_word": np.float64(tot / nwt), "accuracy": np.float64(nct / nwt),
            "ent": np.asarray([ent_d["recall"], ent_d["precision"], ent_d["f1"]], np.float64),
            "ac": np.asarray([ac_d["recall"], ac_d["precision"], ac_d["f1"]], np.float64)}


if __name__ == "__main__":
    out = run_reference()
    path = os.path.join(ROOT, "tests", "golden", "metrics_counts.npz")
    np.savez(path, **out)
    print("wrote", path, out)
