"""Seeded case shared by the metrics tests and tests/golden/make_metrics_golden.py: 3 steps of 3 ragged videos."""
import numpy as np
import torch

N_STEPS, LT, V, A = 3, 25, 61, 40
STEPS_PER_VIDEO = [4, 1, 3]
NI = [5, 3, 7]
OOV = [0, 2, 1]


def make_step(i, seed=23):
    rs = np.random.RandomState(seed * 100 + i)
    preds, labels, ent, al, ac, acts = [], [], [], [], [], []
    for S, ni, oov in zip(STEPS_PER_VIDEO, NI, OOV):
        p = rs.rand(S, LT, V + oov).astype(np.float32)
        g = rs.randint(0, V + oov, size=(S, LT)).astype(np.int64)
        hit = rs.rand(S, LT) < 0.4                               # make ~40 % of the labelled positions correct
        p[hit, g[hit]] = 2.0
        g[rs.rand(S, LT) < 0.3] = -1                             # ignored positions
        preds.append(torch.from_numpy(p)); labels.append(torch.from_numpy(g))
        ent.append(torch.from_numpy(rs.rand(S, ni).astype(np.float32)))
        al.append(torch.from_numpy((rs.rand(S, ni) < 0.3).astype(np.float32)))
        ac.append(torch.from_numpy(rs.rand(S, A).astype(np.float32)))
        acts.append(torch.from_numpy((rs.rand(S, A) < 0.05).astype(np.float32)))
    loss = torch.tensor(float(rs.rand() * 1000 + 100))
    return loss, preds, labels, ent, al, ac, acts
