"""Seeded synthetic case shared by the optimizer tests and tests/golden/make_optim_golden.py: a handful of named
parameters (names chosen so that train.py:339-343's weight-decay grouping has members in both groups), one of which
never receives a gradient (BertAdam skips it, optimization.py:286-287), and per-iteration gradients whose scale
makes the global clip (train.py:141-142) and the per-parameter clip (optimization.py:307-308) active on some
iterations and inactive on others.  Sizes are odd on purpose (unaligned tails, > one chunk)."""
import numpy as np
import torch

SHAPES = [
    ("enc.layer.0.weight", (96, 131)),
    ("enc.layer.0.bias", (131,)),
    ("enc.LayerNorm.weight", (131,)),
    ("enc.LayerNorm.bias", (131,)),
    ("emb.word_embeddings.weight", (257, 77)),      # 19789 elements: more than one 16384-element chunk
    ("dead.weight", (33, 7)),                       # never gets a gradient
    ("head.decoder.weight", (5, 3)),
    ("reasoner.action_embeddings", (40, 30)),
]
NO_GRAD = {"dead.weight"}
N_ITER = 6
GRAD_SCALE = [3.0, 0.5, 1e-3, 2e-4, 40.0, 1e-2]     # iteration i's gradients ~ N(0, scale^2 / numel)
HYPER = dict(lr=1e-2, warmup=0.25, t_total=8, schedule="warmup_linear", b1=0.9, b2=0.999, e=1e-6, max_grad_norm=1.0)
GRAD_CLIP = 1.0
EMA_DECAY = 0.9999
NO_DECAY = ["bias", "LayerNorm.bias", "LayerNorm.weight"]       # train.py:339


def make_params(seed=11):
    rs = np.random.RandomState(seed)
    return {n: torch.from_numpy((rs.randn(*s) * 0.05).astype(np.float32)) for n, s in SHAPES}


def make_grads(it, seed=11):
    rs = np.random.RandomState(seed * 1000 + it)
    out = {}
    for n, s in SHAPES:
        g = (rs.randn(*s) * (GRAD_SCALE[it] / np.sqrt(np.prod(s)))).astype(np.float32)
        out[n] = None if n in NO_GRAD else torch.from_numpy(g)
    return out


def decay_groups(names):
    """train.py:339-343 as index lists."""
    dec = [i for i, n in enumerate(names) if not any(nd in n for nd in NO_DECAY)]
    nod = [i for i, n in enumerate(names) if any(nd in n for nd in NO_DECAY)]
    return [{"params": dec, "weight_decay": 0.01}, {"params": nod, "weight_decay": 0.0}]
