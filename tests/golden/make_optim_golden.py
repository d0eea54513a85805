"""Generates tests/golden/optim_iterations.npz by running the UNMODIFIED reference optimiser classes
(/root/reference/src/rtransformer/optimization.py: BertAdam, EMA) and torch's clip_grad_norm_ in the order of
src/train.py:140-147 on the seeded case of tests/optim_case.py.  Only runs in the build container.

    python tests/golden/make_optim_golden.py
"""
import importlib.util
import os
import sys
import warnings

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import optim_case as OC  # noqa: E402

REF = os.environ.get("SVPC_REFERENCE_ROOT", "/root/reference")


def load_reference_optimization():
    spec = importlib.util.spec_from_file_location("svpc_reference_optimization",
                                                  os.path.join(REF, "src", "rtransformer", "optimization.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class Holder(torch.nn.Module):
    """named_parameters() with the case's names (dots are not allowed in register_parameter, so override)."""

    def __init__(self, params):
        super().__init__()
        self._names = list(params)
        self._plist = torch.nn.ParameterList([torch.nn.Parameter(p.clone()) for p in params.values()])

    def named_parameters(self, *a, **k):
        return list(zip(self._names, self._plist))

    def parameters(self, *a, **k):
        return list(self._plist)


def run_reference(with_ema=True, grad_clip=OC.GRAD_CLIP):
    ref = load_reference_optimization()
    model = Holder(OC.make_params())
    named = model.named_parameters()
    groups = [                                                                      # train.py:339-343
        {"params": [p for n, p in named if not any(nd in n for nd in OC.NO_DECAY)], "weight_decay": 0.01},
        {"params": [p for n, p in named if any(nd in n for nd in OC.NO_DECAY)], "weight_decay": 0.0}]
    ema = None
    if with_ema:
        ema = ref.EMA(OC.EMA_DECAY)                                                 # train.py:344-348
        for n, p in named:
            ema.register(n, p.data)
    opt = ref.BertAdam(groups, **OC.HYPER)
    norms = []
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for it in range(OC.N_ITER):
            grads = OC.make_grads(it)
            opt.zero_grad()
            for n, p in named:
                p.grad = None if grads[n] is None else grads[n].clone()
            if grad_clip != -1:
                norms.append(float(torch.nn.utils.clip_grad_norm_(model.parameters(), grad_clip)))   # train.py:141-142
            opt.step()
            if ema is not None:
                ema(model, it)                                                      # train.py:146-147
    out = {"norms": np.asarray(norms, np.float32)}
    for n, p in named:
        out["p/" + n] = p.data.numpy().copy()
        st = opt.state[p]
        if len(st):
            out["m/" + n] = st["next_m"].numpy().copy()
            out["v/" + n] = st["next_v"].numpy().copy()
            out["step/" + n] = np.asarray(st["step"])
        if ema is not None:
            out["shadow/" + n] = ema.shadow[n].numpy().copy()
    return out


if __name__ == "__main__":
    out = run_reference()
    path = os.path.join(ROOT, "tests", "golden", "optim_iterations.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.shape for k, v in out.items() if k.startswith("p/")}, out["norms"])
