"""Times the fused iteration tail (global clip + BertAdam + EMA, svpc_b200.optim) on the full-size vivt parameter set
and, beside it, the same iteration written with stock torch ops the way the reference runs it (clip_grad_norm_,
per-parameter BertAdam loop, per-parameter EMA loop).  Prints one JSON line."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from svpc_b200 import optim as SO, synth  # noqa: E402
from svpc_b200.config import youcook2_config  # noqa: E402
from svpc_b200.model import StateAwareRecursiveTransformer  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    cfg = youcook2_config("full")
    model = StateAwareRecursiveTransformer(cfg)
    model.load_state_dict(synth.init_state_dict(cfg, seed=3))
    model.to(dev)
    named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
    for _, p in named:
        p.grad = torch.randn_like(p) * 1e-3
    nod = ["bias", "LayerNorm.bias", "LayerNorm.weight"]
    groups = [{"params": [p for n, p in named if not any(x in n for x in nod)], "weight_decay": 0.01},
              {"params": [p for n, p in named if any(x in n for x in nod)], "weight_decay": 0.0}]
    opt = SO.BertAdam(groups, lr=1e-4, warmup=0.1, t_total=100000)
    ema = SO.EMA(0.9999)
    for n, p in named:
        ema.register(n, p.data)
    n_el = sum(p.numel() for _, p in named)

    def fused(i):
        opt.step(grad_clip=1.0, ema=ema, niter=i)

    def stock(i):
        # the reference's sequence with stock torch kernels (optimization.py:284-331, 196-203; train.py:141-147)
        torch.nn.utils.clip_grad_norm_([p for _, p in named], 1.0)
        for g in groups:
            for p in g["params"]:
                st = stock_state.setdefault(p, {"m": torch.zeros_like(p.data), "v": torch.zeros_like(p.data)})
                torch.nn.utils.clip_grad_norm_(p, 1.0)
                st["m"].mul_(0.9).add_(p.grad, alpha=0.1)
                st["v"].mul_(0.999).addcmul_(p.grad, p.grad, value=0.001)
                upd = st["m"] / (st["v"].sqrt() + 1e-6)
                if g["weight_decay"] > 0:
                    upd += g["weight_decay"] * p.data
                p.data.add_(-(1e-4 * upd))
        d = min(0.9999, (1 + i) / (10.0 + i))
        for n, p in named:
            stock_shadow[n] = ((1.0 - d) * p.data + d * stock_shadow[n]).clone()
    stock_state, stock_shadow = {}, {n: p.data.clone() for n, p in named}

    def timeit(fn, reps):
        for i in range(3):
            fn(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(reps):
            fn(3 + i)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps
    ms_f = timeit(fused, 20)
    ms_s = timeit(stock, 5)
    be = SO._be()
    be.start_timing()
    for i in range(10):
        fused(100 + i)
    per = {k: round(t / n, 4) for k, (n, t) in be.stop_timing().items()}
    byts = n_el * 40.0          # sqnorm reads g (4 B); apply reads g,p,m,v,shadow (20 B) and writes p,m,v,shadow (16 B)
    print(json.dumps({"tensors": len(named), "elements": n_el, "fused_ms": round(ms_f, 4), "stock_torch_ms": round(ms_s, 3),
                      "speedup": round(ms_s / ms_f, 1), "algorithmic_bytes": byts, "fused_GBps": round(byts / ms_f / 1e6, 1),
                      "launches_fused": 3, "kernel_ms": per,
                      "apply_GBps": round(n_el * 36.0 / per["svpc_optim_apply"] / 1e6, 1),
                      "sqnorm_GBps": round(n_el * 4.0 / per["svpc_optim_sqnorm"] / 1e6, 1)}))


if __name__ == "__main__":
    main()
