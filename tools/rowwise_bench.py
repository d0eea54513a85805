#!/usr/bin/env python
"""Micro-benchmark of the row-wise backward kernels at the headline shapes (19200 x 768 and 4800 x 768, bf16):
ln_bwd alone, act_bwd (dropout / GELU), and ln_bwd with the fused dz output; GB/s against the algorithmic bytes."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from svpc_b200.cuda_backend import CudaBackend  # noqa: E402

dev = torch.device("cuda")
cb = CudaBackend()
bf = torch.bfloat16


def timeit(fn, reps=30):
    for _ in range(5):
        fn()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ts = []
    for _ in range(reps):
        flush.zero_()                       # evict L2 between launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2] * 1e3


out = {}
for rows in (19200, 4800):
    D = 768
    dy, pre, ref = (torch.randn(rows, D, device=dev).to(bf) for _ in range(3))
    mean, rstd = torch.randn(rows, device=dev), torch.rand(rows, device=dev) + 0.5
    gamma = torch.ones(D, device=dev)
    drop = (torch.tensor([7], dtype=torch.int32, device=dev), 3, 0.1)
    n = rows * D * 2
    r = {}
    t = timeit(lambda: cb.ln_bwd(dy, pre, mean, rstd, gamma))
    r["ln_bwd"] = {"us": round(t, 1), "GBs": round(3 * n / t / 1e3, 0)}
    t = timeit(lambda: cb.act_bwd(dy, None, 0, drop=drop))
    r["act_bwd_drop"] = {"us": round(t, 1), "GBs": round(2 * n / t / 1e3, 0)}
    t = timeit(lambda: cb.act_bwd(dy, ref, 2))
    r["act_bwd_gelu"] = {"us": round(t, 1), "GBs": round(3 * n / t / 1e3, 0)}
    t = timeit(lambda: cb.colsum(dy))
    r["colsum"] = {"us": round(t, 1), "GBs": round(n / t / 1e3, 0)}
    t = timeit(lambda: cb.ln_bwd(dy, pre, mean, rstd, gamma, fuse=(0, None, drop, True)))
    r["ln_bwd_fused_drop"] = {"us": round(t, 1), "GBs": round(4 * n / t / 1e3, 0)}
    t = timeit(lambda: cb.ln_bwd(dy, pre, mean, rstd, gamma, fuse=(2, ref, None, True)))
    r["ln_bwd_fused_gelu"] = {"us": round(t, 1), "GBs": round(5 * n / t / 1e3, 0)}
    out[f"{rows}x{D}"] = r
print(json.dumps(out))
