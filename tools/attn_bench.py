#!/usr/bin/env python
"""Encoder self-attention at the headline shape (192 clips x 12 heads, L = 100, dh = 64, bf16, dropout 0.1): forward and
backward, 20 launches replayed from a CUDA graph."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from svpc_b200.cuda_backend import CudaBackend  # noqa: E402

dev = torch.device("cuda")
cb = CudaBackend()
N, L, H, dh = 192, 100, 12, 64
D = H * dh
bf = torch.bfloat16
qkv = torch.randn(N * L, 3 * D, device=dev).to(bf)
q, k, v = qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:]
off = torch.arange(N, dtype=torch.int32, device=dev) * L
ln = torch.full((N,), L, dtype=torch.int32, device=dev)
km = torch.ones(N * L, device=dev)
drop = (torch.tensor([3], dtype=torch.int32, device=dev), 2, 0.1) if "--no-drop" not in sys.argv else None
dctx = torch.randn(N * L, D, device=dev).to(bf)
dqkv = torch.empty_like(qkv)


def fwd():
    return cb.attn_fwd(q, k, v, off, ln, off, ln, km, False, H, L, L, bf, drop=drop)


def bwd():
    cb.attn_bwd(dctx, q, k, v, off, ln, off, ln, km, False, H, L, L, dqkv[:, :D], dqkv[:, D:2 * D], dqkv[:, 2 * D:], drop=drop)


def graph_time(fn, reps=20):
    for _ in range(3):
        fn()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


flops = 4.0 * N * H * L * L * dh
f, b = graph_time(fwd), graph_time(bwd)
print(json.dumps({"fwd_us": round(f, 1), "bwd_us": round(b, 1), "fwd_tflops": round(flops / f / 1e6, 1),
                  "bwd_tflops": round(2.5 * flops / b / 1e6, 1), "dropout": drop is not None}))
