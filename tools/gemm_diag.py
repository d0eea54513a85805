#!/usr/bin/env python
"""Diagnostic sweep of the tcgen05 GEMM: one work unit per SM, varying K / width / epilogue (per-unit overhead)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from svpc_b200 import backend
dev = torch.device("cuda")
be = backend.get()
def t(M, N, K, lay="NN", od=torch.bfloat16, bias=False, reps=20):
    a = torch.randn((K, M) if lay[0] == "T" else (M, K), device=dev).to(torch.bfloat16)
    b = torch.randn((K, N) if lay[1] == "T" else (N, K), device=dev).to(torch.bfloat16)
    a = a.t() if lay[0] == "T" else a
    b = b.t() if lay[1] == "T" else b
    out = torch.empty(M, N, dtype=od, device=dev)
    bs = torch.randn(N, device=dev) if bias else None
    run = lambda: be.gemm(a, b, od, bs, 0, None, out=out)
    for _ in range(3): run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps): run()
    g.replay()
    e0.record()
    g.replay()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    print(f"{M:6d}x{N:5d}x{K:6d} {lay} out={str(od)[6:]:8s} bias={int(bias)} {us:8.1f} us  {2.0*M*N*K/us/1e6:7.1f} TF/s", flush=True)
M1 = 128 * 148
for K in (64, 256, 768, 3072):
    for N in (64, 128, 256):
        t(M1, N, K)
for N in (256,):
    t(M1, N, 64, od=torch.float32)
    t(M1, N, 64, bias=True)
    t(M1, N, 768, lay="NT")
    t(M1, N, 768, lay="TN")
    t(M1, N, 768, lay="TT")
    t(M1 * 2, N, 768)
    t(M1 * 3, N, 768)
