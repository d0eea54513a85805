#!/usr/bin/env python
"""Micro-benchmark of the tcgen05 GEMM on the shapes of the SVPC step: back-to-back launches between two CUDA events on
the launching stream.  usage: python tools/gemm_bench.py [reps]  -> one line per shape, TFLOP/s and fraction of the
measured sustained bf16 peak (MEASURED_PEAKS.json)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from svpc_b200 import backend  # noqa: E402

# (M, N, K, layout, epilogue)   layout: first letter A, second B; N = K-major (row-major, K contiguous), T = MN-major
SHAPES = [
    (19200, 768, 3072, "NN", "bias+relu"),      # video embedding
    (19200, 2304, 768, "NN", "bias"),           # QKV, encoder layer 1
    (19200, 1536, 768, "NN", "bias"),           # KV, encoder layer 2
    (19200, 768, 768, "NN", "bias"),
    (19200, 768, 768, "NN", "bias+gelu"),       # FFN in
    (19200, 768, 768, "NN", "bias+res"),        # attention output / FFN out with residual
    (19200, 768, 768, "NT", "plain"),           # dgrad
    (19200, 3072, 768, "NT", "plain"),          # dgrad of the video projection (train mode)
    (768, 768, 19200, "TT", "plain-f32"),       # wgrad (split-K)
    (768, 2304, 19200, "TT", "plain-f32"),
    (768, 3072, 19200, "TT", "plain-f32"),
    (4800, 768, 768, "NN", "bias"),             # decoder
    (4800, 2304, 768, "NN", "bias"),
    (4800, 951, 768, "NN", "bias-f32"),         # LM head
    (192, 768, 768, "NN", "bias"),              # step encoder / reasoner
    (192, 768, 768, "NT", "plain"),
    (768, 768, 4800, "TT", "plain-f32"),        # decoder wgrad (split-K)
    (768, 768, 192, "TT", "plain-f32"),         # step-level wgrad
]


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    only = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else None      # indices into SHAPES
    dev = torch.device("cuda")
    be = backend.get()
    peak = 1400.0
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        peak = json.load(open(p)).get("bf16_tflops_sustained", peak)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for si, (M, N, K, lay, epi) in enumerate(SHAPES):
        if only is not None and si not in only:
            continue
        a = torch.randn((K, M) if lay[0] == "T" else (M, K), device=dev).to(torch.bfloat16)
        b = torch.randn((K, N) if lay[1] == "T" else (N, K), device=dev).to(torch.bfloat16)
        a = a.t() if lay[0] == "T" else a
        b = b.t() if lay[1] == "T" else b
        od = torch.float32 if epi.endswith("f32") else torch.bfloat16
        out = torch.empty(M, N, dtype=od, device=dev)
        bias = torch.randn(N, device=dev) if "bias" in epi else None
        res = torch.randn(M, N, device=dev).to(od) if "res" in epi else None
        act = 1 if "relu" in epi else (2 if "gelu" in epi else 0)
        run = lambda: be.gemm(a, b, od, bias, act, res, out=out)
        for _ in range(3):
            run()
        # the launches are replayed from a CUDA graph, as in the training step (no interpreter gap between kernels)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(reps):
                run()
        g.replay()
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / reps * 1e3
        tf = 2.0 * M * N * K / us / 1e6
        ref = (a.float() @ b.float().t())
        if bias is not None:
            ref = ref + bias
        if act == 1:
            ref = ref.relu()
        elif act == 2:
            ref = torch.nn.functional.gelu(ref)
        if res is not None:
            ref = ref + res.float()
        err = float((out.float() - ref).abs().max() / ref.abs().max())
        print(f"{M:6d}x{N:5d}x{K:6d} {lay} {epi:10s} {us:8.1f} us  {tf:7.1f} TFLOP/s  {tf / peak:5.3f} of peak  relerr {err:.1e}", flush=True)


if __name__ == "__main__":
    main()
