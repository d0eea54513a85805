#!/usr/bin/env python
"""Entity-state recurrence kernels at the headline shape (16 videos x 12 clips, Ni = 10, D = 768): 50 launches replayed from
a CUDA graph.  SVPC_REASONER_PIPE = 0 / 1 / 2 selects the plain, pipelined and cluster kernels."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from svpc_b200.cuda_backend import CudaBackend  # noqa: E402

dev = torch.device("cuda")
cb = CudaBackend()
B, S, Ni, D = 16, 12, 10, 768
I32 = lambda x: torch.tensor(x, dtype=torch.int32, device=dev)
cs = [b * S for b in range(B + 1)]
ci = [b * Ni for b in range(B + 1)]
clip_eo = [b * S * Ni + t * Ni for b in range(B) for t in range(S)]
lay = dict(cu_ingr=I32(ci), cu_steps=I32(cs), clip_ent_off=I32(clip_eo), max_ni=Ni, NE=B * S * Ni, B=B, max_steps=S)
Nc = B * S
E0 = torch.randn(B * Ni, D, device=dev) * 0.5
w2v, clog, w4s = torch.randn(Nc, D, device=dev) * 0.2, torch.randn(Nc, 3, device=dev), torch.randn(Nc, device=dev)
ent, ebar, Es, saved = cb.reasoner_fwd(E0, lay, w2v, clog, w4s)
d_ent, d_ebar, d_Es = torch.randn_like(ent), torch.randn_like(ebar), torch.randn_like(Es)


def graph_time(fn, reps=50):
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


out = {"mode": os.environ.get("SVPC_REASONER_PIPE", "1"),
       "fwd_us": round(graph_time(lambda: cb.reasoner_fwd(E0, lay, w2v, clog, w4s)), 1),
       "bwd_us": round(graph_time(lambda: cb.reasoner_bwd(d_ent, d_ebar, d_Es, E0, lay, w2v, clog, w4s, ent, ebar, Es, saved)), 1)}
print(json.dumps(out))
