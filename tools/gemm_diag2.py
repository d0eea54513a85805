import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from svpc_b200 import backend
dev = torch.device("cuda"); be = backend.get()
def t(M, N, K, tag, reps=20):
    a = torch.randn(M, K, device=dev).to(torch.bfloat16); b = torch.randn(N, K, device=dev).to(torch.bfloat16)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    run = lambda: be.gemm(a, b, torch.bfloat16, None, 0, None, out=out)
    for _ in range(3): run()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps): run()
    g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    print(f"dbg={tag:3s} {M}x{N}x{K}: {e0.elapsed_time(e1)/reps*1e3:7.1f} us", flush=True)
for dbg in ("0", "1", "2", "3", "4", "6", "8"):
    os.environ["SVPC_TC_DBG"] = dbg
    t(128 * 148, 256, 64, dbg)
    t(128 * 148 * 3, 256, 768, dbg)
