"""Pointer-generator head and Gumbel front end at the benchmark shape (192 clips x 25 positions, D 768, V 951, 10 ingredients):
CUDA-event times of the forward / backward launches, L2 flushed between iterations."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from svpc_b200 import backend
dev = torch.device("cuda")
cb = backend.get()
I32 = lambda x: torch.tensor(x, dtype=torch.int32, device=dev)
B, S, Lt, D, V, ni = 16, 12, 25, 768, 951, 10
Nc = B * S
g = torch.Generator().manual_seed(0)
ntok = torch.randint(1, 4, (B * ni,), generator=g).tolist()
tok_ptr, tok_ids = [0], []
for n in ntok:
    tok_ids += torch.randint(4, V, (n,), generator=g).tolist(); tok_ptr.append(len(tok_ids))
clip_b = [b for b in range(B) for _ in range(S)]
clip_t = [t for b in range(B) for t in range(S)]
lay = dict(Lt=Lt, V=V, smoothing=0.1, max_ni=ni, clip_src_off=I32([(b * S + t) * ni for b, t in zip(clip_b, clip_t)]),
           clip_ni=I32([ni] * Nc), clip_gi=I32([b * ni for b in clip_b]), clip_vx=I32([V] * Nc),
           clip_P_off=I32([c * Lt * V for c in range(Nc)]), P_numel=Nc * Lt * V, tok_ptr=I32(tok_ptr), tok_ids=I32(tok_ids),
           labels=torch.randint(0, V, (Nc * Lt,), generator=g, dtype=torch.int32).to(dev), max_vx=V, tau=0.5)
rnd = lambda *s, scale=1.0: (torch.randn(*s, generator=g) * scale).to(dev)
dec = rnd(Nc * Lt, D).bfloat16()
G, E = rnd(Nc * ni, D, scale=0.05), rnd(Nc * ni, D)
w, bb, logits = rnd(2 * D, scale=0.05), rnd(1), rnd(Nc * Lt, V)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timed(fn, n=20):
    for _ in range(3): fn()
    t = 0.0
    for _ in range(n):
        flush.zero_(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); t += a.elapsed_time(b)
    return t / n * 1e3
P, rows, saved = cb.pointer_fwd(dec, G, E, w, bb, logits, lay)
dP = rnd(lay["P_numel"]); dloss = torch.tensor(1.0, device=dev)
print(f"pointer_fwd {timed(lambda: cb.pointer_fwd(dec, G, E, w, bb, logits, lay)):8.1f} us")
print(f"pointer_bwd {timed(lambda: cb.pointer_bwd(dP, dloss, dec, G, E, w, bb, logits, lay, P, saved)):8.1f} us (all launches)")
noise = -torch.log(torch.empty(lay['P_numel']).exponential_(generator=g)).to(dev)
emb = rnd(V, 300, scale=0.4); mask = torch.ones(Nc * Lt, device=dev)
x, idx, yv = cb.gumbel_fwd(P, noise, emb, lay, mask)
dx = rnd(Nc, 300)
print(f"gumbel_fwd  {timed(lambda: cb.gumbel_fwd(P, noise, emb, lay, mask)):8.1f} us")
print(f"gumbel_bwd  {timed(lambda: cb.gumbel_bwd(dx, P, noise, emb, lay, mask, idx, yv)):8.1f} us")
