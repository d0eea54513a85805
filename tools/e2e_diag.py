"""Where does the end-to-end step lose time against the device-resident replay?"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from svpc_b200 import synth, backend
from svpc_b200.config import youcook2_config
from svpc_b200.model import StateAwareRecursiveTransformer
from svpc_b200.graph import GraphedTrainStep
dev = torch.device("cuda")
cfg = youcook2_config("full")
sd = synth.init_state_dict(cfg, seed=2019)
model = StateAwareRecursiveTransformer(cfg, compute_dtype=torch.bfloat16); model.load_state_dict(sd); model.to(dev).train()
batch = synth.make_batch(cfg, B=16, S=12, seed=2019, ragged=False, n_ingr=10, n_oov=0)
def pin_all(x):
    if isinstance(x, torch.Tensor): return x.pin_memory()
    if isinstance(x, list) and x and isinstance(x[0], torch.Tensor): return [t.pin_memory() for t in x]
    return x
host = [pin_all(a) for a in synth.forward_args(batch)]
noise = [g.pin_memory() for g in batch["gumbel"]]
to_dev = lambda x: x.to(dev) if isinstance(x, torch.Tensor) else ([t.to(dev) for t in x] if isinstance(x, list) and x and isinstance(x[0], torch.Tensor) else x)
g = GraphedTrainStep(model, [to_dev(a) for a in host], gumbel_noise=[n.to(dev) for n in noise], warmup=2)
def timed(fn, k=10):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(k): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / k * 1e3
print("replay only            %.3f ms" % timed(lambda: g.run()))
print("replay + float(loss)   %.3f ms" % timed(lambda: float(g.run())))
# arena path (what bench.py's e2e uses): collate into the pinned arena once
a_args, a_noise = g.host_inputs()
for dst, src in zip(a_args, host):
    if isinstance(dst, torch.Tensor): dst.copy_(src)
    elif isinstance(dst, list) and dst and isinstance(dst[0], torch.Tensor):
        for d, s_ in zip(dst, src): d.copy_(s_)
for d, s_ in zip(a_noise, noise): d.copy_(s_)
host, noise = a_args, a_noise
print("arena bytes %d, uploaded %d, plan %s" % (g.arena.nbytes, g.arena.h2d_nbytes, g.arena.plan))
g.prefetch(host, noise); torch.cuda.synchronize()
def commit_only():
    g._pf.pairs = g._pf.pairs or g._pf_last
    g._pf_last = g._pf.pairs
    g._pf.commit()
g._pf_last = g._pf.pairs
print("D2D commit only        %.3f ms" % timed(commit_only))
def h2d_only():
    g.prefetch(host, noise)
print("H2D only (copy stream) %.3f ms" % timed(h2d_only))
t0 = time.perf_counter(); g.prefetch(host, noise); t1 = time.perf_counter(); torch.cuda.synchronize()
print("prefetch host enqueue  %.3f ms" % ((t1 - t0) * 1e3))
def e2e(k=10):
    g.prefetch(host, noise)
    for i in range(k):
        loss = g.run_prefetched()
        if i + 1 < k: g.prefetch(host, noise)
        float(loss)
e2e(2); torch.cuda.synchronize()
t0 = time.perf_counter(); e2e(10); torch.cuda.synchronize()
print("e2e pipelined          %.3f ms" % ((time.perf_counter() - t0) / 10 * 1e3))
# replay with a concurrent H2D but no dependency
def conc():
    g.prefetch(host, noise); g._pf.pairs = None if False else g._pf.pairs
    float(g.run())
print("replay || H2D (no dep) %.3f ms" % timed(conc))
