"""Per-kernel CUDA-event timing of one eager training step (all tags), sorted by time."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from svpc_b200 import synth, backend
from svpc_b200.config import youcook2_config
from svpc_b200.model import StateAwareRecursiveTransformer
dev = torch.device("cuda")
cfg = youcook2_config("full")
sd = synth.init_state_dict(cfg, seed=2019)
model = StateAwareRecursiveTransformer(cfg, compute_dtype=torch.bfloat16); model.load_state_dict(sd); model.to(dev).train()
batch = synth.make_batch(cfg, B=16, S=12, seed=2019, ragged=False, n_ingr=10, n_oov=0)
args = synth.forward_args(batch, dev)
noise = [g.to(dev) for g in batch["gumbel"]]
be = backend.get()
def step():
    model.zero_grad(set_to_none=True)
    loss = model(*args, gumbel_noise=noise)[0]
    loss.backward()
for _ in range(3): step()
torch.cuda.synchronize()
be.start_timing()
for _ in range(2): step()
per = be.stop_timing()
tot = sum(t for _, t in per.values()) / 2
print(f"total {tot:.3f} ms/step (per-launch events, eager)")
for k, (n, t) in sorted(per.items(), key=lambda kv: -kv[1][1]):
    print(f"{t/2*1e3:9.1f} us  n={n//2:4d}  avg={t/n*1e3:7.1f} us  {k}")
