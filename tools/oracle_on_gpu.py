"""The reference algorithm as eager PyTorch fp32 ON THE B200 ("same box" figure of SURVEY.md 8d): the oracle restatement
(the same torch ops in the same order as the reference model, oracle/svpc_oracle.py) with every tensor on cuda:0, on the
headline batch (16 videos x 12 clips).  A reported baseline, like the CPU one; prints one JSON line."""
import copy
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import svpc_oracle as O  # noqa: E402
from svpc_b200 import synth  # noqa: E402
from svpc_b200.config import youcook2_config  # noqa: E402


def to_dev(x, dev):
    if isinstance(x, torch.Tensor):
        return x.to(dev)
    if isinstance(x, list):
        return [to_dev(v, dev) for v in x]
    if isinstance(x, tuple):
        return tuple(to_dev(v, dev) for v in x)
    if isinstance(x, dict):
        return {k: to_dev(v, dev) for k, v in x.items()}
    return x


def main():
    videos = int(os.environ.get("VIDEOS", "16"))
    steps = int(os.environ.get("STEPS", "3"))
    dev = torch.device("cuda:0")
    cfg = youcook2_config("full")
    sd = {k: v.to(dev) for k, v in synth.init_state_dict(cfg, seed=2019).items()}
    batch = to_dev(synth.make_batch(cfg, B=videos, S=12, seed=2019, ragged=False, n_ingr=10, n_oov=0), dev)
    times = []
    for i in range(1 + steps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        O.loss_and_grads(sd, cfg, copy.deepcopy(batch))
        torch.cuda.synchronize()
        if i >= 1:
            times.append(time.perf_counter() - t0)
    per = sum(times) / len(times)
    print(json.dumps({"what": "oracle port of the reference fwd+bwd, eager PyTorch fp32 on cuda:0", "videos": videos,
                      "clips": batch["n_clips"], "s_per_step": per, "clips_per_s": batch["n_clips"] / per, "steps": steps}))


if __name__ == "__main__":
    main()
