"""Times the per-step bookkeeping at the headline size (16 videos x 12 clips, 951-word scores): StepMetrics.update (no
host synchronisation, 3 launches) against the reference's sequence of cal_performance / calculate_f1 calls with their
.item() reads (src/train.py:153-173) on the same CUDA tensors.  Prints one JSON line."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from svpc_b200.metrics import StepMetrics  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    B, S, Lt, V, Ni, A = 16, 12, 25, 951, 10, 384
    g = torch.Generator(device=dev).manual_seed(1)
    preds = [torch.rand(S, Lt, V, device=dev, generator=g) for _ in range(B)]
    labels = [torch.randint(0, V, (S, Lt), device=dev, generator=g) for _ in range(B)]
    for lab in labels:
        lab[:, 20:] = -1
    ent = [torch.rand(S, Ni, device=dev, generator=g) for _ in range(B)]
    al = [(torch.rand(S, Ni, device=dev, generator=g) < 0.2).float() for _ in range(B)]
    ac = [torch.rand(S, A, device=dev, generator=g) for _ in range(B)]
    acts = [(torch.rand(S, A, device=dev, generator=g) < 0.01).float() for _ in range(B)]
    loss = torch.tensor(29000.0, device=dev)
    m = StepMetrics(dev)

    def fused():
        m.update(loss, preds, labels, ent, al, ac, acts)

    def stock():                                        # train.py:153-173 with its .item() reads
        n_correct = n_word = 0
        for pred, gold in zip(preds, labels):
            p = pred.max(2)[1].contiguous().view(-1)
            gd = gold.contiguous().view(-1)
            valid = gd.ne(-1)
            n_correct += p.eq(gd).masked_select(valid).sum().item()
            n_word += valid.sum().item()
        out = []
        for probs, golds in ((ent, al), (ac, acts)):
            c = r = q = 0
            for prob, gold in zip(probs, golds):
                pp = (prob > 0.5).flatten()
                gf = gold.flatten()
                c += gf[pp].sum().item(); r += gf.sum().item(); q += pp.sum().item()
            out.append((c, r, q))
        return n_correct, n_word, out, loss.item()

    def wall(fn, reps):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / reps * 1e3
    ms_f, ms_s = wall(fused, 50), wall(stock, 10)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        fused()
    e1.record()
    torch.cuda.synchronize()
    n_correct, n_word, f1, _ = stock()
    m.reset(); fused()
    c = m.counts()
    assert c[0] == n_correct and c[1] == n_word and c[2:5] == list(f1[0]) and c[5:8] == list(f1[1]), (c, n_correct, n_word, f1)
    print(json.dumps({"fused_host_ms_per_step": round(ms_f, 4), "fused_device_ms_per_step": round(e0.elapsed_time(e1) / 50, 4),
                      "stock_ms_per_step": round(ms_s, 3), "host_syncs_fused": 0, "host_syncs_stock": 2 * B + 6 * B + 1,
                      "speedup": round(ms_s / ms_f, 1)}))


if __name__ == "__main__":
    main()
