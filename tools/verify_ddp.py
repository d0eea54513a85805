"""NCCL data-parallel equivalence on real GPUs: N ranks, each running the CAPTURED step (NCCL all-reduces
inside the CUDA graph, gradients left in their buckets) on its own videos, against a 1-rank eager step on the concatenated
batch.  Launched through bench.py so that it runs exactly like the benchmark:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        bench.py --gpus 2 --verify-ddp > profiles/r2_verify_ddp_n2.json

Prints one JSON line (rank 0): loss and per-parameter gradient errors, and `ok`.
"""
import json

import torch
import torch.distributed as dist


def _slice_batch(batch, vids):
    S = max(batch["batch_step_num"][b] for b in vids)
    out = dict(batch)
    idx = torch.tensor(vids)
    for k in ("input_ids_list", "video_features_list", "input_masks_list", "token_type_ids_list", "input_labels_list"):
        out[k] = [t[idx] for t in batch[k][:S]]
    for k in ("ingr_input_ids", "ingr_masks", "ingr_sep_masks"):
        out[k] = batch[k][idx]
    for k in ("batch_step_num", "ingr_id_dict", "oov_word_dict", "extra_zeros", "alignments", "actions", "gumbel"):
        out[k] = [batch[k][b] for b in vids]
    out["n_clips"] = sum(out["batch_step_num"])
    return out


def run(args, dev, rank, world):
    from svpc_b200 import synth
    from svpc_b200.config import youcook2_config
    from svpc_b200.ddp import GradAllReducer
    from svpc_b200.graph import GraphedTrainStep
    from svpc_b200.model import StateAwareRecursiveTransformer
    from svpc_b200.synth import forward_args
    assert world >= 2, "--verify-ddp needs torchrun with at least 2 ranks"
    results = {}
    for dtype, name in ((torch.float32, "fp32"), (torch.bfloat16, "bf16")):
        cfg = youcook2_config(args.mode)
        sd = synth.init_state_dict(cfg, seed=2019)
        per = 2
        full = synth.make_batch(cfg, B=per * world, S=4, seed=77, ragged=False, n_ingr=5, n_oov=0)
        mine = _slice_batch(full, list(range(rank * per, (rank + 1) * per)))
        m = StateAwareRecursiveTransformer(cfg, compute_dtype=dtype)
        m.load_state_dict(sd)
        m.to(dev).eval()
        red = GradAllReducer(m, bucket_bytes=8 << 20)
        assert red.enabled and len(red.buckets) > 1
        noise = [g.to(dev) for g in mine["gumbel"]]
        step = GraphedTrainStep(m, forward_args(mine, dev), gumbel_noise=noise, reducer=red, warmup=2)
        for _ in range(2):
            loss = step.run()
        torch.cuda.synchronize()
        tot = loss.detach().clone()
        dist.all_reduce(tot)
        grads_n = {n: (None if p.grad is None else p.grad.detach().clone()) for n, p in m.named_parameters()}
        # every rank must hold the same summed gradients
        same = True
        for n, g in grads_n.items():
            if g is None:
                continue
            ref = g.clone()
            dist.broadcast(ref, src=0)
            same = same and bool(torch.equal(ref, g))
        # the single-process statement on the concatenated batch (rank 0, no collective)
        if rank == 0:
            red.enabled = False
            m1 = StateAwareRecursiveTransformer(cfg, compute_dtype=dtype)
            m1.load_state_dict(sd)
            m1.to(dev).eval()
            l1 = m1(*forward_args(full, dev), gumbel_noise=[g.to(dev) for g in full["gumbel"]])[0]
            l1.backward()
            torch.cuda.synchronize()
            gscale = max(float(p.grad.abs().max()) for p in m1.parameters() if p.grad is not None)
            worst, worst_n, missing = 0.0, None, []
            worst_l2, worst_l2_n = 0.0, None
            nscale = max(float(p.grad.norm()) for p in m1.parameters() if p.grad is not None)
            for n, p in m1.named_parameters():
                g = grads_n[n]
                if p.grad is None:
                    if g is not None and float(g.abs().max()) != 0.0:
                        missing.append(n)
                    continue
                if g is None:
                    missing.append(n)
                    continue
                err = float((g - p.grad).abs().max()) / max(float(p.grad.abs().max()), 1e-3 * gscale)
                if err > worst:
                    worst, worst_n = err, n
                l2 = float((g.double() - p.grad.double()).norm()) / max(float(p.grad.norm()), 1e-3 * nscale)
                if l2 > worst_l2:
                    worst_l2, worst_l2_n = l2, n
            # fp32: element-wise (max-norm) bound.  bf16: the two runs take different kernel routes (8 rows per rank fall below
            # the tensor-core tile, 64 rows do not), so activations are rounded differently: the bound is the bf16 path's own
            # noise level, relative L2 per tensor (tests/test_headline_gpu.py: 6e-2 against the fp32 path)
            tol_l, tol_g = (1e-5, 1e-3) if dtype == torch.float32 else (2e-3, 6e-2)
            lerr = abs(float(tot) - float(l1)) / abs(float(l1))
            gerr = worst if dtype == torch.float32 else worst_l2
            results[name] = {"loss_n_ranks": float(tot), "loss_1_rank": float(l1), "loss_rel_err": lerr, "worst_grad_max_err": worst,
                             "worst_grad": worst_n, "worst_grad_rel_l2": worst_l2, "worst_grad_l2": worst_l2_n,
                             "params_with_mismatched_presence": missing, "ranks_identical": same,
                             "tol_loss": tol_l, "tol_grad": tol_g, "grad_criterion": "max-norm" if dtype == torch.float32 else "relative L2",
                             "ok": bool(lerr <= tol_l and gerr <= tol_g and not missing and same)}
        dist.barrier()
        red.remove()
    if rank == 0:
        print(json.dumps({"verify_ddp": results, "n_gpus": world, "ok": all(r["ok"] for r in results.values()),
                          "what": "captured step with in-graph NCCL all-reduce on N shards vs 1-rank eager step on the concatenated batch "
                                  "(YouCook2 layer sizes, 2 videos x 4 clips per rank, eval mode)"}))
    import os
    import sys
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(0)
