"""Data-parallel path on CPU (gloo, world size 2): each rank runs the step on its own videos, GradAllReducer SUMS the
gradients with bucketed all-reduces launched from autograd hooks, and the result must equal a single-process run on the
concatenated batch (the reference loss is a sum over videos, model.py:1110-1115).  Host logic only: the operators are
the plain-torch statements of tests/torch_backend.py."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from svpc_b200 import backend, synth
from svpc_b200.config import tiny_config
from svpc_b200.ddp import GradAllReducer
from svpc_b200.model import StateAwareRecursiveTransformer
from svpc_b200.synth import forward_args
from tests.torch_backend import TorchBackend


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _slice_batch(batch, vids):
    """Sub-batch holding the videos `vids` of a collated batch (step-major tensors are indexed along dim 0)."""
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


def _grads(model, batch):
    model.zero_grad(set_to_none=True)
    loss = model(*forward_args(batch), gumbel_noise=batch["gumbel"])[0]
    loss.backward()
    return loss.detach(), {n: (None if p.grad is None else p.grad.clone()) for n, p in model.named_parameters()}


def _worker(rank, world, port, mode, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        backend.set_backend(TorchBackend())
        cfg = tiny_config(mode)
        sd = synth.init_state_dict(cfg, seed=1)
        full = synth.make_batch(cfg, B=4, S=3, seed=5, ragged=True, n_ingr=4, n_oov=1)
        m = StateAwareRecursiveTransformer(cfg)
        m.load_state_dict(sd)
        m.eval()
        red = GradAllReducer(m, bucket_bytes=1 << 16)
        assert red.enabled and len(red.buckets) > 1
        mine = _slice_batch(full, [b for b in range(4) if b % world == rank])
        # two iterations: the first one learns which parameters get no gradient in this mode; in the second every bucket
        # with a live parameter must have been launched from the hooks, before finish()
        for it in range(2):
            m.zero_grad(set_to_none=True)
            loss = m(*forward_args(mine), gumbel_noise=mine["gumbel"])[0]
            loss.backward()
            if it == 1:
                assert red._absent is not None and all(l or n == 0 for l, n in zip(red._launched, red._need)), \
                    (red._launched, red._need)
                if mode == "video":
                    assert len(red._absent) > 0
            red.finish()
            assert all(p.grad is None for p in red._absent)
        if red._absent:
            # a gradient for a parameter that was learned to be dead must not be dropped silently
            dead = next(iter(red._absent))
            try:
                red._on_grad(dead)
                raise AssertionError("expected a RuntimeError for a gradient on a dead parameter")
            except RuntimeError as e:
                assert "reset_absent" in str(e)
            red.reset_absent()
            assert red._absent is None and red._need == [len(b) for b in red.buckets]
        tot = loss.detach().clone()
        dist.all_reduce(tot)
        if rank == 0:
            torch.save((float(tot), {n: (None if p.grad is None else p.grad.clone()) for n, p in m.named_parameters()}), out_path)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ("video", "full"))
def test_two_ranks_equal_one_process(mode, tmp_path):
    ctx = mp.get_context("spawn")
    out_path = str(tmp_path / "rank0.pt")
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mode, out_path)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=300)
        assert p.exitcode == 0
    loss2, grads2 = torch.load(out_path)
    prev = backend._impl
    backend.set_backend(TorchBackend())
    try:
        cfg = tiny_config(mode)
        sd = synth.init_state_dict(cfg, seed=1)
        full = synth.make_batch(cfg, B=4, S=3, seed=5, ragged=True, n_ingr=4, n_oov=1)
        m = StateAwareRecursiveTransformer(cfg)
        m.load_state_dict(sd)
        m.eval()
        loss1, grads1 = _grads(m, full)
    finally:
        backend.set_backend(prev)
    assert abs(loss2 - float(loss1)) <= 1e-5 * abs(float(loss1))
    gscale = max(float(g.abs().max()) for g in grads1.values() if g is not None)
    for n, g1 in grads1.items():
        g2 = grads2[n]
        if g1 is None:
            assert g2 is None, n            # parameters without a gradient stay None (BertAdam skips them)
            continue
        assert g2 is not None, n
        assert float((g1 - g2).abs().max()) <= 1e-4 * max(float(g1.abs().max()), 1e-3 * gscale), n
