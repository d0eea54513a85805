"""Checkpoint layout: files written by svpc_b200.checkpoint load into the UNMODIFIED reference model
and reference-written files load into the drop-in; EMA assign / resume keeps the reference's call-order contract
(train.py:383, 393-406)."""
import copy

import pytest
import torch

from oracle import refshim
from svpc_b200 import checkpoint as CK
from svpc_b200.model import StateAwareRecursiveTransformer
from tests.helpers import load_case


@pytest.mark.parametrize("mode", ["video", "full"])
def test_round_trip_keeps_layout_and_values(tmp_path, mode):
    cfg, sd, _, _ = load_case("ragged", mode)
    m = StateAwareRecursiveTransformer(cfg)
    m.load_state_dict(sd)
    path = str(tmp_path / "m.chkpt")
    ck = CK.save_checkpoint(path, m, opt={"lr": 1e-4}, epoch=3)
    assert sorted(ck) == ["epoch", "model", "model_cfg", "opt"]                      # train.py:402-406
    ck2, m2 = CK.load_checkpoint(path)
    assert ck2["epoch"] == 3 and ck2["opt"] == {"lr": 1e-4} and not m2.training
    assert sorted(ck2["model"].keys()) == sorted(sd.keys())
    for k, v in m2.state_dict().items():
        assert torch.equal(v, sd[k]), k
    # the snapshot does not alias live parameters (an in-place EMA.resume must not change a checkpoint taken before it)
    with torch.no_grad():
        for p in m.parameters():
            p.add_(1.0)
    for k, v in ck["model"].items():
        assert torch.equal(v, sd[k]), k


@pytest.mark.skipif(not refshim.available(), reason="reference tree not mounted (GPU box)")
def test_interchange_with_the_reference_model(tmp_path):
    cfg, sd, _, _ = load_case("ragged", "full")
    m = StateAwareRecursiveTransformer(cfg)
    m.load_state_dict(sd)
    path = str(tmp_path / "ours.chkpt")
    CK.save_checkpoint(path, m, epoch=1)
    ref = refshim.build_model(cfg, sd)                    # reference model in its post-set_pretrained_embedding layout
    with torch.no_grad():
        for p in ref.parameters():
            p.zero_()
    ck = torch.load(path, map_location="cpu", weights_only=False)
    res = ref.load_state_dict(ck["model"])                # translator.py:38, strict
    assert not res.missing_keys and not res.unexpected_keys
    for k, v in ref.state_dict().items():
        assert torch.equal(v, sd[k]), k
    # the other way: a checkpoint dict assembled the reference's way (train.py:402-406)
    path2 = str(tmp_path / "ref.chkpt")
    # (model_cfg as a plain dict: the EasyDict stand-in of oracle/refshim.py is a local class and cannot be pickled)
    torch.save({"model": ref.state_dict(), "model_cfg": dict(ref.config), "opt": None, "epoch": 0}, path2)
    _, m3 = CK.load_checkpoint(path2)
    for k, v in m3.state_dict().items():
        assert torch.equal(v, sd[k]), k


@pytest.mark.gpu
def test_ema_assign_resume_in_place():
    """Values move, addresses stay: after assign the parameters hold the shadows, a checkpoint taken then survives resume,
    after resume the parameters are bit-identical to what they were; shadows are never modified (optimization.py:205-216)."""
    from svpc_b200 import optim as SO
    dev = torch.device("cuda:0")
    cfg, sd, _, _ = load_case("ragged", "full")
    m = StateAwareRecursiveTransformer(cfg)
    m.load_state_dict(sd)
    m.to(dev)
    ema = SO.EMA(0.9999)
    g = torch.Generator(device="cpu").manual_seed(3)
    for n, p in m.named_parameters():
        if p.requires_grad:
            ema.register(n, p.data)
            ema.shadow[n].add_(torch.randn(p.shape, generator=g).to(dev))
    before = {n: p.detach().clone() for n, p in m.named_parameters()}
    shadows = {n: s.clone() for n, s in ema.shadow.items()}
    ptrs = {n: p.data_ptr() for n, p in m.named_parameters()}
    vers = {n: p._version for n, p in m.named_parameters()}
    ema.assign(m)
    ck = CK.make_checkpoint(m, epoch=0)
    for n, p in m.named_parameters():
        assert p.data_ptr() == ptrs[n]
        if p.requires_grad:
            assert torch.equal(p.detach(), shadows[n]), n
            assert p._version > vers[n]
            assert torch.equal(ema.original[n], before[n]), n
    ema.resume(m)
    for n, p in m.named_parameters():
        assert p.data_ptr() == ptrs[n] and torch.equal(p.detach(), before[n]), n
        if p.requires_grad:
            assert torch.equal(ema.shadow[n], shadows[n]), n
            assert torch.equal(ck["model"][n], shadows[n].cpu()), n
    ema.assign(m)                                             # second window re-uses the buffers and the tables
    ema.resume(m)
    for n, p in m.named_parameters():
        assert torch.equal(p.detach(), before[n]), n
