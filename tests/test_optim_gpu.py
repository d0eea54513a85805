"""GPU parity of the fused optimiser tail (svpc_b200.optim -> svpc_optim_* in libsvpc_b200.so) against the golden fixture
of the unmodified reference classes and the CPU oracle.  Tolerance: 2e-6 of each tensor's largest magnitude on parameters / moments / EMA shadows
(fp32 throughout; the only differences are FMA contraction and the summation order of the gradient norms)."""
import os

import numpy as np
import pytest
import torch

from oracle import optim_oracle as OO
from tests import optim_case as OC

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "optim_iterations.npz")
TOL = 2e-6


def build(with_ema=True):
    from svpc_b200 import optim as SO
    dev = torch.device("cuda:0")
    params = OC.make_params()
    names = list(params)
    ps = [torch.nn.Parameter(params[n].to(dev)) for n in names]
    g = OC.decay_groups(names)
    groups = [{"params": [ps[i] for i in grp["params"]], "weight_decay": grp["weight_decay"]} for grp in g]
    opt = SO.BertAdam(groups, **OC.HYPER)
    ema = None
    if with_ema:
        ema = SO.EMA(OC.EMA_DECAY)
        for n, p in zip(names, ps):
            ema.register(n, p.data)
    return names, ps, opt, ema


def set_grads(names, ps, it):
    g = OC.make_grads(it)
    for n, p in zip(names, ps):
        p.grad = None if g[n] is None else g[n].to(p.device)


def close(a, b, what, atol):
    """max |a - b| <= TOL * max |b| (+ atol): the tolerance is relative to the tensor's scale - an update that nearly
    cancels a parameter leaves an element whose own magnitude is far below the rounding step of the operands."""
    a = a.detach().cpu().numpy().astype(np.float64)
    b = np.asarray(b, np.float64)
    err = np.abs(a - b).max()
    assert err <= TOL * np.abs(b).max() + atol, (what, err, np.abs(b).max())


def test_fused_iteration_matches_reference_golden():
    gold = dict(np.load(GOLD))
    names, ps, opt, ema = build()
    norms = []
    for it in range(OC.N_ITER):
        set_grads(names, ps, it)
        opt.step(grad_clip=OC.GRAD_CLIP, ema=ema, niter=it)
        norms.append(float(opt.total_norm()))
    np.testing.assert_allclose(np.asarray(norms, np.float32), gold["norms"], rtol=1e-6)
    for n, p in zip(names, ps):
        close(p.data, gold["p/" + n], n, 1e-8)
        close(ema.shadow[n], gold["shadow/" + n], n, 1e-8)
        st = opt.state[p]
        if n in OC.NO_GRAD:
            assert len(st) == 0
            assert torch.equal(p.data.cpu(), OC.make_params()[n])
        else:
            close(st["next_m"], gold["m/" + n], n, 1e-12)
            close(st["next_v"], gold["v/" + n], n, 1e-20)
            assert st["step"] == int(gold["step/" + n])
    # reference quirk kept: get_lr() answers [0] as soon as it meets a parameter without state (optimization.py:264-266)
    assert opt.get_lr() == [0]
    live = [p for p in opt.param_groups[1]["params"]]
    opt.param_groups[0]["params"] = [p for p in opt.param_groups[0]["params"] if len(opt.state[p])]
    assert opt.get_lr()[0] == pytest.approx(OC.HYPER["lr"] * (1.0 / 3.0))     # warmup_linear at step 6 of 8, warm-up 0.25
    assert len(opt.get_lr()) == len(live) + len(opt.param_groups[0]["params"])


def test_separate_calls_match_oracle():
    """train.py's three separate calls - torch clip_grad_norm_, optimizer.step(), ema(model, niter) - with the drop-in
    classes (the un-fused use) give the oracle's result, and write_grads reproduces the in-place clipping of p.grad."""
    names, ps, opt, ema = build()
    cpu = OC.make_params()
    o = OO.OptimOracle(list(cpu.values()), OC.decay_groups(names), ema_decay=OC.EMA_DECAY, **OC.HYPER)

    class M:
        @staticmethod
        def named_parameters():
            return list(zip(names, ps))
    for it in range(OC.N_ITER):
        set_grads(names, ps, it)
        g = OC.make_grads(it)
        og = [None if g[n] is None else g[n].clone() for n in names]
        o.iteration(og, it, OC.GRAD_CLIP)
        torch.nn.utils.clip_grad_norm_([p for p in ps], OC.GRAD_CLIP)
        opt.step(write_grads=True)
        ema(M, it)
        for n, p, gg in zip(names, ps, og):
            if gg is not None:
                close(p.grad, gg.numpy(), n + ".grad", 1e-12)
    for i, (n, p) in enumerate(zip(names, ps)):
        close(p.data, o.p[i].numpy(), n, 1e-8)
        close(ema.shadow[n], o.shadow[i].numpy(), n, 1e-8)


def test_step_bumps_parameter_versions():
    """The bf16 weight cache of the model is keyed on the parameters' version counters: a fused step must advance them."""
    names, ps, opt, _ = build(with_ema=False)
    set_grads(names, ps, 0)
    before = [p._version for p in ps]
    opt.step(grad_clip=1.0)
    for n, p, b in zip(names, ps, before):
        assert (p._version > b) == (n not in OC.NO_GRAD), n


def test_no_clip_no_ema_and_zero_lr_properties():
    names, ps, opt, _ = build(with_ema=False)
    for grp in opt.param_groups:
        grp["max_grad_norm"] = -1
        grp["lr"] = 0.0
    before = [p.data.clone() for p in ps]
    set_grads(names, ps, 4)                  # large gradients, no clipping anywhere
    opt.step()
    for n, p, b in zip(names, ps, before):
        assert torch.equal(p.data, b), n      # lr 0: parameters untouched
        if p.grad is not None:
            close(opt.state[p]["next_m"], (0.1 * p.grad).cpu().numpy(), n, 1e-12)   # m = (1-b1) g from zero state


def test_model_scale_against_torch_ops():
    """All parameters of the full-size vivt model (42.5 M elements, 180 tensors): one fused iteration against the same
    arithmetic written with plain torch ops on the device."""
    from svpc_b200 import optim as SO, synth
    from svpc_b200.config import youcook2_config
    from svpc_b200.model import StateAwareRecursiveTransformer
    dev = torch.device("cuda:0")
    cfg = youcook2_config("full")
    model = StateAwareRecursiveTransformer(cfg)
    model.load_state_dict(synth.init_state_dict(cfg, seed=3))
    model.to(dev)
    named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
    names, ps = [n for n, _ in named], [p for _, p in named]
    assert sum(p.numel() for p in ps) > 40e6
    gen = torch.Generator(device=dev).manual_seed(5)
    for p in ps:
        p.grad = torch.randn(p.shape, device=dev, generator=gen) * 1e-3
    nod = ["bias", "LayerNorm.bias", "LayerNorm.weight"]
    groups = [{"params": [p for n, p in zip(names, ps) if not any(x in n for x in nod)], "weight_decay": 0.01},
              {"params": [p for n, p in zip(names, ps) if any(x in n for x in nod)], "weight_decay": 0.0}]
    order = [p for g in groups for p in g["params"]]
    wds = [g["weight_decay"] for g in groups for _ in g["params"]]
    opt = SO.BertAdam(groups, lr=1e-4, warmup=0.1, t_total=1000)
    ema = SO.EMA(0.9999)
    for n, p in zip(names, ps):
        ema.register(n, p.data)
    # plain torch statement, fp32, same order of operations as optimization.py / train.py
    ref_p = [p.data.clone() for p in order]
    ref_g = [p.grad.clone() for p in order]
    ref_s = [p.data.clone() for p in order]
    total = torch.linalg.vector_norm(torch.stack([torch.linalg.vector_norm(g) for g in ref_g]))
    cg = torch.clamp(1.0 / (total + 1e-6), max=1.0)
    for i, g in enumerate(ref_g):
        g.mul_(cg)
        g.mul_(torch.clamp(1.0 / (torch.linalg.vector_norm(g) + 1e-6), max=1.0))
        m = g * (1 - 0.9)
        v = g * g * (1 - 0.999)
        upd = m / (v.sqrt() + 1e-6)
        if wds[i] > 0:
            upd += wds[i] * ref_p[i]
        ref_p[i].add_(-(1e-4 * 0.0) * upd)           # warmup_linear at step 0 is 0: first update leaves p unchanged
        d = min(0.9999, 1.0 / 10.0)
        ref_s[i] = (1.0 - d) * ref_p[i] + d * ref_s[i]
    opt.step(grad_clip=1.0, ema=ema, niter=0)
    assert float(opt.total_norm()) == pytest.approx(float(total), rel=1e-5)
    # second step has a non-zero learning rate
    for i, p in enumerate(order):
        assert torch.equal(p.data, ref_p[i])
    m1 = [opt.state[p]["next_m"].clone() for p in order]
    v1 = [opt.state[p]["next_v"].clone() for p in order]
    opt.step(grad_clip=1.0, ema=ema, niter=1)        # same gradients again
    lr1 = 1e-4 * (1.0 / 1000) / 0.1
    worst = 0.0
    for i, p in enumerate(order):
        g = ref_g[i]                                  # clipped gradient (clip coefficients are the same both steps)
        m = m1[i] * 0.9 + g * (1 - 0.9)
        v = v1[i] * 0.999 + g * g * (1 - 0.999)
        upd = m / (v.sqrt() + 1e-6)
        if wds[i] > 0:
            upd += wds[i] * ref_p[i]
        want = ref_p[i] - lr1 * upd
        err = float((p.data - want).abs().max() / (want.abs().max() + 1e-30))
        worst = max(worst, err)
        d = min(0.9999, 2.0 / 11.0)
        want_s = (1.0 - d) * want + d * ref_s[i]
        name = [n for n, q in zip(names, ps) if q is p][0]
        errs = float((ema.shadow[name] - want_s).abs().max() / (want_s.abs().max() + 1e-30))
        worst = max(worst, errs)
    assert worst < TOL, worst


@pytest.mark.parametrize("graph", [False, True])
def test_weight_cache_follows_the_optimizer(graph):
    """After a fused optimiser step the compute-dtype weight copies (functional.WeightCache, refreshed by ONE multi-tensor
    cast at the start of the next forward, inside the captured graph too) must equal a fresh cast of the new parameters:
    the next loss equals that of a new model loaded from the updated state_dict."""
    import copy
    from svpc_b200 import optim as SO, synth, functional as Fn
    from svpc_b200.config import tiny_config
    from svpc_b200.model import StateAwareRecursiveTransformer
    dev = torch.device("cuda:0")
    cfg = tiny_config("full")
    sd = synth.init_state_dict(cfg, seed=1)
    batch = synth.make_batch(cfg, B=3, S=4, seed=5, ragged=True, n_ingr=4, n_oov=2)
    args = synth.forward_args(batch, dev)
    noise = [g.to(dev) for g in batch["gumbel"]]
    model = StateAwareRecursiveTransformer(cfg, compute_dtype=torch.bfloat16)
    model.load_state_dict(sd)
    model.to(dev).eval()
    opt = SO.BertAdam([p for p in model.parameters() if p.requires_grad], lr=5e-3, warmup=-1, t_total=-1, weight_decay=0.0)
    if graph:
        from svpc_b200.graph import GraphedTrainStep
        step = GraphedTrainStep(model, args, gumbel_noise=noise, warmup=1)
        run = lambda: step.run()
    else:
        def run():
            model.zero_grad(set_to_none=True)
            loss = model(*args, gumbel_noise=noise)[0]
            loss.backward()
            return loss
    l0 = float(run())
    opt.step(grad_clip=1.0)
    l1 = float(run())
    assert abs(l1 - l0) > 1e-4 * abs(l0)                         # the update is visible in the next step
    fresh = StateAwareRecursiveTransformer(cfg, compute_dtype=torch.bfloat16)
    fresh.load_state_dict(copy.deepcopy(model.state_dict()))
    fresh.to(dev).eval()
    Fn.WEIGHTS.clear()
    with torch.no_grad():
        want = float(fresh(*args, gumbel_noise=noise)[0])
    assert abs(l1 - want) <= 1e-6 * abs(want), (l0, l1, want)
