"""CPU tests of the optimiser row: the oracle restatement against the golden fixture produced by the
UNMODIFIED reference classes (tests/golden/make_optim_golden.py) and, where /root/reference exists, against the live
classes; the host-side schedule / constructor logic of svpc_b200.optim against the reference's."""
import os

import numpy as np
import pytest
import torch

from oracle import optim_oracle as OO
from svpc_b200 import optim as SO
from tests import optim_case as OC

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "optim_iterations.npz")
HAVE_REF = os.path.isfile("/root/reference/src/rtransformer/optimization.py")


def run_oracle(with_ema=True, grad_clip=OC.GRAD_CLIP):
    params = OC.make_params()
    names = list(params)
    o = OO.OptimOracle(list(params.values()), OC.decay_groups(names), ema_decay=OC.EMA_DECAY if with_ema else None, **OC.HYPER)
    norms = []
    for it in range(OC.N_ITER):
        g = OC.make_grads(it)
        norms.append(o.iteration([None if g[n] is None else g[n].clone() for n in names], it, grad_clip))
    return names, o, norms


def check_against(gold, names, o, norms, tol):
    np.testing.assert_allclose(np.asarray(norms, np.float32), gold["norms"], rtol=1e-6)
    for i, n in enumerate(names):
        np.testing.assert_allclose(o.p[i].numpy(), gold["p/" + n], rtol=tol, atol=tol * 1e-2, err_msg=n)
        if n in OC.NO_GRAD:
            assert "m/" + n not in gold and o.m[i] is None
            np.testing.assert_array_equal(o.p[i].numpy(), OC.make_params()[n].numpy())
        else:
            np.testing.assert_allclose(o.m[i].numpy(), gold["m/" + n], rtol=tol, atol=1e-12, err_msg=n)
            np.testing.assert_allclose(o.v[i].numpy(), gold["v/" + n], rtol=tol, atol=1e-20, err_msg=n)
            assert int(gold["step/" + n]) == o.step[i] == OC.N_ITER
        np.testing.assert_allclose(o.shadow[i].numpy(), gold["shadow/" + n], rtol=tol, atol=tol * 1e-2, err_msg=n)


def test_oracle_matches_reference_golden():
    names, o, norms = run_oracle()
    check_against(dict(np.load(GOLD)), names, o, norms, tol=1e-6)


@pytest.mark.skipif(not HAVE_REF, reason="reference tree not present on this box")
def test_oracle_matches_live_reference():
    from tests.golden.make_optim_golden import run_reference
    live = run_reference()
    gold = dict(np.load(GOLD))
    for k in gold:
        np.testing.assert_array_equal(live[k], gold[k], err_msg=k)          # the committed fixture is what the reference gives
    names, o, norms = run_oracle()
    check_against(live, names, o, norms, tol=1e-6)


@pytest.mark.skipif(not HAVE_REF, reason="reference tree not present on this box")
@pytest.mark.parametrize("name", ["warmup_linear", "warmup_constant", "warmup_cosine", "none", None])
def test_schedules_match_live_reference(name):
    from tests.golden.make_optim_golden import load_reference_optimization
    ref = load_reference_optimization()
    for warmup, t_total in [(0.1, 100), (0.25, 8), (-1, 50), (0.1, -1), (0.0, 10)]:
        r = ref.SCHEDULES[name](warmup=warmup, t_total=t_total)
        s = SO.SCHEDULES[name](warmup=warmup, t_total=t_total)
        for step in range(0, 130, 3):
            a = r.get_lr(step, nowarn=True)
            assert s.get_lr(step) == a, (name, warmup, t_total, step)
            assert OO.schedule_multiplier(name, step, warmup, t_total) == a


def test_schedule_known_answers():
    """warmup_linear (optimization.py:165-168): ramp to 1 over the warm-up fraction, then linearly to 0 at t_total."""
    s = SO.WarmupLinearSchedule(warmup=0.25, t_total=8)
    assert [s.get_lr(i) for i in range(10)] == [0.0, 0.5, 1.0, 0.8333333333333334, 0.6666666666666666, 0.5,
                                                 0.3333333333333333, 0.16666666666666666, 0.0, 0.0]
    assert SO.WarmupLinearSchedule(warmup=0.1, t_total=-1).get_lr(5) == 1.0


def test_constructor_errors_mirror_reference():
    p = [torch.nn.Parameter(torch.zeros(3))]
    for kw in (dict(lr=-1.0), dict(lr=1e-3, schedule="bogus"), dict(lr=1e-3, b1=1.0), dict(lr=1e-3, b2=-0.1), dict(lr=1e-3, e=-1.0),
               dict(lr=1e-3, warmup=1.5)):
        with pytest.raises(ValueError):
            SO.BertAdam(p, **kw)
    opt = SO.BertAdam([{"params": p, "weight_decay": 0.0}], lr=1e-3, warmup=0.1, t_total=10)
    assert opt.get_lr() == [0]                                  # no state yet (optimization.py:264-266)
    assert opt.defaults["max_grad_norm"] == 1.0 and opt.defaults["e"] == 1e-6 and opt.defaults["weight_decay"] == 0.01


def test_no_cpu_path():
    from svpc_b200.cuda_backend import SvpcError
    p = torch.nn.Parameter(torch.zeros(3))
    p.grad = torch.ones(3)
    opt = SO.BertAdam([p], lr=1e-3)
    with pytest.raises(SvpcError):
        opt.step()
