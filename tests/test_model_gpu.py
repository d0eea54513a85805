"""GPU parity tests of the whole step through the drop-in model (svpc_b200.model -> ctypes -> libsvpc_b200.so):
against the committed reference golden fixtures (tiny configuration, all four model modes, ragged + OOV and
full-length batches) and against the CPU oracle at the real YouCook2 layer sizes."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu

from svpc_b200 import backend  # noqa: E402
from svpc_b200.config import youcook2_config  # noqa: E402
from svpc_b200 import synth  # noqa: E402
from svpc_b200.model import StateAwareRecursiveTransformer  # noqa: E402
from svpc_b200.synth import forward_args  # noqa: E402
from tests.helpers import CASES, MODES, load_case, check_against_golden  # noqa: E402

DEV = torch.device("cuda")


@pytest.fixture(autouse=True)
def _cuda_backend():
    backend.set_backend(None)          # product default: the CUDA library
    yield
    backend.set_backend(None)


def build(cfg, sd, dtype=torch.float32):
    m = StateAwareRecursiveTransformer(cfg, compute_dtype=dtype)
    m.load_state_dict(sd, strict=True)
    return m.to(DEV).eval()


def run(m, batch):
    loss, preds, ents, acts = m(*forward_args(batch, DEV), gumbel_noise=batch["gumbel"])
    loss.backward()
    torch.cuda.synchronize()
    return loss.detach().cpu(), preds, ents, acts, {n: p.grad for n, p in m.named_parameters()}


@pytest.mark.parametrize("case", list(CASES))
@pytest.mark.parametrize("mode", MODES)
def test_fp32_matches_reference_golden(case, mode):
    cfg, sd, batch, gold = load_case(case, mode)
    m = build(cfg, sd)
    assert backend.get().name == "cuda-sm100a"
    loss, preds, ents, acts, grads = run(m, batch)
    # north_star: fp32-accumulate path within 1e-4 relative on loss and prediction scores
    check_against_golden(gold, loss, preds, ents, acts, grads, loss_tol=1e-5, out_tol=1e-4, grad_tol=1e-3, what=f"{case}/{mode}")


@pytest.mark.parametrize("mode", MODES)
def test_bf16_loss_within_1e2(mode):
    cfg, sd, batch, gold = load_case("ragged", mode)
    m = build(cfg, sd, torch.bfloat16)
    loss, preds, ents, acts, grads = run(m, batch)
    assert abs(float(loss) - float(gold["loss"])) <= 1e-2 * abs(float(gold["loss"])), (float(loss), float(gold["loss"]))
    for g in grads.values():
        assert g is None or torch.isfinite(g).all()


@pytest.mark.parametrize("mode", ("video", "full"))
def test_fp32_matches_oracle_at_youcook2_sizes(mode):
    from oracle import svpc_oracle as O
    cfg = youcook2_config(mode)
    sd = synth.init_state_dict(cfg, seed=3)
    batch = synth.make_batch(cfg, B=2, S=3, seed=11, ragged=True, n_ingr=6, n_oov=2)
    ref_loss, parts, ref_grads, ref_preds, ref_ents, ref_acts = O.loss_and_grads(sd, cfg, copy.deepcopy(batch))
    m = build(cfg, sd)
    loss, preds, ents, acts, grads = run(m, batch)
    assert abs(float(loss) - float(ref_loss)) <= 1e-4 * abs(float(ref_loss)), (float(loss), float(ref_loss))
    for a, b in zip(preds, ref_preds):
        err = (a.detach().cpu() - b.detach()).abs().max().item()
        assert err <= 1e-4 * b.abs().max().item(), err
    gscale = max(float(g.abs().max()) for g in ref_grads.values() if g is not None)
    for n, rg in ref_grads.items():
        if rg is None:
            continue
        g = grads[n]
        assert g is not None, n
        err = (g.cpu() - rg).abs().max().item()
        assert err <= 2e-3 * max(float(rg.abs().max()), 1e-3 * gscale), (n, err)


@pytest.mark.parametrize("mode", ("full",))
def test_bf16_tcgen05_at_youcook2_sizes(mode):
    """bf16 operands on the tcgen05 route at real layer sizes: loss within 1e-2 relative of the fp32 oracle."""
    from oracle import svpc_oracle as O
    cfg = youcook2_config(mode)
    sd = synth.init_state_dict(cfg, seed=3)
    batch = synth.make_batch(cfg, B=2, S=3, seed=11, ragged=True, n_ingr=6, n_oov=2)
    ref_loss = O.forward(sd, cfg, copy.deepcopy(batch))[0]
    m = build(cfg, sd, torch.bfloat16)
    loss, preds, ents, acts, grads = run(m, batch)
    assert abs(float(loss) - float(ref_loss)) <= 1e-2 * abs(float(ref_loss)), (float(loss), float(ref_loss))


def test_no_cpu_path():
    from svpc_b200.cuda_backend import SvpcError
    be = backend.get()
    with pytest.raises(SvpcError):
        be.colsum(torch.zeros(4, 4))


@pytest.mark.parametrize("mode", ("video", "full"))
@pytest.mark.parametrize("dtype", (torch.float32, torch.bfloat16))
def test_graph_replay_matches_eager(mode, dtype):
    """CUDA-graph replay of fwd+bwd gives the eager result, also after new inputs are loaded into the static buffers."""
    from svpc_b200.graph import GraphedTrainStep
    cfg, sd, batch, gold = load_case("ragged", mode)
    m = build(cfg, sd, dtype)

    def eager():                         # no reference to the eager autograd graph may survive into the capture
        loss, _p, _e, _a, grads = run(m, batch)
        return float(loss), {n: (None if g is None else g.detach().clone()) for n, g in grads.items()}
    loss, grads = eager()
    m.zero_grad(set_to_none=True)
    import gc
    gc.collect()
    # capture on a DIFFERENT batch of the same signature, then load the real one
    other = synth.make_batch(cfg, **{**CASES["ragged"][1], "seed": 99})
    if m.plan_signature(other["batch_step_num"], other["ingr_id_dict"], other["extra_zeros"]) != \
            m.plan_signature(batch["batch_step_num"], batch["ingr_id_dict"], batch["extra_zeros"]):
        other = batch
    step = GraphedTrainStep(m, forward_args(other, DEV), gumbel_noise=[g.to(DEV) for g in other["gumbel"]])
    args = forward_args(batch)           # host tensors
    assert step.matches(args)
    step.load(args, batch["gumbel"])
    for _ in range(2):
        gl = step.run()
    torch.cuda.synchronize()
    tol = 1e-6 if dtype == torch.float32 else 1e-5
    assert abs(float(gl) - float(loss)) <= tol * abs(float(loss)), (float(gl), float(loss))
    gscale = max(float(g.abs().max()) for g in grads.values() if g is not None)
    for n, p in m.named_parameters():
        if grads[n] is None:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, n
        else:          # atomics (split-K, scatter) reorder fp32 sums: compare against the tensor's / the global scale
            err = float((p.grad - grads[n]).abs().max())
            assert err <= 1e-4 * max(float(grads[n].abs().max()), 1e-3 * gscale), (n, err)


def test_prefetch_pipeline_matches_load():
    """The host->device input pipeline (copy stream + staging set, GraphedTrainStep.prefetch / run_prefetched) feeds the
    captured step the same inputs as a synchronous load(): alternating two different batches gives each one's own loss."""
    from svpc_b200.graph import GraphedTrainStep
    # (the full-length case: two ragged batches of different seeds never share a plan signature, and the test would skip)
    cfg, sd, batch, gold = load_case("full_len", "full")
    m = build(cfg, sd, torch.float32)
    other = synth.make_batch(cfg, **{**CASES["full_len"][1], "seed": 99})
    if m.plan_signature(other["batch_step_num"], other["ingr_id_dict"], other["extra_zeros"]) != \
            m.plan_signature(batch["batch_step_num"], batch["ingr_id_dict"], batch["extra_zeros"]):
        pytest.skip("no second batch with the same signature")
    step = GraphedTrainStep(m, forward_args(other, DEV), gumbel_noise=[g.to(DEV) for g in other["gumbel"]])
    pin = lambda x: x.pin_memory() if isinstance(x, torch.Tensor) else ([t.pin_memory() for t in x] if isinstance(x, list) and x and isinstance(x[0], torch.Tensor) else x)
    hosts = [([pin(a) for a in forward_args(b)], [g.pin_memory() for g in b["gumbel"]]) for b in (batch, other)]
    want = []
    for a, g in hosts:
        step.load(a, g)
        want.append(float(step.run()))
    assert abs(want[0] - want[1]) > 1e-3 * abs(want[0])          # the two batches are distinguishable
    got = []
    step.prefetch(*hosts[0])
    for i in range(6):
        loss = step.run_prefetched()
        if i + 1 < 6:
            step.prefetch(*hosts[(i + 1) % 2])                   # upload of the next batch overlaps this replay
        got.append(float(loss))
    for i, v in enumerate(got):
        assert abs(v - want[i % 2]) <= 1e-6 * abs(want[i % 2]), (i, v, want)
    # arena path: the batch is collated into the pinned host views -> one H2D + one D2D memcpy per step
    a_args, a_noise = step.host_inputs()
    for which in (0, 1, 0):
        src_args, src_noise = hosts[which]
        for dst, src in zip(a_args, src_args):
            if isinstance(dst, torch.Tensor):
                dst.copy_(src)
            elif isinstance(dst, list) and dst and isinstance(dst[0], torch.Tensor):
                for d, s in zip(dst, src):
                    d.copy_(s)
        for d, s in zip(a_noise, src_noise):
            d.copy_(s)
        a_args[9] = src_args[9]
        for f in a_args[1]:
            f[:, cfg.max_v_len:] = float("nan")      # text rows of the feature blocks: never uploaded, never read
        assert step.arena.h2d_nbytes < step.arena.nbytes
        assert step.arena.zc_table is not None       # feature blocks: uploaded into alternating buffers, table switched
        before = step.arena.cur
        step.prefetch(a_args, a_noise)
        v = float(step.run_prefetched())
        assert step.arena.cur == 1 - before
        assert abs(v - want[which]) <= 1e-6 * abs(want[which]), (which, v, want)
    # a synchronous load() after the pipeline writes the static arena: the table must point back at it
    step.load(*hosts[1])
    assert abs(float(step.run()) - want[1]) <= 1e-6 * abs(want[1])
    assert step.arena.cur == 0


def test_graph_survives_optimizer_step_and_decode():
    """replay -> optimiser step -> greedy decode -> replay: the decode sees stale cached weight copies (the optimiser
    bumped the parameter versions); they must be refreshed IN PLACE, because the captured graph reads and rewrites them by
    address.  The second replay must equal an eager step on the updated weights."""
    from svpc_b200.graph import GraphedTrainStep
    from svpc_b200.decode import greedy_decode
    from svpc_b200 import optim as SO
    from svpc_b200 import functional as Fn
    from tests.test_decode_host import decode_inputs
    cfg, sd, batch, gold = load_case("ragged", "full")
    m = build(cfg, sd, torch.bfloat16)
    noise = [g.to(DEV) for g in batch["gumbel"]]
    step = GraphedTrainStep(m, forward_args(batch, DEV), gumbel_noise=noise)
    l0 = float(step.run())
    mine = {id(p) for p in m.parameters()}          # (the process-wide cache may still list models of earlier tests)
    ptrs0 = {k: e[1].data_ptr() for k, e in Fn.WEIGHTS.store.items() if set(k[:-1]) <= mine}
    assert len(ptrs0) >= 16
    opt = SO.BertAdam([p for p in m.parameters() if p.requires_grad], lr=5e-3, warmup=-1, t_total=-1, schedule="warmup_constant")
    opt.step()
    with torch.no_grad():
        greedy_decode(m, decode_inputs(batch, DEV))            # stale hits on the decoder / LM-head weights
    for k, ptr in ptrs0.items():          # (the decode adds entries of its own: single-projection copies)
        assert Fn.WEIGHTS.store[k][1].data_ptr() == ptr, "a cached weight copy was re-allocated"
    # a second model moved to the device must not free the first model's cached copies either
    m2 = build(cfg, sd, torch.bfloat16)
    float(m2(*forward_args(batch, DEV), gumbel_noise=noise)[0])
    l1 = float(step.run())
    torch.cuda.synchronize()
    g1 = {n: p.grad.detach().clone() for n, p in m.named_parameters() if p.grad is not None}
    assert abs(l1 - l0) > 1e-4 * abs(l0), "the optimiser step did not change the loss"
    m.zero_grad(set_to_none=True)
    le = m(*forward_args(batch, DEV), gumbel_noise=noise)[0]
    le.backward()
    torch.cuda.synchronize()
    assert abs(l1 - float(le)) <= 1e-5 * abs(float(le)), (l1, float(le))
    gscale = max(float(g.abs().max()) for g in g1.values())
    for n, p in m.named_parameters():
        if n in g1:
            err = float((p.grad - g1[n]).abs().max())
            assert err <= 1e-4 * max(float(g1[n].abs().max()), 1e-3 * gscale), (n, err)


def test_token_csr_updates_do_not_race_the_upload():
    """Several different ingr_id_dicts in flight without any host synchronisation: every replay must scatter the copy mass
    with ITS batch's token ids (StepPlan.update_tokens writes a ring of pinned buffers guarded by events)."""
    from svpc_b200.graph import GraphedTrainStep
    cfg, sd, batch, gold = load_case("full_len", "full")
    m = build(cfg, sd, torch.float32)
    step = GraphedTrainStep(m, forward_args(batch, DEV), gumbel_noise=[g.to(DEV) for g in batch["gumbel"]])
    V = cfg.vocab_size
    dicts = []
    for k in range(6):                        # same lengths, different token ids
        dicts.append([{i: [7 + (t + 3 * k + i) % (V - 7) for t in ids] for i, ids in d.items()} for d in batch["ingr_id_dict"]])
    args = forward_args(batch, DEV)
    want = []
    for d in dicts:                           # synchronous statement
        a = list(args)
        a[9] = d
        step.load(a, batch["gumbel"])
        want.append(float(step.run()))
        torch.cuda.synchronize()
    assert len({round(w, 3) for w in want}) > 1
    losses = []
    for d in dicts:                           # asynchronous: the host runs ahead of the device
        a = list(args)
        a[9] = d
        step.load(a, batch["gumbel"])
        losses.append(step.run().detach().clone())
    torch.cuda.synchronize()
    for w, l in zip(want, losses):
        assert abs(float(l) - w) <= 1e-6 * abs(w), (float(l), w)


def test_bf16_features_in_the_arena():
    """The bf16 path takes the clip features as bf16 (half the upload): same loss as with fp32 features within 2e-3, through
    the eager call and through a captured step whose arena holds bf16 feature blocks; `load()` converts fp32 host features."""
    from svpc_b200.graph import GraphedTrainStep
    cfg = youcook2_config("full")
    sd = synth.init_state_dict(cfg, seed=3)
    batch = synth.make_batch(cfg, B=2, S=3, seed=11, ragged=True, n_ingr=6, n_oov=2)
    m = build(cfg, sd, torch.bfloat16)
    noise = [g.to(DEV) for g in batch["gumbel"]]
    args32 = forward_args(batch, DEV)
    l32 = float(m(*args32, gumbel_noise=noise)[0])
    args16 = list(args32)
    args16[1] = [t.to(torch.bfloat16) for t in args32[1]]
    l16 = float(m(*args16, gumbel_noise=noise)[0])
    assert abs(l16 - l32) <= 2e-3 * abs(l32), (l16, l32)
    step = GraphedTrainStep(m, args16, gumbel_noise=noise)
    assert step.feature_dtype == torch.bfloat16
    lg = float(step.run())
    assert abs(lg - l16) <= 1e-5 * abs(l16), (lg, l16)
    step.load(forward_args(batch), batch["gumbel"])          # fp32 host features: converted on the way in
    lg2 = float(step.run())
    assert abs(lg2 - l16) <= 1e-5 * abs(l16), (lg2, l16)


def test_graph_cache_for_ragged_batches():
    """A stream of ragged batches with three signatures: each is captured the second time it appears and replayed from
    then on; every step - eager, capturing, replayed - gives the eager loss of its batch."""
    from svpc_b200.graph import GraphedStepCache
    cfg, sd, _, _ = load_case("ragged", "full")
    m = build(cfg, sd, torch.float32)
    batches = [synth.make_batch(cfg, B=3, S=4, seed=40 + i, ragged=True, n_ingr=4, n_oov=2) for i in range(3)]
    want = []
    for b in batches:
        want.append(float(m(*forward_args(b, DEV), gumbel_noise=[g.to(DEV) for g in b["gumbel"]])[0]))
    sigs = {m.plan_signature(b["batch_step_num"], b["ingr_id_dict"], b["extra_zeros"]) for b in batches}
    cache = GraphedStepCache(m, max_graphs=3, capture_after=2)
    order = [0, 1, 2, 0, 1, 2, 0, 1, 2, 2, 0]
    for i in order:
        b = batches[i]
        loss = cache.step(forward_args(b, DEV), [g.to(DEV) for g in b["gumbel"]])
        assert abs(float(loss) - want[i]) <= 1e-5 * abs(want[i]), (i, float(loss), want[i])
    st = cache.stats()
    assert st["steps"] == len(order) and st["signatures_seen"] == len(sigs)
    if len(sigs) == 3:
        assert st["captures"] == 3 and st["graph_hits"] == len(order) - 6 and st["graphs_resident"] == 3, st
    small = GraphedStepCache(m, max_graphs=1, capture_after=1)          # least-recently-used eviction
    for i in (0, 1, 1, 0):
        b = batches[i]
        loss = small.step(forward_args(b, DEV), [g.to(DEV) for g in b["gumbel"]])
        assert abs(float(loss) - want[i]) <= 1e-5 * abs(want[i]), (i, float(loss), want[i])
    assert small.stats()["graphs_resident"] == 1
