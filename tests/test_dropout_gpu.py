"""Train-mode dropout (reference nn.Dropout sites model.py:214, 231, 287, 495, 522, 550, 757): every kernel recomputes a
counter-based mask keep(seed, site, element index) instead of storing it.  The contract is stated by
``svpc_dropout_mask`` (scale 0 or 1/(1-p) per element index); these tests check that forward and backward kernels apply
exactly that mask at the documented element indices, and that the model trains with it."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from svpc_b200 import backend  # noqa: E402
from svpc_b200.cuda_backend import CudaBackend  # noqa: E402
from svpc_b200.model import StateAwareRecursiveTransformer  # noqa: E402
from svpc_b200.synth import forward_args  # noqa: E402
from tests.helpers import MODES, load_case  # noqa: E402

DEV = torch.device("cuda")
F32, BF16 = torch.float32, torch.bfloat16
I32 = lambda x: torch.tensor(x, dtype=torch.int32, device=DEV)


@pytest.fixture(scope="module")
def cb():
    return CudaBackend()


def mkdrop(site=3, p=0.1, seed=1234):
    return (torch.tensor([seed], dtype=torch.int32, device=DEV), site, p)


def rnd(*shape, dtype=F32, seed=0):
    g = torch.Generator(device="cpu")
    g.manual_seed(seed + sum(shape))
    return torch.randn(*shape, generator=g).to(DEV).to(dtype)


def close(a, b, tol, name=""):
    a, b = a.double(), b.double()
    err = (a - b).abs().max().item()
    assert err <= tol * max(b.abs().max().item(), 1e-6), (name, err)


def test_mask_statistics(cb):
    for p in (0.1, 0.4):
        m = cb.dropout_mask(1 << 20, mkdrop(p=p))
        keep = (m > 0).float().mean().item()
        assert abs(keep - (1 - p)) < 5e-3, (p, keep)
        assert torch.allclose(m[m > 0], torch.tensor(1 / (1 - p), device=DEV))
    a, b, c = cb.dropout_mask(4096, mkdrop(site=1)), cb.dropout_mask(4096, mkdrop(site=2)), cb.dropout_mask(4096, mkdrop(site=1, seed=99))
    assert not torch.equal(a, b) and not torch.equal(a, c)
    assert torch.equal(a, cb.dropout_mask(4096, mkdrop(site=1)))


@pytest.mark.parametrize("dt,M,N,K", [(BF16, 256, 192, 128), (F32, 70, 50, 33), (BF16, 100, 300, 64)])
def test_gemm_epilogue_dropout(cb, dt, M, N, K):
    a, w, bias, res = rnd(M, K, dtype=dt), rnd(N, K, dtype=dt, seed=1), rnd(N, seed=2), rnd(M, N, dtype=dt, seed=3)
    d = mkdrop()
    mask = cb.dropout_mask(M * N, d).view(M, N)
    for act in (0, 1):
        y0 = cb.gemm(a, w, F32, bias, act)
        y = cb.gemm(a, w, dt, bias, act, residual=res, drop=d)
        close(y, y0 * mask + res.float(), 2e-2 if dt == BF16 else 1e-5, f"gemm act={act}")
    # backward: dz = dy * mask * relu'(out)
    out = cb.gemm(a, w, dt, bias, 1, drop=d)
    dy = rnd(M, N, dtype=dt, seed=4)
    dz, dbias = cb.act_bwd(dy, out, 1, drop=d)
    ref = dy.float() * mask * (out.float() > 0)
    close(dz, ref, 2e-2 if dt == BF16 else 1e-6, "act_bwd dz")
    close(dbias, ref.sum(0), 2e-2, "act_bwd dbias")


def test_embed_ln_dropout(cb):
    V, W, rows = 50, 300, 77
    table, gamma, beta = rnd(V, W), rnd(W, seed=1), rnd(W, seed=2)
    ids = torch.randint(0, V, (rows,), device=DEV, dtype=torch.int32)
    d = mkdrop(site=7)
    mask = cb.dropout_mask(rows * W, d).view(rows, W)
    y0, mean, rstd = cb.embed_ln_fwd(table, ids, gamma, beta, 1e-12, F32)
    y, _, _ = cb.embed_ln_fwd(table, ids, gamma, beta, 1e-12, F32, drop=d)
    close(y, y0 * mask, 1e-6, "embed_ln fwd")
    dy = rnd(rows, W, seed=5)
    got = cb.embed_ln_bwd(dy, table, ids, mean, rstd, gamma, 1e-12, 0, drop=d)
    ref = cb.embed_ln_bwd((dy * mask).contiguous(), table, ids, mean, rstd, gamma, 1e-12, 0)
    for g, r, n in zip(got, ref, ("dtable", "dgamma", "dbeta")):
        close(g, r, 1e-4, n)


def attn_ref(q, k, v, ql, kl, km, causal, H, mask_fn):
    """fp64 statement with an explicit dropout mask; returns ctx and a function giving (dq, dk, dv) for a dctx."""
    q, k, v = (t.detach().double().requires_grad_(True) for t in (q, k, v))
    D = q.shape[1]
    dh = D // H
    outs = []
    qo = ko = 0
    for n, (lq, lk) in enumerate(zip(ql, kl)):
        rows = []
        for h in range(H):
            Q, K, Vv = q[qo:qo + lq, h * dh:(h + 1) * dh], k[ko:ko + lk, h * dh:(h + 1) * dh], v[ko:ko + lk, h * dh:(h + 1) * dh]
            s = Q @ K.t() / math.sqrt(dh)
            m = torch.ones(lq, lk, dtype=torch.float64, device=DEV)
            if km is not None:
                m = m * km[ko:ko + lk].double()[None, :]
            if causal:
                m = m * torch.tril(torch.ones(lq, lk, dtype=torch.float64, device=DEV))
            P = torch.softmax(s + (1 - m) * -10000.0, dim=-1)
            rows.append((P * mask_fn(n, h, lq, lk)) @ Vv)
        outs.append(torch.cat(rows, dim=1))
        qo += lq
        ko += lk
    ctx = torch.cat(outs, dim=0)
    return ctx, lambda dctx: torch.autograd.grad(ctx, (q, k, v), dctx.double())


@pytest.mark.parametrize("dt,ql,kl,causal", [(BF16, [100, 37], [100, 37], False), (BF16, [25, 25], [25, 25], True),
                                              (F32, [12, 7], [12, 7], False), (BF16, [25, 25], [3, 3], False),
                                              (F32, [25, 9], [3, 2], False), (BF16, [1, 1], [100, 70], False)])
def test_attention_dropout(cb, dt, ql, kl, causal):
    H, dh = 12, 64
    D = H * dh
    nq, nk = sum(ql), sum(kl)
    q, k, v = rnd(nq, D, dtype=dt), rnd(nk, D, dtype=dt, seed=1), rnd(nk, D, dtype=dt, seed=2)
    km = torch.ones(nk, device=DEV)
    km[nk // 2::5] = 0
    qo, ko = I32([sum(ql[:i]) for i in range(len(ql))]), I32([sum(kl[:i]) for i in range(len(kl))])
    d = mkdrop(site=11, p=0.1)
    mlq, mlk = max(ql), max(kl)
    lks = (mlk + 1) // 2 * 2          # index-space row stride: max_lk rounded up to even
    full = cb.dropout_mask(len(ql) * H * mlq * lks, d).view(len(ql), H, mlq, lks).double()
    ctx = cb.attn_fwd(q, k, v, qo, I32(ql), ko, I32(kl), km, causal, H, mlq, mlk, dt, drop=d)
    ref, bwd = attn_ref(q, k, v, ql, kl, km, causal, H, lambda n, h, lq, lk: full[n, h, :lq, :lk])
    tol = 3e-2 if dt == BF16 else 3e-5
    close(ctx, ref, tol, "attn ctx")
    dctx = rnd(nq, D, dtype=dt, seed=5)
    dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    cb.attn_bwd(dctx, q, k, v, qo, I32(ql), ko, I32(kl), km, causal, H, mlq, mlk, dq, dk, dv, drop=d)
    rq, rk, rv = bwd(dctx)
    close(dq, rq, tol, "dq"); close(dk, rk, tol, "dk"); close(dv, rv, tol, "dv")


def test_video_ln_affine_dropout(cb):
    B, L, Lv, Fd = 2, 10, 8, 256
    feats = [rnd(B, L, Fd, seed=i) for i in range(3)]
    ct, cbb = I32([0, 1, 2, 0, 1]), I32([0, 0, 0, 1, 1])
    gamma, beta = rnd(Fd, seed=7), rnd(Fd, seed=8)
    d = mkdrop(site=5)
    rows = 5 * Lv
    mask = cb.dropout_mask(rows * Fd, d).view(rows, Fd)
    xhat, mean, rstd = cb.video_ln_pack(feats, ct, cbb, Lv, 1e-12, F32)
    y, _, _ = cb.video_ln_pack(feats, ct, cbb, Lv, 1e-12, F32, gamma=gamma, beta=beta, drop=d)
    close(y, (xhat * gamma + beta) * mask, 1e-5, "video affine+drop")
    dy = rnd(rows, Fd, seed=9)
    dg, db = cb.video_ln_affine_bwd(dy, feats, ct, cbb, Lv, mean, rstd, drop=d)
    close(dg, (dy * mask * xhat).sum(0), 1e-4, "dgamma")
    close(db, (dy * mask).sum(0), 1e-4, "dbeta")


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("dtype", (F32, BF16))
def test_model_trains_with_dropout(mode, dtype):
    backend.set_backend(None)
    cfg, sd, batch, gold = load_case("ragged", mode)
    m = StateAwareRecursiveTransformer(cfg, compute_dtype=dtype)
    m.load_state_dict(sd)
    m.to(DEV).train()

    def run(seed):
        m.zero_grad(set_to_none=True)
        m._seed_t = torch.tensor([seed - 1], dtype=torch.int32, device=DEV)     # forward bumps it by one
        loss = m(*forward_args(batch, DEV), gumbel_noise=batch["gumbel"])[0]
        loss.backward()
        g = torch.cat([p.grad.reshape(-1) for p in m.parameters() if p.grad is not None])
        return float(loss), g
    l1, g1 = run(5)
    l2, g2 = run(5)
    l3, _ = run(6)
    assert math.isfinite(l1) and torch.isfinite(g1).all()
    assert abs(l1 - l2) <= 1e-5 * abs(l1) and (g1 - g2).abs().max() <= 1e-3 * g1.abs().max()   # same seed, same masks
    assert l1 != l3                                                                           # new seed, new masks
    assert abs(l1 - float(gold["loss"])) < 0.5 * abs(float(gold["loss"]))                     # a perturbation, not garbage
    m.eval()
    le = float(m(*forward_args(batch, DEV), gumbel_noise=batch["gumbel"])[0])
    tol = 1e-4 if dtype == F32 else 1e-2
    assert abs(le - float(gold["loss"])) <= tol * abs(float(gold["loss"]))                    # eval mode is dropout-free
