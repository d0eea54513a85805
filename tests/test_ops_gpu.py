"""GPU parity tests, operator by operator: every C-ABI entry point of libsvpc_b200.so (through svpc_b200.cuda_backend)
against the plain-torch statement of the same contract (tests/torch_backend.py, fp64 math) on identical seeded inputs."""
import math
import numpy as np

import pytest
import torch

pytestmark = pytest.mark.gpu

from tests.torch_backend import TorchBackend  # noqa: E402

DEV = "cuda"
F32, BF16 = torch.float32, torch.bfloat16


@pytest.fixture(scope="module")
def cb():
    from svpc_b200.cuda_backend import CudaBackend
    return CudaBackend()


@pytest.fixture(scope="module")
def tb():
    return TorchBackend(math_dtype=torch.float64)


def rnd(*shape, dtype=F32, scale=1.0, seed=None):
    g = torch.Generator(device="cpu")
    g.manual_seed(hash(shape) % 100000 if seed is None else seed)
    return (torch.randn(*shape, generator=g) * scale).to(DEV).to(dtype)


def close(name, got, ref, tol):
    got, ref = got.detach().double().cpu(), ref.detach().double().cpu()
    assert got.shape == ref.shape, (name, got.shape, ref.shape)
    assert torch.isfinite(got).all(), name
    err = (got - ref).abs().max().item()
    scale = max(ref.abs().max().item(), 1e-6)
    assert err <= tol * scale, f"{name}: max err {err:.3e} vs scale {scale:.3e} (tol {tol})"


I32 = lambda x: torch.tensor(x, dtype=torch.int32, device=DEV)


# ---------------------------------------------------------------------------------------------- GEMM
@pytest.mark.parametrize("dt", [F32, BF16])
@pytest.mark.parametrize("M,N,K", [(70, 50, 33), (128, 64, 64), (5, 3, 300), (16, 130, 20), (1, 1, 7),
                                   (192, 3, 768), (192, 1, 300), (3, 768, 192), (192, 768, 3)])   # thin kernels
def test_gemm_simt(cb, tb, dt, M, N, K):
    from svpc_b200.cuda_backend import GEMM_SIMT
    a, b = rnd(M, K, dtype=dt), rnd(N, K, dtype=dt, seed=2)
    bias, res = rnd(N, seed=3), rnd(M, N, dtype=dt, seed=4)
    tol = 1e-5 if dt == F32 else 2e-2
    for act in (0, 1, 2, 3):
        got, pre = cb.gemm(a, b, dt, bias, act, res, want_pre=True, route=GEMM_SIMT)
        ref, rpre = tb.gemm(a, b, torch.float64, bias, act, res, want_pre=True)
        close(f"gemm act{act}", got, ref, tol)
        close(f"gemm pre act{act}", pre, rpre, tol)
    # transposed views (dgrad / wgrad addressing) and accumulate into a strided output
    at, bt = rnd(K, M, dtype=dt, seed=5), rnd(K, N, dtype=dt, seed=6)
    big = torch.zeros(M, N + 9, dtype=F32, device=DEV)
    out = big[:, 4:4 + N]
    cb.gemm(at.t(), bt.t(), F32, out=out, route=GEMM_SIMT)
    cb.gemm(at.t(), bt.t(), F32, out=out, accumulate=True, route=GEMM_SIMT)
    close("gemm T/T accumulate", out, 2 * (at.double().t() @ bt.double()), tol)
    assert float(big[:, :4].abs().max()) == 0 and float(big[:, 4 + N:].abs().max()) == 0


TC_SHAPES = [(128, 128, 64), (256, 128, 128), (200, 951, 768), (1000, 768, 3072), (384, 64, 96), (130, 72, 40),
             (4800, 768, 768),
             # persistent schedule: several rounds per SM + a last round cut into narrower column slices
             (19200, 768, 192), (4800, 2304, 128), (19000, 700, 72), (192, 768, 768), (20000, 300, 64),
             # fewer rows than half a tile (greedy decode of a few clips per GPU)
             (24, 768, 768), (16, 64, 32)]


@pytest.mark.parametrize("a_mn", [False, True])
@pytest.mark.parametrize("b_mn", [False, True])
@pytest.mark.parametrize("M,N,K", TC_SHAPES)
def test_gemm_tc(cb, a_mn, b_mn, M, N, K):
    from svpc_b200.cuda_backend import GEMM_TC
    pad = lambda n: (n + 7) // 8 * 8
    # operands as views of padded buffers so TMA strides are 16-byte multiples even for odd extents
    if a_mn:
        A = rnd(K, pad(M), dtype=BF16)[:, :M].t()
    else:
        A = rnd(M, pad(K), dtype=BF16)[:, :K]
    if b_mn:
        Bm = rnd(K, pad(N), dtype=BF16, seed=2)[:, :N].t()
    else:
        Bm = rnd(N, pad(K), dtype=BF16, seed=2)[:, :K]
    ref = A.double() @ Bm.double().t()
    tol = 2e-3
    got = cb.gemm(A, Bm, F32, route=GEMM_TC)
    close(f"tc plain a_mn={a_mn} b_mn={b_mn}", got, ref, tol)
    bias, res = rnd(N, seed=3), rnd(M, N, dtype=BF16, seed=4)
    got, pre = cb.gemm(A, Bm, BF16, bias, 2, res, want_pre=True, route=GEMM_TC)
    rpre = ref + bias.double()
    rr = rpre * 0.5 * (1 + torch.erf(rpre / math.sqrt(2))) + res.double()
    close("tc gelu+res bf16", got, rr, 2e-2)
    close("tc pre bf16", pre, rpre, 2e-2)
    acc = torch.ones(M, N, dtype=F32, device=DEV)
    cb.gemm(A, Bm, F32, out=acc, accumulate=True, route=GEMM_TC)
    close("tc accumulate", acc, ref + 1, tol)


def test_gemm_tc_splitk(cb):
    from svpc_b200.cuda_backend import GEMM_TC
    M, N, K = 768, 768, 19200          # weight-gradient shape: dW = dY^T X, both operands MN-major
    dY, X = rnd(K, M, dtype=BF16, scale=0.1), rnd(K, N, dtype=BF16, seed=2)
    got = cb.gemm(dY.t(), X.t(), F32, route=GEMM_TC)
    close("tc split-k wgrad", got, dY.double().t() @ X.double(), 2e-3)
    buf = torch.zeros(M, N + 8, dtype=F32, device=DEV)
    cb.gemm(dY.t(), X.t(), F32, out=buf[:, :N], route=GEMM_TC)
    close("tc split-k strided out", buf[:, :N], dY.double().t() @ X.double(), 2e-3)
    assert float(buf[:, N:].abs().max()) == 0


@pytest.mark.parametrize("a_mn", [False, True])
@pytest.mark.parametrize("b_mn", [False, True])
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (200, 951, 768), (1000, 768, 3072), (130, 72, 40), (4800, 768, 300),
                                   (768, 768, 19200), (192, 768, 768)])
def test_gemm_split_fp32(cb, a_mn, b_mn, M, N, K):
    """fp32 operands on the tcgen05 route (split into hi / lo bf16 planes, three passes into one fp32 accumulator): within
    3e-5 of the fp64 product relative to its scale (plain bf16 operands: ~4e-3), all four operand majors, epilogue included."""
    from svpc_b200.cuda_backend import Split2
    A = rnd(K, M).t() if a_mn else rnd(M, K)
    Bm = rnd(K, N, seed=2).t() if b_mn else rnd(N, K, seed=2)
    ref = A.double() @ Bm.double().t()
    assert cb.tc3
    # the tensor core adds every K = 16 slice into the fp32 accumulator with truncation, so the error of ONE accumulator
    # grows linearly with K: forward shapes (K <= 3072) stay below 3e-5; the weight-gradient shape (K = 19200) is only
    # launched without an epilogue, where split-K keeps the chains short (plain case below) - with an epilogue it is 2e-4
    tol = 3e-5
    cb.reset_launch_count()
    got = cb.gemm(A, Bm, F32)
    assert cb.launch_count() == 3, "two split passes + one tensor-core GEMM expected"
    close(f"split plain a_mn={a_mn} b_mn={b_mn}", got, ref, tol)
    if K > 4096:
        tol = 2e-4
    sA, sB = cb.split2(A), cb.split2(Bm)
    assert isinstance(sA, Split2)
    close("hi + lo", sA.hi.double() + sA.lo.double(), A.double(), 2e-5)
    bias, res = rnd(N, seed=3), rnd(M, N, seed=4)
    got, pre = cb.gemm(sA, sB, F32, bias, 2, res, want_pre=True)
    rpre = ref + bias.double()
    rr = rpre * 0.5 * (1 + torch.erf(rpre / math.sqrt(2))) + res.double()
    close("split gelu+res", got, rr, tol)
    close("split pre", pre, rpre, tol)
    acc = torch.ones(M, N, dtype=F32, device=DEV)
    cb.gemm(sA, sB, F32, out=acc, accumulate=True)
    close("split accumulate", acc, ref + 1, tol)
    # a problem that is too small for the tcgen05 kernel falls back to the fp32 FFMA kernel on the fp32 data
    small = cb.gemm(sA[:8], sB, F32)
    close("split -> simt fallback", small, ref[:8], 1e-5)


# ---------------------------------------------------------------------------------------------- row-wise
@pytest.mark.parametrize("dt", [F32, BF16])
@pytest.mark.parametrize("rows,D", [(37, 64), (200, 768), (9, 3072)])
def test_layernorm(cb, tb, dt, rows, D):
    x, res = rnd(rows, D, dtype=dt), rnd(rows, D, dtype=dt, seed=2)
    g, b = 1 + 0.1 * rnd(D, seed=3), 0.1 * rnd(D, seed=4)
    pe, tt = rnd(7, D, seed=5), rnd(4, D, seed=6)
    ids = I32([i % 4 for i in range(rows)])
    tol = 2e-5 if dt == F32 else 2e-2
    for r in (None, res):
        y, pre, mean, rstd = cb.ln_fwd(x, r, g, b, 1e-12, dt, pe, 7, tt, ids)
        ry, rpre, rmean, rrstd = tb.ln_fwd(x, r, g, b, 1e-12, torch.float64, pe, 7, tt, ids)
        close("ln y", y, ry, tol)
        close("ln mean", mean, rmean, tol)
        close("ln rstd", rstd, rrstd, tol)
        dy = rnd(rows, D, dtype=dt, seed=7)
        dpre, dg, db = cb.ln_bwd(dy, pre, mean, rstd, g)
        rdpre, rdg, rdb = tb.ln_bwd(dy.double(), pre.double(), mean.double(), rstd.double(), g)
        close("ln dpre", dpre, rdpre, tol)
        close("ln dgamma", dg, rdg, tol)
        close("ln dbeta", db, rdb, tol)
    close("index_colsum", cb.index_colsum(dy, ids, 4), tb.index_colsum(dy, ids, 4), tol)


@pytest.mark.parametrize("dt", [F32, BF16])
def test_act_bwd_colsum_cast(cb, tb, dt):
    dy, ref = rnd(301, 77, dtype=dt), rnd(301, 77, dtype=dt, seed=2)
    tol = 2e-5 if dt == F32 else 2e-2
    for act in (0, 1, 2, 3):
        r = torch.sigmoid(ref.float()).to(dt) if act == 3 else ref
        dz, dbias = cb.act_bwd(dy, r, act)
        rdz, rdb = tb.act_bwd(dy.double(), r.double(), act)
        close(f"act_bwd dz {act}", dz, rdz, tol)
        close(f"act_bwd dbias {act}", dbias, rdb, tol)
    close("colsum", cb.colsum(dy), dy.double().sum(0), tol)
    assert torch.equal(cb.cast(dy, BF16).float(), dy.to(BF16).float())
    close("sum", cb.sum_all(rnd(5000)), rnd(5000).double().sum(), 1e-5)


@pytest.mark.parametrize("dt", [F32, BF16])
def test_embed_ln(cb, tb, dt):
    V, W, R = 60, 300, 333
    table, g, b = rnd(V, W), 1 + 0.1 * rnd(W, seed=2), 0.1 * rnd(W, seed=3)
    ids = I32([(7 * i) % V for i in range(R)])
    tol = 2e-5 if dt == F32 else 2e-2
    y, mean, rstd = cb.embed_ln_fwd(table, ids, g, b, 1e-12, dt)
    ry, rmean, rrstd = tb.embed_ln_fwd(table, ids, g, b, 1e-12, torch.float64)
    close("embed y", y, ry, tol)
    dy = rnd(R, W, dtype=dt, seed=4)
    dt_, dg, db = cb.embed_ln_bwd(dy, table, ids, mean, rstd, g, 1e-12, 0)
    rdt, rdg, rdb = tb.embed_ln_bwd(dy.double(), table, ids, rmean, rrstd, g, 1e-12, 0)
    close("embed dtable", dt_, rdt, tol)
    close("embed dgamma", dg, rdg, tol)
    close("embed dbeta", db, rdb, tol)
    assert float(dt_[0].abs().max()) == 0


@pytest.mark.parametrize("dt", [F32, BF16])
def test_video_ln_pack(cb, tb, dt):
    S, B, L, Lv, Fd = 3, 2, 10, 8, 96
    feats = [rnd(B, L, Fd, seed=10 + s) for s in range(S)]
    feats[0][0, 0] = 0            # [CLS] rows hold all-zero features
    ct, cbb = I32([0, 1, 2, 0, 1]), I32([0, 0, 0, 1, 1])
    x, mean, rstd = cb.video_ln_pack(feats, ct, cbb, Lv, 1e-12, dt)
    rx, rmean, rrstd = tb.video_ln_pack(feats, ct, cbb, Lv, 1e-12, torch.float64)
    close("video xhat", x, rx, 2e-5 if dt == F32 else 2e-2)
    close("video mean", mean, rmean, 1e-5)
    assert float(x[0].abs().max()) == 0


def test_video_ln_pack_bf16_features(cb, tb):
    """The bf16 upload path: bf16 feature blocks at the real width (3072) through the warp-per-row kernel, eval (plain
    normalisation) and train mode (affine + dropout mask at element index row * F + c), and the affine backward."""
    S, B, L, Lv, Fd = 2, 2, 12, 9, 3072
    feats32 = [rnd(B, L, Fd, seed=20 + s) for s in range(S)]
    feats = [f.to(BF16) for f in feats32]
    ct, cbb = I32([0, 1, 0, 1]), I32([0, 0, 1, 1])
    x, mean, rstd = cb.video_ln_pack(feats, ct, cbb, Lv, 1e-12, BF16)
    rx, rmean, rrstd = tb.video_ln_pack([f.float() for f in feats], ct, cbb, Lv, 1e-12, torch.float64)
    close("video xhat (bf16 in)", x, rx, 1e-2)
    close("video mean (bf16 in)", mean, rmean, 1e-5)
    close("video rstd (bf16 in)", rstd, rrstd, 1e-5)
    gamma, beta = 1 + 0.1 * rnd(Fd, seed=7), 0.1 * rnd(Fd, seed=8)
    drop = (torch.tensor([11], dtype=torch.int32, device=DEV), 5, 0.1)
    rows = 4 * Lv
    mask = cb.dropout_mask(rows * Fd, drop).view(rows, Fd)
    y, _, _ = cb.video_ln_pack(feats, ct, cbb, Lv, 1e-12, BF16, gamma=gamma, beta=beta, drop=drop)
    close("video affine+drop (bf16 in)", y, (rx * gamma.double() + beta.double()) * mask.double(), 1e-2)
    dy = rnd(rows, Fd, dtype=BF16, seed=9)
    dg, db = cb.video_ln_affine_bwd(dy, feats, ct, cbb, Lv, mean, rstd, drop=drop)
    close("video dgamma (bf16 in)", dg, (dy.double() * mask.double() * rx).sum(0), 2e-3)
    close("video dbeta (bf16 in)", db, (dy.double() * mask.double()).sum(0), 2e-3)


@pytest.mark.parametrize("dt", [F32, BF16])
@pytest.mark.parametrize("rows,D", [(200, 768), (37, 64)])
def test_ln_bwd_fused_activation_backward(cb, tb, dt, rows, D):
    """ln_bwd with the dz output: dz = dpre * dropscale * act'(ref) and its column sums, against ln_bwd followed by act_bwd."""
    dy, pre, ref = rnd(rows, D, dtype=dt, seed=1), rnd(rows, D, dtype=dt, seed=2), rnd(rows, D, dtype=dt, seed=3)
    mean, rstd = pre.float().mean(1), 1.0 / (pre.float().var(1, unbiased=False) + 1e-12).sqrt()
    g = 1 + 0.1 * rnd(D, seed=4)
    drop = (torch.tensor([5], dtype=torch.int32, device=DEV), 9, 0.1)
    tol = 1e-5 if dt == F32 else 2e-2
    for act, r, d in ((0, None, drop), (2, ref, None), (1, None, None), (1, ref, drop), (3, torch.sigmoid(ref.float()).to(dt), None)):
        dpre, dg, db, dz, dbias = cb.ln_bwd(dy, pre, mean, rstd, g, fuse=(act, r, d, True))
        dpre0, dg0, db0 = cb.ln_bwd(dy, pre, mean, rstd, g)
        dz0, dbias0 = cb.act_bwd(dpre0, pre if (r is None and act == 1) else r, act, drop=d)
        assert torch.equal(dpre, dpre0)
        close(f"fused dgamma act{act}", dg, dg0, 1e-4); close(f"fused dbeta act{act}", db, db0, 1e-4)
        close(f"fused dz act{act}", dz, dz0, tol)
        close(f"fused dbias act{act}", dbias, dbias0, 2e-2 if dt == BF16 else 1e-4)      # (sums of bf16-rounded vs fp32 terms)


@pytest.mark.parametrize("dt,M,N,K", [(BF16, 256, 192, 128), (F32, 70, 50, 33), (BF16, 4800, 768, 768), (F32, 200, 768, 768)])
def test_gemm_fused_activation_gradient(cb, tb, dt, M, N, K):
    """Input-gradient GEMM whose epilogue adds the skip-connection gradient and multiplies by act'(ref): the previous
    Linear's dz written by the GEMM itself."""
    a, w = rnd(M, K, dtype=dt, scale=0.3), rnd(N, K, dtype=dt, seed=1, scale=0.3)
    res, ref = rnd(M, N, dtype=dt, seed=3), rnd(M, N, dtype=dt, seed=4)
    base = a.double() @ w.double().t()
    tol = 2e-2 if dt == BF16 else 1e-4
    for act in (1, 2, 3):
        r = torch.sigmoid(ref.float()).to(dt) if act == 3 else ref
        got = cb.gemm(a, w, dt, residual=res, grad_act=(act, r))
        want = tb.act_bwd((base + res.double()), r.double(), act)[0]
        close(f"gemm grad_act {act}", got, want, tol)
    got = cb.gemm(a, w, dt, grad_act=(2, ref))
    close("gemm grad_act no residual", got, tb.act_bwd(base, ref.double(), 2)[0], tol)


# ---------------------------------------------------------------------------------------------- attention
ATT_CASES = {
    # name: (q_len list, k_len list, causal, use mask, cross)
    "encoder": ([12, 12, 12], [12, 12, 12], False, True, False),
    "cls_only": ([1, 1, 1], [12, 12, 12], False, True, True),
    "decoder": ([8, 8], [8, 8], True, True, False),
    "cross": ([8, 8], [3, 3], False, False, True),
    "ragged_steps": ([4, 2, 3], [4, 2, 3], False, False, False),
    "full_len": ([100, 100], [100, 100], False, True, False),
    # shapes that take the tensor-core kernel on bf16 / head dim 64 (16 <= L <= 128)
    "tc_ragged": ([100, 37, 64, 16], [100, 37, 64, 16], False, True, False),
    "tc_causal": ([25, 25, 17], [25, 25, 17], True, True, False),
    "tc_cross": ([25, 30], [20, 48], False, True, True),
    "tc_128": ([128], [128], True, False, False),
    # warp-per-(sequence, head) kernels: decoder cross attention (25 x <= 3, ragged context) and the [CLS]-only layer (1 x 100)
    "warp_cross": ([25, 25, 25, 9], [3, 2, 1, 3], False, False, True),
    "warp_cross_masked": ([25, 25], [3, 3], False, True, True),
    "warp_cls": ([1, 1, 1, 1], [100, 100, 37, 70], False, True, True),
}


@pytest.mark.parametrize("dt", [F32, BF16])
@pytest.mark.parametrize("case", list(ATT_CASES))
@pytest.mark.parametrize("H,dh", [(4, 16), (12, 64)])
def test_attention(cb, tb, dt, case, H, dh):
    ql, kl, causal, use_mask, cross = ATT_CASES[case]
    if (case == "full_len" or case.startswith("tc_")) and dh == 16:
        pytest.skip("covered by the other head size")
    if case == "tc_128" and dt == F32:
        pytest.skip("the fp32 short-sequence kernel stages Q/K/V/dO/P/dS in one CTA's shared memory: L <= ~110")
    D = H * dh
    nq, nk = sum(ql), sum(kl)
    qo = I32([sum(ql[:i]) for i in range(len(ql))]); ko = I32([sum(kl[:i]) for i in range(len(kl))])
    qlen, klen = I32(ql), I32(kl)
    if cross:
        qsrc, kvsrc = rnd(nq, D, dtype=dt), rnd(nk, 2 * D, dtype=dt, seed=2)
        q, k, v = qsrc, kvsrc[:, :D], kvsrc[:, D:]
    else:
        qkv = rnd(nq, 3 * D, dtype=dt)
        q, k, v = qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:]
    mask = None
    if use_mask:
        mask = torch.ones(nk, device=DEV)
        mask[nk // 2 + 1::3] = 0
    tol = 3e-5 if dt == F32 else 3e-2
    out = cb.attn_fwd(q, k, v, qo, qlen, ko, klen, mask, causal, H, max(ql), max(kl), dt)
    ref = tb.attn_fwd(q, k, v, qo, qlen, ko, klen, mask, causal, H, max(ql), max(kl), torch.float64)
    close("attn ctx", out, ref, tol)
    dctx = rnd(nq, D, dtype=dt, seed=5)
    dq, dk, dv = torch.empty_like(q.contiguous()), torch.empty_like(k.contiguous()), torch.empty_like(v.contiguous())
    cb.attn_bwd(dctx, q, k, v, qo, qlen, ko, klen, mask, causal, H, max(ql), max(kl), dq, dk, dv)
    rq, rk, rv = (torch.empty(t.shape, dtype=torch.float64, device=DEV) for t in (dq, dk, dv))
    tb.attn_bwd(dctx, q, k, v, qo, qlen, ko, klen, mask, causal, H, max(ql), max(kl), rq, rk, rv)
    close("attn dq", dq, rq, tol); close("attn dk", dk, rk, tol); close("attn dv", dv, rv, tol)


# ---------------------------------------------------------------------------------------------- ingredients
@pytest.mark.parametrize("dt", [F32, BF16])
def test_segmean_ragged_rownorm(cb, tb, dt):
    B, Li, D = 3, 16, 64
    sep = torch.zeros(B, Li, dtype=torch.int32)
    for b, pos in enumerate([[2, 5, 6 + 3], [1, 4], [3, 7, 9, 12]]):
        sep[b, pos] = 1
    sep = sep.to(DEV)
    ni = [3, 2, 4]
    cu = I32([0, 3, 5, 9])
    x, pe = rnd(B * Li, D, dtype=dt), rnd(20, D, seed=2)
    tol = 1e-5 if dt == F32 else 2e-2
    E0 = cb.segmean_fwd(x, sep, cu, pe, sum(ni))
    close("segmean", E0, tb.segmean_fwd(x.double(), sep, cu, pe), tol)
    dE = rnd(sum(ni), D, seed=3)
    close("segmean bwd", cb.segmean_bwd(dE, sep, cu, Li, dt), tb.segmean_bwd(dE.double(), sep, cu, Li, torch.float64), tol)
    close("ragged mean", cb.ragged_mean_fwd(E0, cu), tb.ragged_mean_fwd(E0, cu), 1e-5)
    dy = rnd(B, D, seed=4)
    close("ragged mean bwd", cb.ragged_mean_bwd(dy, cu, sum(ni)), tb.ragged_mean_bwd(dy, cu), 1e-5)
    p = torch.sigmoid(rnd(7, 384, seed=5))
    close("row normalize", cb.row_normalize_fwd(p), tb.row_normalize_fwd(p), 1e-5)
    dn = rnd(7, 384, seed=6)
    close("row normalize bwd", cb.row_normalize_bwd(dn, p), tb.row_normalize_bwd(dn, p), 1e-4)


# ---------------------------------------------------------------------------------------------- reasoner
def _reasoner_case(D):
    steps, ni = [4, 2, 3], [3, 5, 2]
    cs, ci, eo = [0], [0], [0]
    for s, n in zip(steps, ni):
        cs.append(cs[-1] + s); ci.append(ci[-1] + n); eo.append(eo[-1] + s * n)
    clip_eo = [eo[b] + t * ni[b] for b in range(3) for t in range(steps[b])]
    lay = dict(cu_ingr=I32(ci), cu_steps=I32(cs), clip_ent_off=I32(clip_eo), max_ni=max(ni), NE=eo[-1], B=3)
    Nc = cs[-1]
    E0 = rnd(ci[-1], D, scale=0.5)
    w2v, clog, w4s = rnd(Nc, D, scale=0.2, seed=2), rnd(Nc, 3, seed=3), rnd(Nc, seed=4)
    return lay, E0, w2v, clog, w4s, eo[-1], Nc


@pytest.mark.parametrize("D", [64, 768])
@pytest.mark.parametrize("max_steps", [0, 4])      # 0: un-pipelined forward kernel; 4: per-step rows staged up front
def test_reasoner(cb, tb, D, max_steps):
    lay, E0, w2v, clog, w4s, NE, Nc = _reasoner_case(D)
    lay = dict(lay, max_steps=max_steps)
    ent, ebar, Es, saved = cb.reasoner_fwd(E0, lay, w2v, clog, w4s)
    rent, rebar, rEs, _ = tb.reasoner_fwd(E0, lay, w2v, clog, w4s)
    close("reasoner ent", ent, rent, 1e-5); close("reasoner ebar", ebar, rebar, 1e-5); close("reasoner Es", Es, rEs, 1e-5)
    d_ent, d_ebar, d_Es = rnd(NE, seed=5), rnd(Nc, D, seed=6), rnd(NE, D, seed=7)
    got = cb.reasoner_bwd(d_ent, d_ebar, d_Es, E0, lay, w2v, clog, w4s, ent, ebar, Es, saved)
    ref = tb.reasoner_bwd(d_ent, d_ebar, d_Es, E0, lay, w2v, clog, w4s, rent, rebar, rEs, ())
    for n, g, r in zip(("dE0", "dw2v", "dclog", "dw4s"), got, ref):
        close("reasoner " + n, g, r, 2e-4)


# ---------------------------------------------------------------------------------------------- output head
def _pointer_case(dt, copy_mode, big=False):
    # big: enough rows (1250 >= 8 x #SMs) for the multi-row kernels, Lt = 25 -> position groups of 7, 7, 7, 4
    Lt, D, V = (25, 64, 60) if big else (8, 64, 60)
    steps, ni, ez = ([20, 30] if big else [2, 3]), [3, 2], [2, 0]
    B = 2
    vx = [V + z for z in ez]
    clip_b = [b for b in range(B) for _ in range(steps[b])]
    clip_t = [t for b in range(B) for t in range(steps[b])]
    Nc = len(clip_b)
    cu_ingr = [0, 3, 5]
    ent_off = [0, steps[0] * ni[0], steps[0] * ni[0] + steps[1] * ni[1]]
    P_off = [0, steps[0] * Lt * vx[0]]
    P_numel = P_off[1] + steps[1] * Lt * vx[1]
    toks = [[7, 61], [9], [7, 10, 60], [20, 21], [22]]            # per global ingredient (61/60 are OOV columns of video 0)
    tok_ptr, tok_ids = [0], []
    for t in toks:
        tok_ids += t; tok_ptr.append(len(tok_ids))
    n_src = cu_ingr[-1] if copy_mode else ent_off[-1]
    src_off = [cu_ingr[b] for b in clip_b] if copy_mode else [ent_off[b] + t * ni[b] for b, t in zip(clip_b, clip_t)]
    g = torch.Generator().manual_seed(1)
    labels = torch.randint(0, V, (Nc * Lt,), generator=g, dtype=torch.int32)
    labels[::5] = -1
    labels[3] = 61
    labels[4] = 59
    lay = dict(Lt=Lt, V=V, smoothing=0.1, max_ni=max(ni), clip_src_off=I32(src_off), clip_ni=I32([ni[b] for b in clip_b]),
               clip_gi=I32([cu_ingr[b] for b in clip_b]), clip_vx=I32([vx[b] for b in clip_b]),
               clip_P_off=I32([P_off[b] + t * Lt * vx[b] for b, t in zip(clip_b, clip_t)]), P_numel=P_numel,
               tok_ptr=I32(tok_ptr), tok_ids=I32(tok_ids), labels=labels.to(DEV), max_vx=max(vx), tau=0.5)
    dec = rnd(Nc * Lt, D, dtype=dt)
    G, E = rnd(n_src, D, scale=0.3, seed=2), rnd(n_src, D, seed=3)
    w, bb = rnd(2 * D, scale=0.1, seed=4), rnd(1, seed=5)
    logits = rnd(Nc * Lt, V, seed=6)
    return lay, dec, G, E, w, bb, logits, P_numel, Nc, Lt, V


@pytest.mark.parametrize("dt", [F32, BF16])
@pytest.mark.parametrize("copy_mode", [False, True])
@pytest.mark.parametrize("big", [False, True])
def test_pointer(cb, tb, dt, copy_mode, big):
    lay, dec, G, E, w, bb, logits, P_numel, Nc, Lt, V = _pointer_case(dt, copy_mode, big)
    tol = 2e-5 if dt == F32 else 2e-2
    P, rows, saved = cb.pointer_fwd(dec, G, E, w, bb, logits, lay)
    rP, rrows, _ = tb.pointer_fwd(dec, G, E, w, bb, logits, lay)
    close("pointer P", P, rP, tol); close("pointer loss rows", rows, rrows, tol)
    dP = rnd(P_numel, seed=7)
    dloss = torch.tensor(0.7, device=DEV)
    for dp in (None, dP):
        got = cb.pointer_bwd(dp, dloss, dec, G, E, w, bb, logits, lay, P, saved)
        ref = tb.pointer_bwd(dp, dloss, dec, G, E, w, bb, logits, lay, rP, ())
        for n, g, r in zip(("ddec", "dG", "dE", "dw", "db", "dlogits"), got, ref):
            close(f"pointer {n} (dP={'y' if dp is not None else 'n'})", g, r, 5e-4 if dt == F32 else 3e-2)


def test_softmax_ls(cb, tb):
    R, V = 37, 951
    logits = rnd(R, V)
    labels = torch.randint(0, V, (R,), dtype=torch.int32)
    labels[::4] = -1
    labels[1] = V - 1
    labels = labels.to(DEV)
    P, rows = cb.softmax_ls_fwd(logits, labels, V, 0.1)
    rP, rrows = tb.softmax_ls_fwd(logits, labels, V, 0.1)
    close("softmax P", P, rP, 1e-5); close("ls rows", rows, rrows, 1e-5)
    dloss = torch.tensor(1.3, device=DEV)
    dP = rnd(R, V, seed=3)
    for dp in (None, dP):
        close("softmax_ls dlogits", cb.softmax_ls_bwd(dp, dloss, logits, labels, V, 0.1, P),
              tb.softmax_ls_bwd(dp, dloss, logits, labels, V, 0.1, rP), 1e-4)


@pytest.mark.parametrize("big", [False, True])
def test_gumbel(cb, tb, big):
    lay, dec, G, E, w, bb, logits, P_numel, Nc, Lt, V = _pointer_case(F32, False, big)
    P, _, _ = cb.pointer_fwd(dec, G, E, w, bb, logits, lay)
    g = -torch.log(torch.empty(P_numel).exponential_(generator=torch.Generator().manual_seed(3))).to(DEV)
    emb = rnd(V, 24, scale=0.4, seed=8)
    mask = torch.ones(Nc * Lt, device=DEV)
    mask[5::Lt] = 0
    mask[6::Lt] = 0
    x, idx, yv = cb.gumbel_fwd(P, g, emb, lay, mask)
    rx, ridx, ryv = tb.gumbel_fwd(P, g, emb, lay, mask)
    assert torch.equal(idx.cpu(), ridx.cpu())
    close("gumbel xmean", x, rx, 1e-5)
    dx = rnd(Nc, 24, seed=9)
    dP, demb = cb.gumbel_bwd(dx, P, g, emb, lay, mask, idx, yv)
    rdP, rdemb = tb.gumbel_bwd(dx, P, g, emb, lay, mask, ridx, ryv)
    close("gumbel dP", dP, rdP, 1e-4); close("gumbel demb", demb, rdemb, 1e-4)


def test_gumbel_noise_kernel(cb):
    """Device-drawn Gumbel(0,1) noise (what F.gumbel_softmax draws, model.py:1018): moments, reproducibility per seed."""
    n = 1 << 20
    seed = torch.tensor([77], dtype=torch.int32, device=DEV)
    g = cb.gumbel_noise(n, seed, 5)
    assert torch.isfinite(g).all()
    assert abs(float(g.mean()) - 0.5772) < 5e-3 and abs(float(g.var()) - math.pi ** 2 / 6) < 2e-2
    # P(g < x) = exp(-exp(-x))
    for x in (-1.0, 0.0, 1.0, 3.0):
        assert abs(float((g < x).float().mean()) - math.exp(-math.exp(-x))) < 3e-3
    assert torch.equal(g, cb.gumbel_noise(n, seed, 5))
    assert not torch.equal(g, cb.gumbel_noise(n, seed, 6))
    assert not torch.equal(g, cb.gumbel_noise(n, seed + 1, 5))
    big = cb.gumbel_noise(1 << 26, seed, 9)       # 64 M draws: the extreme hash values occur (u must never round to 0 or 1)
    assert torch.isfinite(big).all() and float(big.max()) < 17.0 and float(big.min()) > -3.0


# ---------------------------------------------------------------------------------------------- BiLSTM + losses
@pytest.mark.parametrize("D,W", [(64, 24), (768, 300)])
def test_bilstm(cb, tb, D, W):
    steps = [4, 1, 3]
    cu = I32([0, 4, 5, 8])
    x = rnd(8, W)
    k = 1 / math.sqrt(D)
    params = {}
    for i, key in enumerate(TorchBackend.LSTM_KEYS):
        shape = (4 * D, W) if "weight_ih" in key else (4 * D, D) if "weight_hh" in key else (4 * D,)
        params[key] = (torch.rand(*shape, generator=torch.Generator().manual_seed(i)) * 2 * k - k).to(DEV)
    out, saved = cb.bilstm_fwd(x, cu, steps, params)
    rout, _ = tb.bilstm_fwd(x, cu, steps, params)
    close("bilstm out", out, rout, 1e-5)
    dout = rnd(8, D, seed=5)
    dx, grads = cb.bilstm_bwd(dout, x, cu, steps, params, saved)
    rdx, rgrads = tb.bilstm_bwd(dout, x, cu, steps, params, ())
    close("bilstm dx", dx, rdx, 1e-4)
    for key in TorchBackend.LSTM_KEYS:
        close("bilstm d" + key, grads[key], rgrads[key], 1e-4)


@pytest.mark.parametrize("persistent", [True, False])
@pytest.mark.parametrize("cd", [torch.float32, torch.bfloat16])
def test_bilstm_many_videos(cb, tb, persistent, cd):
    """20 ragged videos (two groups of 16 inside the kernels), full hidden size; the persistent sweep (one launch, grid
    barrier per step) and the per-step launches must both match the torch statement; on the bf16 path only the four
    batched products run on the tensor cores (2e-2)."""
    D, W = 768, 300
    rs = np.random.RandomState(3)
    steps = [int(v) for v in rs.randint(1, 13, size=20)]
    steps[3] = 12
    cu = I32([0] + list(np.cumsum(steps)))
    Nc = sum(steps)
    x = rnd(Nc, W)
    k = 1 / math.sqrt(D)
    params = {}
    for i, key in enumerate(TorchBackend.LSTM_KEYS):
        shape = (4 * D, W) if "weight_ih" in key else (4 * D, D) if "weight_hh" in key else (4 * D,)
        params[key] = (torch.rand(*shape, generator=torch.Generator().manual_seed(i)) * 2 * k - k).to(DEV)
    was = cb.lstm_persistent
    cb.lstm_persistent = persistent
    cb.check_lstm_timeout()
    try:
        out, saved = cb.bilstm_fwd(x, cu, steps, params, cd=cd)
        dout = rnd(Nc, D, seed=5)
        dx, grads = cb.bilstm_bwd(dout, x, cu, steps, params, saved, cd=cd)
    finally:
        cb.lstm_persistent = was
    if persistent:
        assert cb._lstm_use_seq(x.device, D), "the persistent sweep must fit co-resident on a whole B200"
    cb.check_lstm_timeout()          # raises if a grid barrier of the persistent sweep timed out
    rout, _ = tb.bilstm_fwd(x, cu, steps, params)
    rdx, rgrads = tb.bilstm_bwd(dout, x, cu, steps, params, ())
    tol = 1e-4 if cd == torch.float32 else 2e-2
    close("bilstm out", out, rout, tol)
    close("bilstm dx", dx, rdx, tol)
    for key in TorchBackend.LSTM_KEYS:
        close("bilstm d" + key, grads[key], rgrads[key], tol)


def test_lstm_timeout_flag_is_reported(cb):
    """A barrier time-out of the persistent sweep (sticky flag, column 2 of the barrier buffer) raises at the next check."""
    from svpc_b200.cuda_backend import SvpcError
    cb.check_lstm_timeout()
    cb.lstm_sync(DEV)[1, 2] = 1
    with pytest.raises(SvpcError):
        cb.check_lstm_timeout()
    cb.check_lstm_timeout()          # cleared by the failed check


def test_losses(cb, tb):
    p = torch.sigmoid(rnd(50, 16) * 3)
    p[0, 0] = 1.0          # saturated probability: log(1 - p) clamps at -100 like the reference
    p[0, 1] = 0.05
    y = (torch.rand(50, 16, generator=torch.Generator().manual_seed(1)) < 0.2).float().to(DEV)
    y[0, 0] = 0.0
    y[3] = 0               # a row without any action: excluded from the ASL
    g = torch.tensor(0.5, device=DEV)
    close("bce", cb.bce_fwd(p.reshape(-1), y.reshape(-1)), tb.bce_fwd(p.reshape(-1), y.reshape(-1)), 1e-5)
    close("bce bwd", cb.bce_bwd(p.reshape(-1), y.reshape(-1), g), tb.bce_bwd(p.reshape(-1), y.reshape(-1), g), 1e-5)
    close("asl", cb.asl_fwd(p, y), tb.asl_fwd(p, y), 1e-5)
    close("asl bwd", cb.asl_bwd(p, y, g), tb.asl_bwd(p, y, g), 1e-5)
    empty = torch.zeros(4, 16, device=DEV)
    assert float(cb.asl_fwd(p[:4], empty)) == 0.0
