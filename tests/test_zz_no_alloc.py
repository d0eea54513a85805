"""svpc_segmean_fwd_static: the ingredient segment means through the entry point that allocates nothing (scratch in a
file-scope device array of the library) against svpc_segmean_fwd with caller-provided scratch.  Called straight through the
C ABI with ctypes.  Collected last (file name) so that a failure here cannot hide another test under `pytest -x`."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

F32, BF16 = 0, 1


def _lib():
    from svpc_b200.cuda_backend import CudaBackend
    return CudaBackend().lib


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _case(B, Li, D, seed, dtype):
    """Ragged [SEP] layout: video b announces ni[b] ingredients; every announced segment is non-empty."""
    g = torch.Generator().manual_seed(seed)
    sep = torch.zeros(B, Li, dtype=torch.int32)
    ni = []
    for b in range(B):
        n = int(torch.randint(1, 6, (1,), generator=g))
        pos, r = [], -1
        for _ in range(n):
            r += int(torch.randint(2, 4, (1,), generator=g))      # 1-2 tokens, then the [SEP]
            if r >= Li:
                break
            pos.append(r)
        sep[b, pos] = 1
        ni.append(len(pos))
    cu = torch.tensor([0] + list(torch.tensor(ni).cumsum(0)), dtype=torch.int32)
    x = torch.randn(B * Li, D, generator=g).to(dtype)
    pe = torch.randn(8, D, generator=g)
    return sep.cuda(), cu.cuda(), x.cuda(), pe.cuda(), int(cu[-1])


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dynamic(lib, B, Li, D, NI, x, dt, sep, cu, pe):
    E0 = torch.empty(NI, D, dtype=torch.float32, device="cuda")
    row_seg = torch.empty(B * Li, dtype=torch.int32, device="cuda")
    lohi = torch.empty(2 * NI, dtype=torch.int32, device="cuda")
    rc = lib.svpc_segmean_fwd(B, Li, D, NI, _ptr(x), dt, _ptr(sep), _ptr(cu), _ptr(pe), _ptr(E0), _ptr(row_seg), _ptr(lohi),
                              _stream())
    assert rc == 0, lib.svpc_last_error().decode()
    return E0


def _static(lib, B, Li, D, NI, x, dt, sep, cu, pe):
    E0 = torch.full((NI, D), float("nan"), dtype=torch.float32, device="cuda")
    before = torch.cuda.memory_allocated()
    rc = lib.svpc_segmean_fwd_static(B, Li, D, NI, _ptr(x), dt, _ptr(sep), _ptr(cu), _ptr(pe), _ptr(E0), _stream())
    assert rc == 0, lib.svpc_last_error().decode()
    assert torch.cuda.memory_allocated() == before
    return E0


@pytest.mark.parametrize("dtype,dt,tol", [(torch.float32, F32, 1e-5), (torch.bfloat16, BF16, 2e-2)])
def test_static_scratch_matches_caller_scratch_twice_in_a_row(dtype, dt, tol):
    lib = _lib()
    lib.svpc_last_error.restype = C.c_char_p
    # two calls in a row with DIFFERENT shapes: the second must not see the first one's tables in the shared array
    for B, Li, D, seed in [(16, 24, 768, 1), (5, 17, 64, 2), (16, 24, 768, 3)]:
        sep, cu, x, pe, NI = _case(B, Li, D, seed, dtype)
        ref = _dynamic(lib, B, Li, D, NI, x, dt, sep, cu, pe)
        got = _static(lib, B, Li, D, NI, x, dt, sep, cu, pe)
        torch.cuda.synchronize()
        assert torch.isfinite(got).all()
        scale = float(ref.abs().max())
        assert float((got - ref).abs().max()) <= tol * scale, (B, Li, D)
        # same kernels, same inputs, same order of additions: the results are in fact identical
        assert torch.equal(got, ref)


def test_static_scratch_rejects_sizes_beyond_the_built_in_maximum():
    lib = _lib()
    lib.svpc_last_error.restype = C.c_char_p
    sep, cu, x, pe, NI = _case(2, 8, 64, 0, torch.float32)
    E0 = torch.empty(NI, 64, device="cuda")
    rc = lib.svpc_segmean_fwd_static(4096, 17, 64, NI, _ptr(x), F32, _ptr(sep), _ptr(cu), _ptr(pe), _ptr(E0), _stream())   # 69632 rows
    assert rc == -1 and b"65536" in lib.svpc_last_error()
    rc = lib.svpc_segmean_fwd_static(2, 8, 64, 8193, _ptr(x), F32, _ptr(sep), _ptr(cu), _ptr(pe), _ptr(E0), _stream())
    assert rc == -1 and b"8192" in lib.svpc_last_error()
    torch.cuda.synchronize()
