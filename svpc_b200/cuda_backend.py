"""ctypes binding of ``libsvpc_b200.so`` (C ABI in ``include/svpc_b200.h``): the ONLY operator implementation of the
product.  torch supplies device memory and the current CUDA stream; every function below checks its tensors are CUDA
and in the expected layout, allocates outputs with torch and enqueues hand-written sm_100a kernels.  No CPU path:
a missing library or a non-CUDA tensor raises.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SVPC_LIB_PATH") or os.path.join(_HERE, "libsvpc_b200.so")     # (override: A/B runs of two builds)

F32, BF16 = 0, 1
_DT = {torch.float32: F32, torch.bfloat16: BF16}
GEMM_AUTO, GEMM_SIMT, GEMM_TC = 0, 1, 2


class SvpcError(RuntimeError):
    pass


def load_library(path: str = LIB_PATH):
    if not os.path.isfile(path):
        raise SvpcError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                        f"(or `make -C svpc_b200/csrc`); there is no CPU fallback for the SVPC operators")
    lib = C.CDLL(path)
    lib.svpc_last_error.restype = C.c_char_p
    lib.svpc_launch_count.restype = C.c_longlong
    return lib


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _ll(x):
    return C.c_longlong(int(x))


def _f(x):
    return C.c_float(float(x))


def _drop(drop):
    """drop = None or (seed tensor (1,) int32 on the device, site id, p) -> the three C arguments."""
    if drop is None:
        return None, C.c_uint(0), C.c_float(0.0)
    seed, site, p = drop
    assert seed.is_cuda and seed.dtype == torch.int32 and seed.numel() == 1
    return C.c_void_p(seed.data_ptr()), C.c_uint(int(site)), C.c_float(float(p))


def _mat_strides(t):
    """Element strides of a 2-D view with size-1 dimensions normalised so that one of the two is 1."""
    s0, s1 = t.stride()
    if t.shape[1] == 1:
        s1 = 1
    if t.shape[0] == 1 and s1 != 1:
        s0 = 1
    if s0 != 1 and s1 != 1:
        raise SvpcError(f"gemm operand needs a unit stride, got shape {tuple(t.shape)} strides {t.stride()}")
    return s0, s1


class Split2:
    """An fp32 matrix as two bf16 planes, x = hi + lo (hi = bf16(x), lo = bf16(x - hi)): the operand format of the fp32
    path's tensor-core route (``svpc_gemm_split``).  ``src`` is the fp32 matrix itself (the FFMA kernel's operand when a
    problem is not eligible for the tcgen05 kernel).  Views (.t(), slices) apply to all three."""
    __slots__ = ("hi", "lo", "src")
    dtype = torch.float32
    is_cuda = True

    def __init__(self, hi, lo, src):
        self.hi, self.lo, self.src = hi, lo, src

    @property
    def shape(self):
        return self.src.shape

    @property
    def device(self):
        return self.src.device

    def t(self):
        return Split2(self.hi.t(), self.lo.t(), self.src.t())

    def __getitem__(self, idx):
        return Split2(self.hi[idx], self.lo[idx], self.src[idx])


class CudaBackend:
    name = "cuda-sm100a"

    def __init__(self):
        self.lib = load_library()
        self.gemm_route = GEMM_AUTO
        # fp32 (parity) path: contractions that fit the tcgen05 kernel run there on split-bf16 operands (3 passes, fp32
        # accumulation); SVPC_FP32_TC=0 keeps every fp32 GEMM on the FFMA kernel
        self.tc3 = os.environ.get("SVPC_FP32_TC", "1") != "0"
        # ... unless they are tiny (M*N*K below this): the step-level modules (192 rows) gain nothing from the tensor cores
        self.tc3_min_work = float(os.environ.get("SVPC_FP32_TC_MIN_WORK", "0"))
        self._ptr_cache = {}
        self._w_cache = {}
        self.timing = None
        # BiLSTM recurrence: one persistent launch per sweep (default) or one launch per time step (SVPC_LSTM_STEPS=1)
        self.lstm_persistent = os.environ.get("SVPC_LSTM_STEPS", "0") != "1"
        self._lstm_sync = {}          # device -> (2, 4) int32: barrier counters of the fwd / bwd sweep, [:, 2] = sticky time-out flag
        self._lstm_ok = {}            # (device, D) -> the persistent sweep fits co-resident on that device

    # ------------------------------------------------------------------ plumbing
    def _call(self, fn, *args, tag=None):
        timing = self.timing
        if timing is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        rc = getattr(self.lib, fn)(*args, C.c_void_p(torch.cuda.current_stream().cuda_stream))
        if timing is not None:
            e1.record()
            timing.append((tag or fn, e0, e1))
        if rc != 0:
            raise SvpcError(f"{fn} failed ({rc}): {self.lib.svpc_last_error().decode()}")

    def start_timing(self):
        """Per-call CUDA-event timing on the launching stream (bench.py roofline pass)."""
        self.timing = []

    def stop_timing(self):
        torch.cuda.synchronize()
        out = {}
        for tag, e0, e1 in self.timing or []:
            ms = e0.elapsed_time(e1)
            n, t = out.get(tag, (0, 0.0))
            out[tag] = (n + 1, t + ms)
        self.timing = None
        return out

    @staticmethod
    def _chk(*ts):
        for t in ts:
            if t is not None and not t.is_cuda:
                raise SvpcError("svpc_b200 operators need CUDA tensors (no CPU path)")

    @staticmethod
    def _c(t):
        return t if t.is_contiguous() else t.contiguous()

    def memcpy2d_h2d(self, dst_ptr, dpitch, src_ptr, spitch, width, height):
        """Pitched async copy from PINNED host memory to the device on the current stream (raw addresses)."""
        self._call("svpc_memcpy2d_async", C.c_void_p(dst_ptr), _ll(dpitch), C.c_void_p(src_ptr), _ll(spitch), _ll(width),
                   _ll(height), 1)

    def memcpy2d_d2d(self, dst_ptr, dpitch, src_ptr, spitch, width, height):
        """Pitched async device -> device copy on the current stream (raw addresses)."""
        self._call("svpc_memcpy2d_async", C.c_void_p(dst_ptr), _ll(dpitch), C.c_void_p(src_ptr), _ll(spitch), _ll(width),
                   _ll(height), 3)

    def clear_weight_caches(self):
        self._w_cache.clear()

    def launch_count(self):
        return int(self.lib.svpc_launch_count())

    def reset_launch_count(self):
        self.lib.svpc_reset_launch_count()

    # ------------------------------------------------------------------ dense
    # ---- split-fp32 operands (the parity path on the tensor cores)
    def _split_shape_ok(self, M, N, K):
        return M >= 64 and N >= 16 and K >= 32 and float(M) * N * K >= self.tc3_min_work

    def split2(self, x):
        """fp32 2-D view with one unit stride -> Split2 (planes in the same orientation, row pitch padded to 16 bytes)."""
        if isinstance(x, Split2):
            return x
        self._chk(x)
        assert x.dtype == torch.float32 and x.dim() == 2
        if x.shape[1] != 1 and x.stride(1) != 1 and x.stride(0) == 1:
            return self.split2(x.t()).t()
        if x.shape[1] > 1 and x.stride(1) != 1:
            x = x.contiguous()
        R, Cc = x.shape
        ld = (Cc + 7) // 8 * 8
        planes = torch.empty(2, R, ld, dtype=torch.bfloat16, device=x.device) if ld == Cc else \
            torch.zeros(2, R, ld, dtype=torch.bfloat16, device=x.device)
        self.split2_into(x, planes)
        return Split2(planes[0, :, :Cc], planes[1, :, :Cc], x)

    def split2_into(self, x, planes):
        """planes (2, R, ld) bf16 <- hi / lo of x (R, C) fp32 (row pitch x.stride(0))."""
        R, Cc = x.shape
        self._call("svpc_split_bf16", R, Cc, _ptr(x), _ll(x.stride(0) if R > 1 else Cc), _ptr(planes[0]), _ptr(planes[1]),
                   _ll(planes.stride(1)))

    def maybe_split(self, x):
        """Split an fp32 activation that is going to be a tensor-core GEMM operand (rows >= 64, columns >= 32); anything
        else is returned unchanged."""
        if self.tc3 and self.gemm_route != GEMM_SIMT and isinstance(x, torch.Tensor) and x.dtype == torch.float32 and \
                x.dim() == 2 and x.shape[0] >= 64 and x.shape[1] >= 32:
            return self.split2(x)
        return x

    def gemm(self, a, bt, out_dtype, bias=None, act=0, residual=None, out=None, accumulate=False, want_pre=False,
             route=None, drop=None, grad_act=None):
        """out[M,N] = act(a[M,K] @ bt[N,K]^T + bias) + residual.  a / bt may be any 2-D views with one unit stride, or
        Split2 operands (fp32 on the tensor cores).  ``grad_act`` = (act code, ref (M, N) of the output dtype): the result
        is finally multiplied by act'(ref) (an input-gradient GEMM that writes the previous Linear's dz)."""
        route = self.gemm_route if route is None else route
        a_s, b_s = isinstance(a, Split2), isinstance(bt, Split2)
        M, K = a.shape
        N = bt.shape[0]
        assert bt.shape[1] == K, (a.shape, bt.shape)
        both_f32 = a.dtype == torch.float32 and bt.dtype == torch.float32
        if (a_s or b_s or (self.tc3 and both_f32)) and route != GEMM_SIMT and self._split_shape_ok(M, N, K) and M > 0 and N > 0 \
                and both_f32:
            sa_, sb_ = self.split2(a), self.split2(bt)
            sa, sb = _mat_strides(sa_.hi), _mat_strides(sb_.hi)
            if self.lib.svpc_gemm_split_eligible(M, N, K, _ll(sa[0]), _ll(sa[1]), _ll(sb[0]), _ll(sb[1])):
                return self._gemm_split(sa_, sb_, M, N, K, sa, sb, out_dtype, bias, act, residual, out, accumulate, want_pre, drop,
                                        grad_act)
        if a_s:
            a = a.src
        if b_s:
            bt = bt.src
        self._chk(a, bt, bias, residual, out)
        if out is None:
            out = self._out_empty(M, N, out_dtype, a.device)
        assert out.shape == (M, N) and (N == 1 or out.stride(1) == 1)
        pre = torch.empty(M, N, dtype=out.dtype, device=a.device) if want_pre else None
        if residual is not None:
            assert residual.dtype == out.dtype and residual.shape == (M, N) and (N == 1 or residual.stride(1) == 1)
        gref, gact = self._grad_act(grad_act, out, M, N)
        if M > 0 and N > 0:
            sa, sb = _mat_strides(a), _mat_strides(bt)
            self._call("svpc_gemm", M, N, K,
                       _ptr(a), _DT[a.dtype], _ll(sa[0]), _ll(sa[1]),
                       _ptr(bt), _DT[bt.dtype], _ll(sb[0]), _ll(sb[1]),
                       _ptr(out), _DT[out.dtype], _ll(out.stride(0)),
                       _ptr(bias), int(act),
                       _ptr(residual), _ll(residual.stride(0) if residual is not None else 0),
                       _ptr(pre), _ll(N),
                       int(bool(accumulate)), int(route), *_drop(drop),
                       _ptr(gref), _ll(gref.stride(0) if gref is not None else 0), int(gact),
                       tag=None if self.timing is None else self._gemm_tag(M, N, K, a, bt, out, sa, sb))
        return (out, pre) if want_pre else out

    @staticmethod
    def _grad_act(grad_act, out, M, N):
        if grad_act is None:
            return None, 0
        gact, gref = grad_act
        assert gref.dtype == out.dtype and gref.shape == (M, N) and (N == 1 or gref.stride(1) == 1), (gref.dtype, out.dtype)
        return gref, gact

    def _gemm_split(self, a, bt, M, N, K, sa, sb, out_dtype, bias, act, residual, out, accumulate, want_pre, drop,
                    grad_act=None):
        self._chk(a.hi, bt.hi, bias, residual, out)
        if out is None:
            out = self._out_empty(M, N, out_dtype, a.device)
        assert out.shape == (M, N) and (N == 1 or out.stride(1) == 1)
        pre = torch.empty(M, N, dtype=out.dtype, device=a.device) if want_pre else None
        if residual is not None:
            assert residual.dtype == out.dtype and residual.shape == (M, N) and (N == 1 or residual.stride(1) == 1)
        gref, gact = self._grad_act(grad_act, out, M, N)
        self._call("svpc_gemm_split", M, N, K, _ptr(a.hi), _ptr(a.lo), _ll(sa[0]), _ll(sa[1]),
                   _ptr(bt.hi), _ptr(bt.lo), _ll(sb[0]), _ll(sb[1]),
                   _ptr(out), _DT[out.dtype], _ll(out.stride(0)), _ptr(bias), int(act),
                   _ptr(residual), _ll(residual.stride(0) if residual is not None else 0), _ptr(pre), _ll(N),
                   int(bool(accumulate)), *_drop(drop), _ptr(gref), _ll(gref.stride(0) if gref is not None else 0), int(gact),
                   tag=None if self.timing is None else f"gemm_tc3:{M}x{N}x{K}:{'T' if sa[1] != 1 else 'N'}{'T' if sb[1] != 1 else 'N'}")
        return (out, pre) if want_pre else out

    def _gemm_tag(self, M, N, K, a, bt, out, sa, sb):
        r = self.lib.svpc_gemm_route(M, N, K, _DT[a.dtype], _ll(sa[0]), _ll(sa[1]), _DT[bt.dtype], _ll(sb[0]), _ll(sb[1]),
                                     _DT[out.dtype], _ll(out.stride(0)))
        if self.gemm_route == GEMM_SIMT:
            r = GEMM_SIMT
        return f"gemm_{'tc' if r == GEMM_TC else 'simt'}:{M}x{N}x{K}:{'T' if sa[1] != 1 else 'N'}{'T' if sb[1] != 1 else 'N'}"

    def act_bwd(self, dy, ref, act, drop=None):
        self._chk(dy, ref)
        dy = self._c(dy)
        rows, cols = dy.shape
        dz = torch.empty_like(dy)
        dbias = torch.empty(cols, dtype=torch.float32, device=dy.device)
        self._call("svpc_act_bwd", rows, cols, _ptr(dy), _ptr(None if ref is None else self._c(ref)), _DT[dy.dtype], int(act),
                   _ptr(dz), _ptr(dbias), *_drop(drop))
        return dz, dbias

    def dropout_mask(self, n, drop):
        """Scale (0 or 1/(1-p)) of elements [0, n) of a dropout site: the contract the tests check the kernels against."""
        out = torch.empty(int(n), dtype=torch.float32, device=drop[0].device)
        self._call("svpc_dropout_mask", _ll(n), *_drop(drop), _ptr(out))
        return out

    def gumbel_noise(self, n, seed, site):
        """Gumbel(0,1) noise of one step (reference: F.gumbel_softmax draws it on the device, model.py:1018)."""
        out = torch.empty(int(n), dtype=torch.float32, device=seed.device)
        self._call("svpc_gumbel_noise", _ll(n), _ptr(seed), C.c_uint(int(site)), _ptr(out))
        return out

    def colsum(self, x):
        self._chk(x)
        if x.dim() == 2 and not x.is_contiguous() and x.stride(1) == 1 and x.stride(0) > x.shape[1] and x.shape[0] > 0:
            # row-pitched view: sum the pitched rows as they lie (the pad columns are summed into slots nobody reads)
            ld = x.stride(0)
            return self.colsum(x.as_strided((x.shape[0], ld), (ld, 1)))[:x.shape[1]]
        x = self._c(x)
        out = torch.empty(x.shape[1], dtype=torch.float32, device=x.device)
        self._call("svpc_colsum", x.shape[0], x.shape[1], _ptr(x), _DT[x.dtype], _ptr(out))
        return out

    @staticmethod
    def _out_empty(M, N, dtype, device):
        """GEMM output: rows pitched to 16 bytes (the epilogue's vector stores need aligned rows; with the 951-word LM head
        an un-pitched fp32 output drops the whole launch onto the element-wise store path, 47 us against 12)."""
        q = 16 // torch.empty((), dtype=dtype).element_size()
        if N <= 8 or N % q == 0:
            return torch.empty(M, N, dtype=dtype, device=device)
        ld = (N + q - 1) // q * q
        return torch.empty(M, ld, dtype=dtype, device=device)[:, :N]

    @staticmethod
    def _padded_empty(rows, cols, dtype, device):
        """(rows, cols) tensor whose row pitch is a multiple of 16 bytes (a TMA requirement for GEMM operands)."""
        q = 16 // torch.empty((), dtype=dtype).element_size()
        ld = (cols + q - 1) // q * q
        if ld == cols:
            return torch.empty(rows, cols, dtype=dtype, device=device)
        return torch.zeros(rows, ld, dtype=dtype, device=device)[:, :cols]

    def cast(self, x, dtype):
        self._chk(x)
        if x.dtype == dtype:
            return x
        if x.dim() == 2 and dtype == torch.bfloat16 and x.shape[1] % 8 != 0 and x.shape[1] > 8 and x.stride(1) == 1:
            out = self._padded_empty(x.shape[0], x.shape[1], dtype, x.device)
            self._call("svpc_cast2d", x.shape[0], x.shape[1], _ptr(x), _DT[x.dtype], _ll(x.stride(0)), _ptr(out), _DT[dtype],
                       _ll(out.stride(0)))
            return out
        x = self._c(x)
        out = torch.empty(x.shape, dtype=dtype, device=x.device)
        self._call("svpc_cast", _ll(x.numel()), _ptr(x), _DT[x.dtype], _ptr(out), _DT[dtype])
        return out

    def cast_multi_tables(self, srcs, dsts, numel, cols, lds, chunk, device):
        """Device tables of one svpc_cast_multi launch (pointer arrays, element counts, row geometry, chunk table)."""
        ct, cs = [], []
        for i, n in enumerate(numel):
            for s in range(0, max(n, 1), chunk):
                ct.append(i)
                cs.append(s)
        mk = lambda v, dt: torch.tensor(v, dtype=dt, device=device)
        return {"n_chunks": len(ct), "chunk_tensor": mk(ct, torch.int32), "chunk_start": mk(cs, torch.int64),
                "src": mk(srcs, torch.int64), "dst": mk(dsts, torch.int64), "numel": mk(numel, torch.int64),
                "cols": mk(cols, torch.int32), "ld": mk(lds, torch.int64)}

    def cast_multi(self, tab, dtype, chunk):
        self._call("svpc_cast_multi", tab["n_chunks"], _ptr(tab["chunk_tensor"]), _ptr(tab["chunk_start"]), int(chunk),
                   _ptr(tab["src"]), _ptr(tab["dst"]), _ptr(tab["numel"]), _ptr(tab["cols"]), _ptr(tab["ld"]), _DT[dtype])

    def sum_all(self, x):
        self._chk(x)
        x = self._c(x)
        out = torch.empty((), dtype=torch.float32, device=x.device)
        self._call("svpc_sum", _ll(x.numel()), _ptr(x), _ptr(out))
        return out

    # ------------------------------------------------------------------ layer norm / embeddings
    def ln_fwd(self, x, res, gamma, beta, eps, out_dtype, pe=None, pe_period=0, tt=None, tt_ids=None):
        self._chk(x, res, gamma, beta, pe, tt, tt_ids)
        x = self._c(x)
        rows, D = x.shape
        y = torch.empty(rows, D, dtype=out_dtype, device=x.device)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        pre = x
        if res is not None:
            res = self._c(res)
            assert res.dtype == x.dtype and res.shape == x.shape
            pre = torch.empty_like(x)
        self._call("svpc_ln_fwd", rows, D, _ptr(x), _ptr(res), _DT[x.dtype], _ptr(gamma), _ptr(beta), _f(eps),
                   _ptr(pe), int(pe_period), _ptr(tt), _ptr(tt_ids), _ptr(y), _DT[out_dtype],
                   _ptr(pre if res is not None else None), _ptr(mean), _ptr(rstd))
        return y, pre, mean, rstd

    def ln_bwd(self, dy, pre, mean, rstd, gamma, fuse=None):
        """``fuse`` = (act, ref or None, drop or None, want_dbias): also produce dz = dpre * dropscale * act'(ref) and its
        column sums - the activation / dropout backward of the Linear in front of the LayerNorm, in the same pass."""
        self._chk(dy, pre)
        dy = self._c(dy)
        rows, D = dy.shape
        dpre = torch.empty_like(pre)
        dgamma = torch.empty(D, dtype=torch.float32, device=dy.device)
        dbeta = torch.empty(D, dtype=torch.float32, device=dy.device)
        dz = dbias = ref = None
        act, drop = 0, None
        if fuse is not None:
            act, ref, drop, want_dbias = fuse
            dz = torch.empty_like(pre)
            dbias = torch.empty(D, dtype=torch.float32, device=dy.device) if want_dbias else None
            if ref is not None:
                ref = self._c(ref)
                assert ref.dtype == pre.dtype and ref.shape == pre.shape
        self._call("svpc_ln_bwd", rows, D, _ptr(dy), _DT[dy.dtype], _ptr(pre), _DT[pre.dtype], _ptr(mean), _ptr(rstd),
                   _ptr(gamma), _ptr(dpre), _ptr(dgamma), _ptr(dbeta), _ptr(dz), int(act), _ptr(ref), _ptr(dbias), *_drop(drop))
        if fuse is not None:
            return dpre, dgamma, dbeta, dz, dbias
        return dpre, dgamma, dbeta

    def index_colsum(self, dy, ids, nbins):
        self._chk(dy, ids)
        dy = self._c(dy)
        out = torch.empty(nbins, dy.shape[1], dtype=torch.float32, device=dy.device)
        self._call("svpc_index_colsum", dy.shape[0], dy.shape[1], _ptr(dy), _DT[dy.dtype], _ptr(ids), int(nbins), _ptr(out))
        return out

    def embed_ln_fwd(self, table, ids, gamma, beta, eps, out_dtype, drop=None):
        self._chk(table, ids)
        rows, W = ids.numel(), table.shape[1]
        y = self._padded_empty(rows, W, out_dtype, table.device) if out_dtype == torch.bfloat16 else \
            torch.empty(rows, W, dtype=out_dtype, device=table.device)
        mean = torch.empty(rows, dtype=torch.float32, device=table.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=table.device)
        self._call("svpc_embed_ln_fwd", rows, W, _ptr(self._c(table)), _ptr(ids), _ptr(gamma), _ptr(beta), _f(eps), _ptr(y),
                   _ll(y.stride(0)), _DT[out_dtype], _ptr(mean), _ptr(rstd), *_drop(drop))
        return y, mean, rstd

    def embed_ln_bwd(self, dy, table, ids, mean, rstd, gamma, eps, padding_idx, drop=None):
        self._chk(dy, table, ids)
        dy = self._c(dy)
        V, W = table.shape
        dtable = torch.empty(V, W, dtype=torch.float32, device=dy.device)
        dgamma = torch.empty(W, dtype=torch.float32, device=dy.device)
        dbeta = torch.empty(W, dtype=torch.float32, device=dy.device)
        self._call("svpc_embed_ln_bwd", dy.shape[0], W, V, _ptr(dy), _DT[dy.dtype], _ptr(self._c(table)), _ptr(ids), _ptr(mean),
                   _ptr(rstd), _ptr(gamma), int(-1 if padding_idx is None else padding_idx), _ptr(dtable), _ptr(dgamma),
                   _ptr(dbeta), *_drop(drop))
        return dtable, dgamma, dbeta

    def _feat_ptrs(self, feats_list):
        f0 = feats_list[0]
        self._chk(f0)
        B, L, Fdim = f0.shape
        for f in feats_list:
            if not (f.is_cuda and f.dtype == f0.dtype and f.dtype in _DT and f.is_contiguous() and f.shape == f0.shape):
                raise SvpcError("video features must be contiguous fp32 (or bf16) CUDA tensors of one shape (B, L, F)")
        key = tuple(f.data_ptr() for f in feats_list)
        ptrs = self._ptr_cache.get(key)
        if ptrs is None:
            if len(self._ptr_cache) > 64:
                self._ptr_cache.clear()
            ptrs = torch.tensor(list(key), dtype=torch.int64).to(f0.device)
            self._ptr_cache[key] = ptrs
        return ptrs, B, L, Fdim

    def video_ln_pack(self, feats_list, clip_t, clip_b, Lv, eps, out_dtype, gamma=None, beta=None, drop=None):
        """(x - mean) * rstd of the first Lv rows of every clip; with gamma / beta (train mode) the LN affine and the
        dropout between LN and the projection are applied here too (model.py:549-550)."""
        ptrs, B, L, Fdim = self._feat_ptrs(feats_list)
        f0 = feats_list[0]
        n_clips = clip_t.numel()
        xhat = torch.empty(n_clips * Lv, Fdim, dtype=out_dtype, device=f0.device)
        mean = torch.empty(n_clips * Lv, dtype=torch.float32, device=f0.device)
        rstd = torch.empty(n_clips * Lv, dtype=torch.float32, device=f0.device)
        self._call("svpc_video_ln_pack", n_clips, B, L, int(Lv), Fdim, _ptr(ptrs), _DT[f0.dtype], _ptr(clip_t), _ptr(clip_b), _f(eps),
                   _ptr(xhat), _DT[out_dtype], _ptr(mean), _ptr(rstd), _ptr(gamma), _ptr(beta), *_drop(drop))
        return xhat, mean, rstd

    def video_ln_affine_bwd(self, dy, feats_list, clip_t, clip_b, Lv, mean, rstd, drop=None):
        ptrs, B, L, Fdim = self._feat_ptrs(feats_list)
        dy = self._c(dy)
        dgamma = torch.empty(Fdim, dtype=torch.float32, device=dy.device)
        dbeta = torch.empty(Fdim, dtype=torch.float32, device=dy.device)
        self._call("svpc_video_ln_affine_bwd", int(clip_t.numel()), B, L, int(Lv), Fdim, _ptr(ptrs), _DT[feats_list[0].dtype],
                   _ptr(clip_t), _ptr(clip_b),
                   _ptr(mean), _ptr(rstd), _ptr(dy), _DT[dy.dtype], _ptr(dgamma), _ptr(dbeta), *_drop(drop))
        return dgamma, dbeta

    # ------------------------------------------------------------------ attention
    @staticmethod
    def _rowview(t):
        assert t.dim() == 2 and t.stride(1) == 1, (t.shape, t.stride())
        return t

    def attn_fwd(self, q, k, v, q_off, q_len, k_off, k_len, key_mask, causal, n_heads, max_lq, max_lk, out_dtype, drop=None):
        self._chk(q, k, v, q_off, key_mask)
        q, k, v = self._rowview(q), self._rowview(k), self._rowview(v)
        D = q.shape[1]
        assert out_dtype == q.dtype == k.dtype == v.dtype
        ctx = torch.empty(q.shape[0], D, dtype=out_dtype, device=q.device)
        self._call("svpc_attn_fwd", int(q_off.numel()), int(n_heads), D // n_heads, _ptr(q), _ll(q.stride(0)), _ptr(k),
                   _ll(k.stride(0)), _ptr(v), _ll(v.stride(0)), _DT[q.dtype], _ptr(q_off), _ptr(q_len), _ptr(k_off),
                   _ptr(k_len), _ptr(key_mask), int(bool(causal)), int(max_lq), int(max_lk), _ptr(ctx), _ll(D), *_drop(drop))
        return ctx

    def attn_bwd(self, dctx, q, k, v, q_off, q_len, k_off, k_len, key_mask, causal, n_heads, max_lq, max_lk, dq, dk, dv,
                 drop=None):
        self._chk(dctx, q, k, v, dq, dk, dv)
        dctx = self._c(dctx)
        D = q.shape[1]
        for t in (q, k, v, dq, dk, dv):
            self._rowview(t)
        assert dctx.dtype == q.dtype == dq.dtype == dk.dtype == dv.dtype
        self._call("svpc_attn_bwd", int(q_off.numel()), int(n_heads), D // n_heads, _ptr(q), _ll(q.stride(0)), _ptr(k),
                   _ll(k.stride(0)), _ptr(v), _ll(v.stride(0)), _DT[q.dtype], _ptr(q_off), _ptr(q_len), _ptr(k_off),
                   _ptr(k_len), _ptr(key_mask), int(bool(causal)), int(max_lq), int(max_lk), _ptr(dctx), _ll(D),
                   _ptr(dq), _ll(dq.stride(0)), _ptr(dk), _ll(dk.stride(0)), _ptr(dv), _ll(dv.stride(0)), *_drop(drop))

    # ------------------------------------------------------------------ ingredients
    def segmean_fwd(self, x, sep_mask, cu_ingr, pe, NI):
        self._chk(x, sep_mask, cu_ingr, pe)
        x = self._c(x)
        B, Li = sep_mask.shape
        D = x.shape[1]
        NI = int(NI)
        E0 = torch.empty(NI, D, dtype=torch.float32, device=x.device)
        row_seg = torch.empty(B * Li, dtype=torch.int32, device=x.device)
        lohi = torch.empty(max(NI, 1) * 2, dtype=torch.int32, device=x.device)
        self._call("svpc_segmean_fwd", B, Li, D, NI, _ptr(x), _DT[x.dtype], _ptr(sep_mask), _ptr(cu_ingr), _ptr(pe), _ptr(E0),
                   _ptr(row_seg), _ptr(lohi))
        return E0

    def segmean_bwd(self, dE0, sep_mask, cu_ingr, Li, out_dtype):
        self._chk(dE0, sep_mask)
        dE0 = self._c(dE0)
        B = sep_mask.shape[0]
        NI, D = dE0.shape
        dx = torch.empty(B * Li, D, dtype=out_dtype, device=dE0.device)
        row_seg = torch.empty(B * Li, dtype=torch.int32, device=dE0.device)
        lohi = torch.empty(max(NI, 1) * 2, dtype=torch.int32, device=dE0.device)
        self._call("svpc_segmean_bwd", B, Li, D, NI, _ptr(dE0), _ptr(sep_mask), _ptr(cu_ingr), _ptr(dx), _DT[out_dtype],
                   _ptr(row_seg), _ptr(lohi))
        return dx

    def ragged_mean_fwd(self, x, cu):
        self._chk(x, cu)
        x = self._c(x)
        B = cu.numel() - 1
        y = torch.empty(B, x.shape[1], dtype=torch.float32, device=x.device)
        self._call("svpc_ragged_mean_fwd", B, x.shape[1], _ptr(x), _ptr(cu), _ptr(y))
        return y

    def ragged_mean_bwd(self, dy, cu, NI):
        self._chk(dy, cu)
        dy = self._c(dy)
        B = cu.numel() - 1
        dx = torch.empty(int(NI), dy.shape[1], dtype=torch.float32, device=dy.device)
        self._call("svpc_ragged_mean_bwd", B, dy.shape[1], _ptr(dy), _ptr(cu), _ptr(dx))
        return dx

    def row_normalize_fwd(self, x):
        self._chk(x)
        x = self._c(x)
        y = torch.empty_like(x)
        self._call("svpc_row_normalize_fwd", x.shape[0], x.shape[1], _ptr(x), _ptr(y))
        return y

    def row_normalize_bwd(self, dy, x):
        self._chk(dy, x)
        dy, x = self._c(dy), self._c(x)
        dx = torch.empty_like(x)
        self._call("svpc_row_normalize_bwd", x.shape[0], x.shape[1], _ptr(dy), _ptr(x), _ptr(dx))
        return dx

    # ------------------------------------------------------------------ visual simulator recurrence
    def reasoner_fwd(self, E0, lay, w2v, clog, w4s):
        cu_ingr, cu_steps, clip_ent_off = lay["cu_ingr"], lay["cu_steps"], lay["clip_ent_off"]
        self._chk(E0, cu_ingr, cu_steps, clip_ent_off, w2v, clog, w4s)
        E0, w2v, clog, w4s = self._c(E0), self._c(w2v), self._c(clog), self._c(w4s)
        B, max_ni, NE = int(lay["B"]), int(lay["max_ni"]), int(lay["NE"])
        Nc, D = w2v.shape
        ent = torch.empty(NE, dtype=torch.float32, device=E0.device)
        a = torch.empty(NE, dtype=torch.float32, device=E0.device)
        ebar = torch.empty(Nc, D, dtype=torch.float32, device=E0.device)
        Es = torch.empty(NE, D, dtype=torch.float32, device=E0.device)
        self._call("svpc_reasoner_fwd", B, D, max_ni, int(lay.get("max_steps", 0)), _ptr(E0), _ptr(cu_ingr), _ptr(cu_steps),
                   _ptr(clip_ent_off),
                   _ptr(w2v), _ptr(clog), _ptr(w4s), _ptr(ent), _ptr(ebar), _ptr(Es), _ptr(a))
        return ent, ebar, Es, (a,)

    def reasoner_bwd(self, d_ent, d_ebar, d_Es, E0, lay, w2v, clog, w4s, ent, ebar, Es, saved):
        cu_ingr, cu_steps, clip_ent_off = lay["cu_ingr"], lay["cu_steps"], lay["clip_ent_off"]
        (a,) = saved
        B, max_ni = int(lay["B"]), int(lay["max_ni"])
        Nc, D = w2v.shape
        E0, w2v, clog, w4s = self._c(E0), self._c(w2v), self._c(clog), self._c(w4s)
        dE0 = torch.empty_like(E0)
        dw2v = torch.empty_like(w2v)
        dclog = torch.empty_like(clog)
        dw4s = torch.empty_like(w4s)
        self._call("svpc_reasoner_bwd", B, D, max_ni, int(lay.get("max_steps", 0)), _ptr(E0), _ptr(cu_ingr), _ptr(cu_steps),
                   _ptr(clip_ent_off),
                   _ptr(w2v), _ptr(clog), _ptr(w4s), _ptr(ent), _ptr(ebar), _ptr(Es), _ptr(a),
                   _ptr(self._c(d_ent)), _ptr(self._c(d_ebar)), _ptr(self._c(d_Es)), _ptr(dE0), _ptr(dw2v), _ptr(dclog),
                   _ptr(dw4s))
        return dE0, dw2v, dclog, dw4s

    # ------------------------------------------------------------------ pointer-generator + caption loss
    def pointer_fwd(self, dec, G, Esrc, wpgen, bpgen, logits, lay):
        self._chk(dec, G, Esrc, wpgen, bpgen, logits)
        dec, G, Esrc, wpgen = self._c(dec), self._c(G), self._c(Esrc), self._c(wpgen)
        logits = self._rowview(logits)                  # (R, V) fp32, rows may be pitched
        Lt, V = int(lay["Lt"]), int(lay["V"])
        R, D = dec.shape
        Nc = R // Lt
        max_ni, max_vx = max(int(lay["max_ni"]), 1), int(lay["max_vx"])
        dev = dec.device
        P = torch.empty(int(lay["P_numel"]), dtype=torch.float32, device=dev)
        rows = torch.empty(R, dtype=torch.float32, device=dev)
        alpha = torch.empty(R, max_ni, dtype=torch.float32, device=dev)
        pgen = torch.empty(R, dtype=torch.float32, device=dev)
        ctx = torch.empty(R, D, dtype=torch.float32, device=dev)
        self._call("svpc_pointer_fwd", Nc, Lt, D, V, max_ni, max_vx, _ptr(dec), _DT[dec.dtype], _ptr(G), _ptr(Esrc),
                   _ptr(lay["clip_src_off"]), _ptr(lay["clip_ni"]), _ptr(lay["clip_gi"]), _ptr(lay["clip_vx"]),
                   _ptr(lay["clip_P_off"]), _ptr(lay["tok_ptr"]), _ptr(lay["tok_ids"]), _ptr(wpgen), _ptr(bpgen),
                   _ptr(logits), _ll(logits.stride(0)), _ptr(lay["labels"]), _f(lay["smoothing"]), _ptr(P), _ptr(rows),
                   _ptr(alpha), _ptr(pgen), _ptr(ctx))
        return P, rows, (alpha, pgen, ctx)

    def pointer_bwd(self, dP, dloss, dec, G, Esrc, wpgen, bpgen, logits, lay, P, saved):
        alpha, pgen, ctx = saved
        dec, G, Esrc, wpgen = self._c(dec), self._c(G), self._c(Esrc), self._c(wpgen)
        logits = self._rowview(logits)
        Lt, V = int(lay["Lt"]), int(lay["V"])
        R, D = dec.shape
        Nc = R // Lt
        max_ni, max_vx = max(int(lay["max_ni"]), 1), int(lay["max_vx"])
        dev = dec.device
        ddec = torch.empty_like(dec)
        dG = torch.zeros_like(G)
        dE = torch.zeros_like(Esrc)
        du = torch.empty(R, dtype=torch.float32, device=dev)
        dw = torch.empty(2 * D, dtype=torch.float32, device=dev)
        db = torch.empty(1, dtype=torch.float32, device=dev)
        # d logits shares the row pitch of the logits (one ldl argument in the C ABI)
        dlog = torch.empty(R, logits.stride(0), dtype=torch.float32, device=dev)[:, :logits.shape[1]]
        dsc = torch.empty(R, max_ni, dtype=torch.float32, device=dev)
        dloss = dloss.to(torch.float32).reshape(1)
        self._call("svpc_pointer_bwd", Nc, Lt, D, V, max_ni, max_vx, _ptr(dec), _DT[dec.dtype], _ptr(G), _ptr(Esrc),
                   _ptr(lay["clip_src_off"]), _ptr(lay["clip_ni"]), _ptr(lay["clip_gi"]), _ptr(lay["clip_vx"]),
                   _ptr(lay["clip_P_off"]), _ptr(lay["tok_ptr"]), _ptr(lay["tok_ids"]), _ptr(wpgen), _ptr(logits),
                   _ll(logits.stride(0)), _ptr(lay["labels"]), _f(lay["smoothing"]), _ptr(P), _ptr(alpha), _ptr(pgen),
                   _ptr(ctx), _ptr(None if dP is None else self._c(dP)), _ptr(dloss), _ptr(ddec), _ptr(dG), _ptr(dE),
                   _ptr(du), _ptr(dw), _ptr(db), _ptr(dlog), _ptr(dsc))
        return ddec, dG, dE, dw, db, dlog

    def softmax_ls_fwd(self, logits, labels, V, smoothing):
        self._chk(logits, labels)
        logits = self._rowview(logits)
        R = logits.shape[0]
        P = torch.empty(R, V, dtype=torch.float32, device=logits.device)
        rows = torch.empty(R, dtype=torch.float32, device=logits.device)
        self._call("svpc_softmax_ls_fwd", R, int(V), _ptr(logits), _ll(logits.stride(0)), _ptr(labels), _f(smoothing), _ptr(P),
                   _ptr(rows))
        return P, rows

    def softmax_ls_bwd(self, dP, dloss, logits, labels, V, smoothing, P):
        R = logits.shape[0]
        dlog = torch.empty(R, logits.shape[1], dtype=torch.float32, device=logits.device)
        dloss = dloss.to(torch.float32).reshape(1)
        self._call("svpc_softmax_ls_bwd", R, int(V), _ptr(None if dP is None else self._c(dP)), _ptr(dloss), _ptr(P),
                   _ptr(labels), _f(smoothing), _ptr(dlog), _ll(dlog.stride(0)))
        return dlog

    # ------------------------------------------------------------------ textual re-simulator
    def gumbel_fwd(self, P, g, emb, lay, text_mask):
        self._chk(P, g, emb, text_mask)
        Lt, V = int(lay["Lt"]), int(lay["V"])
        Nc = lay["clip_P_off"].numel()
        W = emb.shape[1]
        dev = P.device
        assert g.numel() == P.numel()
        x = torch.empty(Nc, W, dtype=torch.float32, device=dev)
        idx = torch.empty(Nc * Lt, dtype=torch.int32, device=dev)
        yval = torch.empty(Nc * Lt, dtype=torch.float32, device=dev)
        self._call("svpc_gumbel_fwd", Nc, Lt, V, W, _ptr(P), _ptr(self._c(g)), _ptr(lay["clip_P_off"]), _ptr(lay["clip_vx"]),
                   _f(lay["tau"]), _ptr(self._c(emb)), _ptr(text_mask), _ptr(x), _ptr(idx), _ptr(yval))
        return x, idx, yval

    def gumbel_bwd(self, dx, P, g, emb, lay, text_mask, idx, yval):
        Lt, V = int(lay["Lt"]), int(lay["V"])
        Nc = lay["clip_P_off"].numel()
        W = emb.shape[1]
        dP = torch.empty_like(P)
        demb = torch.empty_like(emb, dtype=torch.float32)
        tv = torch.empty(Nc, V, dtype=torch.float32, device=P.device)
        self._call("svpc_gumbel_bwd", Nc, Lt, V, W, _ptr(P), _ptr(self._c(g)), _ptr(lay["clip_P_off"]), _ptr(lay["clip_vx"]),
                   _f(lay["tau"]), _ptr(self._c(emb)), _ptr(text_mask), _ptr(idx), _ptr(yval), _ptr(self._c(dx)), _ptr(dP),
                   _ptr(demb), _ptr(tv))
        return dP, demb

    # ------------------------------------------------------------------ greedy decoding
    def greedy_pick(self, scores, row_off, ld, row_width, width, row_noov, unk, next_in, ext_out, pos):
        self._chk(scores, next_in, ext_out)
        assert ext_out.dtype == torch.int64 and ext_out.stride(1) == 1 and next_in.dtype == torch.int32
        self._call("svpc_greedy_pick", int(next_in.numel()), _ptr(scores), _ptr(row_off), _ll(ld), _ptr(row_width), int(width),
                   _ptr(row_noov), int(unk), _ptr(next_in), _ptr(ext_out), _ll(ext_out.stride(0)), int(pos))

    # ------------------------------------------------------------------ losses
    def bce_fwd(self, p, y):
        self._chk(p, y)
        out = torch.empty((), dtype=torch.float32, device=p.device)
        self._call("svpc_bce_fwd", _ll(p.numel()), _ptr(self._c(p)), _ptr(self._c(y)), _ptr(out))
        return out

    def bce_bwd(self, p, y, g):
        dp = torch.empty_like(p)
        self._call("svpc_bce_bwd", _ll(p.numel()), _ptr(self._c(p)), _ptr(self._c(y)), _ptr(g.to(torch.float32).reshape(1)), _ptr(dp))
        return dp

    def asl_fwd(self, p, y):
        self._chk(p, y)
        out = torch.empty((), dtype=torch.float32, device=p.device)
        self._call("svpc_asl_fwd", p.shape[0], p.shape[1], _ptr(self._c(p)), _ptr(self._c(y)), _ptr(out))
        return out

    def asl_bwd(self, p, y, g):
        dp = torch.empty_like(p)
        self._call("svpc_asl_bwd", p.shape[0], p.shape[1], _ptr(self._c(p)), _ptr(self._c(y)), _ptr(g.to(torch.float32).reshape(1)),
                   _ptr(dp))
        return dp

    # ------------------------------------------------------------------ recipe encoder (BiLSTM)
    _lstm_rows = {}

    def lstm_sync(self, device):
        """Barrier counters of the persistent BiLSTM sweeps on ``device`` (row 0 forward, row 1 backward).  Column 2 is the
        time-out flag: set by a sweep whose grid barrier gave up (results of that step are wrong), never cleared by the
        kernels.  ``check_lstm_timeout`` reads it."""
        t = self._lstm_sync.get(device)
        if t is None:
            t = self._lstm_sync[device] = torch.zeros(2, 4, dtype=torch.int32, device=device)
        return t

    def check_lstm_timeout(self, flags=None):
        """Raise if a persistent BiLSTM sweep timed out at a grid barrier since the last check (synchronises unless
        ``flags``, a host copy of ``lstm_sync(device)``, is given)."""
        for dev, t in list(self._lstm_sync.items()):
            f = t if flags is None else flags
            if int(f[:, 2].sum()) != 0:
                t[:, 2].zero_()
                raise SvpcError("persistent BiLSTM sweep: a grid barrier timed out (the CTAs of a direction were not "
                                "co-resident - MPS / MIG SM limits or a concurrent kernel); the recurrent states of that step "
                                "are invalid.  Set SVPC_LSTM_STEPS=1 to use the per-step kernels.")
            if flags is not None:
                break

    def _lstm_use_seq(self, device, D):
        if not self.lstm_persistent:
            return False
        key = (device, int(D))
        ok = self._lstm_ok.get(key)
        if ok is None:
            with torch.cuda.device(device):
                ok = self._lstm_ok[key] = bool(self.lib.svpc_lstm_seq_supported(int(D)))
        return ok

    def _lstm_row_table(self, steps, device):
        key = (tuple(steps), str(device))
        t = self._lstm_rows.get(key)
        if t is None:
            B, Smax = len(steps), max(steps)
            cu = [0]
            for s in steps:
                cu.append(cu[-1] + s)
            fwd = [[cu[b] + j if j < steps[b] else -1 for b in range(B)] for j in range(Smax)]
            rev = [[cu[b] + steps[b] - 1 - j if j < steps[b] else -1 for b in range(B)] for j in range(Smax)]
            if len(self._lstm_rows) > 64:
                self._lstm_rows.clear()
            t = torch.tensor([fwd, rev], dtype=torch.int32).to(device)
            self._lstm_rows[key] = t
        return t

    def _lstm_whh(self, params, transposed):
        """(2, 4D, D) stack of both directions' W_hh (or (2, D, 4D) of their transposes), rebuilt when the optimiser
        has touched the parameters."""
        ws = [params["weight_hh_l0"], params["weight_hh_l0_reverse"]]
        key = ("whh_t" if transposed else "whh",) + tuple(w.data_ptr() for w in ws)
        ver = tuple(w._version for w in ws)
        hit = self._w_cache.get(key)
        if hit is not None and hit[0] == ver:
            return hit[1]
        G, D = ws[0].shape
        out = torch.empty(2, D, G, dtype=torch.float32, device=ws[0].device) if transposed else \
            torch.empty(2, G, D, dtype=torch.float32, device=ws[0].device)
        for di, w in enumerate(ws):
            w = self._c(w)
            if transposed:
                self._call("svpc_transpose_f32", G, D, _ptr(w), _ptr(out[di]))
            else:
                self._call("svpc_cast", _ll(w.numel()), _ptr(w), F32, _ptr(out[di]), F32)
        self._w_cache[key] = (ver, out)
        return out

    def _lstm_wih_cd(self, params, suf, cd):
        """W_ih of one direction in the compute dtype (row pitch padded for TMA), re-cast when the optimiser touched it."""
        w = params["weight_ih_l0" + suf]
        if cd == torch.float32:
            return w
        key = ("wih", suf, cd, w.data_ptr())
        hit = self._w_cache.get(key)
        if hit is not None and hit[0] == w._version:
            return hit[1]
        out = self.cast(self._c(w), cd)
        self._w_cache[key] = (w._version, out)
        return out

    def bilstm_fwd(self, x, cu_steps, steps, params, cd=torch.float32):
        """x (Nc, W) fp32 -> (Nc, D): both directions of nn.LSTM over every video's clips, summed (model.py:1022-1024).
        The input projection of all clips is one GEMM per direction; the recurrence is Smax fused step kernels
        (both directions, all videos, recurrent product + cell in one launch)."""
        self._chk(x)
        x = self._c(x)
        Nc, W = x.shape
        D = params["weight_hh_l0"].shape[1]
        B, Smax = len(steps), max(steps)
        dev = x.device
        rows = self._lstm_row_table(steps, dev)
        f32 = torch.float32
        xg = torch.empty(2, Nc, 4 * D, dtype=f32, device=dev)
        xc = self.cast(x, cd)         # bf16 path: the four batched products of the LSTM run on the tensor cores
        for di, suf in enumerate(("", "_reverse")):
            bias = params["bias_ih_l0" + suf] + params["bias_hh_l0" + suf]
            self.gemm(xc, self._lstm_wih_cd(params, suf, cd), f32, bias, out=xg[di])
        whh = self._lstm_whh(params, False)
        out = torch.zeros(Nc, D, dtype=f32, device=dev)
        hs = torch.empty(2, Nc, D, dtype=f32, device=dev)
        cs = torch.empty(2, Nc, D, dtype=f32, device=dev)
        hprev = torch.empty(2, Nc, D, dtype=f32, device=dev)
        gates = torch.empty(2, Nc, 4 * D, dtype=f32, device=dev)
        if self._lstm_use_seq(dev, D):
            # one launch for the whole sweep: W_hh slices resident in shared memory, per-direction grid barrier per step
            self._call("svpc_lstm_seq_fwd", B, D, Nc, Smax, _ptr(xg), _ptr(whh), _ptr(rows), _ptr(hs), _ptr(cs),
                       _ptr(gates), _ptr(out), _ptr(hprev), _ptr(self.lstm_sync(dev)[0]))
        else:
            for j in range(Smax):
                self._call("svpc_lstm_step_fwd", B, D, Nc, Smax, j, _ptr(xg), _ptr(whh), _ptr(rows), _ptr(hs), _ptr(cs),
                           _ptr(gates), _ptr(out), _ptr(hprev))
        return out, (rows, cs, gates, hprev)

    def bilstm_bwd(self, dout, x, cu_steps, steps, params, saved, cd=torch.float32):
        dout, x = self._c(dout), self._c(x)
        Nc, W = x.shape
        D = params["weight_hh_l0"].shape[1]
        B, Smax = len(steps), max(steps)
        dev = x.device
        rows, cs, gates, hprev = saved
        f32 = torch.float32
        whh_t = self._lstm_whh(params, True)
        dgates = torch.empty(2, Nc, 4 * D, dtype=f32, device=dev)
        dc = torch.zeros(2, B, D, dtype=f32, device=dev)
        if self._lstm_use_seq(dev, D):
            self._call("svpc_lstm_seq_bwd", B, D, Nc, Smax, _ptr(whh_t), _ptr(rows), _ptr(cs), _ptr(gates), _ptr(dout),
                       _ptr(dgates), _ptr(dc), _ptr(self.lstm_sync(dev)[1]))
        else:
            for j in range(Smax - 1, -1, -1):
                self._call("svpc_lstm_step_bwd", B, D, Nc, Smax, j, _ptr(whh_t), _ptr(rows), _ptr(cs), _ptr(gates),
                           _ptr(dout), _ptr(dgates), _ptr(dc))
        grads = {}
        dx = torch.empty(Nc, W, dtype=f32, device=dev)
        xc = self.cast(x, cd)
        dgc = self.cast(dgates.view(2 * Nc, 4 * D), cd).view(2, Nc, 4 * D)
        hpc = self.cast(hprev.view(2 * Nc, D), cd).view(2, Nc, D)
        for di, suf in enumerate(("", "_reverse")):
            dg = dgc[di]
            grads["weight_hh_l0" + suf] = self.gemm(dg.t(), hpc[di].t(), f32)
            grads["weight_ih_l0" + suf] = self.gemm(dg.t(), xc.t(), f32)
            db = self.colsum(dgates[di])
            grads["bias_ih_l0" + suf] = db
            grads["bias_hh_l0" + suf] = db.clone()
            self.gemm(dg, self._lstm_wih_cd(params, suf, cd).t(), f32, out=dx, accumulate=(di == 1))
        return dx, grads
