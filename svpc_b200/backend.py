"""Operator table of the SVPC hot path.

The model (``svpc_b200/model.py``) and the autograd glue (``svpc_b200/functional.py``) call every
device operator through ``backend.get()``.  The product has exactly ONE implementation:
``svpc_b200.cuda_backend`` - ctypes bindings of the C-ABI library ``libsvpc_b200.so`` (hand-written sm_100a CUDA,
``include/svpc_b200.h``).  There is no CPU implementation in the package: without the built library,
or on non-CUDA tensors, the operators raise.  ``tests/torch_backend.py`` (test infrastructure) holds a plain-torch
statement of each operator's contract; CPU tests install it with ``backend.set_backend`` to check the host-side
logic (batching, ragged layouts, autograd wiring) against the golden fixtures, and the GPU tests use it as the
per-operator checker.
"""
from __future__ import annotations

ACT_NONE, ACT_RELU, ACT_GELU, ACT_SIGMOID = 0, 1, 2, 3

_impl = None


def set_backend(impl) -> None:
    global _impl
    _impl = impl


def get():
    global _impl
    if _impl is None:
        from . import cuda_backend  # raises loudly when libsvpc_b200.so is missing
        _impl = cuda_backend.CudaBackend()
    return _impl
