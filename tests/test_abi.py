"""The C-ABI library loads without a GPU and exports every entry point include/svpc_b200.h declares (no compute calls)."""
import ctypes
import os
import re

from svpc_b200 import cuda_backend

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "svpc_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(svpc_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = declared_symbols()
    assert len(names) >= 46 and "svpc_gemm" in names and "svpc_optim_apply" in names
    lib = cuda_backend.load_library()
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.svpc_abi_version() == 1
    assert isinstance(lib.svpc_last_error(), bytes)


def test_missing_library_fails_loudly(tmp_path):
    try:
        cuda_backend.load_library(str(tmp_path / "nope.so"))
    except cuda_backend.SvpcError as e:
        assert "no CPU fallback" in str(e)
    else:
        raise AssertionError("a missing libsvpc_b200.so must raise")


def test_no_torch_types_in_the_abi():
    src = open(os.path.join(ROOT, "include", "svpc_b200.h")).read()
    assert "at::" not in src and "torch" not in re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    assert ctypes.sizeof(ctypes.c_void_p) == 8


def test_integration_stub_matches_header():
    """The ctypes stub INTEGRATION.md shows for svpc_gemm passes exactly as many arguments as the header declares (a
    maintainer pasting a stale stub would hand the stream where the dropout seed pointer belongs)."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "svpc_b200.h")).read()
    decl = re.search(r"int svpc_gemm\((.*?)\);", hdr, re.S).group(1)
    n_decl = len([a for a in decl.split(",") if a.strip()])
    doc = open(os.path.join(root, "INTEGRATION.md")).read()
    call = doc[doc.index("lib.svpc_gemm(") + len("lib.svpc_gemm("):]
    depth, end = 1, 0
    for end, ch in enumerate(call):
        depth += ch == "("
        depth -= ch == ")"
        if depth == 0:
            break
    body = re.sub(r"#[^\n]*", "", call[:end])
    args, depth, cur = [], 0, ""
    for ch in body:
        if ch == "," and depth == 0:
            args.append(cur)
            cur = ""
            continue
        depth += ch in "(["
        depth -= ch in ")]"
        cur += ch
    args.append(cur)
    n_stub = len([a for a in args if a.strip()])
    assert n_stub == n_decl, (n_stub, n_decl)
