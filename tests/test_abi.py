"""The C-ABI library loads, exports every symbol include/grappa_b200.h declares, and the ctypes prototypes agree with
the header on arity and on pointer-vs-scalar for every argument.  No GPU needed (no compute calls)."""
import ctypes
import os
import re

import pytest

from grappa_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declarations():
    src = open(os.path.join(ROOT, "include", "grappa_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    src = re.sub(r"//[^\n]*", " ", src)
    decls = {}
    for m in re.finditer(r"([A-Za-z_][\w\s\*]*?)\b(grappa_\w+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        arglist = [] if args in ("", "void") else [a.strip() for a in args.split(",")]
        decls[name] = (ret, arglist)
    return decls


def test_header_parses():
    d = _declarations()
    assert len(d) >= 25
    for required in ("grappa_gemm", "grappa_energy_fwd", "grappa_energy_vjp", "grappa_gnn_attn_fwd", "grappa_tuple_attn_fwd"):
        assert required in d


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "libgrappa_b200.so must be built (python grappa_b200/build.py)"
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declarations():
        assert hasattr(handle, name), f"{name} declared in the header but not exported"


def test_prototypes_match_header():
    decls = _declarations()
    assert set(decls) == set(_lib.PROTOTYPES), set(decls) ^ set(_lib.PROTOTYPES)
    for name, (ret, args) in decls.items():
        res, argtypes = _lib.PROTOTYPES[name]
        assert len(args) == len(argtypes), f"{name}: header has {len(args)} args, ctypes has {len(argtypes)}"
        for i, (a, t) in enumerate(zip(args, argtypes)):
            is_ptr = "*" in a
            assert is_ptr == (t is ctypes.c_void_p), f"{name} arg {i} ({a!r}) vs {t}"
            if not is_ptr:
                if "uint64_t" in a:
                    assert t is ctypes.c_uint64, f"{name} arg {i} ({a!r}) vs {t}"
                elif "int64_t" in a:
                    assert t is ctypes.c_int64, f"{name} arg {i} ({a!r}) vs {t}"
                elif "float" in a:
                    assert t is ctypes.c_float, f"{name} arg {i} ({a!r}) vs {t}"
                else:
                    assert t is ctypes.c_int, f"{name} arg {i} ({a!r}) vs {t}"


def test_loader_and_error_string():
    L = _lib.lib()
    assert L.grappa_abi_version() == 2
    # argument validation happens on the host before any launch: usable without a GPU
    rc = L.grappa_gemm(7, 1, 1, 1, None, 1, None, 1, None, 1, None, 0, None, 0, None, 0, 0.0, 0, None, 0, None, 0, None)
    assert rc != 0
    assert b"layout" in L.grappa_last_error()
    with pytest.raises(_lib.GrappaKernelError):
        _lib.check(rc)


def test_no_cpu_fallback():
    import torch
    from grappa_b200 import ops
    with pytest.raises(_lib.GrappaKernelError):
        ops.linear(torch.zeros(2, 4), torch.zeros(3, 4))
