"""ctypes binding of libgrappa_b200.so (the C ABI declared in include/grappa_b200.h).

There is no fallback: if the library is missing or a symbol cannot be resolved, importing the kernels raises.
"""
import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int32, c_int64, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgrappa_b200.so")

P = c_void_p
I64 = c_int64
I = c_int
F = c_float

# name -> (restype, argtypes); mirrors include/grappa_b200.h one to one
PROTOTYPES = {
    "grappa_last_error": (c_char_p, []),
    "grappa_abi_version": (I, []),
    "grappa_launch_count": (I64, []),
    "grappa_gemm": (I, [I, I64, I64, I64, P, I64, P, I64, P, I64, P, I, P, I64, P, I64, F, c_uint64, P, I, P]),
    "grappa_epilogue_bwd": (I, [I64, I64, P, I64, P, I64, P, I64, P, I64, I, F, c_uint64, P, P, I64, P]),
    "grappa_colsum_scratch_floats": (I64, [I64, I64]),
    "grappa_colsum": (I, [I64, I64, P, I64, P, P, P]),
    "grappa_layernorm_fwd": (I, [I64, I64, P, I64, P, P, P, I64, P, P, P]),
    "grappa_layernorm_bwd_scratch_floats": (I64, [I64, I64]),
    "grappa_layernorm_bwd": (I, [I64, I64, P, I64, P, I64, P, P, P, P, I64, P, P, P, P]),
    "grappa_featurize": (I, [I64, I, P, P, P, P, P, I64, P]),
    "grappa_gnn_attn_fwd": (I, [I64, I, I, P, P, P, P, P, P]),
    "grappa_gnn_attn_bwd": (I, [I64, I, I, P, P, P, P, P, P, P, P, P, P, P]),
    "grappa_tuple_gather_fwd": (I, [I64, I, I, I, P, I64, P, P, P, P]),
    "grappa_tuple_gather_bwd": (I, [I64, I, I, P, P, P, P, I64, P]),
    "grappa_tuple_attn_fwd": (I, [I64, I, I, I, P, P, P]),
    "grappa_tuple_attn_bwd": (I, [I64, I, I, I, P, P, P, P]),
    "grappa_token_residual": (I, [I64, I, I, P, P, P, P]),
    "grappa_token_residual_bwd": (I, [I64, I, I, P, P, P]),
    "grappa_symm_permute_fwd": (I, [I64, I, I, I, P, P, P, P]),
    "grappa_symm_permute_bwd": (I, [I64, I, I, I, P, P, P, P]),
    "grappa_readout_finalize_fwd": (I, [I, I64, I, I, P, P, P, P, P, I, I, P, P, P, P]),
    "grappa_readout_finalize_bwd": (I, [I, I64, I, I, P, P, P, P, P, I, I, P, P, P, P]),
    "grappa_energy_plan_build": (I, [I64, I64, P, P, I64, P, I64, P, I64, P, I64, I, P, P, P, P, P, P, P, P]),
    "grappa_energy_fwd": (I, [I64, I64, I64, P, I64, P, P, P, P, P, P, I, I, I, P, P, P, P, P, I, P, I,
                              P, P, P, P, P, P, P, P, P, P, P, P, P]),
    "grappa_energy_vjp": (I, [I64, P, P, P, P, P, I64, P, I64, P, I64, P, I64, P, P, P, P, P, I, P, I,
                              P, P, P, P, P, P, P]),
    "grappa_csr_build": (I, [I64, I64, P, P, P]),
    "grappa_molwise_loss": (I, [I64, I64, P, P, P, P, P, P, P, P, P, I, P, I, F, F, F, F, P, P, P, P, P, P, P, P]),
    "grappa_adamw": (I, [I64, P, P, P, P, F, F, F, F, F, I64, F, P]),
}

_lib = None


class GrappaKernelError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GrappaKernelError(
                f"{LIB_PATH} not found: the CUDA library is required (no CPU fallback). "
                f"Build it with `python -c 'import __graft_entry__ as g; g.build()'` or `python grappa_b200/build.py`.")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(handle, name)     # AttributeError if the header and the library disagree
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int):
    if rc != 0:
        msg = lib().grappa_last_error().decode("utf-8", "replace")
        raise GrappaKernelError(f"libgrappa_b200 error {rc}: {msg}")


def launch_count() -> int:
    return int(lib().grappa_launch_count())
