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
    "grappa_gemm": (I, [I, I64, I64, I64, P, I64, P, I64, P, I64, P, I, P, I64, P, I64, F, c_uint64, P, I, P, I64, P]),
    "grappa_gemm_actgrad": (I, [I, I64, I64, I64, P, I64, P, I64, P, I64, P, I64, I, P, I64, P]),
    "grappa_gemm_ex": (I, [I, I64, I64, I64, P, I64, P, I64, P, I64, P, I, P, I64, P, I64, F, c_uint64, P, I, P, I64, P, I64, P, P]),
    "grappa_split_lo": (I, [I64, P, P, P]),
    "grappa_gemm_bf16x3": (I, [I64, I64, I64, P, I64, P, P, I64, P, I64, P, I, P, I64, P, I64, F, c_uint64, P, I, P, I64, P, I64, P, P]),
    "grappa_split_bf16_elems": (I64, [I64, I64]),
    "grappa_split_bf16": (I, [I64, P, I64, P, P, P]),
    "grappa_epilogue_bwd": (I, [I64, I64, P, I64, P, I64, P, I64, P, I64, I, F, c_uint64, P, P, I64, P, P, P]),
    "grappa_colsum_scratch_floats": (I64, [I64, I64]),
    "grappa_colsum": (I, [I64, I64, P, I64, P, P, P]),
    "grappa_colsum_noalloc": (I, [I64, I64, P, I64, P, P]),
    "grappa_layernorm_fwd": (I, [I64, I64, P, I64, P, P, P, I64, P, P, P]),
    "grappa_layernorm_bwd_scratch_floats": (I64, [I64, I64]),
    "grappa_layernorm_bwd": (I, [I64, I64, P, I64, P, I64, P, P, P, P, I64, P, P, P, P]),
    "grappa_featurize": (I, [I64, I, P, P, P, P, P, I64, P]),
    "grappa_gnn_attn_fwd": (I, [I64, I, I, P, P, P, P, P, P]),
    "grappa_gnn_attn_bwd": (I, [I64, I, I, P, P, P, P, P, P, P, P, P, P, P]),
    "grappa_tuple_gather_fwd": (I, [I64, I, I, I, P, I64, P, P, P, P]),
    "grappa_tuple_gather_bwd": (I, [I64, I, I, P, P, P, P, I64, P]),
    "grappa_tuple_attn_fwd": (I, [I64, I, I, I, P, P, P, P]),
    "grappa_tuple_attn_bwd": (I, [I64, I, I, I, P, P, P, P, P]),
    "grappa_token_residual": (I, [I64, I, I, P, P, P, P, P]),
    "grappa_token_residual_bwd": (I, [I64, I, I, P, P, P]),
    "grappa_symm_permute_fwd": (I, [I64, I, I, I, P, P, P, P]),
    "grappa_symm_permute_bwd": (I, [I64, I, I, I, P, P, P, P]),
    "grappa_symm_ln_fwd": (I, [I64, I, I, I, P, P, P, P, P, P, P, P]),
    "grappa_symm_ln_bwd_scratch_floats": (I64, [I64, I, I]),
    "grappa_symm_ln_bwd": (I, [I64, I, I, I, P, P, P, P, P, P, P, P, P, P, P]),
    "grappa_readout_finalize_fwd": (I, [I, I64, I, I, P, P, P, P, P, I, I, P, P, P, P]),
    "grappa_readout_finalize_bwd": (I, [I, I64, I, I, P, P, P, P, P, I, I, P, P, P, P]),
    "grappa_energy_plan_build": (I, [I64, I64, P, P, I64, P, I64, P, I64, P, I64, I, I, P, P, P, P, P]),
    "grappa_energy_fwd": (I, [I64, I64, I64, P, I64, P, P, P, P, I, P, P, P, P, P, I, P, I,
                              P, P, P, P, P, P, P, P, P, P, P, P, P]),
    "grappa_energy_vjp": (I, [I64, P, P, P, P, P, I64, P, I64, P, I64, P, I64, P, P, P, P, P, I, P, I,
                              P, P, P, P, P, P, P]),
    "grappa_csr_build": (I, [I64, I64, P, P, P]),
    "grappa_plan_build": (I, [I64, I64, P, P, P, P, P, I64, P, P, I, I, I, P, P, P, P]),
    "grappa_plan_fill": (I, [P, P, I64]),
    "grappa_plan_free": (None, [P]),
    "grappa_enumerate_tuples": (I, [I64, P, I64, P, P, P]),
    "grappa_pe_table_fwd": (I, [I64, I, I, P, I64, P, P]),
    "grappa_pe_table_bwd": (I, [I64, I, I, P, P, I64, P]),
    "grappa_row_gather_fwd": (I, [I64, I, P, I64, P, P, I64, P]),
    "grappa_row_gather_bwd": (I, [I64, I, P, I64, P, P, P, I64, P]),
    "grappa_collate_int": (I, [I64, P, P, P, P, P]),
    "grappa_collate_float": (I, [I64, P, P, P, P, P]),
    "grappa_molwise_loss": (I, [I64, I64, P, P, P, P, P, P, P, P, P, I, P, I, F, F, F, F, P, P, P, P, P, P, P, P]),
    "grappa_adamw": (I, [I64, P, P, P, P, F, F, F, F, F, I64, F, P]),
    "grappa_dp_shard": (I, [I64, I64, I, I, P, P]),
    "grappa_dp_exchange_adamw": (I, [I64, I64, I, I, P, P, P, P, P, P, I64, I64, I, F, F, F, F, F, I64, F, F, P]),
}

_lib = None


class GrappaKernelError(RuntimeError):
    pass


class Profiler:
    """Per-entry-point CUDA-event timing on the launching stream (used by bench.py for the roofline numbers).
    Enabled with `_lib.profiler = Profiler()`; every C-ABI call is then bracketed by two events."""

    def __init__(self):
        self.records = []          # (name, start_event, end_event, args)
        self.keepalive = []        # tensors whose pointers appear in the records

    def summary(self):
        import torch
        torch.cuda.synchronize()
        out = {}
        for name, a, b, args in self.records:
            d = out.setdefault(name, {"calls": 0, "ms": 0.0, "flops": 0.0, "bytes": 0.0})
            d["calls"] += 1
            d["ms"] += a.elapsed_time(b)
            if name in ("grappa_gemm", "grappa_gemm_actgrad", "grappa_gemm_ex"):
                d["flops"] += 2.0 * args[1] * args[2] * args[3]
            elif name == "grappa_gemm_bf16x3":
                d["flops"] += 2.0 * args[0] * args[1] * args[2]
            elif name == "grappa_energy_fwd":                      # 24 B per atom-conformation + 20 B per molecule-conformation
                d["bytes"] += args[2] * (24.0 * args[0] + 20.0 * args[1])
            elif name == "grappa_energy_vjp":
                d["bytes"] += args[0] * 24.0 * PROFILE_HINTS.get("n_atoms", 0)
            elif name == "grappa_gnn_attn_fwd":                    # (2N + E) rows of H*D floats
                d["bytes"] += (2.0 * args[0] + PROFILE_HINTS.get("n_edges", 0)) * args[1] * args[2] * 4.0
        return out


    def replay_ms(self, names, iters: int = 5, keep=None) -> float:
        """Re-issues the recorded calls of the given entry points (same arguments, same buffers) inside ONE CUDA graph
        and times its replay with CUDA events: back-to-back device time of those kernels, free of host launch gaps.
        The recorded buffers may have been recycled by the allocator; they are still mapped, which is all timing needs."""
        import ctypes
        import torch
        raw = _raw_lib()
        calls = [(getattr(raw, n), args) for n, _, _, args in self.records if n in names and (keep is None or keep(n, args))]
        if not calls:
            return 0.0
        torch.cuda.synchronize()
        stream = torch.cuda.Stream()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            sp = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            for fn, args in calls:
                rc = fn(*args[:-1], sp)
                if rc != 0:
                    raise GrappaKernelError(raw.grappa_last_error().decode())
        graph.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            graph.replay()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / iters


PROFILE_HINTS = {}
profiler = None


class _ProfiledLib:
    def __init__(self, handle, prof):
        self._h, self._p = handle, prof

    def __getattr__(self, name):
        fn = getattr(self._h, name)
        if not name.startswith("grappa_") or name in ("grappa_last_error", "grappa_launch_count", "grappa_abi_version",
                                                       "grappa_colsum_scratch_floats", "grappa_layernorm_bwd_scratch_floats", "grappa_symm_ln_bwd_scratch_floats",
                                                       "grappa_csr_build", "grappa_energy_plan_build", "grappa_enumerate_tuples", "grappa_plan_build",
                                                       "grappa_plan_fill", "grappa_plan_free"):
            return fn
        prof = self._p

        def timed(*args):
            import torch
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            rc = fn(*args)
            b.record()
            prof.records.append((name, a, b, args))
            return rc
        return timed


def lib():
    if profiler is not None:
        return _ProfiledLib(_raw_lib(), profiler)
    return _raw_lib()


def _raw_lib():
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
    return int(_raw_lib().grappa_launch_count())
