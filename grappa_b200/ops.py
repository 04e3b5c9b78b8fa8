"""
torch.autograd.Function wrappers around the C ABI (include/grappa_b200.h).  PyTorch supplies device memory, the
current stream and the autograd tape; every arithmetic step is a kernel of libgrappa_b200.so.  There is no fallback:
CPU tensors are rejected.
"""
import ctypes
import weakref
from typing import Optional

import numpy as np
import torch

from . import _lib

import os as _os
# 0 = fp32 FFMA GEMM tiles, 1 = tcgen05 3xTF32 tiles (see grappa_gemm), 2 = tcgen05 bf16x3 tiles for the GEMMs whose B
# operand is a weight (forward and dX; grappa_gemm_bf16x3) and 3xTF32 for the rest (weight gradients)
MATH_MODE = int(_os.environ.get("GRAPPA_MATH_MODE", "2"))
ACT_NONE, ACT_ELU = 0, 1
# Optional {id(parameter): gradient buffer}: backward kernels write parameter gradients straight into these buffers
# (views of the trainer's flat gradient buffer) and autograd sees None, i.e. no per-parameter accumulate kernels.
# Valid because every parameter of the model is used exactly once per forward.
GRAD_SINK = None
FUSE_SKIP_GRAD = _os.environ.get("GRAPPA_FUSE_SKIP_GRAD", "1") != "0"    # Linear-ELU-Linear + skip: the skip's gradient added in the first layer's dX GEMM epilogue
FUSE_ELU_GRAD = _os.environ.get("GRAPPA_FUSE_ELU_GRAD", "1") != "0"      # Linear-ELU-Linear blocks: ELU' of the hidden tensor in the dX GEMM epilogue (see LinearFn)
# With a gradient sink nothing downstream in autograd consumes parameter gradients, so the weight-gradient GEMM, the
# bias column sum and the LayerNorm affine reductions are issued on a side stream and overlap with the activation-
# gradient chain (most of these kernels have far fewer tiles than SMs).  The step driver joins the side stream after
# backward; tensors read by the side work are parked in _DEFERRED until then.
_SIDE_STREAMS = {}
_DEFERRED = []


def norm_device(device) -> torch.device:
    """'cuda' and 'cuda:0' must name the same per-device state (streams, workspaces): always carry the index."""
    device = torch.device(device)
    if device.type == "cuda" and device.index is None:
        device = torch.device("cuda", torch.cuda.current_device())
    return device


SIDE_STREAM_PER_STREAM = _os.environ.get("GRAPPA_SIDE_PER_STREAM", "1") != "0"
SIDE_POOL = int(_os.environ.get("GRAPPA_SIDE_POOL", "2"))      # side streams per compute stream, used round-robin (measured: 1 -> 2: -1.5 %)
_SIDE_RR = {}


def _side_stream(device):
    """The side stream paired with the CURRENT stream: the four writers run on four streams, and one shared side stream
    would serialise their parameter-gradient work (and delay the GNN's behind it)."""
    device = norm_device(device)
    cs = torch.cuda.current_stream(device).cuda_stream if SIDE_STREAM_PER_STREAM else 0
    slot = 0
    if SIDE_POOL > 1:
        slot = _SIDE_RR.get((device, cs), 0)
        _SIDE_RR[(device, cs)] = (slot + 1) % SIDE_POOL
    key = (device, (cs, slot))
    st = _SIDE_STREAMS.get(key)
    if st is None:
        st = torch.cuda.Stream(device=device)
        _SIDE_STREAMS[key] = st
    return st


_ACTIVE_SIDE = {}          # side streams that received work since the last join (stale ones must not be waited on
                           # under graph capture: "dependency on uncaptured work")


def side_streams(device):
    """Side streams of `device` with gradient work pending since the last join."""
    device = norm_device(device)
    return [st for (d, _), st in _ACTIVE_SIDE.items() if d == device]


def join_side_work(device):
    """Make the current stream wait for all side-stream gradient work; called once after backward."""
    device = norm_device(device)
    cur = torch.cuda.current_stream(device)
    for key in [k for k in _ACTIVE_SIDE if k[0] == device]:
        cur.wait_stream(_ACTIVE_SIDE.pop(key))
    _DEFERRED.clear()


class _on_side_stream:
    """Context: fork the side stream from the current one (everything issued so far is visible to it)."""
    def __init__(self, device, keep):
        self.device, self.keep = device, keep

    def __enter__(self):
        cur = torch.cuda.current_stream(self.device)
        side = _side_stream(self.device)
        _ACTIVE_SIDE[(norm_device(self.device), side.cuda_stream)] = side
        ev = torch.cuda.Event()
        ev.record(cur)
        side.wait_event(ev)
        _DEFERRED.append(self.keep)
        self.ctx = torch.cuda.stream(side)
        self.ctx.__enter__()
        _IN_SIDE[0] += 1

    def __exit__(self, *a):
        _IN_SIDE[0] -= 1
        self.ctx.__exit__(*a)


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t: Optional[torch.Tensor]):
    if t is None:
        return None
    prof = _lib.profiler
    if prof is not None:
        prof.keepalive.append(t)      # recorded launches are replayed later: their buffers must stay allocated
    return ctypes.c_void_p(t.data_ptr())


def _req(t: torch.Tensor, name: str, dtype=torch.float32):
    if not t.is_cuda:
        raise _lib.GrappaKernelError(f"{name} must be a CUDA tensor (grappa_b200 has no CPU path)")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    return t


def _rows2d(t: torch.Tensor, name: str):
    """2-D view with unit inner stride; returns (tensor, leading dimension)."""
    _req(t, name)
    if t.dim() != 2:
        raise ValueError(f"{name} must be 2-D, got shape {tuple(t.shape)}")
    if t.stride(1) != 1 or (t.shape[0] > 1 and t.stride(0) < t.shape[1]):
        t = t.contiguous()
    return t, (t.stride(0) if t.shape[0] > 1 else max(t.shape[1], 1))


class DropoutSeeds:
    """Per-process dropout seed stream.  `base` is a host counter advanced per call; `dev` is an optional device
    uint64 that the step driver bumps between CUDA-graph replays so that replays draw fresh masks."""
    base = 0x5EED
    dev: Optional[torch.Tensor] = None

    @classmethod
    def next(cls) -> int:
        cls.base = (cls.base * 6364136223846793005 + 1442695040888963407) & 0xFFFFFFFFFFFFFFFF
        return cls.base

    @classmethod
    def manual_seed(cls, seed: int):
        cls.base = seed & 0xFFFFFFFFFFFFFFFF


_WORKSPACE = {}
WORKSPACE_FLOATS = 8 * 1024 * 1024       # 32 MB split-K scratch per device, allocated on first use


def _workspace(device):
    """Split-K scratch, one per (device, stream): GEMMs of the four parameter writers run on parallel streams."""
    device = norm_device(device)
    key = (device, torch.cuda.current_stream(device).cuda_stream)
    ws = _WORKSPACE.get(key)
    if ws is None:
        ws = torch.empty(WORKSPACE_FLOATS, device=device, dtype=torch.float32)
        _WORKSPACE[key] = ws
    return ws


# Split-K grid budget (CTAs) of the weight-gradient GEMMs issued on side streams.  They are background work: sized for
# the whole machine (0) they occupy every SM with a 192 KB CTA and each kernel of the critical chain has to wait for
# one of them to retire; un-split (budget 8: below any of their tile counts) they leave SMs free and need no reduce
# pass.  Measured on the cfg-2 step: 4.06 -> 3.93 ms.
SIDE_CTA_BUDGET = int(_os.environ.get("GRAPPA_SIDE_CTA_BUDGET", "8"))
_IN_SIDE = [0]
# streams on which the machine is saturated by several concurrent chains (the four writers): split-K there trades SM time
# for latency, which does not pay when throughput is the limit
WRITER_CTA_BUDGET = int(_os.environ.get("GRAPPA_WRITER_CTA_BUDGET", "8"))       # measured: 3.915 -> 3.893 ms
THROUGHPUT_STREAMS = set()


def _gemm_mode():
    budget = SIDE_CTA_BUDGET if _IN_SIDE[0] else 0
    if not budget and WRITER_CTA_BUDGET and torch.cuda.current_stream().cuda_stream in THROUGHPUT_STREAMS:
        budget = WRITER_CTA_BUDGET
    return min(MATH_MODE, 1) | (budget << 8) | (0x10000 if _IN_SIDE[0] else 0)     # bit 16: bookkeeping only (ignored by the library)


def gemm(layout, M, N, K, A, lda, B, ldb, C, ldc, bias=None, act=0, res1=None, ldr1=0, res2=None, ldr2=0,
         drop_p=0.0, seed=0, seed_dev=None, elu_out=None, ld_eo=0, b_lo=None):
    ws = _workspace(C.device)
    wsf = WORKSPACE_FLOATS if ws is not None else 0
    L = _lib.lib()
    if elu_out is None and b_lo is None:
        _lib.check(L.grappa_gemm(layout, M, N, K, _p(A), lda, _p(B), ldb, _p(C), ldc, _p(bias), act, _p(res1), ldr1, _p(res2), ldr2,
                                 float(drop_p), seed, _p(seed_dev), _gemm_mode(), _p(ws), wsf, _stream()))
    else:
        _lib.check(L.grappa_gemm_ex(layout, M, N, K, _p(A), lda, _p(B), ldb, _p(C), ldc, _p(bias), act, _p(res1), ldr1, _p(res2), ldr2,
                                    float(drop_p), seed, _p(seed_dev), _gemm_mode(), _p(ws), wsf,
                                    _p(elu_out), ld_eo, _p(b_lo), _stream()))


# Pre-split weights (see grappa_gemm_ex / grappa_split_lo): {id(parameter): lo plane}.  The training-step driver keeps a
# flat lo buffer next to its flat parameter buffer and refreshes it after every AdamW launch; without a driver (inference)
# lo planes are made on first use and cached per weight version.
WEIGHT_LO = None
WEIGHT_EPOCH = [0]        # bumped by every optimiser launch of this library (its kernels do not touch tensor versions)
USE_WEIGHT_LO = _os.environ.get("GRAPPA_WEIGHT_LO", "1") != "0"
_LO_CACHE = {}


def split_lo(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    x = _req(x, "x").contiguous()
    out = torch.empty_like(x) if out is None else out
    _lib.check(_lib.lib().grappa_split_lo(x.numel(), _p(x), _p(out), _stream()))
    return out


def weight_lo(w: torch.Tensor):
    """lo plane of a weight matrix for the B operand of the NT / NN GEMMs, or None."""
    if not USE_WEIGHT_LO or MATH_MODE < 1 or w.dim() != 2 or w.shape[1] % 4 or w.shape[1] < 32 or w.shape[0] < 32 or not w.is_contiguous():
        return None
    if WEIGHT_LO is not None:
        return WEIGHT_LO.get(id(w))
    key = id(w)
    hit = _LO_CACHE.get(key)
    # ids and device addresses are recycled once a tensor dies: an entry is only valid for the very object it was made for
    stamp = (w._version, w.data_ptr(), WEIGHT_EPOCH[0])
    if hit is not None and hit[0]() is w and hit[1] == stamp:
        return hit[2]
    lo = split_lo(w.detach())
    _LO_CACHE[key] = (weakref.ref(w, lambda _r, k=key: _LO_CACHE.pop(k, None)), stamp, lo)
    return lo


class WeightPlanes:
    """bf16 hi / lo planes of one weight matrix W (N, K) in both orientations (grappa_split_bf16): `nt` = planes of W for
    y = x W^T, `nn` = planes of W^T for dx = dy W; each (hi pointer, lo pointer, leading dimension)."""
    __slots__ = ("buf", "nt", "nn", "shape")

    def __init__(self, buf: torch.Tensor, base_elems: int, N: int, K: int):
        ldk, ldn = (K + 7) // 8 * 8, (N + 7) // 8 * 8
        p0 = buf.data_ptr() + 2 * base_elems
        self.buf, self.shape = buf, (N, K)
        self.nt = (p0, p0 + 2 * N * ldk, ldk)
        self.nn = (p0 + 4 * N * ldk, p0 + 4 * N * ldk + 2 * K * ldn, ldn)


WEIGHT_PLANES = None      # {id(parameter): WeightPlanes}, set by the training-step driver (refreshed after every AdamW)
_PLANE_CACHE = {}


def make_planes(w: torch.Tensor) -> WeightPlanes:
    """Planes of one (N, K) fp32 matrix, made now (one launch)."""
    w = _req(w, "weight").contiguous()
    N, K = w.shape
    L = _lib.lib()
    buf = torch.zeros(int(L.grappa_split_bf16_elems(N, K)), dtype=torch.int16, device=w.device)
    table = torch.tensor([0, N, K, 0], dtype=torch.int64, device=w.device)
    _lib.check(L.grappa_split_bf16(1, _p(table), ((N + 31) // 32) * ((K + 31) // 32), _p(w), _p(buf), _stream()))
    return WeightPlanes(buf, 0, N, K)


def weight_planes(w: torch.Tensor):
    """bf16 planes of a weight matrix for grappa_gemm_bf16x3, or None (mode off / shape not eligible)."""
    if MATH_MODE != 2 or w.dim() != 2 or w.shape[0] < 32 or w.shape[1] < 32 or not w.is_contiguous():
        return None
    if WEIGHT_PLANES is not None:
        return WEIGHT_PLANES.get(id(w))
    key = id(w)
    hit = _PLANE_CACHE.get(key)
    stamp = (w._version, w.data_ptr(), WEIGHT_EPOCH[0])
    if hit is not None and hit[0]() is w and hit[1] == stamp:
        return hit[2]
    pl = make_planes(w.detach())
    _PLANE_CACHE[key] = (weakref.ref(w, lambda _r, k=key: _PLANE_CACHE.pop(k, None)), stamp, pl)
    return pl


def _bf16_ok(A: torch.Tensor, lda: int, M: int, N: int, K: int) -> bool:
    return M >= 32 and N >= 32 and K >= 32 and lda % 4 == 0 and A.data_ptr() % 16 == 0


def gemm_bf16(M, N, K, A, lda, planes3, C, ldc, bias=None, act=0, res1=None, ldr1=0, res2=None, ldr2=0,
              drop_p=0.0, seed=0, seed_dev=None, elu_out=None, ld_eo=0, res1_idx=None):
    """C[M,N] = epi(A[M,K] . Bw[N,K]^T) with Bw given as bf16 planes (hi pointer, lo pointer, ld)."""
    ws = _workspace(C.device)
    wsf = WORKSPACE_FLOATS if ws is not None else 0
    _lib.check(_lib.lib().grappa_gemm_bf16x3(M, N, K, _p(A), lda, planes3[0], planes3[1], planes3[2], _p(C), ldc, _p(bias), act,
                                             _p(res1), ldr1, _p(res2), ldr2, float(drop_p), seed, _p(seed_dev), _gemm_mode(),
                                             _p(ws), wsf, _p(elu_out), ld_eo, res1_idx, _stream()))


def gemm_actgrad(layout, M, N, K, A, lda, B, ldb, C, ldc, elu_out, ld_eo, b_lo=None):
    gemm(layout, M, N, K, A, lda, B, ldb, C, ldc, elu_out=elu_out, ld_eo=ld_eo, b_lo=b_lo)


_SKIP_DW = _os.environ.get("GRAPPA_SKIP_DW", "0") == "1"      # timing experiments only: wrong gradients


class LinearFn(torch.autograd.Function):
    """y = dropout(act(x W^T + b)) + res1 + res2, columns padded to `out_ld` (pad columns are zero).

    Two flags let a Linear-ELU-Linear block skip the elementwise ELU backward pass over its hidden tensor:
      x_is_elu_out  (second layer): x is the plain ELU output of the preceding layer and feeds nothing else, so the
                    backward returns dL/d(pre-activation of that layer) -- ELU' is fused into the dX GEMM epilogue;
      grad_is_preact (first layer, act = ELU, no dropout / residual): the incoming gradient is already w.r.t. the
                    pre-activation."""

    @staticmethod
    def forward(ctx, x, weight, bias, res1, res2, act, drop_p, out_ld, x_is_elu_out=False, grad_is_preact=False, res1_gather=None,
                dx_add=None, res1_grad_to=None):
        # dx_add / res1_grad_to (a shared _GradHolder): in y = W2 f(W1 x) + x the skip's gradient reaches x twice; the second
        # layer hands dL/dy to the holder instead of returning it for res1, and the first layer's dX GEMM adds it in its
        # epilogue -- one fused add instead of an autograd accumulation kernel over the (rows, features) tensor
        # res1_gather = (plan, names): res1 is a TABLE of distinct token contents; row m of the residual is
        # res1[idx[m]], gathered inside the GEMM epilogue (its gradient is pull-reduced per table row in backward)
        x, lda = _rows2d(x, "x")
        w, ldb = _rows2d(weight, "weight")
        M, K = x.shape
        N = w.shape[0]
        if w.shape[1] != K:
            raise ValueError(f"weight {tuple(w.shape)} does not match input features {K}")
        ldc = int(out_ld) if out_ld else N
        y = torch.empty((M, ldc), device=x.device, dtype=torch.float32)
        if ldc > N:
            y[:, N:].zero_()
        r1 = r2 = None
        l1 = l2 = 0
        if res1 is not None:
            r1, l1 = _rows2d(res1, "res1")
        if res2 is not None:
            r2, l2 = _rows2d(res2, "res2")
        seed = DropoutSeeds.next() if drop_p > 0 else 0
        seed_dev = DropoutSeeds.dev if drop_p > 0 else None
        planes = weight_planes(weight) if ldb == K else None
        w_lo = None
        use_bf16 = planes is not None and _bf16_ok(x, lda, M, N, K)
        if res1_gather is not None and not use_bf16:        # other GEMM paths take a materialised residual
            r1 = torch.empty((M, r1.shape[1]), device=x.device, dtype=torch.float32)
            _lib.check(_lib.lib().grappa_row_gather_fwd(M, r1.shape[1], _p(res1), l1, res1_gather[0].ptr(res1_gather[1][0]), _p(r1),
                                                        r1.shape[1], _stream()))
            l1 = r1.shape[1]
        if use_bf16:
            gemm_bf16(M, N, K, x, lda, planes.nt, y, ldc, bias, act, r1, l1, r2, l2, drop_p, seed, seed_dev,
                      res1_idx=res1_gather[0].ptr(res1_gather[1][0]) if res1_gather is not None else None)
        else:
            w_lo = weight_lo(weight) if ldb == K else None
            gemm(0, M, N, K, x, lda, w, ldb, y, ldc, bias, act, r1, l1, r2, l2, drop_p, seed, seed_dev, b_lo=w_lo)
        ctx.w_lo = w_lo
        ctx.planes = planes
        ctx.res1_gather = (res1_gather, res1.shape[0]) if res1_gather is not None else None
        if res1_gather is not None and act != ACT_NONE:
            raise ValueError("a gathered residual needs act = none (the epilogue backward would need it per token)")
        ctx.save_for_backward(x, w, y if act != ACT_NONE else None, r1 if act != ACT_NONE else None,
                              r2 if act != ACT_NONE else None)
        if grad_is_preact and (act != ACT_ELU or drop_p > 0 or res1 is not None or res2 is not None):
            raise ValueError("grad_is_preact needs a plain ELU layer (no dropout, no residual)")
        ctx.cfg = (act, float(drop_p), seed, seed_dev, N, ldc, bias is not None, res1 is not None, res2 is not None)
        ctx.flags = (bool(x_is_elu_out), bool(grad_is_preact))
        ctx.dx_add, ctx.res1_grad_to = dx_add, res1_grad_to
        ctx.sink_ids = (id(weight), id(bias))
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, y, r1, r2 = ctx.saved_tensors
        act, drop_p, seed, seed_dev, N, ldc, has_bias, has_r1, has_r2 = ctx.cfg
        x_is_elu_out, grad_is_preact = ctx.flags
        if grad_is_preact:
            act = ACT_NONE                   # dy already is dL/d(pre-activation): the next layer's dX GEMM applied ELU'
        L = _lib.lib()
        dy, lddy = _rows2d(dy, "grad")
        M, K = x.shape
        lda = x.stride(0) if M > 1 else K
        sink = GRAD_SINK
        want_db = has_bias and ctx.needs_input_grad[2]
        db = None
        db_done = False
        if want_db:
            tgt = sink.get(ctx.sink_ids[1]) if sink is not None else None
            db = tgt if tgt is not None else torch.empty((N,), device=dy.device, dtype=torch.float32)
            scratch = torch.empty((int(L.grappa_colsum_scratch_floats(M, N)),), device=dy.device, dtype=torch.float32)
        if act != ACT_NONE or drop_p > 0:
            lddz = (N + 3) // 4 * 4
            dz = torch.empty((M, lddz), device=dy.device, dtype=torch.float32)
            _lib.check(L.grappa_epilogue_bwd(M, N, _p(dy), lddy, _p(y), ldc, _p(r1), r1.stride(0) if r1 is not None and M > 1 else ldc,
                                             _p(r2), r2.stride(0) if r2 is not None and M > 1 else ldc, act, drop_p, seed,
                                             _p(seed_dev), _p(dz), lddz, _p(db) if want_db else None,
                                             _p(scratch) if want_db else None, _stream()))
            db_done = want_db
        else:
            dz, lddz = dy, lddy
        dx = dw = None
        w_tgt = sink.get(ctx.sink_ids[0]) if (sink is not None and ctx.needs_input_grad[1]) else None
        b_sunk = want_db and sink is not None and sink.get(ctx.sink_ids[1]) is not None
        side_ok = (not ctx.needs_input_grad[1] or w_tgt is not None) and (not want_db or b_sunk) and M > 0

        def param_grads():
            nonlocal dw
            if ctx.needs_input_grad[1]:
                dw = w_tgt if w_tgt is not None else torch.empty((N, K), device=dy.device, dtype=torch.float32)
                if not _SKIP_DW:
                    gemm(2, N, K, M, dz, lddz, x, lda, dw, K)
                if w_tgt is not None:
                    dw = None
            if want_db and not db_done:
                _lib.check(L.grappa_colsum(M, N, _p(dz), lddz, _p(db), _p(scratch), _stream()))

        if side_ok and (ctx.needs_input_grad[1] or (want_db and not db_done)):
            with _on_side_stream(dy.device, (dz, x, scratch if want_db else None)):
                param_grads()
        else:
            param_grads()
        extra = ldx_extra = None
        if ctx.dx_add is not None and ctx.dx_add.value is not None:      # the skip's gradient, handed over by the second layer
            extra, ldx_extra = _rows2d(ctx.dx_add.value, "skip gradient")
            ctx.dx_add.value = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((M, K), device=dy.device, dtype=torch.float32)
            if ctx.planes is not None and _bf16_ok(dz, lddz, M, K, N):     # dx = dz W: the planes of W^T are K-major for it
                gemm_bf16(M, K, N, dz, lddz, ctx.planes.nn, dx, K, res1=extra, ldr1=ldx_extra or 0,
                          elu_out=x if x_is_elu_out else None, ld_eo=lda if x_is_elu_out else 0)
            elif x_is_elu_out:
                gemm_actgrad(1, M, K, N, dz, lddz, w, w.stride(0), dx, K, x, lda, b_lo=ctx.w_lo)
                if extra is not None:
                    dx = dx + extra
            else:
                gemm(1, M, K, N, dz, lddz, w, w.stride(0), dx, K, res1=extra, ldr1=ldx_extra or 0, b_lo=ctx.w_lo)
        if b_sunk:
            db = None
        dr = dy if ldc == N else dy[:, :N]
        dr1 = dr if (has_r1 and ctx.needs_input_grad[3]) else None
        if dr1 is not None and ctx.res1_gather is not None:        # gradient of the table: fixed-order pull over its CSR
            (plan, names), R = ctx.res1_gather
            dr1c = dr1.contiguous()
            dtab = torch.empty((R, N), device=dy.device, dtype=torch.float32)
            _lib.check(L.grappa_row_gather_bwd(R, N, _p(dr1c), N, plan.ptr(names[1]), plan.ptr(names[2]), _p(dtab), N, _stream()))
            dr1 = dtab
        if ctx.res1_grad_to is not None and dr1 is not None and ctx.res1_gather is None:
            ctx.res1_grad_to.value = dr1          # consumed by the first layer's dX GEMM (runs next in this backward pass)
            dr1 = None
        return (dx, dw, db, dr1, dr if (has_r2 and ctx.needs_input_grad[4]) else None, None, None, None, None, None, None, None, None)


class _GradHolder:
    __slots__ = ("value",)

    def __init__(self):
        self.value = None


def skip_grad_holder(x):
    """A holder for the gradient of a skip connection that starts at x, or None when the fusion does not apply."""
    return _GradHolder() if (FUSE_SKIP_GRAD and torch.is_grad_enabled() and x.requires_grad) else None


def linear(x, weight, bias=None, act=ACT_NONE, res1=None, res2=None, drop_p=0.0, out_ld=0, x_is_elu_out=False,
           grad_is_preact=False, res1_gather=None, dx_add=None, res1_grad_to=None):
    return LinearFn.apply(x, weight, bias, res1, res2, act, drop_p, out_ld, x_is_elu_out, grad_is_preact, res1_gather,
                          dx_add, res1_grad_to)


def elu_mlp(x, w1, b1, w2, b2, act2=ACT_NONE, res1=None, res2=None, drop_p=0.0):
    """Linear-ELU-Linear with the hidden tensor's ELU backward fused into the second layer's dX GEMM."""
    fuse = FUSE_ELU_GRAD and torch.is_grad_enabled()
    # the block's input is also its skip connection: its two gradient contributions are summed inside the first layer's dX GEMM
    holder = skip_grad_holder(x) if res1 is x else None
    z = linear(x, w1, b1, act=ACT_ELU, grad_is_preact=fuse, dx_add=holder)
    return linear(z, w2, b2, act=act2, res1=res1, res2=res2, drop_p=drop_p, x_is_elu_out=fuse, res1_grad_to=holder)


class LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta):
        x, ldx = _rows2d(x, "x")
        M, D = x.shape
        y = torch.empty((M, D), device=x.device, dtype=torch.float32)
        stats = torch.empty((2, max(M, 1)), device=x.device, dtype=torch.float32)
        _lib.check(_lib.lib().grappa_layernorm_fwd(M, D, _p(x), ldx, _p(gamma), _p(beta), _p(y), D, _p(stats[0]), _p(stats[1]),
                                                   _stream()))
        ctx.save_for_backward(x, gamma, stats)
        ctx.sink_ids = (id(gamma), id(beta))
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gamma, stats = ctx.saved_tensors
        L = _lib.lib()
        dy, lddy = _rows2d(dy, "grad")
        M, D = x.shape
        ldx = x.stride(0) if M > 1 else D
        dx = torch.empty((M, D), device=dy.device, dtype=torch.float32)
        need_affine = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
        dg = db = scratch = None
        sunk = False
        if need_affine:
            sink = GRAD_SINK
            tg = sink.get(ctx.sink_ids[0]) if sink is not None else None
            tb = sink.get(ctx.sink_ids[1]) if sink is not None else None
            sunk = tg is not None and tb is not None
            dg = tg if sunk else torch.empty((D,), device=dy.device, dtype=torch.float32)
            db = tb if sunk else torch.empty((D,), device=dy.device, dtype=torch.float32)
            scratch = torch.empty((int(L.grappa_layernorm_bwd_scratch_floats(M, D)),), device=dy.device, dtype=torch.float32)
        # (measured: issuing the affine reductions on a side stream instead costs 2 % of the step: they are short, and the
        #  fork / join events outweigh what they free on the chain)
        _lib.check(L.grappa_layernorm_bwd(M, D, _p(dy), lddy, _p(x), ldx, _p(gamma), _p(stats[0]), _p(stats[1]), _p(dx), D,
                                          _p(dg), _p(db), _p(scratch), _stream()))
        if sunk:
            dg = db = None
        return dx, dg, db


def layer_norm(x, gamma, beta):
    return LayerNormFn.apply(x, gamma, beta)


def featurize(blocks, charge: torch.Tensor) -> torch.Tensor:
    """blocks: list of (tensor or None, width).  Returns (N, sum(width) + 16) features (no gradient: inputs)."""
    n = charge.shape[0]
    ptrs = (ctypes.c_void_p * 8)()
    widths = (ctypes.c_int32 * 8)()
    lds = (ctypes.c_int32 * 8)()
    keep = []
    total = 0
    for i, (t, wdt) in enumerate(blocks):
        if t is not None:
            t = _req(t, "feature").reshape(n, -1)
            t, ld = _rows2d(t, "feature")
            keep.append(t)
            ptrs[i] = t.data_ptr()
            lds[i] = ld
        else:
            ptrs[i] = None
            lds[i] = 0
        widths[i] = wdt
        total += wdt
    out = torch.empty((n, total + 16), device=charge.device, dtype=torch.float32)
    charge = _req(charge, "partial_charge").contiguous()
    _lib.check(_lib.lib().grappa_featurize(n, len(blocks), ptrs, widths, lds, _p(charge), _p(out), total + 16, _stream()))
    return out


class GnnAttnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ft, plan, n_heads):
        ft, ld = _rows2d(ft, "ft")
        if ld != ft.shape[1]:
            ft = ft.contiguous()
        N, HD = ft.shape
        D = HD // n_heads
        out = torch.empty_like(ft)
        lse = torch.empty((N, n_heads), device=ft.device, dtype=torch.float32)
        _lib.check(_lib.lib().grappa_gnn_attn_fwd(N, n_heads, D, _p(ft), plan.ptr("in_ptr"), plan.ptr("in_nbr"), _p(out), _p(lse),
                                                  _stream()))
        ctx.save_for_backward(ft, out, lse)
        ctx.plan, ctx.n_heads = plan, n_heads
        return out

    @staticmethod
    def backward(ctx, dout):
        ft, out, lse = ctx.saved_tensors
        plan, H = ctx.plan, ctx.n_heads
        dout = dout.contiguous()
        N, HD = ft.shape
        delta = torch.empty((N, H), device=ft.device, dtype=torch.float32)
        dft = torch.empty_like(ft)
        _lib.check(_lib.lib().grappa_gnn_attn_bwd(N, H, HD // H, _p(ft), _p(out), _p(lse), _p(dout), plan.ptr("in_ptr"),
                                                  plan.ptr("in_nbr"), plan.ptr("out_ptr"), plan.ptr("out_nbr"), _p(delta), _p(dft),
                                                  _stream()))
        return dft, None, None


def gnn_attention(ft, plan, n_heads):
    return GnnAttnFn.apply(ft, plan, n_heads)


class TupleGatherFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, plan, level, n_seq, F, W, pe):
        a, lda = _rows2d(a, "atom_feats")
        T = plan.T[level]
        x = torch.empty((T * n_seq, W), device=a.device, dtype=torch.float32)
        _lib.check(_lib.lib().grappa_tuple_gather_fwd(T, n_seq, F, W, _p(a), lda, plan.ptr("idx_" + level), _p(pe), _p(x), _stream()))
        ctx.cfg = (plan, level, F, W, a.shape[0], lda)
        return x

    @staticmethod
    def backward(ctx, dx):
        plan, level, F, W, N, lda = ctx.cfg
        dx = dx.contiguous()
        ldda = max(lda, W)
        da = torch.empty((N, ldda), device=dx.device, dtype=torch.float32)
        if ldda > W:
            da[:, W:].zero_()
        _lib.check(_lib.lib().grappa_tuple_gather_bwd(N, F, W, _p(dx), plan.ptr("inc_ptr_" + level), plan.ptr("inc_tok_" + level),
                                                      _p(da), ldda, _stream()))
        return da, None, None, None, None, None, None


def tuple_gather(a, plan, level, n_seq, F, W, pe=None):
    return TupleGatherFn.apply(a, plan, level, n_seq, F, W, pe)


class PeTableFn(torch.autograd.Function):
    """U[2a + pe] = [P[a, :F], pe] for pe in {0, 1}: the distinct transformer inputs of a level with a PE column."""

    @staticmethod
    def forward(ctx, a, F, W):
        a, lda = _rows2d(a, "atom_feats")
        N = a.shape[0]
        u = torch.empty((2 * N, W), device=a.device, dtype=torch.float32)
        _lib.check(_lib.lib().grappa_pe_table_fwd(N, F, W, _p(a), lda, _p(u), _stream()))
        ctx.cfg = (N, F, W, lda)
        return u

    @staticmethod
    def backward(ctx, du):
        N, F, W, lda = ctx.cfg
        du = du.contiguous()
        ldda = max(lda, W)
        da = torch.empty((N, ldda), device=du.device, dtype=torch.float32)
        _lib.check(_lib.lib().grappa_pe_table_bwd(N, F, W, _p(du), _p(da), ldda, _stream()))
        return da, None, None


def pe_table(a, F, W):
    return PeTableFn.apply(a, F, W)


class RowGatherFn(torch.autograd.Function):
    """X[r] = table[idx[r]] with a pull-style, fixed-order backward over the CSR (ptr, tok) of idx."""

    @staticmethod
    def forward(ctx, table, plan, names, rows):
        table, ldt = _rows2d(table, "table")
        D = table.shape[1]
        x = torch.empty((rows, D), device=table.device, dtype=torch.float32)
        _lib.check(_lib.lib().grappa_row_gather_fwd(rows, D, _p(table), ldt, plan.ptr(names[0]), _p(x), D, _stream()))
        ctx.cfg = (plan, names, table.shape[0], D)
        return x

    @staticmethod
    def backward(ctx, dx):
        plan, names, R, D = ctx.cfg
        dx = dx.contiguous()
        dt = torch.empty((R, D), device=dx.device, dtype=torch.float32)
        _lib.check(_lib.lib().grappa_row_gather_bwd(R, D, _p(dx), D, plan.ptr(names[1]), plan.ptr(names[2]), _p(dt), D, _stream()))
        return dt, None, None, None


def row_gather(table, plan, names, rows):
    """names = (idx, ptr, tok) array names of the batch plan."""
    return RowGatherFn.apply(table, plan, names, rows)


class TupleAttnFn(torch.autograd.Function):
    """Attention over the tokens of every tuple.  With `gather = (plan, names)` the input is the QKV TABLE of distinct
    token contents and the per-token gather is fused into the kernels (forward reads table rows through the token index;
    backward writes per-token dqkv and pull-reduces it into the table in fixed order)."""

    @staticmethod
    def forward(ctx, qkv, T, n_seq, n_heads, gather=None):
        qkv = _req(qkv, "qkv").contiguous()
        E = qkv.shape[1] // 3
        out = torch.empty((T * n_seq, E), device=qkv.device, dtype=torch.float32)
        idx = gather[0].ptr(gather[1][0]) if gather is not None else None
        _lib.check(_lib.lib().grappa_tuple_attn_fwd(T, n_seq, n_heads, E // n_heads, _p(qkv), idx, _p(out), _stream()))
        ctx.save_for_backward(qkv)
        ctx.cfg = (T, n_seq, n_heads, E, gather)
        return out

    @staticmethod
    def backward(ctx, dout):
        (qkv,) = ctx.saved_tensors
        T, n_seq, n_heads, E, gather = ctx.cfg
        dout = dout.contiguous()
        L = _lib.lib()
        if gather is None:
            dqkv = torch.empty_like(qkv)
            _lib.check(L.grappa_tuple_attn_bwd(T, n_seq, n_heads, E // n_heads, _p(qkv), None, _p(dout), _p(dqkv), _stream()))
            return dqkv, None, None, None, None
        plan, names = gather
        dtok = torch.empty((T * n_seq, 3 * E), device=qkv.device, dtype=torch.float32)
        _lib.check(L.grappa_tuple_attn_bwd(T, n_seq, n_heads, E // n_heads, _p(qkv), plan.ptr(names[0]), _p(dout), _p(dtok), _stream()))
        dtab = torch.empty_like(qkv)
        _lib.check(L.grappa_row_gather_bwd(qkv.shape[0], 3 * E, _p(dtok), 3 * E, plan.ptr(names[1]), plan.ptr(names[2]), _p(dtab), 3 * E,
                                           _stream()))
        return dtab, None, None, None, None


def tuple_attention(qkv, T, n_seq, n_heads, gather=None):
    return TupleAttnFn.apply(qkv, T, n_seq, n_heads, gather)


class TokenResidualFn(torch.autograd.Function):
    """r = x3 + x * (col < F ? 2 : 1).  With `gather = (plan, names)`, `x` is the table of distinct token contents, read
    through the token index inside the kernel; its gradient is pull-reduced per table row (fixed order) and scaled there,
    so neither the gathered input nor its per-token gradient is ever written."""

    @staticmethod
    def forward(ctx, x3, x, F, gather=None):
        x3 = _req(x3, "x3").contiguous()
        x = _req(x, "x").contiguous()
        M, W = x3.shape
        r = torch.empty_like(x3)
        idx = gather[0].ptr(gather[1][0]) if gather is not None else None
        _lib.check(_lib.lib().grappa_token_residual(M, W, F, _p(x3), _p(x), idx, _p(r), _stream()))
        ctx.cfg = (M, W, F, gather, x.shape[0])
        return r

    @staticmethod
    def backward(ctx, dr):
        M, W, F, gather, R = ctx.cfg
        dr = dr.contiguous()
        dx = None
        L = _lib.lib()
        if ctx.needs_input_grad[1]:
            if gather is None:
                dx = torch.empty_like(dr)
                _lib.check(L.grappa_token_residual_bwd(M, W, F, _p(dr), _p(dx), _stream()))
            else:
                plan, names = gather
                tmp = torch.empty((R, W), device=dr.device, dtype=torch.float32)
                _lib.check(L.grappa_row_gather_bwd(R, W, _p(dr), W, plan.ptr(names[1]), plan.ptr(names[2]), _p(tmp), W, _stream()))
                dx = torch.empty_like(tmp)
                _lib.check(L.grappa_token_residual_bwd(R, W, F, _p(tmp), _p(dx), _stream()))
        return dr, dx, None, None


def token_residual(x3, x, F, gather=None):
    return TokenResidualFn.apply(x3, x, F, gather)


class SymmPermuteFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, T, n_seq, perms):
        x = _req(x, "x").contiguous()
        W = x.shape[1]
        P = len(perms)
        flat = (ctypes.c_int32 * (P * n_seq))(*[int(v) for p in perms for v in p])
        y = torch.empty((P * T, n_seq * W), device=x.device, dtype=torch.float32)
        _lib.check(_lib.lib().grappa_symm_permute_fwd(T, n_seq, W, P, flat, _p(x), _p(y), _stream()))
        ctx.cfg = (T, n_seq, W, P, flat)
        return y

    @staticmethod
    def backward(ctx, dy):
        T, n_seq, W, P, flat = ctx.cfg
        dy = dy.contiguous()
        dx = torch.empty((T * n_seq, W), device=dy.device, dtype=torch.float32)
        _lib.check(_lib.lib().grappa_symm_permute_bwd(T, n_seq, W, P, flat, _p(dy), _p(dx), _stream()))
        return dx, None, None, None


class SymmLayerNormFn(torch.autograd.Function):
    """The symmetriser's permuted copies with the first LayerNorm fused in (grappa_symm_ln_fwd / _bwd): statistics once per
    tuple, the un-normalised permuted copy is never written."""

    @staticmethod
    def forward(ctx, x, gamma, beta, T, n_seq, perms):
        x = _req(x, "x").contiguous()
        W = x.shape[1]
        P = len(perms)
        flat = (ctypes.c_int32 * (P * n_seq))(*[int(v) for p in perms for v in p])
        y = torch.empty((P * T, n_seq * W), device=x.device, dtype=torch.float32)
        stats = torch.empty((2, max(T, 1)), device=x.device, dtype=torch.float32)
        _lib.check(_lib.lib().grappa_symm_ln_fwd(T, n_seq, W, P, flat, _p(x), _p(gamma), _p(beta), _p(y), _p(stats[0]), _p(stats[1]),
                                                 _stream()))
        ctx.save_for_backward(x, gamma, stats)
        ctx.cfg = (T, n_seq, W, P, flat)
        ctx.sink_ids = (id(gamma), id(beta))
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gamma, stats = ctx.saved_tensors
        T, n_seq, W, P, flat = ctx.cfg
        L = _lib.lib()
        dy = dy.contiguous()
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        need_affine = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
        dg = db = scratch = None
        sunk = False
        if need_affine:
            sink = GRAD_SINK
            tg = sink.get(ctx.sink_ids[0]) if sink is not None else None
            tb = sink.get(ctx.sink_ids[1]) if sink is not None else None
            sunk = tg is not None and tb is not None
            D = n_seq * W
            dg = tg if sunk else torch.empty((D,), device=dy.device, dtype=torch.float32)
            db = tb if sunk else torch.empty((D,), device=dy.device, dtype=torch.float32)
            scratch = torch.empty((int(L.grappa_symm_ln_bwd_scratch_floats(T, n_seq, W)),), device=dy.device, dtype=torch.float32)
        _lib.check(L.grappa_symm_ln_bwd(T, n_seq, W, P, flat, _p(dy), _p(x), _p(gamma), _p(stats[0]), _p(stats[1]), _p(dx), _p(dg), _p(db),
                                        _p(scratch), _stream()))
        if sunk:
            dg = db = None
        return dx, dg, db, None, None, None


def symm_layer_norm(x, gamma, beta, T, n_seq, perms):
    return SymmLayerNormFn.apply(x, gamma, beta, T, n_seq, perms)


def symm_permute(x, T, n_seq, perms):
    return SymmPermuteFn.apply(x, T, n_seq, perms)


class ReadoutFinalizeFn(torch.autograd.Function):
    """kind 0 bond / 1 angle -> (eq, k); kind 2 gated torsion / 3 ungated torsion -> k (T, n_per)."""

    @staticmethod
    def forward(ctx, s, kind, T, prefactors, consts, kstd, kmean, n_per, col0):
        s = _req(s, "s").contiguous()
        P = len(prefactors)
        n_out = s.shape[1]
        pref = (ctypes.c_float * P)(*[float(v) for v in prefactors])
        cst = (ctypes.c_float * 8)(*([float(v) for v in consts] + [0.0] * (8 - len(consts))))
        c = torch.empty((T, n_out), device=s.device, dtype=torch.float32)
        if kind <= 1:
            eq = torch.empty((T,), device=s.device, dtype=torch.float32)
            k = torch.empty((T,), device=s.device, dtype=torch.float32)
        else:
            eq = None
            k = torch.empty((T, n_per), device=s.device, dtype=torch.float32)
        _lib.check(_lib.lib().grappa_readout_finalize_fwd(kind, T, P, n_out, _p(s), pref, cst, _p(kstd), _p(kmean), n_per, col0,
                                                          _p(c), _p(eq), _p(k), _stream()))
        ctx.save_for_backward(c, kstd, kmean)
        ctx.cfg = (kind, T, P, n_out, pref, cst, n_per, col0)
        if kind <= 1:
            return eq, k
        return k

    @staticmethod
    def backward(ctx, *grads):
        c, kstd, kmean = ctx.saved_tensors
        kind, T, P, n_out, pref, cst, n_per, col0 = ctx.cfg
        if kind <= 1:
            d_eq, d_k = grads
        else:
            d_eq, d_k = None, grads[0]
        d_eq = d_eq.contiguous() if d_eq is not None else None
        d_k = d_k.contiguous() if d_k is not None else None
        ds = torch.empty((P * T, n_out), device=c.device, dtype=torch.float32)
        _lib.check(_lib.lib().grappa_readout_finalize_bwd(kind, T, P, n_out, _p(c), pref, cst, _p(kstd), _p(kmean), n_per, col0,
                                                          _p(d_eq), _p(d_k), _p(ds), _stream()))
        return ds, None, None, None, None, None, None, None, None


def readout_finalize(s, kind, T, prefactors, consts, kstd=None, kmean=None, n_per=0, col0=0):
    return ReadoutFinalizeFn.apply(s, kind, T, prefactors, consts, kstd, kmean, n_per, col0)


class EnergyFn(torch.autograd.Function):
    """(k2, eq2, k3, eq3, k4, k4i, xyz) -> (energy (n_mol,C), gradient (N,C,3), term_energy (4,n_mol,C), x..., e...)."""

    @staticmethod
    def forward(ctx, k2, eq2, k3, eq3, k4, k4i, xyz, plan, term_mask, want_coords):
        xyz = _req(xyz, "xyz").contiguous()
        N, C, _ = xyz.shape
        dev = xyz.device
        T = plan.T

        def prep(t, shape):
            if t is None:
                return torch.zeros(shape, device=dev, dtype=torch.float32)
            return _req(t, "parameter").contiguous()
        P4 = k4.shape[1] if k4 is not None and k4.dim() == 2 else 1
        P4i = k4i.shape[1] if k4i is not None and k4i.dim() == 2 else 1
        k2, eq2 = prep(k2, (T["n2"],)), prep(eq2, (T["n2"],))
        k3, eq3 = prep(k3, (T["n3"],)), prep(eq3, (T["n3"],))
        k4, k4i = prep(k4, (T["n4"], P4)), prep(k4i, (T["n4_improper"], P4i))
        es = plan.energy_sizes
        n_mol = plan.n_mol
        energy = torch.empty((n_mol, C), device=dev, dtype=torch.float32)
        term = torch.empty((4, n_mol, C), device=dev, dtype=torch.float32)
        grad = torch.empty((N, C, 3), device=dev, dtype=torch.float32)
        lv = ("n2", "n3", "n4", "n4_improper")
        xs = [torch.empty((T[l], C), device=dev, dtype=torch.float32) if want_coords else None for l in lv]
        ts = [torch.empty((T[l], C), device=dev, dtype=torch.float32) if want_coords else None for l in lv]
        partial = torch.empty((es["n_blocks"], 4, C), device=dev, dtype=torch.float32) if es["any_multi"] else None
        _lib.check(_lib.lib().grappa_energy_fwd(
            N, n_mol, C, _p(xyz), es["n_blocks"], plan.ptr("e_blk"), plan.ptr("e_prog"), plan.ptr("e_mol_toff"),
            plan.energy_sizes_array().ctypes.data, term_mask,
            _p(k2), _p(eq2), _p(k3), _p(eq3), _p(k4), P4, _p(k4i), P4i, _p(energy), _p(term), _p(grad),
            _p(xs[0]), _p(xs[1]), _p(xs[2]), _p(xs[3]), _p(ts[0]), _p(ts[1]), _p(ts[2]), _p(ts[3]), _p(partial), _stream()))
        ctx.save_for_backward(xyz, k2, eq2, k3, eq3, k4, k4i)
        ctx.cfg = (plan, P4, P4i, term_mask)
        outs = [energy, grad, term] + [x if x is not None else torch.empty(0, device=dev) for x in xs] + \
               [t if t is not None else torch.empty(0, device=dev) for t in ts]
        ctx.mark_non_differentiable(*outs[2:])
        return tuple(outs)

    @staticmethod
    def backward(ctx, g_energy, g_grad, *unused):
        xyz, k2, eq2, k3, eq3, k4, k4i = ctx.saved_tensors
        plan, P4, P4i, term_mask = ctx.cfg
        if term_mask != 15:
            raise NotImplementedError("backward through a term-masked Energy evaluation is not supported")
        T = plan.T
        dev = xyz.device
        C = xyz.shape[1]
        g_energy = g_energy.contiguous() if g_energy is not None else None
        g_grad = g_grad.contiguous() if g_grad is not None else None
        dk2, deq2 = torch.empty_like(k2), torch.empty_like(eq2)
        dk3, deq3 = torch.empty_like(k3), torch.empty_like(eq3)
        dk4, dk4i = torch.empty_like(k4), torch.empty_like(k4i)
        s, n = plan.offsets["tuple_mol"]
        _lib.check(_lib.lib().grappa_energy_vjp(
            C, _p(xyz), _p(g_energy), _p(g_grad), plan.ptr("tuple_mol"),
            plan.ptr("idx_n2"), T["n2"], plan.ptr("idx_n3"), T["n3"], plan.ptr("idx_n4"), T["n4"],
            plan.ptr("idx_n4_improper"), T["n4_improper"],
            _p(k2), _p(eq2), _p(k3), _p(eq3), _p(k4), P4, _p(k4i), P4i,
            _p(dk2), _p(deq2), _p(dk3), _p(deq3), _p(dk4), _p(dk4i), _stream()))
        ng = ctx.needs_input_grad
        return (dk2 if ng[0] else None, deq2 if ng[1] else None, dk3 if ng[2] else None, deq3 if ng[3] else None,
                dk4 if ng[4] else None, dk4i if ng[5] else None, None, None, None, None)


class MolwiseLossFn(torch.autograd.Function):
    """Fused molecule-wise loss; the backward re-runs the kernel with gradient outputs scaled by the upstream scalar."""

    @staticmethod
    def _call(plan, weights, energy, gradient, k4, k4i, energy_ref, gradient_ref, is_dummy, upstream, parts, loss, dE, dG, dk4, dk4i):
        ew, gw, preg, ireg = weights
        n_mol, C = energy.shape
        _lib.check(_lib.lib().grappa_molwise_loss(
            n_mol, C, plan.ptr("mol_atom_off"), plan.ptr("mol_off_n4"), plan.ptr("mol_off_n4_improper"),
            _p(energy), _p(energy_ref), _p(gradient), _p(gradient_ref), _p(is_dummy),
            _p(k4), k4.shape[1] if k4.dim() == 2 else 1, _p(k4i), k4i.shape[1] if k4i.dim() == 2 else 1,
            ew, gw, preg, ireg, _p(upstream), _p(parts), _p(loss), _p(dE), _p(dG), _p(dk4), _p(dk4i), _stream()))

    @staticmethod
    def forward(ctx, energy, gradient, k4, k4i, energy_ref, gradient_ref, is_dummy, plan, weights):
        energy = _req(energy, "energy").contiguous()
        gradient = _req(gradient, "gradient").contiguous()
        k4, k4i = k4.contiguous(), k4i.contiguous()
        energy_ref, gradient_ref = energy_ref.contiguous(), gradient_ref.contiguous()
        is_dummy = is_dummy.contiguous() if is_dummy is not None else None
        dev = energy.device
        parts = torch.empty((energy.shape[0],), device=dev, dtype=torch.float32)
        loss = torch.empty((), device=dev, dtype=torch.float32)
        MolwiseLossFn._call(plan, weights, energy, gradient, k4, k4i, energy_ref, gradient_ref, is_dummy, None, parts, loss,
                            None, None, None, None)
        ctx.save_for_backward(energy, gradient, k4, k4i, energy_ref, gradient_ref, is_dummy)
        ctx.cfg = (plan, weights)
        return loss

    @staticmethod
    def backward(ctx, g):
        energy, gradient, k4, k4i, energy_ref, gradient_ref, is_dummy = ctx.saved_tensors
        plan, weights = ctx.cfg
        g = g.contiguous().float()
        dE, dG = torch.empty_like(energy), torch.empty_like(gradient)
        dk4, dk4i = torch.empty_like(k4), torch.empty_like(k4i)
        MolwiseLossFn._call(plan, weights, energy, gradient, k4, k4i, energy_ref, gradient_ref, is_dummy, g, None, None,
                            dE, dG, dk4, dk4i)
        return dE, dG, dk4, dk4i, None, None, None, None, None
