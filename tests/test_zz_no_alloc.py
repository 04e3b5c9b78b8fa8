"""grappa_colsum_noalloc: the column sum whose scratch lives inside the library instead of in a caller buffer.

Who needs it: code that holds device pointers it did not allocate and has no allocator of its own, e.g. a force-field
plugin called from an MD engine's step loop, or a bare ctypes / cgo / JNI host.  Such a caller cannot size and
provide the partial-sum scratch that grappa_colsum wants, and must not trigger cudaMalloc in the middle of a step.

This file is collected last on purpose: if it fails under `pytest -x`, every other test has already run.
It goes through the C ABI directly (ctypes), twice in a row, and compares with the caller-scratch entry point.
"""
import ctypes

import pytest
import torch

from grappa_b200 import _lib

pytestmark = pytest.mark.gpu


def _both(x):
    L = _lib.lib()
    M, N = x.shape
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    ref = torch.empty(N, device=x.device)
    scratch = torch.empty(int(L.grappa_colsum_scratch_floats(M, N)), device=x.device)
    _lib.check(L.grappa_colsum(M, N, x.data_ptr(), x.stride(0), ref.data_ptr(), scratch.data_ptr(), stream))
    out = torch.full((N,), float("nan"), device=x.device)
    _lib.check(L.grappa_colsum_noalloc(M, N, x.data_ptr(), x.stride(0), out.data_ptr(), stream))
    torch.cuda.synchronize()
    return out, ref


def test_noalloc_colsum_matches_caller_scratch_variant_twice_in_a_row():
    gen = torch.Generator().manual_seed(0)
    # second call: other shape and other values, so stale partials of the first call would show up
    for M, N in ((6080, 512), (1280, 2048)):
        x = torch.randn(M, N, generator=gen).cuda()
        out, ref = _both(x)
        torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-5 * float(ref.abs().max()))
        torch.testing.assert_close(out.cpu(), x.cpu().double().sum(0).float(), rtol=1e-4, atol=1e-3)


def test_noalloc_colsum_odd_width_and_single_row():
    gen = torch.Generator().manual_seed(1)
    for M, N in ((37, 511), (1, 256)):
        x = torch.randn(M, N, generator=gen).cuda()
        out, ref = _both(x)
        torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-5)


def test_noalloc_colsum_rejects_problems_beyond_the_builtin_scratch():
    L = _lib.lib()
    M, N = 4096, 200000                      # needs far more than 65536 floats of partial sums
    assert L.grappa_colsum_scratch_floats(M, N) > 65536
    x = torch.zeros(8, device="cuda")        # never dereferenced: the size check comes first
    rc = L.grappa_colsum_noalloc(M, N, x.data_ptr(), N, x.data_ptr(), None)
    assert rc != 0
    msg = L.grappa_last_error().decode()
    assert "65536" in msg and "grappa_colsum" in msg
