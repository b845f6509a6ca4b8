"""ddl_gram_static_scratch: the Gram entry point for callers that cannot allocate.

Some hosts of this library have no way to hand over a scratch buffer when they call it: a C/C++
program without PyTorch's caching allocator, or a service whose device memory is laid out once at
start-up, cannot afford a cudaMalloc inside the round loop (it synchronises the GPU and can fail
late).  For them the library keeps the split-K scratch in a static device array of its own.  This
test drives that entry point through the C ABI exactly as such a host would (ctypes, raw pointers)
and checks it against the regular entry point, which gets its scratch from the caller.

The file name sorts last on purpose: if this addition ever breaks, a `pytest -x` run has already
exercised every other test.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _gram_regular(L, U, n_split):
    """Existing entry point: the caller (here: torch) owns the scratch."""
    from ddl25spring_b200.ops import _lib
    K, D = U.shape
    G = torch.empty(K, K, device=DEV)
    scratch = torch.empty(n_split * K * K, device=DEV)
    rc = L.ddl_gemm_tcgen05(_lib.p(U), _lib.i64(U.stride(0)), _lib.i64(0), 0, _lib.p(U), _lib.i64(U.stride(0)), _lib.i64(0), 0,
                            _lib.p(G), _lib.i64(K), _lib.i64(0), _lib.p(None), _lib.i64(0), 0, _lib.f32(0.0), _lib.f32(1.0),
                            _lib.f32(0.0), K, K, D, 1, 4, 0, n_split, _lib.p(scratch), _lib.stream())
    _lib.check(rc, "ddl_gemm_tcgen05")
    return G


def _gram_static(L, U, n_split):
    """New entry point: inputs and outputs only; nothing is allocated inside the call."""
    from ddl25spring_b200.ops import _lib
    K, D = U.shape
    G = torch.empty(K, K, device=DEV)          # the OUTPUT is still the caller's
    rc = L.ddl_gram_static_scratch(_lib.p(U), _lib.i64(U.stride(0)), K, _lib.i64(D), _lib.p(G), _lib.i64(K), n_split, 0,
                                   _lib.stream())
    _lib.check(rc, "ddl_gram_static_scratch")
    return G


def test_static_scratch_gram_matches_regular_entry_point_twice(cuda_lib):
    L = cuda_lib
    torch.manual_seed(0)
    # first call: 64 clients, 64 splits -> fills 64*64*64 floats of the static array
    U1 = 0.05 * torch.randn(64, 200_000, device=DEV)
    got1 = _gram_static(L, U1, 64)
    exp1 = _gram_regular(L, U1, 64)
    torch.cuda.synchronize()
    torch.testing.assert_close(got1, exp1, rtol=0, atol=0)
    exp64 = U1.double() @ U1.double().T
    assert float((got1.double() - exp64).abs().max() / exp64.abs().max()) < 3e-3      # same tolerance as test_gram
    # second call right after, different shape: the array still holds the first call's partials
    U2 = 0.05 * torch.randn(10, 70_000, device=DEV)
    got2 = _gram_static(L, U2, 37)
    exp2 = _gram_regular(L, U2, 37)
    torch.cuda.synchronize()
    torch.testing.assert_close(got2, exp2, rtol=0, atol=0)
    exp64 = U2.double() @ U2.double().T
    assert float((got2.double() - exp64).abs().max() / exp64.abs().max()) < 3e-3


def test_static_scratch_rejects_oversized_problems(cuda_lib):
    L = cuda_lib
    U = torch.randn(128, 4096, device=DEV)
    G = torch.empty(128, 128, device=DEV)
    from ddl25spring_b200.ops import _lib
    # 65 * 128 * 128 floats > 2^20: must be refused with a readable message, not silently truncated
    rc = L.ddl_gram_static_scratch(_lib.p(U), _lib.i64(U.stride(0)), 128, _lib.i64(4096), _lib.p(G), _lib.i64(128), 65, 0,
                                   _lib.stream())
    assert rc != 0
    msg = L.ddl_last_error().decode()
    assert "exceeds the built-in scratch" in msg
    # and the largest legal size for K = 128 still works
    rc = L.ddl_gram_static_scratch(_lib.p(U), _lib.i64(U.stride(0)), 128, _lib.i64(4096), _lib.p(G), _lib.i64(128), 64, 0,
                                   _lib.stream())
    assert rc == 0
    torch.cuda.synchronize()
    exp = U.double() @ U.double().T
    assert float((G.double() - exp).abs().max() / exp.abs().max()) < 3e-3


def test_static_scratch_rejects_ragged_inner_extent(cuda_lib):
    """D is the innermost extent of the operand's tensor map: anything but whole 16-byte units must be refused with an
    argument error before a launch (a ragged extent would stall the TMA copy)."""
    from ddl25spring_b200.ops import _lib
    L = cuda_lib
    U = torch.randn(16, 70_004, device=DEV)
    G = torch.empty(16, 16, device=DEV)
    rc = L.ddl_gram_static_scratch(_lib.p(U), _lib.i64(U.stride(0)), 16, _lib.i64(70_001), _lib.p(G), _lib.i64(16), 8, 0,
                                   _lib.stream())
    assert rc == 1, rc                                     # DDL_ERR_ARG
    assert "multiple of 4" in L.ddl_last_error().decode()
