"""tcgen05 GEMM front-end (K10) and the Krum Gram matrix (K3).

``gemm`` computes ``C[g] = act(alpha * A[g] @ B[g]^T + bias[g])`` with fp32 tensors consumed as
TF32 (or exactly, "3xTF32", with ``precise=True``) on the Blackwell tensor cores.  Operands may
be K-major (``[rows, K]``) or MN-major (``[K, rows]``), which is what lets the three GEMMs of a
Linear layer (forward, input gradient, weight gradient) share one kernel with no transposes:

===========  =======================  ======================  =====================
GEMM         A                        B                       C
===========  =======================  ======================  =====================
forward      X   [B, in]  K-major     W  [out, in] K-major    Y  [B, out]
dgrad        dY  [B, out] K-major     W  [out, in] MN-major   dX [B, in]
wgrad        dY  [B, out] MN-major    X  [B, in]   MN-major   dW [out, in]
===========  =======================  ======================  =====================

Reference call sites replaced: every ``nn.Linear`` listed in SURVEY §2.5.
"""
from __future__ import annotations

import os

import torch

from . import _lib

ACT = {"none": 0, None: 0, "relu": 1, "leaky_relu": 2}
PRECISE_CHAIN = 1024
MAX_FIXUP_TILES = 8192
# In-kernel split-K fix-up (the split CTAs of a tile meet at a counter and each reduces its share of the rows; cooperative
# launch).  Opt-in: on the LLaMA step the separate reduce kernel behind a programmatic dependent launch measured faster
# (2.74 ms vs 3.30 ms cooperative / 3.20 ms programmatic launch of the fix-up), see profiles/r2_llama_fixup_modes.txt.
USE_FIXUP = os.environ.get("DDL_GEMM_FIXUP", "0") == "1"
AUTO_SPLIT = True           # spread the K loop of small grids over more SMs (see gemm()); benchmarks switch it off to isolate it
_TILE_COUNTERS: dict = {}


def _tile_counters(device: torch.device) -> torch.Tensor:
    """Zeroed per-(device, stream) tile counters of the in-kernel split-K fix-up (the kernel leaves them zero; launches on one
    stream are ordered, so they can share the buffer).  The fix-up makes the CTAs of a tile wait for each other, which is safe
    for ONE such grid at a time (it is a single wave); code that runs GEMMs concurrently on several streams
    (``ops.conv.workspace_scope``) gets the separate reduce kernel instead."""
    from .conv import _SCOPE                 # concurrent client slots (fl/server.py) each work inside their own scope
    key = (device.index, torch.cuda.current_stream(device).cuda_stream, _SCOPE["id"])
    t = _TILE_COUNTERS.get(key)
    if t is None:
        t = _TILE_COUNTERS[key] = torch.zeros(MAX_FIXUP_TILES, dtype=torch.int32, device=device)
    return t


def _grouped(t: torch.Tensor):
    if t.dim() == 2:
        return t.unsqueeze(0)
    assert t.dim() == 3
    return t


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False, bias: torch.Tensor | None = None,
         act: str | None = None, slope: float = 0.01, alpha: float = 1.0, beta: float = 0.0, precise: bool = False,
         n_split: int = 1,
         out: torch.Tensor | None = None) -> torch.Tensor:
    """``C = act(alpha * op(A) @ op(B)^T + bias)``.

    ``a``: ``[G?, M, K]`` (K-major) or ``[G?, K, M]`` (``a_mn``); ``b``: ``[G?, N, K]`` or ``[G?, K, N]``
    (``b_mn``).  fp32 (TF32 math; ``precise`` = 3xTF32 error-compensated) or bf16 operands, fp32 output.
    """
    squeeze = a.dim() == 2
    A, B = _grouped(a), _grouped(b)
    G = A.shape[0]
    M, K = (A.shape[2], A.shape[1]) if a_mn else (A.shape[1], A.shape[2])
    N, Kb = (B.shape[2], B.shape[1]) if b_mn else (B.shape[1], B.shape[2])
    assert K == Kb and B.shape[0] == G, (A.shape, B.shape)
    assert A.is_cuda and A.dtype == B.dtype and A.dtype in (torch.float32, torch.bfloat16)
    assert A.stride(2) == 1 and B.stride(2) == 1
    eb = A.element_size()
    if out is None:
        out = torch.empty((G, M, N), device=a.device, dtype=torch.float32)
    O = _grouped(out)
    assert O.stride(2) == 1 and O.dtype == torch.float32
    bias_stride = 0
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.shape[-1] == N
        bias_stride = bias.stride(0) if bias.dim() == 2 else 0
    chain_min = 1
    if precise:
        # The tensor core's fp32 accumulator truncates on every add, so the error of a K-long chain
        # grows ~linearly in K (measured on B200: 2e-6 at K=128, 8e-5 at K=9216).  Cap each
        # accumulation chain at PRECISE_CHAIN products and finish the sum in fp32 on the CUDA cores.
        chain_min = -(-K // PRECISE_CHAIN)
    if AUTO_SPLIT and n_split == 1 and eb == 4:
        # A grid that covers less than half the SMs with a long K loop is latency-bound per CTA: spread the reduction over more
        # SMs, partials summed in a fixed order by splitk_reduce_kernel.  Tuned inside a CUDA graph on the LLaMA lab's shapes
        # (scripts/dev/gemm_split_tune.py, profiles/r2_gemm_split_tune.txt): a k-block costs ~0.75 us (3xTF32), the extra
        # reduce launch 3-5 us depending on the output size, so 9 k-blocks are only worth splitting on very small grids.
        bn = 32 if N <= 32 else (64 if N <= 64 else 128)
        ctas = -(-M // 128) * -(-N // bn) * G
        kb = -(-K // 32)
        if 2 * ctas <= 148 and (kb >= 16 or (kb >= 8 and ctas <= 20)):
            n_split = max(1, min(kb // 3, 148 // ctas))
    n_split = max(n_split, chain_min)
    L = _lib.lib()
    scratch = counters = None
    if n_split > 1:
        scratch = torch.empty(n_split * G * M * N, device=a.device, dtype=torch.float32)
        from .conv import _SCOPE
        if USE_FIXUP and beta == 0.0 and L.ddl_gemm_tile_count(int(M), int(N), int(G)) <= MAX_FIXUP_TILES and not _SCOPE["concurrent"]:
            counters = _tile_counters(a.device)
    args = (_lib.p(A), _lib.i64(A.stride(1)), _lib.i64(A.stride(0)), int(a_mn),
            _lib.p(B), _lib.i64(B.stride(1)), _lib.i64(B.stride(0)), int(b_mn),
            _lib.p(O), _lib.i64(O.stride(1)), _lib.i64(O.stride(0)),
            _lib.p(bias), _lib.i64(bias_stride), ACT[act], _lib.f32(slope), _lib.f32(alpha), _lib.f32(beta),
            int(M), int(N), int(K), int(G), int(eb), int(precise), int(n_split), _lib.p(scratch))
    if counters is not None:
        rc = L.ddl_gemm_tcgen05_fixup(*args, _lib.p(counters), _lib.stream())
    else:
        rc = L.ddl_gemm_tcgen05(*args, _lib.stream())
    _lib.check(rc, "ddl_gemm_tcgen05")
    return out[0] if (squeeze and out.dim() == 3) else out


def linear(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None, act: str | None = None,
           precise: bool = True, slope: float = 0.01) -> torch.Tensor:
    """``act(x @ w.T + bias)`` – nn.Linear forward on tcgen05 (fp32-grade by default)."""
    return gemm(x, w, bias=bias, act=act, precise=precise, slope=slope)


def linear_dgrad(dy: torch.Tensor, w: torch.Tensor, precise: bool = True) -> torch.Tensor:
    """``dX = dY @ W``  (W [out, in] consumed MN-major; needs ``in % 32 == 0``)."""
    return gemm(dy, w, b_mn=True, precise=precise)


def linear_wgrad(dy: torch.Tensor, x: torch.Tensor, precise: bool = True, n_split: int = 1) -> torch.Tensor:
    """``dW = dY^T @ X`` (both consumed MN-major; needs ``out % 32 == 0`` and ``in % 32 == 0``)."""
    return gemm(dy, x, a_mn=True, b_mn=True, precise=precise, n_split=n_split)


def default_gram_splits(K: int, D: int) -> int:
    """Enough split-K slices to put one CTA on each of the 148 SMs."""
    kb = (D + 31) // 32
    return max(1, min(148, kb))


def gram_tcgen05(updates: torch.Tensor, out: torch.Tensor | None = None, n_split: int | None = None,
                 precise: bool = False, center: torch.Tensor | None = None) -> torch.Tensor:
    """``G = U U^T`` for ``U`` = ``[K <= 128, D]`` client updates: split-K over D across all SMs with
    TMEM accumulators, deterministic partial reduction.  TF32 inputs (pass update *deltas*, whose
    norms are small, to keep cancellation in ``G_ii + G_jj - 2 G_ij`` benign) or ``precise``.

    ``center`` ([D], e.g. the global model): the Gram matrix of ``U - center`` is computed WITHOUT forming the
    K x D delta matrix - the centre is subtracted from each staged tile in shared memory (3xTF32 path)."""
    K, D = updates.shape
    assert K <= 128, "one M tile: at most 128 clients per Gram launch"
    if n_split is None:
        n_split = default_gram_splits(K, D)
    if center is None:
        return gemm(updates, updates, n_split=n_split, precise=precise, out=out)
    assert updates.is_cuda and updates.dtype == torch.float32 and updates.stride(1) == 1 and center.is_contiguous()
    assert center.numel() >= D and center.dtype == torch.float32
    if out is None:
        out = torch.empty(K, K, device=updates.device, dtype=torch.float32)
    n_split = max(1, min(n_split, (D + 31) // 32))
    scratch = torch.empty(n_split * K * K, device=updates.device, dtype=torch.float32) if n_split > 1 else None
    rc = _lib.lib().ddl_gram_tcgen05(_lib.p(updates), _lib.i64(updates.stride(0)), int(K), _lib.i64(D), _lib.p(center), _lib.p(out),
                                     _lib.i64(out.stride(0)), 1, int(n_split), _lib.p(scratch), _lib.stream())
    _lib.check(rc, "ddl_gram_tcgen05")
    return out


def gather_linear_ptrs(part_ptrs, part_dims, part_lds, M: int, w: torch.Tensor, bias: torch.Tensor | None = None,
                       act: str | None = None, precise: bool = True, slope: float = 0.01, out: torch.Tensor | None = None):
    """K7: ``act(cat_p(parts[p]) @ w.T + bias)`` where ``parts[p]`` are raw device pointers (local or peer-mapped)
    to ``[M, dims[p]]`` fp32 matrices with row strides ``part_lds[p]``."""
    import ctypes as C
    n = len(part_ptrs)
    N = w.shape[0]
    assert w.is_cuda and w.dtype == torch.float32 and w.shape[1] == sum(part_dims)
    # re-pack W so that every party's slab starts on a 16-byte boundary (TMA box origins must be aligned)
    pads = [(d + 3) // 4 * 4 for d in part_dims]
    if pads == list(part_dims) and w.stride(1) == 1 and (w.stride(0) * 4) % 16 == 0:
        wp = w
    else:
        wp = torch.zeros(N, sum(pads), device=w.device, dtype=torch.float32)
        so = do = 0
        for d, pd in zip(part_dims, pads):
            wp[:, do:do + d] = w[:, so:so + d]
            so, do = so + d, do + pd
    w = wp
    if out is None:
        out = torch.empty(M, N, device=w.device, dtype=torch.float32)
    rc = _lib.lib().ddl_gather_gemm_tcgen05(_lib.host_ptrs(part_ptrs), (C.c_int * n)(*[int(d) for d in part_dims]),
                                            (C.c_int64 * n)(*[int(x) for x in part_lds]), n, _lib.p(w), _lib.i64(w.stride(0)),
                                            _lib.p(out), _lib.i64(out.stride(0)), _lib.p(bias), ACT[act], _lib.f32(slope), int(M),
                                            int(N), int(precise), 1, _lib.p(None), _lib.stream())
    _lib.check(rc, "ddl_gather_gemm_tcgen05")
    return out


def gram_rowgather_ptrs(part_ptrs, rows_per_part: int, ld: int, d0: int, d1: int, out: torch.Tensor, n_split: int = 148,
                        precise: bool = True, center: torch.Tensor | None = None):
    """K3 fused all-gather + Gram: ``out = U[:, d0:d1] @ U[:, d0:d1].T`` with the rows of ``U`` on different GPUs
    (``part_ptrs[p]`` = device pointer, possibly peer-mapped, of GPU p's ``rows_per_part`` rows).  Row order p-major."""
    Kc = len(part_ptrs) * rows_per_part
    kb = (d1 - d0 + 31) // 32
    if precise and center is None:                 # raw weights: keep accumulation chains short; re-centred deltas do not need it
        n_split = max(n_split, -(-(d1 - d0) // PRECISE_CHAIN))
    n_split = max(1, min(n_split, kb))
    scratch = torch.empty(n_split * Kc * Kc, device=out.device, dtype=torch.float32) if n_split > 1 else None
    rc = _lib.lib().ddl_gram_rowgather_tcgen05(_lib.host_ptrs(part_ptrs), len(part_ptrs), int(rows_per_part), _lib.i64(ld), _lib.i64(d0),
                                               _lib.i64(d1), _lib.p(center), _lib.p(out), _lib.i64(out.stride(0)), int(precise), int(n_split),
                                               _lib.p(scratch), _lib.stream())
    _lib.check(rc, "ddl_gram_rowgather_tcgen05")
    return out
