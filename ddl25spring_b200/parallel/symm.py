"""Public entry points of the symmetric-memory (NVLink peer memory) collectives.

The implementation lives in :mod:`parallel.symm_impl` (arena, channels, the fused aggregation / gradient-reduce / VFL
gather hooks); this module re-exports it and provides :func:`make_collective`, the factory the servers and ``bench.py``
use.  On a multi-GPU process group a failure to set the peer-memory path up is an ERROR: silently running the NCCL
baseline instead would hide a performance regression.  Pass ``allow_fallback=True`` to get a loud warning and ``None``.
"""
from __future__ import annotations

from .symm_impl import (Channel, FusedVFLGather, OptArgs, SymmetricAggregator, SymmetricArena, SymmetricGradReducer,  # noqa: F401
                        shard_bounds)


def make_collective(allow_fallback: bool = False, **kw):
    """The fused peer-memory aggregator for the current process group, or ``None`` on a single GPU / CPU (the server then
    uses its local path).  ``kw`` goes to :class:`SymmetricAggregator` (``max_rows_per_rank``, ``row_elems``, ``use_nvls``)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1 and torch.cuda.is_available()):
        return None
    try:
        return SymmetricAggregator(**kw)
    except Exception as e:
        if not allow_fallback:
            raise RuntimeError(f"peer-memory aggregation could not be set up ({type(e).__name__}: {e}); pass "
                               "allow_fallback=True to run on NCCL collectives instead") from e
        import sys
        import warnings
        msg = f"peer-memory aggregation unavailable ({type(e).__name__}: {e}); FALLING BACK to NCCL collectives"
        warnings.warn(msg)
        print("WARNING: " + msg, file=sys.stderr, flush=True)
        return None
