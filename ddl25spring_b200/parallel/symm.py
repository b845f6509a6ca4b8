"""Symmetric peer memory over NVLink and the fused cross-GPU aggregation built on it.

(placeholder entry point; the symmetric arena and peer-pointer kernels are filled in below as the
multi-GPU path is brought up)
"""
from __future__ import annotations


def make_collective():
    """Return the fused peer-memory aggregator for the current process group, or ``None`` when
    running on a single GPU / CPU (the server then uses its local or NCCL/gloo path)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1 and torch.cuda.is_available()):
        return None
    try:
        from .symm_impl import SymmetricAggregator
        return SymmetricAggregator()
    except Exception as e:      # pragma: no cover - reported once, NCCL path still correct
        import warnings
        warnings.warn(f"peer-memory aggregation unavailable ({type(e).__name__}: {e}); using NCCL collectives")
        return None
