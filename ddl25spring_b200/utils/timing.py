"""Device timing helpers: CUDA events, L2 flush, max-over-ranks."""
from __future__ import annotations

import torch
import torch.distributed as dist

L2_BYTES = 126 * 1024 * 1024


class L2Flusher:
    """Writes a buffer twice the size of the 126 MB L2 so the next timed step starts cold."""

    def __init__(self, device):
        self.buf = torch.empty(2 * L2_BYTES // 4, dtype=torch.float32, device=device)
        self.v = 0.0

    def flush(self):
        self.v += 1.0
        self.buf.fill_(self.v)


def max_over_ranks(value: float, device) -> float:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        t = torch.tensor([value], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return value


def barrier_sync(device) -> None:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()
    if torch.device(device).type == "cuda":
        torch.cuda.synchronize()
