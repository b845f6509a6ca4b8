"""NVLS (in-switch) all-reduce over a multicast mapping of symmetric memory.

The multicast object and the per-rank mappings are set up by ``torch.distributed._symmetric_memory`` (it does the
``cuMulticastCreate`` / ``cuMulticastBindMem`` dance and the handle exchange); the reduction itself is this
framework's kernel (``csrc/nvls.cu``): every GPU issues ``multimem.ld_reduce`` on ITS coordinate shard — the NVSwitch
sums the W replicas — and ``multimem.st`` to broadcast the result.  Optional: ``NvlsAllReduce.create`` returns ``None``
when the fabric / driver does not expose multicast, and callers fall back to the peer-pointer kernel of ``symm_impl``.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..ops import _lib
from .symm_impl import SymmetricArena, shard_bounds


class NvlsAllReduce:
    def __init__(self, buf: torch.Tensor, handle, arena: SymmetricArena):
        self.buf, self.hdl, self.arena = buf, handle, arena
        self.mc = int(handle.multicast_ptr)
        self.world, self.rank = arena.world, arena.rank
        self.stream = torch.cuda.Stream()

    @classmethod
    def create(cls, n_elems: int, arena: SymmetricArena, group=None):
        try:
            import torch.distributed._symmetric_memory as symm_mem
            grp = group or dist.group.WORLD
            n = (n_elems + 127) // 128 * 128
            buf = symm_mem.empty(n, dtype=torch.float32, device=torch.device("cuda", torch.cuda.current_device()))
            hdl = symm_mem.rendezvous(buf, grp.group_name if hasattr(grp, "group_name") else grp)
            if not int(getattr(hdl, "multicast_ptr", 0)):
                return None
            buf.zero_()
            return cls(buf, hdl, arena)
        except Exception:
            return None

    def all_reduce_(self, lo: int = 0, hi: int | None = None, scale: float = 1.0, max_ctas: int = 0) -> None:
        """In-place sum over ranks of ``buf[lo:hi]`` (float4-aligned bounds), on the current stream."""
        hi = self.buf.numel() if hi is None else hi
        s0, s1 = shard_bounds(hi - lo, self.world, self.rank)
        self.arena.barrier()                                   # every rank's inputs are written
        _lib.check(_lib.lib().ddl_nvls_allreduce(_lib.C.c_void_p(self.mc), _lib.i64(lo + s0), _lib.i64(lo + s1), _lib.f32(scale),
                                                 int(max_ctas), _lib.stream()), "ddl_nvls_allreduce")
        self.arena.barrier()                                   # every shard has been broadcast

    # ---- reducer protocol of parallel.dp.DataParallel
    @property
    def grad(self) -> torch.Tensor:
        return self.buf

    class _Work:
        def __init__(self, ev):
            self.ev = ev

        def wait(self):
            torch.cuda.current_stream().wait_event(self.ev)

    def __call__(self, view: torch.Tensor, lo: int, hi: int):
        ready = torch.cuda.Event()
        ready.record()
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(ready)
            self.all_reduce_(lo, hi, max_ctas=32)
            done = torch.cuda.Event()
            done.record()
        return self._Work(done)
