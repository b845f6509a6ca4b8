"""Symmetric peer memory over NVLink (one process per GPU) and the fused collectives built on it.

:class:`SymmetricArena` allocates same-sized device buffers on every rank (``cudaMalloc`` in
``csrc/symm.cu``), exchanges CUDA-IPC handles through ``torch.distributed`` once, and exposes

* ``local(name)``      – this rank's buffer as a torch tensor (zero-copy, ``__cuda_array_interface__``);
* ``peer_ptrs(name)``  – device pointers of the same buffer on every rank (index = rank);
* a signal pad + error flag for the in-kernel cross-GPU barriers.

:class:`SymmetricAggregator` is the multi-GPU server-aggregate hook of :class:`fl.packed_server.
PackedFedAvgServer`: each rank's clients' weight rows live in a symmetric buffer, and ONE kernel per
GPU (``ddl_symm_reduce``) pulls this GPU's coordinate shard of every row on every GPU, reduces it
(weighted mean / median / trimmed mean) and stores the result into every rank's replica.
:class:`SymmetricGradReducer` plugs the same kernel into :class:`parallel.dp.DataParallel` as the
bucketed gradient all-reduce (in place over the symmetric gradient arena, on a side stream).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from ..ops import _lib


class _DevBuf:
    """Raw device allocation viewed as a torch tensor through the CUDA array interface."""

    def __init__(self, ptr: int, nbytes: int, device_index: int):
        self.ptr, self.nbytes = ptr, nbytes
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}
        self.tensor = torch.as_tensor(self, device=torch.device("cuda", device_index))


class OptArgs(C.Structure):
    """``ddl_opt_args`` of include/ddl25_kernels.h: the optimizer epilogue of the fused bucket reduce (K2)."""
    _fields_ = [("kind", C.c_int), ("lr", C.c_float), ("weight_decay", C.c_float), ("momentum", C.c_float),
                ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float), ("m", C.c_void_p), ("v", C.c_void_p),
                ("state_off4", C.c_int64), ("step_dev", C.c_void_p), ("bump_step", C.c_int)]


class Channel:
    """One ordered sequence of collectives: signal pads on every rank + a device-resident sequence number.

    The kernels advance the sequence number themselves, so a launch carries no per-call host value and can be captured
    into a CUDA graph.  Collectives of one channel must be stream-ordered on every rank and issued in the same order
    by all ranks; independent streams (the DP bucket reducer next to the main stream) use separate channels."""

    def __init__(self, pad_ptrs: list[int], state: torch.Tensor):
        self.pad_ptrs, self.state = pad_ptrs, state


class SymmetricArena:
    def __init__(self, group=None):
        assert dist.is_initialized() and torch.cuda.is_available()
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.dev = torch.cuda.current_device()
        self.L = _lib.lib()
        self._bufs: dict[str, _DevBuf] = {}
        self._peers: dict[str, list[int]] = {}
        self._sizes: dict[str, int] = {}
        self._n_channels = 0
        self.err = torch.zeros(1, dtype=torch.int32, device="cuda")
        self.main = self.new_channel()
        self.barrier()

    # ------------------------------------------------------------------ allocation + exchange
    def alloc(self, name: str, nbytes: int, dtype=torch.float32) -> torch.Tensor:
        """Collective: every rank must call with the same ``name`` / size, in the same order."""
        if name in self._bufs:
            raise KeyError(f"symmetric buffer {name!r} already exists")
        want = int(nbytes)
        nbytes = (want + 255) // 256 * 256
        ptr = C.c_void_p()
        _lib.check(self.L.ddl_symm_alloc(C.byref(ptr), C.c_int64(nbytes)), "ddl_symm_alloc")
        buf = _DevBuf(ptr.value, nbytes, self.dev)
        handle = (C.c_ubyte * 64)()
        _lib.check(self.L.ddl_ipc_export(C.c_void_p(ptr.value), handle), "ddl_ipc_export")
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle), group=self.group)
        peers = []
        for r, h in enumerate(handles):
            if r == self.rank:
                peers.append(ptr.value)
                continue
            out = C.c_void_p()
            _lib.check(self.L.ddl_ipc_import((C.c_ubyte * 64).from_buffer_copy(h), C.byref(out)), "ddl_ipc_import")
            peers.append(out.value)
        self._bufs[name], self._peers[name] = buf, peers
        self._sizes[name] = want
        return buf.tensor[:want].view(dtype)

    def local(self, name: str, dtype=torch.float32) -> torch.Tensor:
        return self._bufs[name].tensor[:self._sizes[name]].view(dtype)

    def peer_ptrs(self, name: str) -> list[int]:
        return self._peers[name]

    def new_channel(self) -> Channel:
        """Collective: a fresh set of signal pads + sequence state (see :class:`Channel`)."""
        name = f"__pads{self._n_channels}__"
        self._n_channels += 1
        self.alloc(name, 4 * 2 * 16 + 64, torch.int32)                # int[2*world] (+ slack), zeroed by the allocator
        return Channel(self.peer_ptrs(name), torch.zeros(4, dtype=torch.int32, device="cuda"))

    # ------------------------------------------------------------------ barriers / errors
    def barrier(self, channel: Channel | None = None) -> None:
        """Cross-GPU barrier executed BY the GPUs (in stream order), not by the hosts."""
        ch = channel or self.main
        _lib.check(self.L.ddl_symm_barrier(_lib.host_ptrs(ch.pad_ptrs), self.rank, self.world, _lib.p(ch.state), _lib.p(self.err),
                                           _lib.stream()), "ddl_symm_barrier")

    def check(self) -> None:
        """Raise if any collective on this arena timed out (a rank did not launch its kernel within 5 s).  Reads one
        device int: call it where the host synchronises anyway (after an accuracy read-back, every N optimizer steps)."""
        e = int(self.err.item())
        if e:
            raise RuntimeError(f"peer-memory collective timed out on rank {self.rank} (code {e}): a rank did not launch its "
                               "kernel; every later collective on this arena skipped its stores")

    def reduce(self, mode: int, row_ptrs, weights, out_ptrs, d0: int, d1: int, *, base=None, alpha=0.0, beta=1.0,
               rank_lo=0, rank_hi=1, max_ctas: int = 0, dev_weights: torch.Tensor | None = None, channel: Channel | None = None,
               opt: OptArgs | None = None, mc_in: int = 0, mc_out: int = 0) -> None:
        ch = channel or self.main
        K = len(row_ptrs) if row_ptrs else (len(weights) if weights is not None else self.world)
        w = None if (weights is None or dev_weights is not None) else _lib.host_floats(weights)
        _lib.check(self.L.ddl_symm_reduce(int(mode), _lib.host_ptrs(row_ptrs) if row_ptrs else None, w, _lib.p(dev_weights), K,
                                          _lib.p(base), _lib.f32(alpha), _lib.f32(beta), int(rank_lo), int(rank_hi),
                                          _lib.host_ptrs(out_ptrs) if out_ptrs else None, len(out_ptrs) if out_ptrs else 0,
                                          _lib.i64(d0), _lib.i64(d1), _lib.host_ptrs(ch.pad_ptrs), self.rank, self.world,
                                          _lib.p(ch.state), _lib.p(self.err), int(max_ctas),
                                          C.byref(opt) if opt is not None else None, C.c_void_p(mc_in or None),
                                          C.c_void_p(mc_out or None), _lib.stream()), "ddl_symm_reduce")


def shard_bounds(n: int, world: int, rank: int, align: int = 4) -> tuple[int, int]:
    """Contiguous, ``align``-aligned coordinate shard of rank ``rank``."""
    per = -(-n // world)
    per = (per + align - 1) // align * align
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


class SymmetricAggregator:
    """Fused cross-GPU aggregation for the packed FedAvg server (weighted mean, median, trimmed mean; Krum
    gathers the rows it needs with peer copies and scores locally)."""

    def __init__(self, arena: SymmetricArena | None = None, max_rows_per_rank: int = 16, row_elems: int = 1_200_000,
                 use_nvls: bool = False, tag: str = "", two_phase_select: bool | None = None):
        self.arena = arena or SymmetricArena()
        # median / trimmed mean as gather + local selection + peer stores instead of the one-kernel form.  Measured with 8 rows
        # of 11.2 M floats per GPU: 2 GPUs 0.46 vs 0.40 ms (one kernel wins), 8 GPUs 0.82 vs 0.96 ms (two phases win) - the
        # default follows the group size.
        self.two_phase_select = (self.arena.world >= 8) if two_phase_select is None else two_phase_select
        self.tag = tag                     # several aggregators may share one arena: buffer names carry the tag
        self.nvls = None
        if use_nvls:                       # in-switch reduction of the per-GPU partial sums (falls back when unsupported)
            from .nvls import NvlsAllReduce
            self.nvls = NvlsAllReduce.create(row_elems, self.arena)
        self.G, self.P = max_rows_per_rank, row_elems
        self.rows = self.arena.alloc("client_rows" + tag, self.G * self.P * 4).view(self.G, self.P)
        self.result = self.arena.alloc("agg_result" + tag, self.P * 4)       # every rank's replica of the aggregate; a server may use it
                                                                          # directly as its global parameter arena (global_storage())
        self.partial = self.arena.alloc("agg_partial" + tag, self.P * 4)
        self.partial_ptrs = self.arena.peer_ptrs("agg_partial" + tag)
        self.row_ptrs = self.arena.peer_ptrs("client_rows" + tag)
        self.res_ptrs = self.arena.peer_ptrs("agg_result" + tag)

    def global_storage(self) -> torch.Tensor:
        """Symmetric buffer the fused kernels write the aggregate into on EVERY rank: a server that keeps its global model
        here (``FlatArena.rebase``) gets the new model without a further copy."""
        return self.result

    def row_buffer(self, G: int) -> torch.Tensor:
        """Symmetric storage for this rank's ``G`` client weight rows (the engine trains in place in it)."""
        assert G <= self.G, f"{G} rows per rank requested, arena was built for {self.G}"
        return self.rows[:G]

    def aggregate(self, aggregator, W_local: torch.Tensor, mine_pos, weights: torch.Tensor, gflat: torch.Tensor) -> None:
        from ..fl.aggregators import CoordinateMedian, Krum, TrimmedMean, WeightedMean

        a = self.arena
        world, K = a.world, len(weights)
        if W_local.numel() and W_local.data_ptr() != self.rows.data_ptr():
            self.rows[:W_local.shape[0]].copy_(W_local)
        # row k of the round lives on rank k % world at local index k // world (round-robin deal)
        ptrs = [self.row_ptrs[k % world] + (k // world) * self.P * 4 for k in range(K)]
        d0, d1 = shard_bounds(self.P, world, a.rank)
        w = [float(x) for x in weights.tolist()]
        if isinstance(aggregator, WeightedMean):
            # hierarchical: a linear rule lets every GPU pre-reduce ITS clients locally (HBM speed) so that only one
            # partial row per GPU crosses NVLink; the fused kernel then sums the W partials shard by shard.
            from .. import ops
            n_local = W_local.shape[0]
            part = self.nvls.buf[:self.P] if self.nvls is not None else self.partial
            if n_local:
                ops.weighted_sum(self.rows[:n_local], [w[p] for p in mine_pos], part)
            else:
                part.zero_()
            if self.nvls is not None:          # NVSwitch sums the W partials (multimem.ld_reduce) and multicasts the result
                self.nvls.all_reduce_(0, self.P)
                gflat.copy_(part)
                return
            a.reduce(0, self.partial_ptrs, [1.0] * world, self.res_ptrs, d0, d1)
        elif isinstance(aggregator, (CoordinateMedian, TrimmedMean)) and K <= 64:
            if isinstance(aggregator, CoordinateMedian):
                lo, hi = ((K - 1) // 2, (K - 1) // 2 + 1) if (K % 2 or aggregator.lower) else (K // 2 - 1, K // 2 + 1)
            else:
                t = aggregator.trim_count(K)
                lo, hi = t, K - t
            if self.two_phase_select and d1 > d0:
                # Two phases: (1) the coordinate-sharded exchange as 128-bit peer loads into a local [K, shard] matrix (the same
                # gather kernel Krum uses), (2) the selection network over LOCAL memory whose result rows are stored straight
                # into every replica (peer stores), (3) a GPU-side barrier.  The one-kernel form reads 4 bytes per thread per
                # row from the peers; here the NVLink side moves float4s and the sort never waits on a remote load.
                shard = d1 - d0
                pad = (-shard) % 32
                Us = getattr(self, "_Usel", None)
                if Us is None or Us.shape != (K, shard + pad):
                    Us = self._Usel = torch.zeros(K, shard + pad, device="cuda")
                a.reduce(2, ptrs, w, [Us.data_ptr()], d0, d1, rank_lo=pad // 4)
                lptrs = [Us.data_ptr() + (k * Us.stride(0) - d0) * 4 for k in range(K)]
                _lib.check(a.L.ddl_rank_select_mean_rows(_lib.host_ptrs(lptrs), K, int(lo), int(hi), _lib.host_ptrs(self.res_ptrs),
                                                         len(self.res_ptrs), _lib.i64(d0), _lib.i64(d1), _lib.stream()),
                           "ddl_rank_select_mean_rows")
                a.barrier()
            else:
                a.reduce(1, ptrs, w, self.res_ptrs, d0, d1, rank_lo=lo, rank_hi=hi)
        else:
            # Krum: coordinate-sharded.  One fused kernel pulls THIS GPU's slice of every client's update delta
            # over NVLink into a local [K, shard] matrix; the tcgen05 Gram GEMM (K3) runs on that slice; the
            # K x K partial Grams are summed with one more fused reduce; scores / selection are computed
            # redundantly (identically) on every rank and the selected rows are averaged by the K1 kernel.
            assert isinstance(aggregator, Krum) and K <= 128
            from ..ops.gemm import gram_tcgen05
            from .. import ops
            shard = d1 - d0
            pad = (-shard) % 32                                   # keep local rows 128-byte aligned for TMA
            U = getattr(self, "_U", None)
            if U is None or U.shape != (K, shard + pad):
                U = self._U = torch.zeros(K, shard + pad, device="cuda")
            kb = getattr(self, "_gram_bufs", None)
            if kb is None:
                kb = self._gram_bufs = {}
            if K not in kb:                                       # K x K partial / summed Gram buffers, one pair per round size
                n1, n2 = f"gram_part_{K}{self.tag}", f"gram_sum_{K}{self.tag}"
                kb[K] = (a.alloc(n1, K * K * 4).view(K, K), a.alloc(n2, K * K * 4).view(K, K), a.peer_ptrs(n1), a.peer_ptrs(n2))
            self._gpart, self._gsum, self._gpart_ptrs, self._gsum_ptrs = kb[K]
            rows_per_rank = K // world
            if getattr(aggregator, "fused_gather", False) and K == rows_per_rank * world and rows_per_rank % 8 == 0:
                # ONE kernel: every k-block of U is TMA-loaded from the owning GPUs and used as both MMA operands (3xTF32,
                # with the global model subtracted in shared memory: the Gram of the deltas); result rows are rank-major.
                from ..ops.gemm import gram_rowgather_ptrs
                g0, g1 = shard_bounds(self.P, world, a.rank, align=32)
                a.barrier()
                gram_rowgather_ptrs(self.row_ptrs, rows_per_rank, self.P, g0, g1, self._gpart, precise=True,
                                    center=gflat.detach())      # re-centred in shared memory: Gram of the deltas
                order = torch.tensor([(k % world) * rows_per_rank + k // world for k in range(K)], device="cuda")
            else:
                order = None
                a.reduce(2, ptrs, w, [U.data_ptr()], d0, d1, base=gflat.detach(), rank_lo=pad // 4)
                gram_tcgen05(U, out=self._gpart, precise=aggregator.precise)
            if (K * K) % 4 == 0:
                a.reduce(0, self._gpart_ptrs, [1.0] * world, self._gsum_ptrs, 0, K * K)
                G = self._gsum
            else:                                                 # odd K: K*K is not float4-divisible
                dist.all_reduce(self._gpart)
                G = self._gpart
            if order is not None:
                G = G[order][:, order].contiguous()                # back to client order
            sel, selw, scores = ops.krum_select(U, aggregator.f, aggregator.multi, gram_matrix=G)
            aggregator.last_selected, aggregator.last_scores = sel, scores
            # selection weights stay on the device (no host round trip) and zero-weight rows are not read
            if order is None:
                # the deltas of THIS shard are already local (U, gathered above): new = global + sum_k w_k delta_k is an
                # HBM-speed pass over local memory + the broadcast of the shard - the selected rows do not cross NVLink twice
                lptrs = [U.data_ptr() + (k * U.stride(0) - d0) * 4 for k in range(K)]
                a.reduce(0, lptrs, None, self.res_ptrs, d0, d1, base=gflat.detach(), alpha=1.0, beta=1.0, dev_weights=selw)
            else:
                a.reduce(0, ptrs, None, self.res_ptrs, d0, d1, dev_weights=selw)
        if gflat.data_ptr() != self.result.data_ptr():          # servers that adopt `result` as their global arena skip this copy
            gflat.copy_(self.result)


class _DevView:
    def __init__(self, ptr: int, n_floats: int):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (ptr, False), "version": 3, "strides": None}
        self.tensor = torch.as_tensor(self, device="cuda")


class _Work:
    def __init__(self, ev):
        self.ev = ev

    def wait(self):
        torch.cuda.current_stream().wait_event(self.ev)


class SymmetricGradReducer:
    """``reducer`` for :class:`parallel.dp.DataParallel` (K2): the bucketed gradient all-reduce as ONE fused peer-memory
    kernel per bucket, on a side stream and its own channel so it runs under the rest of backward.

    ``fuse_optimizer=True`` (default) makes that kernel the whole tail of the training step for its bucket: every rank sums
    ITS coordinate shard of the bucket over all ranks (reduce-scatter), applies 1/W and the optimizer update to that shard
    of the parameters - momentum / Adam moments exist only for the owned shard (ZeRO-1) - and stores the NEW PARAMETERS into
    every rank's replica (all-gather).  Gradients and parameters therefore both live in symmetric memory.  The launches
    carry no per-step host state, so DataParallel(graph=True) records them as side branches of the step graph.

    ``use_nvls=True`` puts both arenas behind NVLS multicast addresses (torch symmetric memory does the binding): the W
    peer loads become one ``multimem.ld_reduce`` (the switch adds) and the W peer stores one ``multimem.st``.

    Replaces lab/tutorial_1b/DP/gradient_aggr/intro_DP_GA.py:53-67 (barrier, flatten, CPU copy, gloo all-reduce, copy back,
    divide, Adam step - strictly one after the other)."""

    capturable = True

    def __init__(self, arena: SymmetricArena, n_elems: int, max_ctas: int = 32, fuse_optimizer: bool = True,
                 use_nvls: bool = False, tag: str = "", tail_ctas: int = 148):
        self.arena = arena
        # buckets reduced WHILE the backward pass still runs take `max_ctas` CTAs (they share the SMs with its GEMMs); the
        # step's last bucket has nothing left to overlap with and takes `tail_ctas` (LLaMA DP, 2 GPUs: 3.11 -> 2.87 ms / step)
        self.tail_ctas = tail_ctas
        n = (n_elems + 127) // 128 * 128
        self.n = n
        self.mc_grad = self.mc_param = 0
        self.param = None
        if use_nvls:
            import torch.distributed._symmetric_memory as symm_mem
            grp = arena.group or dist.group.WORLD
            dev = torch.device("cuda", torch.cuda.current_device())
            bufs = []
            for _ in range(2 if fuse_optimizer else 1):
                b = symm_mem.empty(n, dtype=torch.float32, device=dev)
                h = symm_mem.rendezvous(b, grp.group_name)
                if not int(getattr(h, "multicast_ptr", 0)):
                    raise RuntimeError("NVLS multicast is not available for this group")
                b.zero_()
                bufs.append((b, h))
            self._nvls_keep = bufs
            self.grad, self.mc_grad = bufs[0][0], int(bufs[0][1].multicast_ptr)
            self.ptrs = None
            if fuse_optimizer:
                self.param, self.mc_param = bufs[1][0], int(bufs[1][1].multicast_ptr)
            self.param_ptrs = None
        else:
            self.grad = arena.alloc("dp_grad" + tag, n * 4)
            self.ptrs = arena.peer_ptrs("dp_grad" + tag)
            if fuse_optimizer:
                self.param = arena.alloc("dp_param" + tag, n * 4)
                self.param_ptrs = arena.peer_ptrs("dp_param" + tag)
        self.fuses_optimizer = fuse_optimizer
        self.channel = arena.new_channel()
        self.stream = torch.cuda.Stream()
        self.max_ctas = max_ctas
        self.step_dev = torch.zeros(1, dtype=torch.int32, device="cuda")       # completed optimizer steps (Adam bias correction)
        self._opt = None
        self._state: dict[tuple[int, int], tuple] = {}
        arena.barrier()

    # ---- optimizer epilogue
    def configure(self, optimizer: str, lr: float, weight_decay: float = 0.0, betas=(0.9, 0.999), eps: float = 1e-8,
                  momentum: float = 0.0) -> None:
        kind = {"sgd": 1, "adam": 2, "adamw": 3}[optimizer]
        self._opt = dict(kind=kind, lr=lr, weight_decay=weight_decay, momentum=momentum, beta1=betas[0], beta2=betas[1], eps=eps)

    def prepare(self, lo: int, hi: int) -> None:
        """Allocate the optimizer moments of the OWNED shard of bucket [lo, hi) (call before any graph capture)."""
        o = self._opt
        if (lo, hi) in self._state:
            return
        s0, s1 = shard_bounds(hi - lo, self.arena.world, self.arena.rank)
        need_m = o["kind"] >= 2 or o["momentum"] != 0.0
        m = torch.zeros(max(s1 - s0, 4), device="cuda") if need_m else None
        v = torch.zeros(max(s1 - s0, 4), device="cuda") if o["kind"] >= 2 else None
        self._state[(lo, hi)] = (m, v)

    def _opt_args(self, lo: int, hi: int, s0: int, s1: int) -> OptArgs:
        o = self._opt
        self.prepare(lo, hi)
        m, v = self._state[(lo, hi)]
        return OptArgs(kind=o["kind"], lr=o["lr"], weight_decay=o["weight_decay"], momentum=o["momentum"], beta1=o["beta1"],
                       beta2=o["beta2"], eps=o["eps"], m=None if m is None else m.data_ptr(), v=None if v is None else v.data_ptr(),
                       state_off4=(lo + s0) // 4, step_dev=self.step_dev.data_ptr(), bump_step=0)

    def reset_state(self) -> None:
        for m, v in self._state.values():
            if m is not None:
                m.zero_()
            if v is not None:
                v.zero_()
        self.step_dev.zero_()

    def end_of_step(self) -> None:
        """Called once per optimizer step after the last bucket (on the reducer's stream): advance the step counter."""
        if self.fuses_optimizer and self._opt is not None and self._opt["kind"] >= 2:
            with torch.cuda.stream(self.stream):
                self.step_dev.add_(1)

    # ---- reducer protocol
    def __call__(self, view: torch.Tensor, lo: int, hi: int, last: bool = False):
        a = self.arena
        ready = torch.cuda.Event()
        ready.record()
        s0, s1 = shard_bounds(hi - lo, a.world, a.rank)
        fused = self.fuses_optimizer and self._opt is not None
        ctas = max(self.max_ctas, self.tail_ctas) if last else self.max_ctas
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(ready)
            if fused:
                a.reduce(0, self.ptrs, [1.0] * a.world, self.param_ptrs, lo + s0, lo + s1, base=self.param, beta=1.0 / a.world,
                         max_ctas=ctas, channel=self.channel, opt=self._opt_args(lo, hi, s0, s1),
                         mc_in=self.mc_grad, mc_out=self.mc_param)
                view.zero_()          # every peer has read this bucket (exit barrier): ready for the next backward
            else:
                a.reduce(0, self.ptrs, [1.0] * a.world, self.ptrs, lo + s0, lo + s1, max_ctas=ctas, channel=self.channel,
                         mc_in=self.mc_grad, mc_out=self.mc_grad)
            done = torch.cuda.Event()
            done.record()
        return _Work(done)


class SymmetricWeightAverager:
    """K2b: in-place average of the replicas' parameters by the fused peer-memory kernel (each rank averages its coordinate
    shard over all replicas and stores it back into all of them).  The intended semantics of
    lab/tutorial_1b/DP/weight_aggr/intro_DP_WA.py:54-67, whose own write-back is a no-op (:67)."""

    def __init__(self, arena: SymmetricArena, n_elems: int, max_ctas: int = 0, tag: str = ""):
        self.arena = arena
        self.n = (n_elems + 127) // 128 * 128
        self.param = arena.alloc("wa_param" + tag, self.n * 4)
        self.ptrs = arena.peer_ptrs("wa_param" + tag)
        self.max_ctas = max_ctas

    def average_(self) -> None:
        a = self.arena
        s0, s1 = shard_bounds(self.n, a.world, a.rank)
        a.reduce(0, self.ptrs, [1.0 / a.world] * a.world, self.ptrs, s0, s1, max_ctas=self.max_ctas)


class FusedVFLGather:
    """``gather_linear`` hook for :class:`parallel.vfl.DistributedVFL` (K7).

    Forward: every party writes its embedding into a (double-buffered) symmetric buffer, a GPU-side barrier
    orders the writes, and ONE tcgen05 GEMM per rank computes ``cat(E_0..E_{W-1}) @ fc1.weight.T + bias`` with
    its k-blocks TMA-loaded directly from the owning GPUs — the concatenated embedding matrix never exists.
    Backward (all on the tcgen05 GEMM, no library matmul): the embedding gradient needs only this party's weight slab
    (``dY W[:, slab]``, the scatter of lab/tutorial_2b/vfl.py:81; the slab is consumed MN-major from a 32-column padded
    copy); the weight gradient ``dY^T E_p`` reads the party's embedding straight out of the symmetric buffer (MN-major,
    zero-padded to 32 columns), lands in this rank's slab of a full-width partial that is zero elsewhere, and the partials
    are summed with one fused peer-memory reduce so every replica of the top model takes the same step.
    """

    def __init__(self, arena: SymmetricArena, max_batch: int, dims, n_out: int):
        self.arena, self.dims = arena, [int(d) for d in dims]
        self.ld = (max(self.dims) + 31) // 32 * 32          # 32-column multiples: the buffer doubles as an MN-major GEMM operand
        self.max_batch = max_batch
        self.bufs = [arena.alloc(f"vfl_emb_{i}", max_batch * self.ld * 4).view(max_batch, self.ld) for i in range(2)]
        self.ptrs = [arena.peer_ptrs(f"vfl_emb_{i}") for i in range(2)]
        ktot = sum(self.dims)
        self.nw = (n_out * ktot + 127) // 128 * 128
        self.dw_part = arena.alloc("vfl_dw_part", self.nw * 4)
        self.dw_sum = arena.alloc("vfl_dw_sum", self.nw * 4)
        self.dw_part_ptrs, self.dw_sum_ptrs = arena.peer_ptrs("vfl_dw_part"), arena.peer_ptrs("vfl_dw_sum")
        self.step = 0

    def __call__(self, emb, dims, rank, weight, bias):
        return _FusedGatherLinear.apply(emb, weight, bias, self, rank)


class _FusedGatherLinear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, emb, weight, bias, h: FusedVFLGather, rank: int):
        from ..ops.gemm import gather_linear_ptrs
        B, d = emb.shape
        slot = h.step % 2
        h.step += 1
        h.bufs[slot][:B, :d].copy_(emb.detach())
        h.arena.barrier()                                   # all parties' embeddings are in place
        out = gather_linear_ptrs(h.ptrs[slot], h.dims, [h.ld] * len(h.dims), B, weight.detach(), bias.detach(), precise=True)
        ctx.save_for_backward(weight.detach())
        ctx.h, ctx.rank, ctx.slot, ctx.B = h, rank, slot, B
        return out

    @staticmethod
    def backward(ctx, gout):
        from ..ops.gemm import gemm
        from ..ops.nn import col_sum, pad2d
        (weight,) = ctx.saved_tensors
        h, rank, B = ctx.h, ctx.rank, ctx.B
        a = h.arena
        off = sum(h.dims[:rank])
        d = h.dims[rank]
        d32 = (d + 31) // 32 * 32
        n_out, ktot = weight.shape
        gout = gout.contiguous()
        if n_out % 32:                                       # top models here have 128 outputs; anything else: pad dY once
            gout = pad2d(gout, B, (n_out + 31) // 32 * 32)
        no32 = gout.shape[1]
        # scatter: only this party's slab of the weight is needed -> dE_p = dY W[:, slab]
        wslab = pad2d(weight[:, off:off + d], no32, d32)
        demb = gemm(gout, wslab, b_mn=True, precise=True)[:, :d]
        # weight gradient of the slab from the embedding that still sits in the symmetric buffer (columns >= d are zero)
        emb_p = h.bufs[ctx.slot][:B, :d32]
        dws = gemm(gout, emb_p, a_mn=True, b_mn=True, precise=True)            # [no32, d32]
        part = h.dw_part[:n_out * ktot].view(n_out, ktot)                      # zero outside this rank's slab, always
        _lib.check(_lib.lib().ddl_pad2d(_lib.p(dws), _lib.i64(dws.stride(0)), int(n_out), int(d), _lib.p(part[:, off:]), _lib.i64(ktot),
                                        int(n_out), int(d), -1, None, _lib.f32(0.0), _lib.stream()), "ddl_pad2d")
        d0, d1 = shard_bounds(h.nw, a.world, a.rank)
        a.reduce(0, h.dw_part_ptrs, [1.0] * a.world, h.dw_sum_ptrs, d0, d1)
        dw = h.dw_sum[:n_out * ktot].view(n_out, ktot).clone()
        return demb, dw, col_sum(gout)[:n_out], None, None
