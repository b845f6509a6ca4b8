"""Data parallelism: bucketed gradient all-reduce overlapped with backward, and weight averaging.

Reference behaviour being replaced (lab/tutorial_1b/DP):

* gradient aggregation (gradient_aggr/intro_DP_GA.py:53-67): after the WHOLE backward, barrier, flatten
  every gradient into one vector, copy it to the CPU, one gloo all-reduce, copy back, divide by W, step.
  Nothing overlaps.
* weight aggregation (weight_aggr/intro_DP_WA.py:52-67): local step, then all-reduce of the weights — but
  the average is assigned to the loop variable (``param = ...``, :67) and never written back, so replicas
  silently diverge.  :class:`WeightAveraging` implements the intended semantics.

Here the gradients already live in one flat arena (no flatten/copy), the arena is cut into buckets in
REVERSE parameter order (the order autograd produces them), and a post-accumulate-grad hook launches the
reduction of a bucket the moment its last gradient lands, so communication runs under the rest of the
backward.  The reduction is pluggable:

* ``torch.distributed`` (NCCL / gloo) all-reduce per bucket, then ONE fused optimizer kernel over the arena with the 1/W
  folded in (``grad_scale``) - the baseline;
* :class:`parallel.symm_impl.SymmetricGradReducer` (K2): one peer-memory kernel per bucket that is reduce-scatter + 1/W +
  optimizer update of the owned parameter shard + all-gather of the new parameters.  Gradients and parameters live in
  symmetric memory, optimizer moments are sharded over the ranks, there is no separate optimizer or zero-grad pass, and
  with ``graph=True`` the bucket kernels are captured as side branches of the step graph, so the whole step - forward,
  backward, overlapped communication, update - is a single ``cudaGraphLaunch`` per rank.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch
import torch.distributed as dist
from torch import nn

from .. import ops
from ..core.arena import FlatArena


@dataclass
class Bucket:
    lo: int                      # element range [lo, hi) of the flat gradient arena
    hi: int
    n_params: int
    ready: int = 0
    work: object = None
    names: list[str] = field(default_factory=list)


def make_buckets(arena: FlatArena, bucket_bytes: int) -> tuple[list[Bucket], dict[str, int]]:
    """Greedy buckets over the parameter list walked backwards; returns (buckets, param name -> bucket)."""
    specs = [s for s in arena.specs if not s.is_buffer]
    buckets: list[Bucket] = []
    owner: dict[str, int] = {}
    cur_hi = arena.size if specs else 0          # include the arena's zero padding: bucket ends stay float4-aligned
    cur_names: list[str] = []
    cur_lo = cur_hi
    for s in reversed(specs):
        cur_lo = s.offset
        cur_names.append(s.name)
        if (cur_hi - cur_lo) * 4 >= bucket_bytes:
            buckets.append(Bucket(cur_lo, cur_hi, len(cur_names), names=cur_names))
            cur_hi, cur_names = cur_lo, []
    if cur_names:
        buckets.append(Bucket(cur_lo, cur_hi, len(cur_names), names=cur_names))
    for i, b in enumerate(buckets):
        for n in b.names:
            owner[n] = i
    return buckets, owner


class DataParallel:
    """Wraps ``module`` for synchronous data-parallel training.

    >>> dp = DataParallel(model, optimizer="adam", lr=8e-4)
    >>> loss = loss_fn(dp.module(x), y); dp.backward(loss); dp.step()
    """

    def __init__(self, module: nn.Module, *, bucket_mb: float = 4.0, optimizer: str = "sgd", lr: float = 0.01,
                 weight_decay: float = 0.0, betas=(0.9, 0.999), eps: float = 1e-8, momentum: float = 0.0, group=None,
                 reducer=None, overlap: bool = True, graph: bool = False, check_every: int = 50):
        self.module = module
        self.check_every = check_every
        # graph=True: forward + backward of train_step() are captured once into a CUDA graph and replayed (utils/graphs.py).
        # With the fused peer-memory reducer the bucket kernels (reduce + optimizer + broadcast) are part of that graph;
        # with NCCL the bucket reductions and the optimizer kernel are issued right after the replay.
        self.graph = graph and torch.cuda.is_available()
        self._gstep = None
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.arena = FlatArena.from_module(module, with_grad=True)
        self.buckets, self._owner = make_buckets(self.arena, int(bucket_mb * 2 ** 20))
        self.reducer = reducer                   # callable(flat_grad_view, lo, hi) -> handle with .wait(), or None
        self.optimizer, self.lr, self.wd, self.betas, self.eps, self.momentum = optimizer, lr, weight_decay, betas, eps, momentum
        # fused path: the reducer's kernel also applies the optimizer to the owned shard and broadcasts the new parameters
        self.fused_step = bool(reducer is not None and getattr(reducer, "fuses_optimizer", False) and self.world > 1)
        if reducer is not None and hasattr(reducer, "grad"):
            # the fused peer-memory reducer works in place on a symmetric buffer: make it the gradient arena
            assert reducer.grad.numel() >= self.arena.size, "reducer buffers are smaller than the model"
            self.arena.grad = reducer.grad[:self.arena.size]
            self.arena.grad.zero_()
            self.arena.bind_grads()
        if self.fused_step:
            self.arena.rebase(reducer.param)     # parameters move to symmetric memory: peers store the updated shards here
            reducer.configure(optimizer, lr, weight_decay, betas, eps, momentum)
            for b in self.buckets:               # optimizer state of the owned shards, allocated before any graph capture
                reducer.prepare(b.lo, b.hi)
        self.overlap = overlap
        self.step_count = 0
        self._m = self._v = None
        if not self.fused_step:
            if optimizer in ("adam", "adamw"):
                self._m, self._v = torch.zeros_like(self.arena.flat), torch.zeros_like(self.arena.flat)
            elif momentum:
                self._m = torch.zeros_like(self.arena.flat)
        self._launched: list[int] = []
        # hooks launch a bucket as soon as its last gradient has landed; inside a graph capture only reducers whose launches
        # are capturable (no per-call host state) may do that - NCCL buckets are then reduced after the replay
        self._hooks_in_graph = bool(reducer is not None and getattr(reducer, "capturable", False))
        if self.world > 1 and overlap:
            for name, p in module.named_parameters():
                if p.requires_grad:
                    p.register_post_accumulate_grad_hook(self._make_hook(self._owner[name]))
        self.sync_params()

    # ------------------------------------------------------------------ set-up
    def sync_params(self) -> None:
        """Rank 0's initial weights everywhere (the labs rely on identical seeds instead)."""
        if self.world > 1:
            dist.broadcast(self.arena.flat, src=dist.get_global_rank(self.group, 0) if self.group else 0, group=self.group)

    def _make_hook(self, bi: int):
        def hook(_param):
            if self.graph and not self._hooks_in_graph:   # captured backward + NCCL: buckets are reduced after the replay
                return
            b = self.buckets[bi]
            b.ready += 1
            if b.ready == b.n_params:
                self._launch(bi)
        return hook

    def _launch(self, bi: int) -> None:
        b = self.buckets[bi]
        view = self.arena.grad[b.lo:b.hi]
        if self.reducer is not None:
            if hasattr(self.reducer, "tail_ctas"):   # the step's last bucket overlaps nothing: the reducer may use the whole GPU for it
                b.work = self.reducer(view, b.lo, b.hi, last=len(self._launched) == len(self.buckets) - 1)
            else:
                b.work = self.reducer(view, b.lo, b.hi)
        else:
            b.work = dist.all_reduce(view, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        self._launched.append(bi)

    # ------------------------------------------------------------------ training step
    def zero_grad(self) -> None:
        self.arena.grad.zero_()
        for b in self.buckets:
            b.ready, b.work = 0, None
        self._launched = []

    def backward(self, loss: torch.Tensor) -> None:
        """Backward with bucket reductions launched from the autograd hooks; returns once all buckets are
        reduced (gradients hold the SUM over ranks; the 1/W is applied in :meth:`step`)."""
        loss.backward()
        self.finish()

    def finish(self) -> None:
        if self.world == 1:
            return
        for bi, b in enumerate(self.buckets):
            if b.work is None:            # parameters without gradient this step, or overlap disabled
                self._launch(bi)
        if self.fused_step:
            self.reducer.end_of_step()
        for b in self.buckets:
            if b.work is not None and hasattr(b.work, "wait"):
                b.work.wait()
        if self.fused_step:               # also joins end_of_step() issued on the reducer's stream
            torch.cuda.current_stream().wait_stream(self.reducer.stream)

    def step(self) -> None:
        """Fused optimizer step over the flat arena (K15) with the DP mean folded in.  With the fused reducer the update
        has already been applied by the bucket kernels (and the gradients re-zeroed): only bookkeeping is left."""
        self.step_count += 1
        if self.fused_step:
            for b in self.buckets:
                b.ready, b.work = 0, None
            self._launched = []
            if self.check_every and self.step_count % self.check_every == 0:
                self.reducer.arena.check()       # one device read every `check_every` steps: a missed collective raises
            return
        scale = 1.0 / self.world
        with torch.no_grad():
            if self.optimizer == "sgd":
                ops.sgd_step(self.arena.flat, self.arena.grad, self.lr, self.wd, self.momentum, self._m, grad_scale=scale)
            else:
                ops.adam_step(self.arena.flat, self.arena.grad, self._m, self._v, self.step_count, self.lr, self.betas[0],
                              self.betas[1], self.eps, self.wd, decoupled=self.optimizer == "adamw", grad_scale=scale)
        self.zero_grad()

    def train_step(self, loss_fn, *batch) -> torch.Tensor:
        if self.graph:
            in_graph = self._hooks_in_graph and self.world > 1
            if self._gstep is None:
                from ..utils.graphs import GraphedStep

                def fwd_bwd(*b):
                    self._launched = []
                    loss = loss_fn(self.module, *b)
                    loss.backward()                      # accumulates in place into the (static) flat gradient arena
                    if in_graph:
                        self.finish()                    # leftover buckets + join of the reducer's stream: part of the graph
                        for bk in self.buckets:
                            bk.ready, bk.work = 0, None
                    return loss.detach()

                buffers = [(b, b.clone()) for b in self.module.buffers()]
                params = self.arena.flat.clone() if self.fused_step else None
                self._gstep = GraphedStep(fwd_bwd, batch)
                torch.cuda.synchronize()
                self.arena.grad.zero_()                  # warm-up and capture runs left gradients behind ...
                with torch.no_grad():
                    for b, saved in buffers:             # ... advanced the BatchNorm running statistics ...
                        b.copy_(saved)
                    if params is not None:               # ... and (fused reducer) took real optimizer steps
                        self.arena.flat.copy_(params)
                        self.reducer.reset_state()
                        torch.cuda.synchronize()
                        self.reducer.arena.barrier()     # nobody replays before every rank has restored its replica
                for bk in self.buckets:
                    bk.ready, bk.work = 0, None
            loss = self._gstep(*batch)
            if not in_graph:
                self.finish()
            self.step()
            return loss
        loss = loss_fn(self.module, *batch)
        self.backward(loss)
        self.step()
        return loss.detach()


class WeightAveraging:
    """DP by weight aggregation: every rank takes its own optimizer step, then the weights are averaged
    and WRITTEN BACK (the reference computes the average and drops it: intro_DP_WA.py:67)."""

    def __init__(self, module: nn.Module, optimizer: torch.optim.Optimizer | None = None, *, lr: float = 0.01,
                 group=None, every: int = 1, averager=None):
        self.module, self.group, self.every = module, group, every
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.arena = FlatArena.from_module(module, with_grad=True)
        # averager: parallel.symm_impl.SymmetricWeightAverager - the parameters move into symmetric memory and ONE fused
        # peer-memory kernel per rank averages them in place (K2b); None: torch.distributed all_reduce + div (baseline)
        self.averager = averager if (averager is not None and self.world > 1) else None
        if self.averager is not None:
            self.arena.rebase(self.averager.param)
        self.optimizer = optimizer
        self.lr = lr
        self.iters = 0
        if self.world > 1:
            dist.broadcast(self.arena.flat, src=0, group=group)

    def step(self) -> None:
        if self.optimizer is not None:
            self.optimizer.step()
            self.optimizer.zero_grad(set_to_none=False)
        else:
            with torch.no_grad():
                ops.sgd_step(self.arena.flat, self.arena.grad, self.lr)
            self.arena.grad.zero_()
        self.iters += 1
        if self.world > 1 and self.iters % self.every == 0:
            with torch.no_grad():
                if self.averager is not None:
                    self.averager.average_()            # fused: reduce-scatter + 1/W + all-gather over peer memory, in place
                else:
                    dist.all_reduce(self.arena.flat, group=self.group)
                    self.arena.flat.div_(self.world)    # in place: every parameter view sees the average
