"""Servers — the *server-aggregate* hook and the federated round loop.

API parity with lab/tutorial_1a/hfl_complete.py:159-390 and lab/homework-1.ipynb cell 12:

* ``Server.run(nr_rounds) -> RunResult`` and ``Server.test() -> float`` (top-1 accuracy in %);
* :class:`CentralizedServer`, :class:`FedSgdGradientServer`, :class:`FedSgdWeightServer`,
  :class:`FedAvgServer` with the reference's constructor argument order;
* client sampling ``rng.choice(N, max(1, round(C N)), replace=False)``, per-client seed
  ``seed + k + 1 + round K`` and message accounting ``2 (round+1) K`` are kept bit-for-bit.

What changes on a B200:

* the global model is a flat device arena; the per-round "snapshot to CPU / copy back"
  (hfl_complete.py:277,356,380-383) disappears;
* client messages land in rows of one ``[K, D]`` device matrix and the aggregation rule
  (:mod:`fl.aggregators`) is one fused kernel over it; an :class:`~fl.attacks.Attack` can rewrite
  the rows of malicious clients first (attack-and-defense hook);
* with ``torch.distributed`` initialised the sampled clients of a round are dealt round-robin to
  the ranks (several simulated clients per GPU) and the rows are combined across ranks — by the
  fused peer-memory kernels of :mod:`parallel.symm` on NVLink, or by gloo/NCCL collectives
  (CPU plumbing and the baseline);
* ``wall_time`` keeps the reference's *simulated-parallel* meaning (setup + slowest client +
  aggregation, hfl_complete.py:283-296); ``device_time_ms`` is the real device time of the round.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from time import perf_counter
from typing import Callable, Sequence

import numpy as np
import numpy.random as npr
import torch
import torch.distributed as dist
from torch import nn

from .. import ops
from ..core.results import RunResult, message_count
from ..core.seeds import client_round_seed, clients_per_round
from ..models.mnist_cnn import MnistCnn
from ..utils.trace import span
from .aggregators import Aggregator, WeightedMean
from .attacks import Attack
from .client import (Client, GradientClient, ModelSlot, WeightClient, default_device, materialize_client_data,
                     nll_of_log_probs)


def _dist_ready() -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


class _RoundTimer:
    """CUDA-event timing on GPU, perf_counter on CPU."""

    def __init__(self, device: torch.device):
        self.cuda = device.type == "cuda"

    def start(self):
        if self.cuda:
            self.e0, self.e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            self.e0.record()
        self.t0 = perf_counter()

    def stop_ms(self) -> float:
        if self.cuda:
            self.e1.record()
            self.e1.synchronize()
            return self.e0.elapsed_time(self.e1)
        return (perf_counter() - self.t0) * 1e3


class Server(ABC):
    def __init__(self, lr: float, batch_size: int, seed: int, *, model_fn: Callable[[], nn.Module] = MnistCnn,
                 test_set=None, device=None, include_buffers: bool = False, test_batch_size: int = 10_000,
                 loss_fn=nll_of_log_probs, graph: bool = False) -> None:
        self.clients: list[Client] = []
        self.lr, self.batch_size, self.seed = lr, batch_size, seed
        self.device = torch.device(device) if device is not None else default_device()
        self.model_fn = model_fn
        self.loss_fn = loss_fn                       # nll for log-prob models (MnistCnn), cross-entropy for logit models
        self.include_buffers = include_buffers
        self.graph = graph                           # replay each local SGD step from a CUDA graph (ModelSlot.sgd_step_graphed)
        self._eval_graphs: dict = {}
        torch.manual_seed(seed)                      # global-model init is seeded (hfl_complete.py:165)
        self.slot = ModelSlot(model_fn, self.device, include_buffers=include_buffers)
        self.model = self.slot.model
        self.arena = self.slot.arena
        self.test_batch_size = test_batch_size
        self._test_xy = None
        if test_set is not None:
            self.set_test_data(test_set)

    # ------------------------------------------------------------------ evaluation
    def set_test_data(self, test_set) -> None:
        if hasattr(test_set, "images"):
            x, y = test_set.images, test_set.targets
        else:
            x, y = materialize_client_data(test_set)
        self._test_xy = (x.to(self.device), y.to(self.device))

    @abstractmethod
    def run(self, nr_rounds: int, *, checkpointer=None, logger=None) -> RunResult:
        """Run rounds up to index ``nr_rounds``.  ``checkpointer`` (core.checkpoint.RoundCheckpointer): state is saved after
        every ``checkpointer.every`` rounds and, if a saved state exists, the run CONTINUES from it (same RunResult and
        weights as an uninterrupted run).  ``logger`` (utils.trace.JsonlLogger): one record per round."""

    def _resume(self, checkpointer, rng, res: RunResult) -> tuple[int, RunResult]:
        if checkpointer is None or not checkpointer.exists():
            return 0, res
        start, saved = checkpointer.resume(self.arena, rng)
        return start, (saved if saved is not None else res)

    @staticmethod
    def _log_round(logger, res: RunResult, nr_round: int) -> None:
        if logger is not None:
            logger.log(event="round", algorithm=res.algorithm, round=nr_round + 1, test_accuracy=res.test_accuracy[-1],
                       device_ms=res.device_time_ms[-1], samples=res.samples[-1], messages=res.message_count[-1])

    @torch.no_grad()
    def test(self) -> float:
        """Top-1 accuracy in percent on the held-out set (hfl_complete.py:172-183).  With several
        ranks each evaluates a strided shard and the correct-counts are summed."""
        if self._test_xy is None:
            return float("nan")
        x, y = self._test_xy
        rank, world = (dist.get_rank(), dist.get_world_size()) if _dist_ready() else (0, 1)
        xs, ys = x[rank::world].contiguous(), y[rank::world].contiguous()   # the eval kernel reads raw pointers
        self.model.eval()
        correct = torch.zeros((), dtype=torch.int32, device=self.device)
        for lo in range(0, len(ys), self.test_batch_size):
            xb = xs[lo:lo + self.test_batch_size]
            if self.graph and xb.is_cuda:
                # inference forward replayed from a CUDA graph per batch shape (weights / BN statistics live in the static arena)
                key = tuple(xb.shape)
                g = self._eval_graphs.get(key)
                if g is None:
                    from ..utils.graphs import GraphedStep
                    g = self._eval_graphs[key] = GraphedStep(lambda t: self.model(t).float().contiguous(), (xb,))
                out = g(xb)
            else:
                out = self.model(xb).float().contiguous()
            ops.eval_top1(out, ys[lo:lo + self.test_batch_size], correct)
        if world > 1:
            c64 = correct.to(torch.int64)
            dist.all_reduce(c64)
            correct = c64
        return 100.0 * int(correct) / len(y)


class CentralizedServer(Server):
    """Plain minibatch SGD over the whole training set; one "round" = one epoch
    (hfl_complete.py:193-216)."""

    def __init__(self, lr: float, batch_size: int, seed: int, *, train_set, **kw) -> None:
        super().__init__(lr, batch_size, seed, **kw)
        self.trainer = WeightClient(train_set, lr, batch_size, 1, model_fn=self.model_fn, device=self.device,
                                    slot=self.slot, loss_fn=self.loss_fn)
        self.clients = []

    def run(self, nr_rounds: int, *, checkpointer=None, logger=None) -> RunResult:
        res = RunResult("Centralized", 1, 1, self.batch_size, 1, self.lr, self.seed)
        rng = npr.default_rng(self.seed)                       # unused by the epochs; kept so checkpoints have one format
        start, res = self._resume(checkpointer, rng, res)
        elapsed = res.wall_time[-1] if res.wall_time else 0.0
        timer = _RoundTimer(self.device)
        for epoch in range(start, nr_rounds):
            t0 = perf_counter()
            timer.start()
            # same reseeding rule as the reference: generator.manual_seed(seed + epoch + 1)
            with span("centralized/epoch"):
                self.trainer.update_flat(self.slot.flat.detach().clone(), self.seed + epoch + 1, out=None)
            res.device_time_ms.append(timer.stop_ms())
            elapsed += perf_counter() - t0
            res.wall_time.append(round(elapsed, 1))
            res.message_count.append(0)
            res.samples.append(self.trainer.n)
            with span("server/test"):
                res.test_accuracy.append(self.test())
            self._log_round(logger, res, epoch)
            if checkpointer is not None:
                checkpointer.maybe_save(epoch, self.arena, rng, res)
        return res


class DecentralizedServer(Server):
    """Shared federated set-up (hfl_complete.py:220-229) plus the attack/defense hooks."""

    algorithm = "Decentralized"
    message_is_gradient = False

    def __init__(self, lr: float, batch_size: int, client_subsets: Sequence, client_fraction: float, seed: int, *,
                 aggregator: Aggregator | None = None, attack: Attack | None = None,
                 malicious: np.ndarray | None = None, collective=None, concurrent_clients: int = 1, **kw) -> None:
        super().__init__(lr, batch_size, seed, **kw)
        self.nr_clients = len(client_subsets)
        self.client_fraction = client_fraction
        self.client_sample_counts = [len(s[1]) if isinstance(s, (tuple, list)) else len(s) for s in client_subsets]
        self.nr_clients_per_round = clients_per_round(self.nr_clients, client_fraction)
        self.rng = npr.default_rng(seed)
        self.aggregator = aggregator or WeightedMean()
        self.attack = attack
        self.malicious = malicious if malicious is not None else np.zeros(self.nr_clients, dtype=bool)
        self.rank, self.world = (dist.get_rank(), dist.get_world_size()) if _dist_ready() else (0, 1)
        # one borrowed model slot for all of this rank's clients (they train one after another here;
        # the packed engine in fl/packed.py trains them concurrently with grouped kernels)
        self.worker_slot = ModelSlot(self.model_fn, self.device, include_buffers=self.include_buffers, graph=self.graph)
        # concurrent_clients = S > 1 (CUDA): S model slots, each with its own stream, convolution workspace and step graph; the
        # rank's clients of a round are dealt to the slots and train SIDE BY SIDE.  A ResNet-18 client step at the lab batch
        # sizes is ~200 kernels that each fill a fraction of the 148 SMs, so several clients' graphs overlap on the GPU - the
        # reference trains them strictly one after another (hfl_complete.py:365-375) and merely *reports* parallel time (:373-375).
        self.worker_slots = [self.worker_slot]
        self._slot_streams = None
        if concurrent_clients > 1 and self.device.type == "cuda":
            self.worker_slots += [ModelSlot(self.model_fn, self.device, include_buffers=self.include_buffers, graph=self.graph)
                                  for _ in range(concurrent_clients - 1)]
            self._slot_streams = [torch.cuda.Stream(device=self.device) for _ in self.worker_slots]
        self._updates: torch.Tensor | None = None
        # parallel.symm_impl.SymmetricAggregator(row_elems=arena.size): fused peer-memory aggregation across ranks
        self.collective = collective

    # ------------------------------------------------------------------ helpers
    def _update_matrix(self, k: int) -> torch.Tensor:
        if self._updates is None or self._updates.shape[0] != k:
            self._updates = torch.zeros(k, self.arena.size, device=self.device)
        return self._updates

    def _collect_updates(self, nr_round: int, chosen: np.ndarray, gflat: torch.Tensor):
        """Run the chosen clients (this rank's share) and return ``(updates[K, D], slowest_client_s)``."""
        K = len(chosen)
        fused = self.collective is not None and self.world > 1
        U = self.collective.row_buffer(-(-K // self.world)) if fused else self._update_matrix(K)
        if self.world > 1 and not fused:
            U.zero_()
        slowest = 0.0
        streams = self._slot_streams
        main = torch.cuda.current_stream(self.device) if streams else None
        if streams:
            for st in streams:
                st.wait_stream(main)                     # the global model (and U) are ready
        n_local = 0
        for pos, c_i in enumerate(chosen):
            if pos % self.world != self.rank:
                continue
            t0 = perf_counter()
            ind = int(c_i)
            seed = client_round_seed(self.seed, ind, nr_round, self.nr_clients_per_round)
            client = self.clients[ind]
            si = n_local % len(self.worker_slots)
            n_local += 1
            client.bind(self.worker_slots[si])
            row = U[pos // self.world] if fused else U[pos]      # fused: this rank's rows only, in symmetric memory
            if streams:
                from ..ops.conv import workspace_scope
                with torch.cuda.stream(streams[si]), workspace_scope(si):
                    client.update_flat(gflat, seed, out=row)
                    if self.attack is not None and self.malicious[ind]:
                        self.attack(row, None if self.message_is_gradient else gflat, client=ind, nr_round=nr_round, out=row)
            else:
                client.update_flat(gflat, seed, out=row)
                if self.attack is not None and self.malicious[ind]:
                    self.attack(row, None if self.message_is_gradient else gflat, client=ind, nr_round=nr_round, out=row)
            slowest = max(slowest, perf_counter() - t0)
        if streams:
            for st in streams:
                main.wait_stream(st)
        if self.world > 1 and not fused:
            # rows trained elsewhere are zero here: a SUM all-reduce assembles the full matrix on
            # every rank (gloo/NCCL path; parallel/symm.py replaces this with peer-memory kernels)
            dist.all_reduce(U)
        return U, slowest

    def _round_weights(self, chosen: np.ndarray) -> torch.Tensor:
        counts = np.asarray([self.client_sample_counts[int(i)] for i in chosen], dtype=np.float64)
        return torch.tensor(counts / counts.sum(), dtype=torch.float32)

    @abstractmethod
    def _apply(self, aggregated: torch.Tensor) -> None:
        """Fold the aggregated message into the global model."""

    def _result(self) -> RunResult:
        return RunResult(self.algorithm, self.nr_clients, self.client_fraction, self.batch_size,
                         getattr(self, "nr_local_epochs", 1), self.lr, self.seed)

    def samples_per_round(self, chosen) -> int:
        return sum(self.client_sample_counts[int(i)] for i in chosen) * getattr(self, "nr_local_epochs", 1)

    # ------------------------------------------------------------------ the round loop
    def run(self, nr_rounds: int, *, checkpointer=None, logger=None) -> RunResult:
        res = self._result()
        start, res = self._resume(checkpointer, self.rng, res)
        elapsed = res.wall_time[-1] if res.wall_time else 0.0
        timer = _RoundTimer(self.device)
        agg_out = torch.empty(self.arena.size, device=self.device)
        for nr_round in range(start, nr_rounds):
            t_setup = perf_counter()
            timer.start()
            self.model.train()
            gflat = self.arena.flat.detach()
            chosen = self.rng.choice(self.nr_clients, self.nr_clients_per_round, replace=False)
            weights = self._round_weights(chosen)
            elapsed += perf_counter() - t_setup
            with span("clients/local_update"):
                U, slowest = self._collect_updates(nr_round, chosen, gflat)
            elapsed += slowest
            t_agg = perf_counter()
            with torch.no_grad(), span("server/aggregate"):
                if self.collective is not None and self.world > 1:
                    mine_pos = [p for p in range(len(chosen)) if p % self.world == self.rank]
                    # gradient messages are aggregated around 0, weight messages around the current global model (Krum scores
                    # deltas); the fused kernels leave the aggregate in agg_out on every rank
                    if self.message_is_gradient:
                        agg_out.zero_()
                    else:
                        agg_out.copy_(gflat)
                    self.collective.aggregate(self.aggregator, U[:len(mine_pos)], mine_pos, weights, agg_out)
                    agg = agg_out
                else:
                    agg = self.aggregator(U, weights, out=agg_out,
                                          global_flat=None if self.message_is_gradient else gflat)
                self._apply(agg)
            res.device_time_ms.append(timer.stop_ms())
            elapsed += perf_counter() - t_agg
            res.wall_time.append(round(elapsed, 1))
            res.message_count.append(message_count(nr_round, self.nr_clients_per_round))
            res.samples.append(self.samples_per_round(chosen))
            with span("server/test"):
                res.test_accuracy.append(self.test())
            if self.collective is not None and self.world > 1:
                self.collective.arena.check()      # test() synchronised the host: raise on a peer time-out instead of going on
            self._log_round(logger, res, nr_round)
            if checkpointer is not None and self.rank == 0:
                checkpointer.maybe_save(nr_round, self.arena, self.rng, res)
        return res


class FedSgdGradientServer(DecentralizedServer):
    """FedSGD: clients send full-batch gradients, the server takes one SGD step on their
    sample-weighted mean (hfl_complete.py:260-312)."""
    algorithm = "FedSGDGradient"
    message_is_gradient = True

    def __init__(self, lr: float, client_subsets, client_fraction: float, seed: int, **kw) -> None:
        super().__init__(lr, -1, client_subsets, client_fraction, seed, **kw)
        ckw = dict(model_fn=self.model_fn, device=self.device, slot=self.worker_slot, loss_fn=self.loss_fn)
        self.clients = [GradientClient(subset, **ckw) for subset in client_subsets]

    def _apply(self, aggregated):
        ops.sgd_step(self.arena.flat, aggregated, self.lr)


class FedSgdWeightServer(DecentralizedServer):
    """Homework A1 variant (lab/homework-1.ipynb cell 12): same gradients, routed through the
    ``.grad`` arena and an optimizer step.  Kept as a separate class because the reference shows
    it is round-for-round identical to :class:`FedSgdGradientServer` (cells 17-18) and the test
    suite re-checks that invariant."""
    algorithm = "FedSGDWeight"
    message_is_gradient = True

    def __init__(self, lr: float, client_subsets, client_fraction: float, seed: int, **kw) -> None:
        super().__init__(lr, -1, client_subsets, client_fraction, seed, **kw)
        ckw = dict(model_fn=self.model_fn, device=self.device, slot=self.worker_slot, loss_fn=self.loss_fn)
        self.clients = [GradientClient(subset, **ckw) for subset in client_subsets]

    def _apply(self, aggregated):
        self.arena.grad.copy_(aggregated)
        ops.sgd_step(self.arena.flat, self.arena.grad, self.lr)


class FedAvgServer(DecentralizedServer):
    """FedAvg: clients run E local epochs and send weights; the server overwrites the global model
    with their sample-weighted mean — or a robust rule (hfl_complete.py:336-390)."""
    algorithm = "FedAvg"

    def __init__(self, lr: float, batch_size: int, client_subsets, client_fraction: float, nr_local_epochs: int,
                 seed: int, **kw) -> None:
        super().__init__(lr, batch_size, client_subsets, client_fraction, seed, **kw)
        self.name = "FedAvg"
        self.nr_local_epochs = nr_local_epochs
        ckw = dict(model_fn=self.model_fn, device=self.device, slot=self.worker_slot, loss_fn=self.loss_fn)
        self.clients = [WeightClient(subset, lr, batch_size, nr_local_epochs, **ckw) for subset in client_subsets]

    def _apply(self, aggregated):
        self.arena.flat.copy_(aggregated)
