"""FedAvg / robust FedAvg over MnistCnn clients packed on the GPUs (the flagship path).

Same round semantics and public API as :class:`fl.server.FedAvgServer` (reference:
lab/tutorial_1a/hfl_complete.py:336-390) — client sampling, per-client shuffle seeds, sample-count
weights, message accounting, ``run(nr_rounds) -> RunResult``, ``test() -> float`` — but the sampled
clients of a round are dealt to the ranks (one rank per GPU) and each rank trains ITS clients
concurrently with :class:`fl.packed.PackedMnistEngine` instead of one after another through
autograd.  The clients' weights never leave HBM: row ``i`` of the engine's ``[G, P]`` arena *is*
client ``i``'s message, and the aggregator kernel reads those rows in place.

Data residency: ``data_on_device=True`` keeps every client's shard in HBM (60 000 MNIST-shaped
samples are 188 MB of 180 GB); ``False`` keeps them in pinned host memory and copies the sampled
clients' shards host->device every round (the end-to-end benchmark path).
"""
from __future__ import annotations

from time import perf_counter

import numpy as np
import numpy.random as npr
import torch
import torch.distributed as dist

from .. import ops
from ..core.arena import FlatArena
from ..core.results import RunResult, message_count
from ..core.seeds import client_round_seed, clients_per_round, epoch_permutations
from ..models.mnist_cnn import MnistCnn
from .aggregators import Aggregator, WeightedMean
from .client import materialize_client_data
from .packed import MAX_G, P, PackedMnistEngine
from ..utils.trace import span
from .server import Server, _RoundTimer, _dist_ready


class _NullCtx:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class PackedFedAvgServer:
    algorithm = "FedAvg"

    def __init__(self, lr: float, batch_size: int, client_subsets, client_fraction: float, nr_local_epochs: int,
                 seed: int, *, test_set=None, device="cuda", aggregator: Aggregator | None = None, attack=None,
                 malicious: np.ndarray | None = None, data_on_device: bool = True, p_drop1: float = 0.25,
                 p_drop2: float = 0.5, eval_chunk: int = 2048, collective=None, prefetch_next: bool = True) -> None:
        self.prefetch_next = prefetch_next
        self.lr, self.batch_size, self.seed = lr, batch_size, seed
        self.nr_local_epochs = nr_local_epochs
        self.device = torch.device(device)
        self.nr_clients = len(client_subsets)
        self.client_fraction = client_fraction
        self.nr_clients_per_round = clients_per_round(self.nr_clients, client_fraction)
        self.rng = npr.default_rng(seed)
        self.aggregator = aggregator or WeightedMean()
        self.attack = attack
        self.malicious = malicious if malicious is not None else np.zeros(self.nr_clients, dtype=bool)
        self.rank, self.world = (dist.get_rank(), dist.get_world_size()) if _dist_ready() else (0, 1)
        self.p1, self.p2 = p_drop1, p_drop2
        self.collective = collective          # parallel.symm.SymmetricAggregator or None (NCCL/gloo fallback)

        # global model: seeded init, flat arena (hfl_complete.py:165-166)
        torch.manual_seed(seed)
        self.model = MnistCnn(p_drop1, p_drop2).to(self.device)
        self.arena = FlatArena.from_module(self.model)
        assert self.arena.size == P
        if collective is not None and self.world > 1 and hasattr(collective, "global_storage"):
            # keep the global model in the symmetric buffer the fused kernels write the aggregate into (on every rank):
            # the new global model needs no further copy
            self.arena.rebase(collective.global_storage())

        # client shards: [N, n_max, 784] + [N, n_max] (+ counts), on device or in pinned host memory
        xs, ys = zip(*[materialize_client_data(s) for s in client_subsets])
        self.client_sample_counts = [len(y) for y in ys]
        self.n_max = max(self.client_sample_counts)
        X = torch.zeros(self.nr_clients, self.n_max, 784)
        Y = torch.zeros(self.nr_clients, self.n_max, dtype=torch.int64)
        for i, (x, y) in enumerate(zip(xs, ys)):
            X[i, :len(y)] = x.reshape(len(y), 784)
            Y[i, :len(y)] = y
        self.data_on_device = data_on_device
        self.X_u8 = None
        if data_on_device:
            self.X_all, self.Y_all = X.to(self.device), Y.to(self.device)
        else:
            # Host-resident shards cross PCIe as the raw uint8 pixels they were made from (1 byte per pixel instead of 4) when
            # the float images are exactly the MNIST transform of 8-bit pixels; the transform then runs on the device
            # (csrc/data.cu), bit-identical to the host one.  Anything else is shipped as fp32.
            from ..core.data import MNIST_MEAN, MNIST_STD
            u8 = torch.round((X * MNIST_STD + MNIST_MEAN) * 255.0).clamp_(0, 255).to(torch.uint8)
            if self.device.type == "cuda" and torch.equal(u8.float().div_(255.0).sub_(MNIST_MEAN).div_(MNIST_STD), X):
                self.X_u8 = u8.pin_memory()
                self._norm = (MNIST_MEAN, MNIST_STD)
            self.X_all, self.Y_all = X.pin_memory(), Y.pin_memory()
        self.h2d_bytes_last_round = 0
        self.test_ms_last_round = 0.0

        self._engines: dict[int, PackedMnistEngine] = {}
        self._bufs: dict[int, tuple] = {}
        self._eval = PackedMnistEngine(1, eval_chunk, device=self.device, train=False)
        self._test_xy = None
        if test_set is not None:
            self.set_test_data(test_set)
        self.clients = list(range(self.nr_clients))    # ids; data lives in X_all / Y_all
        self._agg_out = torch.empty(P, device=self.device)
        self._copy_stream = torch.cuda.Stream(device=self.device) if self.device.type == "cuda" else None
        self._slot_free = [None, None]
        self._next_slot = 0
        self._prefetch = None            # (chosen, staged) of the round that will run next

    # ------------------------------------------------------------------ evaluation
    def set_test_data(self, test_set) -> None:
        x, y = (test_set.images, test_set.targets) if hasattr(test_set, "images") else materialize_client_data(test_set)
        self.n_test = len(y)
        xs, ys = x[self.rank::self.world], y[self.rank::self.world]      # each rank evaluates a strided shard
        self._test_xy = (xs.reshape(len(ys), 784).to(self.device).contiguous(), ys.to(self.device).contiguous())
        self._correct = torch.zeros((), dtype=torch.int32, device=self.device)
        self._correct_host = torch.zeros((), dtype=torch.int32).pin_memory() if self.device.type == "cuda" else None

    @torch.no_grad()
    def test(self) -> float:
        """Top-1 accuracy (%) of the global model (hfl_complete.py:172-183), sharded over ranks."""
        if self._test_xy is None:
            return float("nan")
        x, y = self._test_xy
        self._correct.zero_()
        self._eval.eval_logits(self.arena.flat, x, y, self._correct, want_logits=False)
        if self.world > 1:
            c = self._correct.to(torch.int64)
            dist.all_reduce(c)
            return 100.0 * int(c) / self.n_test
        return 100.0 * int(self._correct) / self.n_test       # device -> host read of the round's metric

    # ------------------------------------------------------------------ one round
    def _engine(self, G: int) -> PackedMnistEngine:
        if G not in self._engines:
            self._engines[G] = PackedMnistEngine(G, self.batch_size, device=self.device, p_drop1=self.p1, p_drop2=self.p2)
            w_rows = self.collective.row_buffer(G) if self.collective is not None else torch.empty(G, P, device=self.device)
            cuda = self.device.type == "cuda"
            n_perm = self.nr_local_epochs * self.n_max

            def staging():
                ph = torch.zeros(G, n_perm, dtype=torch.int32)
                return (torch.zeros(G, self.n_max, 784, device=self.device),
                        torch.zeros(G, self.n_max, dtype=torch.int64, device=self.device),
                        ph.pin_memory() if cuda else ph,
                        torch.zeros(G, n_perm, dtype=torch.int32, device=self.device),
                        torch.zeros(G, self.n_max, 784, dtype=torch.uint8, device=self.device) if self.X_u8 is not None else None)
            # two staging sets: while round r trains out of one, round r+1's shards / permutations are staged into the other
            self._bufs[G] = (w_rows, [staging(), staging()])
        return self._engines[G]

    def _stage(self, nr_round: int, mine: list[int], slot: int):
        """Host permutations + data movement for one round into staging set ``slot`` (on the copy stream on GPU)."""
        G = len(mine)
        self._engine(G)
        Xg, Yg, perm_host, perm_dev, Ug = self._bufs[G][1][slot]
        counts = [self.client_sample_counts[c] for c in mine]
        h2d = 0
        cuda = self.device.type == "cuda"
        ctx = torch.cuda.stream(self._copy_stream) if cuda else _NullCtx()
        if cuda and self._slot_free[slot] is not None:
            self._copy_stream.wait_event(self._slot_free[slot])     # the round that last trained out of this set is done
        with ctx:
            # shard copies first: they run on the copy engine while the host draws this round's permutations below
            for i, c in enumerate(mine):
                if Ug is not None:
                    Ug[i].copy_(self.X_u8[c], non_blocking=True)     # pinned host -> device: raw pixels, 1 byte each
                else:
                    Xg[i].copy_(self.X_all[c], non_blocking=True)    # pinned host -> device in the end-to-end mode
                Yg[i].copy_(self.Y_all[c], non_blocking=True)
                if not self.data_on_device:
                    h2d += self.X_all[c].numel() * (1 if Ug is not None else 4) + self.Y_all[c].numel() * 8
            if Ug is not None:
                ops.u8_normalize(Ug, self._norm[0], self._norm[1], out=Xg)   # on the copy stream, behind the copies
        for i, c in enumerate(mine):
            seed = client_round_seed(self.seed, c, nr_round, self.nr_clients_per_round)
            for e, pm in enumerate(epoch_permutations(counts[i], seed, self.nr_local_epochs)):
                perm_host[i, e * self.n_max:e * self.n_max + counts[i]] = pm.to(torch.int32)
        with ctx:
            perm_dev.copy_(perm_host, non_blocking=True)
            h2d += perm_host.numel() * 4
            ready = None
            if cuda:
                ready = torch.cuda.Event()
                ready.record(self._copy_stream)
        return {"round": nr_round, "mine": list(mine), "slot": slot, "ready": ready, "h2d": h2d, "counts": counts}

    def _train_local_clients(self, nr_round: int, mine: list[int], staged=None):
        """Train this rank's clients of the round; returns their ``[G, P]`` weight rows."""
        G = len(mine)
        eng = self._engine(G)
        if staged is None or staged["round"] != nr_round or staged["mine"] != list(mine):
            staged = self._stage(nr_round, mine, self._next_slot)
            self._next_slot ^= 1
        W = self._bufs[G][0]
        Xg, Yg, _perm_host, perm_dev, _Ug = self._bufs[G][1][staged["slot"]]
        gflat = self.arena.flat.detach()
        W.copy_(gflat.unsqueeze(0).expand(G, P))                 # server -> clients broadcast (hfl_complete.py:356)
        self.h2d_bytes_last_round = staged["h2d"]
        if staged["ready"] is not None:
            torch.cuda.current_stream().wait_event(staged["ready"])
        eng.local_train(W, Xg, Yg, perm_dev, staged["counts"], self.batch_size, self.nr_local_epochs, self.lr,
                        seed=self.seed * 1_000_003 + nr_round)
        if self.device.type == "cuda":
            ev = torch.cuda.Event()
            ev.record()
            self._slot_free[staged["slot"]] = ev
        if self.attack is not None:
            for i, c in enumerate(mine):
                if self.malicious[c]:
                    self.attack(W[i], gflat, client=c, nr_round=nr_round, out=W[i])
        return W

    def _aggregate(self, W_local: torch.Tensor, mine_pos: list[int], weights: torch.Tensor) -> None:
        gflat = self.arena.flat
        if self.world == 1:
            agg = self.aggregator(W_local, weights, out=self._agg_out, global_flat=gflat.detach())
            gflat.copy_(agg)
            return
        if self.collective is not None:
            self.collective.aggregate(self.aggregator, W_local, mine_pos, weights, gflat)
            return
        K = len(weights)
        if isinstance(self.aggregator, WeightedMean):
            # baseline collective path: local partial sum, then a SUM all-reduce of one model
            part = self._agg_out
            if len(mine_pos):
                ops.weighted_sum(W_local, [float(weights[p]) for p in mine_pos], part)
            else:
                part.zero_()
            dist.all_reduce(part)
            gflat.copy_(part)
        else:
            U = torch.zeros(K, P, device=self.device)
            for i, p in enumerate(mine_pos):
                U[p].copy_(W_local[i])
            dist.all_reduce(U)
            agg = self.aggregator(U, weights, out=self._agg_out, global_flat=gflat.detach())
            gflat.copy_(agg)

    def run(self, nr_rounds: int, *, checkpointer=None, logger=None) -> RunResult:
        """Rounds up to index ``nr_rounds``; ``checkpointer`` / ``logger`` as in :meth:`fl.server.Server.run` (a saved state
        is resumed: the sampling RNG is stored as it was BEFORE the next round's prefetch draw)."""
        res = RunResult(self.algorithm, self.nr_clients, self.client_fraction, self.batch_size, self.nr_local_epochs,
                        self.lr, self.seed)
        start = 0
        if checkpointer is not None and checkpointer.exists():
            start, saved = checkpointer.resume(self.arena, self.rng)
            res = saved if saved is not None else res
            self._prefetch = None                               # staged data of the interrupted run is gone
        timer = _RoundTimer(self.device)
        elapsed = res.wall_time[-1] if res.wall_time else 0.0
        for nr_round in range(start, nr_rounds):
            t0 = perf_counter()
            timer.start()
            if self._prefetch is not None:
                chosen, staged = self._prefetch                 # drawn (in rng order) and staged during the previous round
                self._prefetch = None
            else:
                chosen, staged = self.rng.choice(self.nr_clients, self.nr_clients_per_round, replace=False), None
            counts = np.asarray([self.client_sample_counts[int(c)] for c in chosen], dtype=np.float64)
            weights = torch.tensor(counts / counts.sum(), dtype=torch.float32)
            mine_pos = [p for p in range(len(chosen)) if p % self.world == self.rank]
            mine = [int(chosen[p]) for p in mine_pos]
            W_local = None
            assert len(mine) <= MAX_G, "more than MAX_G clients per GPU per round: raise MAX_G or add ranks"
            if mine:
                with span("clients/packed_local_train"):
                    W_local = self._train_local_clients(nr_round, mine, staged)
            # While the GPU trains: draw the NEXT round's clients (same rng order as drawing them later) and stage their
            # shards + permutations into the other staging set.  The next round is nr_round + 1 of this run(), or round 0
            # of a following run() call; a wrong guess only costs a re-stage.
            rng_state = self.rng.bit_generator.state            # what a resumed run must start from
            if self.prefetch_next:
                nxt_round = nr_round + 1 if nr_round + 1 < nr_rounds else 0
                nxt = self.rng.choice(self.nr_clients, self.nr_clients_per_round, replace=False)
                nxt_mine = [int(nxt[p]) for p in range(len(nxt)) if p % self.world == self.rank]
                st = None
                if nxt_mine:
                    st = self._stage(nxt_round, nxt_mine, self._next_slot)
                    self._next_slot ^= 1
                self._prefetch = (nxt, st)
            if W_local is None:
                W_local = torch.empty(0, P, device=self.device)
            with torch.no_grad(), span("server/aggregate"):
                self._aggregate(W_local, mine_pos, weights)
            ev_t = None
            if self.device.type == "cuda":
                ev_t = torch.cuda.Event(enable_timing=True)
                ev_t.record()
            with span("server/test"):
                acc = self.test()
            res.device_time_ms.append(timer.stop_ms())
            if ev_t is not None:                     # share of the round spent in the (rank-sharded) evaluation
                self.test_ms_last_round = ev_t.elapsed_time(timer.e1)
            if self.collective is not None and self.world > 1:
                self.collective.arena.check()          # the host is synchronised here anyway: a peer time-out raises instead of
                                                       # silently leaving a stale global model
            elapsed += perf_counter() - t0
            res.wall_time.append(round(elapsed, 1))
            res.message_count.append(message_count(nr_round, self.nr_clients_per_round))
            res.samples.append(int(counts.sum()) * self.nr_local_epochs)
            res.test_accuracy.append(acc)
            Server._log_round(logger, res, nr_round)
            if checkpointer is not None and self.rank == 0:
                checkpointer.maybe_save(nr_round, self.arena, rng_state, res)
        return res
