"""Clients — the *client-local-train* hook.

API parity with lab/tutorial_1a/hfl_complete.py:145-155,233-256,316-332:

* ``Client.update(weights: list[Tensor], seed: int) -> list[Tensor]`` is the message-level
  contract (a list of per-parameter tensors in ``model.parameters()`` order);
* :class:`GradientClient` does one full-batch forward/backward and returns gradients (FedSGD);
* :class:`WeightClient` runs ``E`` local epochs of minibatch SGD and returns weights (FedAvg).

Differences that matter on a B200:

* messages never leave the device: ``update_flat(global_flat, seed, out=row)`` reads the global
  flat arena and writes the client's message straight into its row of the server's ``[K, D]``
  update matrix.  ``update()`` is a thin list-of-tensors wrapper around it;
* the reference gives every one of the N clients its own model on the device; here a client owns
  only its data and borrows a :class:`ModelSlot` (model + flat arena + grad arena) from the
  server, so memory scales with clients-per-round, not with N;
* the sample order is produced by :func:`core.seeds.epoch_permutations`, which reproduces the
  reference's ``DataLoader(shuffle=True, generator=g)`` order for the same seed without a
  DataLoader in the hot loop;
* the local SGD step is one fused launch over the flat arena (K15).
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Callable

import numpy as np
import torch
import torch.nn.functional as F
from torch import nn
from torch.utils.data import Subset

from .. import ops
from ..core.arena import FlatArena
from ..core.seeds import epoch_permutations
from ..models.mnist_cnn import MnistCnn


def default_device() -> torch.device:
    """cuda -> cpu (the reference also tries mps: hfl_complete.py:10-15; this framework is B200-only)."""
    return torch.device("cuda" if torch.cuda.is_available() else "cpu")


def nll_of_log_probs(output: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    return F.nll_loss(output, target)


def cross_entropy(output: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    return F.cross_entropy(output, target)


class ModelSlot:
    """A model whose parameters (and optionally float buffers) live in one flat arena."""

    def __init__(self, model_fn: Callable[[], nn.Module], device, include_buffers: bool = False, graph: bool = False):
        self.model = model_fn().to(device)
        self.arena = FlatArena.from_module(self.model, include_buffers=include_buffers, with_grad=True)
        self.device = torch.device(device)
        # graph=True: a full local SGD step (zero grads, forward, backward, fused SGD) is captured once per batch shape
        # into a CUDA graph and replayed for every client that borrows this slot (utils/graphs.py)
        self.graph = graph and self.device.type == "cuda"
        self._graphs: dict = {}

    def sgd_step_graphed(self, loss_fn, lr: float, data: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        from ..utils.graphs import GraphedStep
        key = (tuple(data.shape), tuple(target.shape), float(lr), loss_fn)
        g = self._graphs.get(key)
        if g is None:
            saved = self.arena.flat.detach().clone()
            buffers = [(b, b.clone()) for b in self.model.buffers()]
            n = self.arena.size

            def step(x, y):
                self.zero_grad()
                loss = loss_fn(self.model(x), y)
                loss.backward()
                with torch.no_grad():
                    ops.sgd_step(self.arena.flat[:n], self.arena.grad[:n], lr)
                return loss.detach()

            g = GraphedStep(step, (data, target))
            with torch.no_grad():                 # undo the warm-up / capture steps
                self.arena.flat.copy_(saved)
                for b, v in buffers:
                    b.copy_(v)
            self._graphs[key] = g
        return g(data, target)

    @property
    def flat(self) -> torch.Tensor:
        return self.arena.flat

    @property
    def grad(self) -> torch.Tensor:
        return self.arena.grad

    def zero_grad(self) -> None:
        # autograd accumulates in place into the pre-bound .grad views, so zeroing the arena is enough
        self.arena.grad.zero_()


def materialize_client_data(client_data, device=None, pin: bool = False):
    """``(x, y)`` tensors for a client from a ``Subset`` of an in-memory set, a ``(x, y)`` pair or any
    ``Dataset``."""
    if isinstance(client_data, (tuple, list)) and len(client_data) == 2 and torch.is_tensor(client_data[0]):
        x, y = client_data
    elif isinstance(client_data, Subset) and hasattr(client_data.dataset, "images"):
        idx = torch.as_tensor(np.asarray(client_data.indices), dtype=torch.long)
        x, y = client_data.dataset.images[idx], client_data.dataset.targets[idx]
    else:
        xs, ys = zip(*[client_data[i] for i in range(len(client_data))])
        x, y = torch.stack([torch.as_tensor(v) for v in xs]), torch.as_tensor(ys, dtype=torch.long)
    if device is not None:
        x, y = x.to(device), y.to(device)
    elif pin and torch.cuda.is_available():
        x, y = x.pin_memory(), y.pin_memory()
    return x, y


class Client(ABC):
    """Owns a private data shard and shuffling seed stream; borrows a model slot."""

    def __init__(self, client_data, batch_size: int, *, model_fn: Callable[[], nn.Module] = MnistCnn,
                 device=None, loss_fn=nll_of_log_probs, slot: ModelSlot | None = None, data_on_device: bool = True):
        self.device = torch.device(device) if device is not None else default_device()
        self.model_fn = model_fn
        self.loss_fn = loss_fn
        self._slot = slot
        self.x, self.y = materialize_client_data(client_data, self.device if data_on_device else None,
                                                 pin=not data_on_device)
        self.n = len(self.y)
        self.batch_size = self.n if batch_size in (-1, None) else batch_size

    # ---- model slot
    @property
    def slot(self) -> ModelSlot:
        if self._slot is None:
            self._slot = ModelSlot(self.model_fn, self.device)
        return self._slot

    def bind(self, slot: ModelSlot) -> None:
        self._slot = slot

    @property
    def model(self) -> nn.Module:
        return self.slot.model

    def batches(self, perm: torch.Tensor):
        for lo in range(0, self.n, self.batch_size):
            idx = perm[lo:lo + self.batch_size].to(self.x.device)
            yield (self.x[idx].to(self.device, non_blocking=True), self.y[idx].to(self.device, non_blocking=True))

    # ---- the hook
    @abstractmethod
    def update_flat(self, global_flat: torch.Tensor, seed: int, out: torch.Tensor | None = None) -> torch.Tensor:
        """Train from ``global_flat`` and write this client's flat message into ``out``."""

    def update(self, weights: list[torch.Tensor], seed: int) -> list[torch.Tensor]:
        """Reference message format (hfl_complete.py:154): list of tensors in, list of tensors out."""
        arena = self.slot.arena
        gflat = arena.flatten_list([w.to(self.device) for w in weights])
        msg = self.update_flat(gflat, seed)
        return arena.to_list(msg, clone=True, device="cpu")


class GradientClient(Client):
    """FedSGD client: one full-batch fwd/bwd, message = gradients (hfl_complete.py:233-256)."""

    def __init__(self, client_data, **kw):
        super().__init__(client_data, -1, **kw)

    def update_flat(self, global_flat, seed, out=None):
        slot = self.slot
        with torch.no_grad():
            slot.flat.copy_(global_flat)
        slot.zero_grad()
        slot.model.train()
        perm = epoch_permutations(self.n, seed, 1)[0]       # seeding kept for parity (hfl_complete.py:243-244)
        for data, target in self.batches(perm):             # always exactly one batch
            self.loss_fn(slot.model(data), target).backward()
        if out is None:
            return slot.grad.clone()
        out.copy_(slot.grad)
        return out


class WeightClient(Client):
    """FedAvg client: E local epochs of minibatch SGD, message = weights (hfl_complete.py:316-332)."""

    def __init__(self, client_data, lr: float, batch_size: int, nr_epochs: int, **kw):
        super().__init__(client_data, batch_size, **kw)
        self.lr, self.nr_epochs = lr, nr_epochs

    def update_flat(self, global_flat, seed, out=None):
        slot = self.slot
        with torch.no_grad():
            slot.flat.copy_(global_flat)
        slot.model.train()
        n_params = slot.arena.size
        for perm in epoch_permutations(self.n, seed, self.nr_epochs):
            for data, target in self.batches(perm):
                if slot.graph and len(target) == self.batch_size:
                    slot.sgd_step_graphed(self.loss_fn, self.lr, data, target)
                    continue
                slot.zero_grad()
                self.loss_fn(slot.model(data), target).backward()
                with torch.no_grad():
                    ops.sgd_step(slot.flat[:n_params], slot.grad[:n_params], self.lr)
        if out is None:
            return slot.flat.detach().clone()
        out.copy_(slot.flat.detach())
        return out
