"""Seeding conventions (hfl_complete.py:165,209,229,289,327,368)."""
from __future__ import annotations

import numpy as np
import torch


def client_round_seed(seed: int, client_index: int, nr_round: int, clients_per_round: int) -> int:
    """``seed + k + 1 + round * K`` – the reference's per-client-per-round shuffle seed."""
    return seed + int(client_index) + 1 + nr_round * clients_per_round


def clients_per_round(nr_clients: int, client_fraction: float) -> int:
    """``max(1, round(C * N))`` (hfl_complete.py:228)."""
    return max(1, round(client_fraction * nr_clients))


def loader_permutation(n: int, seed: int) -> torch.Tensor:
    """Sample order a ``DataLoader(shuffle=True, generator=g)`` yields after ``g.manual_seed(seed)``.

    The DataLoader iterator first draws one int64 from the generator (its worker base seed)
    and then ``RandomSampler`` draws ``randperm(n)``.  Reproducing that on the host lets the
    packed GPU engine visit samples in exactly the reference's order without a DataLoader.
    """
    return epoch_permutations(n, seed, 1)[0]


def epoch_permutations(n: int, seed: int, epochs: int) -> list[torch.Tensor]:
    """Sample orders of ``epochs`` consecutive passes over a ``DataLoader(shuffle=True,
    generator=g)`` after one ``g.manual_seed(seed)`` – what ``WeightClient.update`` sees
    (hfl_complete.py:327-330): the generator state carries over from epoch to epoch."""
    g = torch.Generator()
    g.manual_seed(seed)
    out = []
    for _ in range(epochs):
        torch.empty((), dtype=torch.int64).random_(generator=g)   # DataLoader iterator base seed
        out.append(torch.randperm(n, generator=g))
        torch.randperm(n, generator=g)   # RandomSampler's trailing (empty-slice) randperm draw
    return out


def sample_clients(rng: np.random.Generator, nr_clients: int, k: int) -> np.ndarray:
    """Without-replacement client sampling (hfl_complete.py:278-279,357-358)."""
    return rng.choice(nr_clients, k, replace=False)


def seed_everything(seed: int) -> None:
    torch.manual_seed(seed)
    np.random.seed(seed % (2 ** 32))
