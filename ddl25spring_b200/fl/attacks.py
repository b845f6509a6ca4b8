"""Byzantine client behaviours — the *attack* hook on the ``Client.update`` return path.

The reference lists labs 3A/3B (attacks, defenses) but ships no code for them;
the north star asks for a sign-flip attacker fraction of 20 % (BASELINE.json config 4).  An
:class:`Attack` rewrites the flat message of a malicious client before the server sees it; on
CUDA that is one launch of ``attack_perturb`` (K6) writing straight into the client's row of
the update matrix.
"""
from __future__ import annotations

from abc import ABC, abstractmethod

import numpy as np
import torch

from .. import ops


class Attack(ABC):
    name = "attack"

    @abstractmethod
    def __call__(self, update: torch.Tensor, global_flat: torch.Tensor | None, *, client: int, nr_round: int,
                 out: torch.Tensor | None = None) -> torch.Tensor:
        """Return the poisoned flat message (may write in place into ``out``)."""


class SignFlip(Attack):
    """``w_g - s (w_k - w_g)`` for weight messages, ``-s g`` for gradient messages (pass
    ``global_flat=None``)."""
    name = "sign_flip"

    def __init__(self, scale: float = 1.0):
        self.scale = scale

    def __call__(self, update, global_flat, *, client, nr_round, out=None):
        return ops.attack_perturb(update, global_flat, self.scale, 0.0, 0, out)


class ScaledUpdate(Attack):
    """Model-replacement style boosting: ``w_g + s (w_k - w_g)`` (sign-flip with negative scale)."""
    name = "scaling"

    def __init__(self, scale: float = 10.0):
        self.scale = scale

    def __call__(self, update, global_flat, *, client, nr_round, out=None):
        return ops.attack_perturb(update, global_flat, -self.scale, 0.0, 0, out)


class NoiseAttack(Attack):
    """Honest update plus uniform noise in ``[-amp, amp]`` (counter-based RNG on device)."""
    name = "noise"

    def __init__(self, amplitude: float = 1.0, seed: int = 0):
        self.amplitude, self.seed = amplitude, seed

    def __call__(self, update, global_flat, *, client, nr_round, out=None):
        seed = (self.seed * 1_000_003 + client * 7919 + nr_round) & 0xFFFFFFFF
        # scale = -1 keeps the honest update (g - (-1)(u - g) = u) and only adds the noise
        return ops.attack_perturb(update, global_flat, -1.0, self.amplitude, seed, out)


def choose_malicious(nr_clients: int, fraction: float, seed: int) -> np.ndarray:
    """Fixed set of ``round(fraction * N)`` malicious client ids (boolean mask)."""
    rng = np.random.default_rng(seed + 7777)
    n_bad = int(round(fraction * nr_clients))
    mask = np.zeros(nr_clients, dtype=bool)
    mask[rng.choice(nr_clients, n_bad, replace=False)] = True
    return mask


def make_attack(name: str, **kw) -> Attack:
    name = name.lower().replace("-", "_")
    if name in ("sign_flip", "signflip"):
        return SignFlip(**kw)
    if name in ("scaling", "scaled"):
        return ScaledUpdate(**kw)
    if name == "noise":
        return NoiseAttack(**kw)
    raise ValueError(f"unknown attack {name!r}")
