"""Plain ``state_dict`` checkpoints (BASELINE.json: "plain torch state_dict checkpoints").

The reference only ever holds a ``state_dict`` in memory, and by reference rather than by copy, so its
"best" parameters are silently the last ones (lab/tutorial_2a/centralized.py:51,67-70).
Here: :func:`snapshot` deep-copies, :func:`save` / :func:`load` write ordinary ``torch.save`` files keyed
exactly like ``nn.Module.state_dict()`` (a :class:`FlatArena` is serialised by name, never as a blob), and
:class:`RoundCheckpointer` adds resume for federated runs (round index, RNG state, RunResult so far).
"""
from __future__ import annotations

import os
from dataclasses import asdict

import numpy as np
import torch
from torch import nn

from .arena import FlatArena
from .results import RunResult


def snapshot(module_or_arena) -> dict[str, torch.Tensor]:
    """Deep copy of a state dict (fixes the aliasing bug of centralized.py:51)."""
    sd = module_or_arena.state_dict()
    return {k: v.detach().clone() for k, v in sd.items()}


def save(obj, path: str) -> None:
    sd = obj if isinstance(obj, dict) else obj.state_dict()
    tmp = f"{path}.tmp"
    torch.save({k: v.detach().cpu() for k, v in sd.items()}, tmp)
    os.replace(tmp, path)                       # atomic: a crash never leaves a half-written checkpoint


def load(target, path: str, strict: bool = True, map_location="cpu") -> None:
    sd = torch.load(path, map_location=map_location, weights_only=True)
    if isinstance(target, FlatArena):
        target.load_state_dict(sd, strict=strict)
    else:
        target.load_state_dict(sd, strict=strict)


class RoundCheckpointer:
    """Save / resume a federated run every ``every`` rounds."""

    def __init__(self, directory: str, every: int = 1):
        self.dir, self.every = directory, every
        os.makedirs(directory, exist_ok=True)

    def path(self) -> str:
        return os.path.join(self.dir, "round_state.pt")

    def maybe_save(self, nr_round: int, model: nn.Module | FlatArena, rng: np.random.Generator, result: RunResult) -> None:
        if (nr_round + 1) % self.every:
            return
        save(model, os.path.join(self.dir, "model.pt"))
        tmp = self.path() + ".tmp"
        torch.save({"round": nr_round + 1, "rng": rng.bit_generator.state, "result": asdict(result)}, tmp)
        os.replace(tmp, self.path())

    def resume(self, model, rng: np.random.Generator) -> tuple[int, RunResult | None]:
        if not os.path.exists(self.path()):
            return 0, None
        st = torch.load(self.path(), weights_only=False)
        load(model, os.path.join(self.dir, "model.pt"))
        rng.bit_generator.state = st["rng"]
        return st["round"], RunResult(**st["result"])
