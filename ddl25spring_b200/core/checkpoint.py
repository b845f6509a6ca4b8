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
    """Save / resume a federated run every ``every`` rounds.

    ONE file (``round_state.pt``) holds everything a resumed run needs - the model ``state_dict``, the index of the next
    round, the server's numpy sampling RNG state, torch's CPU and CUDA RNG states (dropout, initialisation) and the
    ``RunResult`` so far - and it is replaced atomically, so a crash can never pair a newer model with an older round index.
    ``Server.run(nr_rounds, checkpointer=ck)`` (fl/server.py, fl/packed_server.py) saves after each round and, when the file
    exists, continues from it; the resumed run's ``RunResult`` and final weights equal the uninterrupted run's."""

    def __init__(self, directory: str, every: int = 1):
        self.dir, self.every = directory, every
        os.makedirs(directory, exist_ok=True)

    def path(self) -> str:
        return os.path.join(self.dir, "round_state.pt")

    def exists(self) -> bool:
        return os.path.exists(self.path())

    def save(self, next_round: int, model: nn.Module | FlatArena, rng_state: dict, result: RunResult, extra: dict | None = None) -> None:
        sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
        state = {"round": int(next_round), "model": sd, "rng": rng_state, "torch_rng": torch.get_rng_state(),
                 "cuda_rng": torch.cuda.get_rng_state_all() if torch.cuda.is_available() else None,
                 "result": asdict(result), "extra": extra or {}}
        tmp = self.path() + ".tmp"
        torch.save(state, tmp)
        os.replace(tmp, self.path())                # atomic

    def maybe_save(self, nr_round: int, model: nn.Module | FlatArena, rng, result: RunResult, extra: dict | None = None) -> None:
        """``rng``: a numpy Generator (its current state is saved) or an already captured ``bit_generator.state`` dict."""
        if (nr_round + 1) % self.every:
            return
        self.save(nr_round + 1, model, rng if isinstance(rng, dict) else rng.bit_generator.state, result, extra)

    def resume(self, model, rng: np.random.Generator) -> tuple[int, RunResult | None]:
        """Load the latest state into ``model`` / ``rng`` / torch's RNGs; returns (next round, RunResult so far)."""
        if not self.exists():
            return 0, None
        st = torch.load(self.path(), weights_only=False)
        if isinstance(model, FlatArena):
            model.load_state_dict(st["model"], strict=True)
        else:
            model.load_state_dict(st["model"], strict=True)
        rng.bit_generator.state = st["rng"]
        torch.set_rng_state(st["torch_rng"])
        if st.get("cuda_rng") is not None and torch.cuda.is_available() and len(st["cuda_rng"]) == torch.cuda.device_count():
            torch.cuda.set_rng_state_all(st["cuda_rng"])
        self.extra = st.get("extra", {})
        return st["round"], RunResult(**st["result"])
