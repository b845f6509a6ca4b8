"""Run configuration (the reference has module-level literals and a notebook cell of
defaults — lab/homework-1.ipynb cell 6: N=100, lr=0.01, C=0.1, E=1, B=100, rounds=10, iid, seed=10)."""
from __future__ import annotations

import argparse
import json
from dataclasses import asdict, dataclass, fields


@dataclass
class FLConfig:
    nr_clients: int = 100
    lr: float = 0.01
    client_fraction: float = 0.1
    nr_local_epochs: int = 1
    batch_size: int = 100
    nr_rounds: int = 10
    iid: bool = True
    seed: int = 10
    algorithm: str = "fedavg"          # fedavg | fedsgd | fedsgd_weight | centralized
    aggregator: str = "mean"           # mean | krum | multi_krum | median | trimmed_mean
    attack: str = "none"               # none | sign_flip | scaling | noise
    malicious_fraction: float = 0.0
    model: str = "mnist_cnn"
    engine: str = "auto"               # auto | packed | generic

    def to_json(self) -> str:
        return json.dumps(asdict(self))

    @classmethod
    def from_args(cls, argv=None) -> "FLConfig":
        ap = argparse.ArgumentParser()
        for f in fields(cls):
            if f.type in ("bool", bool):
                ap.add_argument(f"--{f.name}", type=lambda s: s.lower() in ("1", "true", "yes"), default=f.default)
            else:
                ap.add_argument(f"--{f.name}", type=type(f.default), default=f.default)
        return cls(**vars(ap.parse_args(argv)))
