"""Run the UNMODIFIED reference (baseline/_ref, installed by baseline/install_reference.sh) on the
benchmark workload through its own public API: ``FedAvgServer(...).run(nr_rounds)``
(lab/tutorial_1a/hfl_complete.py:336-390).

Nothing of this framework runs on that path.  The only thing supplied from outside is DATA: the
reference downloads MNIST at import time (hfl_complete.py:26), there is no network here, so the same
synthetic MNIST-shaped set the product arm trains on is written to disk in MNIST's IDX format where
torchvision looks for it; ``datasets.MNIST(download=True)`` then finds the files and skips the
download.  The reference is single-process and single-device: with N > 1 ranks, rank 0 runs the whole
job (all clients of the round) on its GPU and the other ranks wait — that *is* how the reference
scales.
"""
from __future__ import annotations

import os
import sys
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
REF_ROOT = REPO / "baseline" / "_ref"


def reference_available() -> str | None:
    """None if usable, else a one-line reason."""
    if not (REF_ROOT / "lab" / "tutorial_1a" / "hfl_complete.py").exists():
        return "baseline/_ref is missing: run baseline/install_reference.sh (needs /root/reference)"
    try:
        import torchvision  # noqa: F401
    except Exception as e:  # pragma: no cover
        return f"torchvision unavailable ({type(e).__name__})"
    return None


def prepare_data(seed: int = 0) -> None:
    """Write the synthetic MNIST IDX files into the reference's data directory (idempotent)."""
    sys.path.insert(0, str(REPO))
    from ddl25spring_b200.core.data import SyntheticMNIST

    raw = REF_ROOT / "lab" / "tutorial_1a" / "data" / "MNIST" / "raw"
    marker = raw / f".synthetic_seed_{seed}"
    if marker.exists():
        return
    SyntheticMNIST(seed=seed).write_idx(raw)
    marker.touch()


def run_reference(nr_clients: int, client_fraction: float, batch_size: int, epochs: int, lr: float, seed: int,
                  warmup: int, steps: int, data_seed: int = 0) -> dict:
    """Returns {"seconds": timed wall seconds for `steps` rounds, "samples": samples trained, "acc": last accuracy}."""
    import torch

    prepare_data(data_seed)
    sys.path.insert(0, str(REF_ROOT))
    import importlib

    hfl = importlib.import_module("lab.tutorial_1a.hfl_complete")     # stock module, stock code path
    subsets = hfl.split(nr_clients, True, seed)
    server = hfl.FedAvgServer(lr, batch_size, subsets, client_fraction, epochs, seed)
    k = server.nr_clients_per_round
    if warmup > 0:
        server.run(warmup)
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = server.run(steps)
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    per_client = len(subsets[0])
    return {"seconds": dt, "samples": steps * k * per_client * epochs, "acc": res.test_accuracy[-1],
            "clients_per_round": k, "device": str(hfl.device)}


if __name__ == "__main__":
    why = reference_available()
    print(why or run_reference(100, 0.08, 100, 1, 0.01, 10, 1, 2))
