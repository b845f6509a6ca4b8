"""Run records (parity with ``RunResult`` in lab/tutorial_1a/hfl_complete.py:113-138)."""
from __future__ import annotations

from dataclasses import asdict, dataclass, field

ETA = "\N{GREEK SMALL LETTER ETA}"


@dataclass
class RunResult:
    """Per-round record of a federated run.

    Same fields and ``as_df`` contract as the reference.  Two additions that the reference
    collects but never shows: ``device_time_ms`` (CUDA-event time of each round, max over
    ranks) and ``samples`` (training samples consumed per round) so rounds/s and samples/s can
    be reported; both are dropped from ``as_df`` unless asked for.
    """
    algorithm: str
    n: int      # number of clients
    c: float    # client_fraction
    b: int      # batch size, -1 == full batch (printed as infinity)
    e: int      # nr_local_epochs
    lr: float   # printed as lowercase eta
    seed: int
    wall_time: list[float] = field(default_factory=list)
    message_count: list[int] = field(default_factory=list)
    test_accuracy: list[float] = field(default_factory=list)
    device_time_ms: list[float] = field(default_factory=list)
    samples: list[int] = field(default_factory=list)

    def as_df(self, skip_wtime: bool = True, perf: bool = False):
        from pandas import DataFrame

        d = {k.capitalize().replace("_", " "): v for k, v in asdict(self).items()}
        if d["B"] == -1:
            d["B"] = "\N{INFINITY}"
        for k in ("Device time ms", "Samples"):
            if not perf or len(d[k]) != len(self.wall_time):
                d.pop(k)
        df = DataFrame({"Round": range(1, len(self.wall_time) + 1), **d})
        df = df.rename(columns={"Lr": ETA})
        if skip_wtime:
            df = df.drop(columns=["Wall time"])
        return df

    # throughput helpers -------------------------------------------------------------
    def rounds_per_sec(self) -> float:
        t = sum(self.device_time_ms) / 1e3
        return len(self.device_time_ms) / t if t > 0 else float("nan")

    def samples_per_sec(self) -> float:
        t = sum(self.device_time_ms) / 1e3
        return sum(self.samples) / t if t > 0 else float("nan")


def message_count(nr_round: int, clients_per_round: int) -> int:
    """Cumulative messages after round ``nr_round`` (0-based): hfl_complete.py:309,387."""
    return 2 * (nr_round + 1) * clients_per_round
