"""Multi-process test helper: the package's own launcher (ddl25spring_b200/utils/launch.py) - W spawned ranks with a
process group on 127.0.0.1 (the reference's way of 'testing' distributed code, lab/tutorial_1b/DP/gradient_aggr/run.sh:4-8)
under a timeout, so a deadlock is a test failure instead of a hang."""
from ddl25spring_b200.utils.launch import free_port, spawn  # noqa: F401


def run_distributed(fn, world: int, args=(), timeout: float = 120.0, backend: str = "gloo"):
    """Run ``fn(rank, world, *args)`` on ``world`` ranks; returns the list of per-rank results."""
    return spawn(fn, world, args, timeout=timeout, backend=backend)
