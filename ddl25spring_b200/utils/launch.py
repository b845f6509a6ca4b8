"""Process launchers: N ranks on this machine, one per GPU (or CPU ranks over gloo).

The labs start their ranks with bash loops of background ``python script.py $rank`` processes
(lab/tutorial_1b/DP/gradient_aggr/run.sh:4-8) or, inside a notebook, with ``multiprocessing`` "fork" and stdout redirected to
``out{rank}.txt`` (lab/tutorial_1b/DP/gradient_aggr/intro_DP_GA_notebook.ipynb cells 0-1); a rank that dies or hangs leaves
the others waiting forever.  :func:`spawn` is the in-process equivalent used by the notebooks, examples and tests here:
``world`` spawned processes, rendezvous on 127.0.0.1, NCCL with one GPU per rank or gloo on CPU, per-rank results returned
to the caller, a failure in any rank re-raised with its traceback, and a timeout so a deadlock is an error, not a hang.
:func:`launch_script` runs a script under ``torch.distributed.run`` (what scripts/run_dp.sh / run_pp.sh do).
"""
from __future__ import annotations

import os
import socket
import subprocess
import sys
import traceback

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(rank, world, port, fn, args, q, release, backend, log_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    if log_dir:                                   # the notebook launcher's `out{rank}.txt`
        f = open(os.path.join(log_dir, f"out{rank}.txt"), "w", buffering=1)
        os.dup2(f.fileno(), 1)
        sys.stdout = f
    try:
        torch.set_num_threads(1)
        if backend == "nccl":
            torch.cuda.set_device(rank)
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        else:
            dist.init_process_group("gloo", rank=rank, world_size=world)
        out = fn(rank, world, *args)
        q.put((rank, "ok", out))
    except Exception:
        q.put((rank, "err", traceback.format_exc()))
    finally:
        try:
            dist.destroy_process_group()
        except Exception:
            pass
        release.wait(60)      # tensors in the queue are shared through this process: stay alive until they are read


def spawn(fn, world: int, args=(), timeout: float = 120.0, backend: str | None = None, log_dir: str | None = None):
    """Run ``fn(rank, world, *args)`` on ``world`` ranks and return the list of per-rank results.

    ``backend``: "nccl" (one GPU per rank), "gloo", or None = NCCL when this machine has at least ``world`` GPUs."""
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() and torch.cuda.device_count() >= world else "gloo"
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    release = ctx.Event()
    procs = [ctx.Process(target=_entry, args=(r, world, port, fn, args, q, release, backend, log_dir)) for r in range(world)]
    for p in procs:
        p.start()
    results = {}
    try:
        for _ in range(world):
            rank, status, payload = q.get(timeout=timeout)
            if status != "ok":
                raise RuntimeError(f"rank {rank} failed:\n{payload}")
            results[rank] = payload
    finally:
        release.set()
        for p in procs:
            p.join(timeout=10)
            if p.is_alive():
                p.terminate()
    return [results[r] for r in range(world)]


def launch_script(script: str, nproc: int, *script_args: str, timeout: float | None = None) -> subprocess.CompletedProcess:
    """``python -m torch.distributed.run --nproc-per-node nproc script ...`` on 127.0.0.1 (a free port is chosen)."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", str(free_port()), script, *script_args]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
