"""Spawn W processes on this machine with a gloo group on 127.0.0.1 (the reference's own way of
'testing' distributed code, lab/tutorial_1b/DP/gradient_aggr/run.sh:4-8) - but under a timeout, so
a deadlock is a test failure instead of a hang."""
import os
import socket
import traceback

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(rank, world, port, fn, args, q, release, backend="gloo"):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
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


def run_distributed(fn, world: int, args=(), timeout: float = 120.0, backend: str = "gloo"):
    """Run ``fn(rank, world, *args)`` on ``world`` gloo ranks; returns the list of per-rank results."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    release = ctx.Event()
    procs = [ctx.Process(target=_entry, args=(r, world, port, fn, args, q, release, backend)) for r in range(world)]
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
