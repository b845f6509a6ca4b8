#!/usr/bin/env python
"""Headline benchmark: federated rounds/s and samples/s of FedAvg on MnistCnn (BASELINE.json metric
"FL rounds/sec + samples/sec at 1/2/4/8 B200").

Workload (identical for both arms): FedAvg as in lab/tutorial_1a/hfl_complete.py:336-390 — 100
simulated clients holding an IID split of a 60 000-sample MNIST-shaped synthetic set, 8 sampled
clients per GPU per round (client_fraction = 0.08 * n_gpus: weak scaling), E = 1 local epoch of
minibatch SGD (B = 100, lr = 0.01), sample-count-weighted aggregation, and the top-1 evaluation of
the new global model on the 10 000-sample test set that ``run()`` performs every round.
One "step" = one federated round.

  python bench.py [--gpus N --steps K --warmup W]           product arm (this framework)
  python bench.py --impl reference [...]                    unmodified reference from baseline/_ref

Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

NR_CLIENTS, CLIENTS_PER_GPU, BATCH, EPOCHS, LR, SEED = 100, 8, 100, 1, 0.01, 10
SAMPLES_PER_CLIENT = 600
METRIC = "fl_samples_per_sec"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--aggregator", default="mean")
    return ap.parse_args()


def init_dist(n_gpus: int):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if torch.cuda.is_available():
        torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        dist.init_process_group("nccl" if torch.cuda.is_available() else "gloo", rank=rank, world_size=world,
                                device_id=torch.device("cuda", local) if torch.cuda.is_available() else None)
    return rank, world, local


def config_dict(n_gpus: int, extra: dict | None = None) -> dict:
    cfg = {"model": "MnistCnn (LeNet-style, 1 199 882 params)", "algorithm": "FedAvg",
           "nr_clients": NR_CLIENTS, "clients_per_round": CLIENTS_PER_GPU * n_gpus,
           "client_fraction": round(CLIENTS_PER_GPU * n_gpus / NR_CLIENTS, 4), "local_epochs": EPOCHS,
           "batch_size": BATCH, "lr": LR, "global_batch": CLIENTS_PER_GPU * n_gpus * BATCH, "seq_len": None,
           "samples_per_round": CLIENTS_PER_GPU * n_gpus * SAMPLES_PER_CLIENT * EPOCHS, "test_samples_per_round": 10_000,
           "parallelism": f"clients packed {CLIENTS_PER_GPU}/GPU x {n_gpus} GPU(s), one rank per GPU",
           "step": "one federated round: broadcast, local training of all sampled clients, aggregation, test()"}
    if extra:
        cfg.update(extra)
    return cfg


# ----------------------------------------------------------------------------------------- reference arm
def run_reference_arm(args):
    sys.path.insert(0, os.path.join(REPO, "baseline"))
    import reference_runner as rr

    why = rr.reference_available()
    if why is not None:
        print(json.dumps({"impl": "reference", "unavailable": why}))
        return
    rank, world, local = init_dist(args.gpus)
    import torch
    import torch.distributed as dist

    out = None
    if rank == 0:
        # the reference is single-process / single-device: rank 0 runs the whole job
        r = rr.run_reference(NR_CLIENTS, CLIENTS_PER_GPU * args.gpus / NR_CLIENTS, BATCH, EPOCHS, LR, SEED, args.warmup,
                             args.steps)
        ms = r["seconds"] * 1e3 / args.steps
        out = {"impl": "reference", "metric": METRIC, "value": r["samples"] / r["seconds"], "unit": "samples/s",
               "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
               "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32 (cuDNN TF32 convs, torch default)",
               "data": "synthetic", "rounds_per_sec": 1e3 / ms, "final_test_accuracy": r["acc"],
               "config": config_dict(args.gpus, {"reference_device": r["device"],
                                                 "note": "stock FedAvgServer.run(); single process, uses one GPU whatever N is; "
                                                         "wall-clock timed (the reference has no device timing)"})}
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if out is not None:
        print(json.dumps(out))


# ----------------------------------------------------------------------------------------- product arm
def build_server(n_gpus: int, data_on_device: bool, aggregator: str, dataset, prefetch: bool = True):
    from ddl25spring_b200.core.data import split
    from ddl25spring_b200.fl.aggregators import make_aggregator
    from ddl25spring_b200.fl.packed_server import PackedFedAvgServer
    from ddl25spring_b200.parallel.symm import make_collective

    subsets = split(dataset.train, NR_CLIENTS, True, SEED)
    kw = {}
    if aggregator == "krum":
        kw = dict(n_byzantine=max(1, int(0.2 * CLIENTS_PER_GPU * n_gpus)))
    return PackedFedAvgServer(LR, BATCH, subsets, CLIENTS_PER_GPU * n_gpus / NR_CLIENTS, EPOCHS, SEED,
                              test_set=dataset.test, device="cuda", aggregator=make_aggregator(aggregator, **kw),
                              data_on_device=data_on_device, collective=make_collective(), prefetch_next=prefetch)


def timed_rounds(server, steps: int, warmup: int, device, flusher, sampler=None):
    """W untimed rounds, then K rounds each bracketed by CUDA events (L2 flushed in between, outside the
    events); returns (sum of device ms over the K rounds — max over ranks, accuracies)."""
    import torch
    from ddl25spring_b200.utils.timing import barrier_sync, max_over_ranks

    for _ in range(warmup):
        server.run(1)
    barrier_sync(device)
    total = 0.0
    accs = []
    if sampler is not None:
        sampler.__enter__()
    for _ in range(steps):
        flusher.flush()
        torch.cuda.synchronize()
        res = server.run(1)              # run() brackets the round with CUDA events on the launching stream
        total += res.device_time_ms[0]
        accs.append(res.test_accuracy[0])
    barrier_sync(device)
    if sampler is not None:
        sampler.__exit__(None, None, None)
    return max_over_ranks(total, device), accs


def run_ours(args):
    rank, world, local = init_dist(args.gpus)
    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        print(json.dumps({"metric": METRIC, "error": "no CUDA device: bench.py needs a B200"}))
        return
    from ddl25spring_b200 import ops
    from ddl25spring_b200.core.data import SyntheticMNIST
    from ddl25spring_b200.utils.clocks import ClockSampler
    from ddl25spring_b200.utils.timing import L2Flusher

    device = torch.device("cuda", local)
    n_gpus = world
    dataset = SyntheticMNIST(seed=0)
    flusher = L2Flusher(device)

    # ---- kernel-side number: data resident in HBM
    server = build_server(n_gpus, True, args.aggregator, dataset)
    sampler = ClockSampler(local)
    for _ in range(args.warmup):
        server.run(1)
    ops.reset_launch_count()
    ms_total, accs = timed_rounds(server, args.steps, 0, device, flusher, sampler)
    launches = ops.launch_count()
    del server
    torch.cuda.empty_cache()

    # ---- end-to-end number through the public API: shards in pinned host memory, copied H2D every round,
    #      accuracy read back D2H every round
    # (no cross-round prefetch here: every round's own H2D copy and D2H read are inside that round's timed region)
    server = build_server(n_gpus, False, args.aggregator, dataset, prefetch=False)
    e2e_ms, _ = timed_rounds(server, args.steps, args.warmup, device, flusher)
    h2d = server.h2d_bytes_last_round
    del server

    samples_round = CLIENTS_PER_GPU * n_gpus * SAMPLES_PER_CLIENT * EPOCHS
    ms = ms_total / args.steps
    e2e = e2e_ms / args.steps
    if rank == 0:
        out = {"metric": METRIC, "value": samples_round / (ms / 1e3), "unit": "samples/s", "n_gpus": n_gpus,
               "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
               "scaling": "weak", "vs_baseline": None, "dtype": "fp32 (TF32 tensor-core convs like the reference's cuDNN default; "
                                                               "3xTF32 fp32-grade linears)",
               "data": "synthetic", "rounds_per_sec": 1e3 / ms, "final_test_accuracy": accs[-1],
               "clocks": sampler.summary(),
               "e2e": {"value": samples_round / (e2e / 1e3), "unit": "samples/s", "ms_per_step": e2e,
                       "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4,
                       "api": "PackedFedAvgServer(...).run(1) with data_on_device=False"},
               "gpu_launches": int(launches),
               "config": config_dict(n_gpus, {"l2": "L2 flushed (252 MB write) between timed rounds",
                                              "aggregator": args.aggregator,
                                              "timing": "CUDA events per round on the launching stream, summed, max over ranks"})}
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference_arm(a)
    else:
        run_ours(a)
