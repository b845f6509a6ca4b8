#!/usr/bin/env python
"""Federated DCGAN on synthetic 64x64 images (BASELINE.json config 5): rounds/s at N GPUs.

  [torchrun --nproc-per-node N] python benchmarks/bench_fed_dcgan.py [--clients-per-gpu 4 --local-steps 4 --batch 64 --rounds 5]

Every round each rank trains its clients' G and D (channels-last tcgen05 convolution stack, the whole D+G iteration incl. fused
Adam replayed from one CUDA graph), then G and D are averaged over all clients: `fused` = local pre-reduce + ONE peer-memory
kernel per rank that writes the new global [G | D] into every replica; `nccl` = the same with torch.distributed all_reduce.
Device-timed (CUDA events around the rounds), max over ranks."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--clients-per-gpu", type=int, default=4)
    ap.add_argument("--local-steps", type=int, default=4)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--rounds", type=int, default=5)
    a = ap.parse_args()
    from ddl25spring_b200.gen.fed_dcgan import FederatedDCGAN
    from ddl25spring_b200.utils.timing import max_over_ranks
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    K = a.clients_per_gpu * world
    g = torch.Generator().manual_seed(0)
    clients = [torch.rand(128, 3, 64, 64, generator=g) * 2 - 1 for _ in range(K)]
    out = {"world": world, "clients_per_round": K, "local_steps": a.local_steps, "batch": a.batch, "model": "DCGAN 64x64 (ngf = ndf = 64)",
           "images_per_round": K * a.local_steps * a.batch, "ms_per_round": {}, "rounds_per_s": {}, "images_per_s": {}}
    arms = ["fused", "nccl"] if world > 1 else ["single_gpu"]
    for arm in arms:
        coll = None
        if arm == "fused":
            from ddl25spring_b200.parallel.symm_impl import SymmetricArena
            coll = SymmetricArena()
        fed = FederatedDCGAN(clients, 1.0, local_steps=a.local_steps, batch_size=a.batch, seed=1, device="cuda", collective=coll)
        fed.run(2)                                             # warm-up (graph capture)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fed.run(a.rounds)
        e1.record(); torch.cuda.synchronize()
        ms = max_over_ranks(e0.elapsed_time(e1) / a.rounds, "cuda")
        out["ms_per_round"][arm] = round(ms, 3)
        out["rounds_per_s"][arm] = round(1e3 / ms, 3)
        out["images_per_s"][arm] = round(out["images_per_round"] / ms * 1e3, 1)
        del fed
        torch.cuda.empty_cache()
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
