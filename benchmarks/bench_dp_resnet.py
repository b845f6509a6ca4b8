"""Data-parallel ResNet-18 on synthetic 32x32 CIFAR-shaped batches (BASELINE.json config 2).

  torchrun --nproc-per-node N benchmarks/bench_dp_resnet.py [--batch 256 --steps 20]

Arms (same model, same data, same fused optimizer kernel):
  fused : DataParallel + SymmetricGradReducer  - bucketed all-reduce by the peer-memory kernel, overlapped with backward
  nccl  : DataParallel with torch.distributed all_reduce per bucket (async)              - "the baseline, not the product"
  ddp   : torch.nn.parallel.DistributedDataParallel + torch.optim.SGD                    - stock PyTorch
  --model nhwc  : models/resnet_nhwc.py - channels-last tcgen05 tap-GEMM convolutions, fused NHWC BatchNorm (default)
  --model aten  : models/resnet.py      - ATen/cuDNN (TF32) convolutions
  --graph       : forward + backward replayed from a CUDA graph (fused / nccl arms)
Device-timed, max over ranks."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn.functional as F
from ddl25spring_b200.ops.nn import kernel_classification_loss


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--model", default="nhwc", choices=["nhwc", "aten"])
    ap.add_argument("--graph", action="store_true")
    a = ap.parse_args()
    if a.model == "nhwc":
        from ddl25spring_b200.models.resnet_nhwc import resnet18_nhwc as resnet18
    else:
        from ddl25spring_b200.models.resnet import resnet18
    from ddl25spring_b200.parallel.dp import DataParallel
    from ddl25spring_b200.utils.timing import max_over_ranks
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    g = torch.Generator().manual_seed(rank)
    x = torch.randn(a.batch, 3, 32, 32, generator=g).cuda()
    y = torch.randint(0, 10, (a.batch,), generator=g).cuda()
    loss_fn = lambda m, xb, yb: kernel_classification_loss(m(xb), yb)
    out = {"world": world, "per_gpu_batch": a.batch, "model": "ResNet-18 (11.18M params)", "convs": a.model, "graph": a.graph,
           "samples_per_s": {}, "ms_per_step": {}}

    def run(step):
        for _ in range(a.warmup):
            step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            step()
        e1.record(); torch.cuda.synchronize()
        return max_over_ranks(e0.elapsed_time(e1) / a.steps, "cuda")

    arms = ["nccl"] + (["fused", "ddp"] if world > 1 else [])
    for arm in arms:
        torch.manual_seed(0)
        model = resnet18().cuda()
        if arm == "ddp":
            ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[torch.cuda.current_device()])
            opt = torch.optim.SGD(ddp.parameters(), lr=0.05)

            def step():
                opt.zero_grad(set_to_none=True)
                loss_fn(ddp, x, y).backward()
                opt.step()
        else:
            reducer = None
            if arm == "fused":
                from ddl25spring_b200.parallel.symm_impl import SymmetricArena, SymmetricGradReducer
                arena = SymmetricArena()
                reducer = SymmetricGradReducer(arena, sum(p.numel() for p in model.parameters()) + 4096)
            dp = DataParallel(model, bucket_mb=8.0, optimizer="sgd", lr=0.05, reducer=reducer, graph=a.graph)
            step = lambda: dp.train_step(loss_fn, x, y)
        ms = run(step)
        out["ms_per_step"][arm] = round(ms, 3)
        out["samples_per_s"][arm] = round(a.batch * world / ms * 1e3, 1)
        if arm == "fused":
            arena.check()
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
