"""Vertical FL, one party per GPU (BASELINE.json config 3): embedding gather fused into the top model's first GEMM.

  torchrun --nproc-per-node N benchmarks/bench_vfl.py [--batch 4096 --emb 64]

Lab shapes (B = 64, 60 features) are microseconds; the bench states the scaled shapes it uses.
fused = FusedVFLGather (barrier kernel + ONE tcgen05 GEMM pulling k-blocks from the owner GPUs with TMA);
nccl  = all_gather + cat + F.linear."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--emb", type=int, default=64)
    ap.add_argument("--out", type=int, default=128)
    ap.add_argument("--iters", type=int, default=50)
    a = ap.parse_args()
    from ddl25spring_b200.ops.gemm import gather_linear_ptrs
    from ddl25spring_b200.parallel.symm_impl import SymmetricArena
    from ddl25spring_b200.utils.timing import max_over_ranks
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    arena = SymmetricArena()
    dims = [a.emb] * world
    emb = arena.alloc("emb", a.batch * a.emb * 4).view(a.batch, a.emb)
    emb.copy_(torch.randn(a.batch, a.emb, device="cuda"))
    torch.manual_seed(0)
    W = torch.randn(a.out, a.emb * world, device="cuda")
    b = torch.randn(a.out, device="cuda")
    ptrs = arena.peer_ptrs("emb")
    out = torch.empty(a.batch, a.out, device="cuda")
    parts = [torch.empty(a.batch, a.emb, device="cuda") for _ in range(world)]

    def fused():
        arena.barrier()
        gather_linear_ptrs(ptrs, dims, [a.emb] * world, a.batch, W, b, precise=True, out=out)

    def nccl():
        dist.all_gather(parts, emb)
        return torch.nn.functional.linear(torch.cat(parts, 1), W, b)

    def t(fn):
        for _ in range(5):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            fn()
        e1.record(); torch.cuda.synchronize()
        return max_over_ranks(e0.elapsed_time(e1) / a.iters * 1e3, "cuda")

    # ---- whole training step of the split network (bottom MLP per party, gather, top model, CE, backward, AdamW)
    from ddl25spring_b200.ops.nn import kernel_cross_entropy, use_kernels
    from ddl25spring_b200.ops.optim import ArenaOptimizer
    from ddl25spring_b200.parallel import vfl
    from ddl25spring_b200.parallel.symm_impl import FusedVFLGather
    g = torch.Generator().manual_seed(rank)
    x_local = torch.randn(a.batch, a.emb // 2, generator=g).cuda()
    y = torch.randint(0, 2, (a.batch,), generator=torch.Generator().manual_seed(0)).cuda()
    step_ms = {}
    from ddl25spring_b200.utils.graphs import GraphedStep
    hook = None
    for arm in ("fused_kernels", "fused_kernels_graphed", "allgather_aten", "allgather_aten_graphed"):
        try:
            ours, graphed = arm.startswith("fused"), arm.endswith("graphed")
            if ours and hook is None:
                hook = FusedVFLGather(arena, a.batch, dims, a.out)
            net = vfl.DistributedVFL(a.emb // 2, dims, act_on_logits=False, seed=3, gather_linear=hook if ours else None).cuda()
            net.bottom.dropout.p = 0.0
            if ours:
                use_kernels(net)
                opt, lf = ArenaOptimizer(net, "adamw", capturable=graphed), kernel_cross_entropy
            else:
                opt, lf = torch.optim.AdamW(net.parameters(), capturable=graphed), torch.nn.functional.cross_entropy

            def one(xb, yb):
                opt.zero_grad()
                loss = lf(net(xb), yb)
                loss.backward()
                opt.step()
                return loss.detach()
            if graphed:                       # forward, gather, backward, fused reduce and optimizer replayed as ONE graph per rank
                gs = GraphedStep(one, (x_local, y))
                step = lambda: gs(x_local, y)
            else:
                step = lambda: one(x_local, y)
            step_ms[arm] = round(t(step) / 1e3, 4)
        except Exception as e:
            step_ms[arm] = None
            r_err = f"{type(e).__name__}: {e}"[:200]
            print(f"[bench_vfl] arm {arm} failed: {r_err}", file=sys.stderr)
    r = {"world": world, "train_step_ms": step_ms, "train_samples_per_s": {k: (round(a.batch / v * 1e3, 1) if v else None) for k, v in step_ms.items()}, "batch": a.batch, "emb_per_party": a.emb, "top_fc1_out": a.out, "us": {"fused": round(t(fused), 2), "nccl_allgather_cat_linear": round(t(nccl), 2)}}
    ref = nccl()
    fused()
    r["max_abs_err_vs_fp32"] = float((out - ref).abs().max())
    arena.check()
    if rank == 0:
        print(json.dumps(r), flush=True)
    torch.cuda.synchronize()
    dist.barrier()
    os._exit(0)          # graphs that captured collectives keep communicator work alive: leave without the slow teardown


if __name__ == "__main__":
    main()
