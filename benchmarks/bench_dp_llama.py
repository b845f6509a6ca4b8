#!/usr/bin/env python
"""Data-parallel LLaMA (the DP lab: 6 x 288, ctx 256, vocab 32 000, batch 3 per rank, Adam; lab/tutorial_1b/DP/gradient_aggr/
intro_DP_GA.py:41-68) on N GPUs with this framework: kernels for every op, the whole step - forward, backward, bucket reduce +
Adam on the owned shard + parameter broadcast in one peer-memory kernel per bucket - replayed as ONE CUDA graph per rank.

  torchrun --nproc-per-node N --master-addr 127.0.0.1 benchmarks/bench_dp_llama.py

The matching baseline row ("reference loop on NCCL": flat cat -> all_reduce -> split -> Adam) is `baseline/nccl_baselines.py --only dp_ga`.
Device-timed, max over ranks."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main():
    from ddl25spring_b200.core.data import SyntheticTokens
    from ddl25spring_b200.models.llama import LLama
    from ddl25spring_b200.ops.nn import kernel_causal_lm_loss, use_kernels
    from ddl25spring_b200.parallel.dp import DataParallel
    from ddl25spring_b200.utils.timing import max_over_ranks
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    V, B, T = 32_000, 3, 256
    out = {"world": world, "model": "LLaMA 6x288, vocab 32000, batch 3 x 256 per rank, Adam", "ms_per_step": {}, "tokens_per_s": {}}
    tok = next(iter(SyntheticTokens(V, B, T, skip=rank * 5000))).cuda()
    ap = argparse.ArgumentParser()
    ap.add_argument("--arms", default="")       # e.g. fused_in_graph:64:16:nofix = max_ctas 64, 16 MB buckets, no in-GEMM fix-up
    a = ap.parse_args()
    arms = a.arms.split(",") if a.arms else (["kernels_graph"] if world == 1 else ["fused_in_graph", "nccl_graph"])
    from ddl25spring_b200.ops import gemm as _gemm
    fix0 = _gemm.MAX_FIXUP_TILES
    for spec in arms:
        arm, *opts = spec.split(":")
        max_ctas = int(opts[0]) if len(opts) > 0 else 32
        bucket_mb = float(opts[1]) if len(opts) > 1 else 16.0
        _gemm.MAX_FIXUP_TILES = 0 if "nofix" in opts else fix0
        torch.manual_seed(0)
        net = use_kernels(LLama(None, V, 288, 6, "cuda", 6, T))
        reducer = None
        if arm.startswith("fused_in_graph"):
            from ddl25spring_b200.parallel.symm_impl import SymmetricArena, SymmetricGradReducer
            reducer = SymmetricGradReducer(SymmetricArena(), sum(p.numel() for p in net.parameters()) + 8192,
                                           max_ctas=max_ctas, use_nvls="nvls" in arm, tag=spec.replace(":", "_"))
        dp = DataParallel(net, optimizer="adam", lr=8e-4, reducer=reducer, graph=True, bucket_mb=bucket_mb)
        loss_fn = lambda m, t: kernel_causal_lm_loss(m(t), t)
        for _ in range(5):
            dp.train_step(loss_fn, tok)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            loss = dp.train_step(loss_fn, tok)
        e1.record(); torch.cuda.synchronize()
        ms = max_over_ranks(e0.elapsed_time(e1) / 20, "cuda")
        out["ms_per_step"][spec] = round(ms, 3)
        out["tokens_per_s"][spec] = round(world * B * T / ms * 1e3, 1)
        out.setdefault("loss", {})[spec] = round(float(loss), 4)
        if reducer is not None:
            reducer.arena.check()
        del dp, net
        torch.cuda.empty_cache()
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        torch.cuda.synchronize(); dist.barrier()
        os._exit(0)


if __name__ == "__main__":
    main()
