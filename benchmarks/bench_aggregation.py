"""Robust-aggregation latency per round (BASELINE.json metric, config 4): K client updates of a ResNet-18-sized
model (11 181 642 params), single GPU or one rank per GPU with the fused peer-memory kernels.

  python benchmarks/bench_aggregation.py                      # 1 GPU, K = 64 rows local
  torchrun --nproc-per-node N benchmarks/bench_aggregation.py # K = 8 rows per GPU, fused vs NCCL+torch

Every timing: CUDA events, >= 3 warm-ups, L2 flushed between iterations, max over ranks.
Roofline: bytes that must be read once (K * D * 4) / measured HBM copy bandwidth (MEASURED_PEAKS.json)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

D = 11_181_642
DP = (D + 127) // 128 * 128


def timeit(fn, flusher, iters=5, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flusher.flush(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


def main():
    from ddl25spring_b200 import ops
    from ddl25spring_b200.ops.gemm import gram_tcgen05
    from ddl25spring_b200.utils.timing import L2Flusher, max_over_ranks
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = peaks.get("hbm_gbs", 6650.0) * 1e9
    fl = L2Flusher("cuda")
    out = {"D": D, "world": world}
    if world == 1:
        K, f = 64, 12
        U = 0.01 * torch.randn(K, DP, device="cuda")
        w = torch.full((K,), 1.0 / K)
        res = torch.empty(DP, device="cuda")
        bytes_in = K * DP * 4
        rows = {}
        rows["mean_ours"] = timeit(lambda: ops.weighted_sum(U, w, res), fl)
        rows["mean_torch"] = timeit(lambda: torch.mv(U.t(), w.cuda(), out=res), fl)
        rows["median_ours"] = timeit(lambda: ops.coordinate_median(U, out=res), fl)
        rows["median_torch"] = timeit(lambda: torch.median(U, dim=0), fl, iters=3, warm=1)
        rows["trimmed_ours"] = timeit(lambda: ops.trimmed_mean(U, 13, res), fl)
        rows["trimmed_torch"] = timeit(lambda: torch.sort(U, dim=0).values[13:K - 13].mean(0), fl, iters=3, warm=1)
        rows["krum_gram_ours_tf32"] = timeit(lambda: gram_tcgen05(U), fl)
        G = gram_tcgen05(U)
        rows["krum_select_ours"] = timeit(lambda: ops.krum_select(U, f, 1, gram_matrix=G), fl)
        rows["krum_gram_torch_fp32"] = timeit(lambda: U @ U.t(), fl)
        rows["krum_cdist_torch"] = timeit(lambda: torch.cdist(U, U), fl, iters=3, warm=1)
        out["K"] = K
        out["ms"] = {k: round(v, 4) for k, v in rows.items()}
        out["roofline_ms_read_once"] = round(bytes_in / hbm * 1e3, 4)
        out["frac_of_measured_hbm"] = {k: round(bytes_in / hbm * 1e3 / v, 3) for k, v in rows.items() if "ours" in k and "select" not in k}
    else:
        from ddl25spring_b200.parallel.symm_impl import SymmetricArena, shard_bounds
        G_, K = 8, 8 * world
        a = SymmetricArena()
        rows_t = a.alloc("rows", G_ * DP * 4).view(G_, DP)
        res = a.alloc("res", DP * 4)
        rows_t.copy_(0.01 * torch.randn(G_, DP, device="cuda"))
        ptrs = [a.peer_ptrs("rows")[k % world] + (k // world) * DP * 4 for k in range(K)]
        w = [1.0 / K] * K
        d0, d1 = shard_bounds(DP, world, rank)
        t = {}
        t["mean_fused"] = timeit(lambda: a.reduce(0, ptrs, w, a.peer_ptrs("res"), d0, d1), fl)
        part = a.alloc("part", DP * 4)
        lw = [1.0 / K] * G_

        def mean_hier():
            ops.weighted_sum(rows_t, lw, part)
            a.reduce(0, a.peer_ptrs("part"), [1.0] * world, a.peer_ptrs("res"), d0, d1)
        t["mean_fused_hierarchical"] = timeit(mean_hier, fl)
        t["median_fused"] = timeit(lambda: a.reduce(1, ptrs, w, a.peer_ptrs("res"), d0, d1, rank_lo=K // 2 - 1, rank_hi=K // 2 + 1), fl)
        part2 = torch.empty(DP, device="cuda")
        wl = torch.full((G_,), 1.0 / K, device="cuda")

        def nccl_mean():
            torch.mv(rows_t.t(), wl, out=part2)
            dist.all_reduce(part2)
        t["mean_nccl"] = timeit(nccl_mean, fl)
        allrows = torch.empty(K, DP, device="cuda") if K * DP * 4 < 60e9 else None

        def nccl_median():
            dist.all_gather_into_tensor(allrows.view(-1), rows_t.view(-1))
            torch.median(allrows, dim=0)
        if allrows is not None:
            t["median_nccl_allgather_torch"] = timeit(nccl_median, fl, iters=3, warm=1)
        a.check()
        out["K"] = K
        out["ms"] = {k: round(max_over_ranks(v, "cuda"), 4) for k, v in t.items()}
        nv = 770e9
        out["nvlink_roofline_ms"] = round((world - 1) / world * G_ * DP * 4 / nv * 1e3, 4)
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
