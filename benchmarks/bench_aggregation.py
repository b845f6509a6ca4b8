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
        # one rank per GPU, 8 client rows per GPU: the product path (SymmetricAggregator.aggregate) for every rule, next to
        # NCCL all_gather + torch for the same rule
        from ddl25spring_b200.fl.aggregators import CoordinateMedian, Krum, TrimmedMean, WeightedMean
        from ddl25spring_b200.parallel.symm_impl import SymmetricAggregator
        G_, K = 8, 8 * world
        f = max(1, int(0.2 * K))
        coll = SymmetricAggregator(max_rows_per_rank=G_, row_elems=DP)
        rows_t = coll.row_buffer(G_)
        gen = torch.Generator(device="cuda").manual_seed(rank)
        gflat = 0.05 * torch.randn(DP, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1234))
        rows_t.copy_(gflat[None, :] + 0.01 * torch.randn(G_, DP, device="cuda", generator=gen))
        mine_pos = [p for p in range(K) if p % world == rank]
        w = torch.full((K,), 1.0 / K)
        g = gflat.clone()
        t = {}
        rules = {"mean": WeightedMean(), "median": CoordinateMedian(), "trimmed_mean": TrimmedMean(0.2),
                 "krum": Krum(f, multi=K - 2 * f), "krum_fused_gather": Krum(f, multi=K - 2 * f, fused_gather=True)}
        sel = {}
        for name, agg in rules.items():
            t[name + "_fused"] = timeit(lambda: coll.aggregate(agg, rows_t, mine_pos, w, g), fl)
            if isinstance(agg, Krum):
                sel[name] = sorted(agg.last_selected.cpu().tolist())
        coll.two_phase_select = True                  # gather (float4 peer loads) + local selection + peer stores
        for name in ("median", "trimmed_mean"):
            t[name + "_two_phase_fused"] = timeit(lambda: coll.aggregate(rules[name], rows_t, mine_pos, w, g), fl)
        coll.two_phase_select = False
        # stages of the Krum path, timed separately (each is a collective: every rank runs the same sequence)
        from ddl25spring_b200.ops.gemm import gram_tcgen05
        from ddl25spring_b200.parallel.symm_impl import shard_bounds
        d0, d1 = shard_bounds(DP, world, rank)
        shard = d1 - d0
        pad = (-shard) % 32
        Ub = torch.zeros(K, shard + pad, device="cuda")
        ptrs_k = [coll.row_ptrs[k % world] + (k // world) * DP * 4 for k in range(K)]
        t["stage_gather_shard"] = timeit(lambda: coll.arena.reduce(2, ptrs_k, [1.0] * K, [Ub.data_ptr()], d0, d1, base=g, rank_lo=pad // 4), fl)
        t["stage_gram_local_tf32"] = timeit(lambda: gram_tcgen05(Ub), fl)
        coll.arena.check()
        allrows = torch.empty(world, G_, DP, device="cuda")

        def gather():
            dist.all_gather_into_tensor(allrows.view(-1), rows_t.reshape(-1))
            return allrows.view(K, DP)
        tt = int(-(-0.2 * K // 1))

        def base(name):
            A = gather()
            if name == "mean":
                return A.mean(0)
            if name == "median":
                return torch.median(A, dim=0).values
            if name == "trimmed_mean":
                return torch.sort(A, dim=0).values[tt:K - tt].mean(0)
            Ad = A - gflat[None, :]
            dd = torch.cdist(Ad, Ad).pow(2)
            dd.fill_diagonal_(float("inf"))
            sc = torch.sort(dd, dim=1).values[:, :K - f - 2].sum(1)
            idx = torch.argsort(sc)[:K - 2 * f]
            # the gathered matrix is rank-major (row = rank * 8 + local); report client indices k (rank k % W, local k // W)
            sel["krum_torch_fp32"] = sorted(int(i % G_) * world + int(i // G_) for i in idx.cpu().tolist())
            return A[idx].mean(0)
        for name in ("mean", "median", "trimmed_mean", "krum"):
            t[name + "_nccl_allgather_torch"] = timeit(lambda: base(name), fl, iters=2, warm=1)
        out["K"] = K
        out["ms"] = {k: round(max_over_ranks(v, "cuda"), 4) for k, v in t.items()}
        local_bytes = G_ * DP * 4
        roof = max(local_bytes / hbm, (world - 1) / world * local_bytes / 900e9) * 1e3
        out["roofline_ms"] = round(roof, 4)
        out["roofline"] = "max(8 local rows read once from HBM at the measured copy bandwidth, (W-1)/W of them over NVLink at 900 GB/s)"
        out["frac_of_roofline"] = {k: round(roof / v, 3) for k, v in out["ms"].items() if k.endswith("_fused")}
        out["frac_of_measured_nvlink_770"] = {k: round((world - 1) / world * local_bytes / 770e9 * 1e3 / v, 3)
                                              for k, v in out["ms"].items() if k.endswith("_fused") or k.startswith("stage_gather")}
        out["krum_selected"] = sel
        out["krum_selection_agrees"] = sel.get("krum") == sel.get("krum_torch_fp32") == sel.get("krum_fused_gather")
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
