"""Byzantine-robust FedAvg (BASELINE.json config 4): 8 ResNet-18 clients per GPU per round (64 at 8 GPUs), 20 % sign-flip
attackers, defended by Krum / coordinate-wise median / trimmed mean.

  [torchrun --nproc-per-node N] python benchmarks/bench_robust_fedavg.py [--rounds 3 --samples 256 --batch 64]

Reports rounds/s, samples/s and the aggregation latency per round (device-timed, max over ranks).  Client-local training
of ResNet-18 runs on the channels-last tcgen05 convolution stack (models/resnet_nhwc.py) with every local SGD step replayed
from a CUDA graph (--model aten --no-graph: the autograd/ATen/cuDNN path); the attack, the robust
aggregation (incl. BN running statistics, `include_buffers=True`), the optimizer step and the evaluation count are this
framework's kernels; across GPUs the aggregation is the fused peer-memory kernel."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn.functional as F
from ddl25spring_b200.ops.nn import kernel_classification_loss


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--samples", type=int, default=256, help="samples per client")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--aggregators", default="mean,krum,median,trimmed_mean")
    ap.add_argument("--model", default="nhwc", choices=["nhwc", "aten"])
    ap.add_argument("--no-graph", action="store_true")
    a = ap.parse_args()
    from ddl25spring_b200.core.data import SyntheticCIFAR, split
    from ddl25spring_b200.fl.aggregators import make_aggregator
    from ddl25spring_b200.fl.attacks import SignFlip, choose_malicious
    from ddl25spring_b200.fl.server import FedAvgServer
    if a.model == "nhwc":
        from ddl25spring_b200.models.resnet_nhwc import resnet18_nhwc as resnet18
    else:
        from ddl25spring_b200.models.resnet import resnet18
    from ddl25spring_b200.utils.timing import max_over_ranks
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    K = 8 * world
    ds = SyntheticCIFAR(n_train=K * a.samples, n_test=2000, seed=0)
    subsets = split(ds.train, K, True, 1)
    mal = choose_malicious(K, 0.2, 0)
    coll = None
    out = {"world": world, "clients_per_round": K, "attackers": int(mal.sum()), "model": "ResNet-18", "convs": a.model, "graph": not a.no_graph, "samples_per_client": a.samples, "results": {}}
    for name in a.aggregators.split(","):
        kw = dict(n_byzantine=int(mal.sum()), multi=K - 2 * int(mal.sum())) if name == "krum" else {}
        srv = FedAvgServer(0.05, a.batch, subsets, 1.0, 1, 10, model_fn=resnet18, test_set=ds.test, device="cuda",
                           loss_fn=kernel_classification_loss, include_buffers=True, graph=not a.no_graph,
                           aggregator=make_aggregator(name, **kw), attack=SignFlip(4.0), malicious=mal, test_batch_size=1000)
        if world > 1:
            if coll is None:
                from ddl25spring_b200.parallel.symm_impl import SymmetricAggregator
                coll = SymmetricAggregator(max_rows_per_rank=8, row_elems=srv.arena.size)
            srv.collective = coll
        srv.run(1)                                           # warm-up round
        # aggregation latency alone, on the last update matrix
        U = srv._updates if srv._updates is not None else coll.row_buffer(8)
        w = torch.full((K,), 1.0 / K)
        g = srv.arena.flat.detach().clone()
        ts = []
        for _ in range(5):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if world > 1:
                coll.aggregate(srv.aggregator, U, [p for p in range(K) if p % world == rank], w, g)
            else:
                srv.aggregator(U, w, global_flat=g)
            e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        agg_ms = max_over_ranks(sorted(ts)[2], "cuda")
        res = srv.run(a.rounds)
        ms = max_over_ranks(sum(res.device_time_ms) / a.rounds, "cuda")
        out["results"][name] = {"ms_per_round": round(ms, 2), "rounds_per_s": round(1e3 / ms, 3),
                                "samples_per_s": round(K * a.samples / ms * 1e3, 1), "aggregation_ms": round(agg_ms, 3),
                                "test_acc": [round(x, 2) for x in res.test_accuracy]}
        del srv
        torch.cuda.empty_cache()
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
