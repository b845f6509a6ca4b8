import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist, torch.nn.functional as F
from ddl25spring_b200.core.data import SyntheticCIFAR, split
from ddl25spring_b200.fl.aggregators import make_aggregator
from ddl25spring_b200.fl.attacks import SignFlip, choose_malicious
from ddl25spring_b200.fl.server import FedAvgServer
from ddl25spring_b200.models.resnet import resnet18
world = int(os.environ["WORLD_SIZE"]); rank = int(os.environ["RANK"])
torch.cuda.set_device(rank); dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
K = 8 * world
ds = SyntheticCIFAR(n_train=K * 128, n_test=500, seed=0)
subsets = split(ds.train, K, True, 1)
mal = choose_malicious(K, 0.2, 0)
from ddl25spring_b200.parallel.symm_impl import SymmetricAggregator
coll = None
for agg in ("median",):
    res = {}
    for path in ("nccl", "fused"):
        srv = FedAvgServer(0.05, 64, subsets, 1.0, 1, 10, model_fn=resnet18, test_set=ds.test, device="cuda", loss_fn=lambda o, t: F.cross_entropy(o, t),
                           include_buffers=True, aggregator=make_aggregator(agg), attack=SignFlip(4.0), malicious=mal, test_batch_size=500)
        if path == "fused":
            coll = coll or SymmetricAggregator(max_rows_per_rank=8, row_elems=srv.arena.size)
            srv.collective = coll
        r = srv.run(1)
        torch.cuda.synchronize()
        f = srv.arena.flat
        res[path] = f.clone()
        print(rank, agg, path, "finite", bool(torch.isfinite(f).all()), "acc", r.test_accuracy, "absmax", float(f.abs().max()), flush=True)
    d = (res["nccl"] - res["fused"]).abs()
    print(rank, "max diff", float(d.max()), "at", int(d.argmax()), "size", f.numel(), flush=True)
dist.destroy_process_group()
