"""Check: concurrent client slots produce the same update rows as sequential training (debug helper)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.nn.functional as F
from ddl25spring_b200.core.data import SyntheticCIFAR, split
from ddl25spring_b200.fl.server import FedAvgServer
from ddl25spring_b200.models.resnet_nhwc import resnet18_nhwc

K = 8
ds = SyntheticCIFAR(n_train=K * 128, n_test=500, seed=0)
subsets = split(ds.train, K, True, 1)
rows = {}
for cc in (1, 4, 8):
    srv = FedAvgServer(0.05, 64, subsets, 1.0, 1, 10, model_fn=resnet18_nhwc, test_set=ds.test, device="cuda", loss_fn=F.cross_entropy,
                       include_buffers=True, graph=True, concurrent_clients=cc, test_batch_size=500)
    g = srv.arena.flat.detach().clone()
    for rnd in range(2):
        U, _ = srv._collect_updates(rnd, np.arange(K), g)
        torch.cuda.synchronize()
        rows[(cc, rnd)] = U.clone()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for rnd in range(3):
        srv._collect_updates(2 + rnd, np.arange(K), g)
    t1.record(); torch.cuda.synchronize()
    print("concurrent", cc, "ms per round of local training", t0.elapsed_time(t1) / 3)
for rnd in range(2):
    a, b = rows[(1, rnd)], rows[(8, rnd)]
    print("round", rnd, "max abs diff per row", [(float((a[i] - b[i]).abs().max())) for i in range(K)], "finite", bool(torch.isfinite(b).all()))
    print("  delta norms seq", [round(float((a[i] - g).norm()), 4) for i in range(K)])
    print("  delta norms con", [round(float((b[i] - g).norm()), 4) for i in range(K)])
