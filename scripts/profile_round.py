"""A short fixed workload for ncu: 2 warm rounds + 1 profiled-size round of the flagship packed FedAvg path."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl25spring_b200.core.data import SyntheticMNIST, split
from ddl25spring_b200.fl.packed_server import PackedFedAvgServer

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 3
ds = SyntheticMNIST(seed=0)
srv = PackedFedAvgServer(0.01, 100, split(ds.train, 100, True, 10), 0.08, 1, 10, test_set=ds.test, device="cuda")
res = srv.run(rounds)
torch.cuda.synchronize()
print("ms/round", [round(x, 3) for x in res.device_time_ms], "acc", res.test_accuracy[-1])
import time
t0 = time.perf_counter(); srv.run(1); t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print("host issue+test sync ms %.3f, total ms %.3f" % ((t1 - t0) * 1e3, (t2 - t0) * 1e3))
