import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl25spring_b200.core.data import SyntheticMNIST, split
from ddl25spring_b200.fl.packed_server import PackedFedAvgServer
ds = SyntheticMNIST(seed=0)
srv = PackedFedAvgServer(0.01, 100, split(ds.train, 100, True, 10), 0.08, 1, 10, test_set=ds.test, device="cuda")
srv.run(3)
torch.cuda.synchronize()
G = 8
eng = srv._engine(G); W, stag = srv._bufs[G]; Xg, Yg, ph, pd, _ = stag[0]
for name, fn in [("local_train", lambda: eng.local_train(W, Xg, Yg, pd, [600] * G, 100, 1, 0.01, seed=1)),
                 ("test", lambda: srv._eval.eval_logits(srv.arena.flat, srv._test_xy[0], srv._test_xy[1], srv._correct, want_logits=False)),
                 ("train_local_clients", lambda: srv._train_local_clients(5, list(range(8))))]:
    for _ in range(2):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); e0.record(); fn(); e1.record(); t1 = time.perf_counter()
        torch.cuda.synchronize()
    print(f"{name}: host enqueue {1e3*(t1-t0):.3f} ms, device {e0.elapsed_time(e1):.3f} ms")
