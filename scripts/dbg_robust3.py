import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist, torch.nn.functional as F
from ddl25spring_b200.core.data import SyntheticCIFAR, split
from ddl25spring_b200.fl.aggregators import make_aggregator
from ddl25spring_b200.fl.attacks import SignFlip, choose_malicious
from ddl25spring_b200.fl.server import FedAvgServer
from ddl25spring_b200.models.resnet import resnet18
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
torch.cuda.set_device(rank)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
K = int(sys.argv[1])
ds = SyntheticCIFAR(n_train=K * 256, n_test=500, seed=0)
subsets = split(ds.train, K, True, 1)
mal = choose_malicious(K, 0.2, 0)
for agg, attack in (("mean", None), ("median", SignFlip(4.0))):
    srv = FedAvgServer(0.05, 64, subsets, 1.0, 1, 10, model_fn=resnet18, test_set=ds.test, device="cuda", loss_fn=lambda o, t: F.cross_entropy(o, t),
                       include_buffers=True, aggregator=make_aggregator(agg), attack=attack, malicious=mal, test_batch_size=500)
    r = srv.run(2)
    if rank == 0:
        print("RES world", world, "K", K, agg, "attack" if attack else "clean", r.test_accuracy, flush=True)
if world > 1:
    dist.destroy_process_group()
