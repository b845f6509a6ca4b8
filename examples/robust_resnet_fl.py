"""Byzantine-robust FedAvg of ResNet-18 (BASELINE.json config 4): sign-flip attackers vs Krum / median / trimmed mean.

  python examples/robust_resnet_fl.py --aggregator krum --rounds 5 [--clients 8 --samples 256 --batch 64 --no-graph]
  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/robust_resnet_fl.py --clients 64

On a B200 the clients train on the channels-last tcgen05 convolution stack (models/resnet_nhwc.py), each local SGD step
replayed from a CUDA graph, and the aggregation (BatchNorm statistics included) runs on the robust-aggregation kernels —
across ranks through the fused peer-memory kernel.  On CPU the same code runs on plain PyTorch ops."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn.functional as F
from ddl25spring_b200.ops.nn import kernel_classification_loss

from ddl25spring_b200.core.data import SyntheticCIFAR, split
from ddl25spring_b200.fl.aggregators import make_aggregator
from ddl25spring_b200.fl.attacks import SignFlip, choose_malicious
from ddl25spring_b200.fl.server import FedAvgServer
from ddl25spring_b200.models.resnet_nhwc import resnet18_nhwc


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--aggregator", default="krum", choices=["mean", "krum", "multi_krum", "median", "trimmed_mean"])
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--clients", type=int, default=8)
    ap.add_argument("--samples", type=int, default=256, help="samples per client")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--malicious_fraction", type=float, default=0.2)
    ap.add_argument("--scale", type=float, default=4.0, help="sign-flip scale")
    ap.add_argument("--no-graph", action="store_true")
    a = ap.parse_args(argv)
    cuda = torch.cuda.is_available()
    if "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        if cuda:
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group("nccl" if cuda else "gloo")
    ds = SyntheticCIFAR(n_train=a.clients * a.samples, n_test=max(64, a.clients * 16), seed=0)
    subsets = split(ds.train, a.clients, True, 1)
    mal = choose_malicious(a.clients, a.malicious_fraction, 0)
    n_byz = int(mal.sum())
    kw = dict(n_byzantine=n_byz, multi=max(1, a.clients - 2 * n_byz)) if "krum" in a.aggregator else {}
    srv = FedAvgServer(0.05, a.batch, subsets, 1.0, 1, 10, model_fn=resnet18_nhwc, test_set=ds.test, device="cuda" if cuda else "cpu",
                       loss_fn=kernel_classification_loss, include_buffers=True, graph=cuda and not a.no_graph,
                       aggregator=make_aggregator(a.aggregator, **kw), attack=SignFlip(a.scale) if n_byz else None, malicious=mal,
                       test_batch_size=min(1000, len(ds.test)))
    if dist.is_initialized() and dist.get_world_size() > 1 and cuda:
        from ddl25spring_b200.parallel.symm_impl import SymmetricAggregator
        srv.collective = SymmetricAggregator(max_rows_per_rank=-(-a.clients // dist.get_world_size()), row_elems=srv.arena.size)
    res = srv.run(a.rounds)
    if not dist.is_initialized() or dist.get_rank() == 0:
        print(res.as_df(perf=True).to_string(index=False))
    if dist.is_initialized():
        dist.destroy_process_group()
    return res


if __name__ == "__main__":
    main()
