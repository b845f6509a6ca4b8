"""Data-parallel LLaMA training (lab/tutorial_1b/DP): gradient aggregation (intro_DP_GA.py) or weight aggregation
(intro_DP_WA.py, with the average actually written back).

  python -m torch.distributed.run --nproc-per-node 3 --master-addr 127.0.0.1 examples/dp_train.py [--mode ga|wa] [--iters 20]
Each rank reads its own slice of the token stream (skip = rank * 5000, intro_DP_GA.py:29)."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from ddl25spring_b200.core.data import SyntheticTokens
from ddl25spring_b200.models.llama import LLama, causalLLMLoss
from ddl25spring_b200.parallel.dp import DataParallel, WeightAveraging


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="ga", choices=["ga", "wa"])
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--dmodel", type=int, default=288)
    ap.add_argument("--vocab", type=int, default=32_000)
    ap.add_argument("--fused", action="store_true", help="peer-memory gradient reducer instead of NCCL (GPU only)")
    ap.add_argument("--kernels", action="store_true", help="run the model on the hand-written kernels (ops/nn.use_kernels)")
    ap.add_argument("--graph", action="store_true", help="replay forward + backward from a CUDA graph (GPU only)")
    a = ap.parse_args()
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dist.init_process_group("nccl" if cuda else "gloo")
    rank, dev = dist.get_rank(), ("cuda" if cuda else "cpu")
    torch.manual_seed(0)
    net = LLama(None, a.vocab, a.dmodel, 6, dev, 6, 256)
    loss_fn = lambda m, t: causalLLMLoss(m(t), t, a.vocab)
    if a.kernels and cuda:
        from ddl25spring_b200.ops.nn import kernel_causal_lm_loss, use_kernels
        use_kernels(net)
        loss_fn = lambda m, t: kernel_causal_lm_loss(m(t), t)
    ds = iter(SyntheticTokens(a.vocab, 1, 256, skip=rank * 5000))
    if a.mode == "ga":
        reducer = None
        if a.fused and cuda:
            from ddl25spring_b200.parallel.symm_impl import SymmetricArena, SymmetricGradReducer
            reducer = SymmetricGradReducer(SymmetricArena(), sum(p.numel() for p in net.parameters()) + 4096)
        dp = DataParallel(net, optimizer="adam", lr=8e-4, reducer=reducer, graph=a.graph)
        for itr in range(a.iters):
            x = next(ds).to(dev)
            loss = dp.train_step(loss_fn, x)
            print(itr, float(loss), flush=True)
    else:
        averager = None
        if a.fused and cuda:            # the replicas' weights are averaged in place by ONE peer-memory kernel per rank (K2b)
            from ddl25spring_b200.parallel.symm_impl import SymmetricArena, SymmetricWeightAverager
            averager = SymmetricWeightAverager(SymmetricArena(), sum(p.numel() for p in net.parameters()) + 4096)
        wa = WeightAveraging(net, torch.optim.Adam(net.parameters(), lr=8e-4), averager=averager)
        for itr in range(a.iters):
            x = next(ds).to(dev)
            loss = loss_fn(net, x)
            loss.backward()
            wa.step()
            print(itr, float(loss), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
