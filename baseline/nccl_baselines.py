#!/usr/bin/env python
"""The reference's algorithms re-stated on torch.distributed NCCL + ATen: "the baseline, not the product".

  [torchrun --nproc-per-node N --master-addr 127.0.0.1] python baseline/nccl_baselines.py [--only dp_ga,dp_wa,fedavg,robust,vfl]

None of this framework's kernels, arenas or graphs runs here - only stock PyTorch (cuBLAS / cuDNN / ATen) and
``torch.distributed`` with the nccl backend, one rank per GPU.  What each section keeps from the reference and what it changes:

* ``dp_ga``   lab/tutorial_1b/DP/gradient_aggr/intro_DP_GA.py:41-68 - forward, backward, barrier, flatten every gradient into
              one vector (``torch.cat``), ONE all-reduce, split, divide by W, Adam step, ``empty_cache``.  Changed: backend
              "nccl" and no ``.to("cpu")`` round trip (the tensors stay on the device); LLaMA from models/llama.py in plain
              PyTorch (simplellm is not installable), synthetic tokens.
* ``dp_wa``   weight_aggr/intro_DP_WA.py:52-67 with the average WRITTEN BACK (the reference drops it, :67).
* ``fedavg``  lab/tutorial_1a/hfl_complete.py:336-390 with the sampled clients dealt round-robin to the ranks, stock autograd
              training per client (MnistCnn, B = 100, E = 1), scaled weights summed with ``dist.all_reduce``.
* ``robust``  [north star] Krum / coordinate median / trimmed mean as ``all_gather`` of the update rows + torch ops.
* ``vfl``     lab/tutorial_2b/vfl.py:36-37,79-82 - party embeddings ``all_gather``-ed, ``torch.cat``, ``nn.Linear`` top model,
              autograd backward, AdamW.

The pipeline baseline (``dist.isend`` / ``irecv`` of CUDA tensors between plain-PyTorch LLaMA stages, lab/tutorial_1b/PP/1F1B/
intro_PP_1F1B_MB.py) is the ``torch`` arm of ``benchmarks/bench_pp_llama.py``; the DCGAN ``dist.all_reduce`` baseline is the
``nccl`` arm of ``benchmarks/bench_fed_dcgan.py``.

Timing: CUDA events around the timed iterations on every rank, max over ranks; rank 0 prints one JSON object whose rows are
the "reference build on NCCL" columns of BASELINE.md section 4.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)

import torch
import torch.distributed as dist
import torch.nn.functional as F


def _max_ms(ms: float) -> float:
    if dist.is_initialized() and dist.get_world_size() > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)
    return ms


def _timed(step, warm: int, iters: int) -> float:
    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    if dist.is_initialized():
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        step()
    e1.record()
    torch.cuda.synchronize()
    return _max_ms(e0.elapsed_time(e1) / iters)


def dp_llama(world: int, rank: int, mode: str) -> dict:
    from ddl25spring_b200.core.data import SyntheticTokens
    from ddl25spring_b200.models.llama import LLama, causalLLMLoss
    V, B, T = 32_000, 3, 256
    torch.manual_seed(0)
    net = LLama(None, V, 288, 6, "cuda", 6, T)
    optim = torch.optim.Adam(net.parameters(), lr=8e-4)
    ds = iter(SyntheticTokens(V, B, T, skip=rank * 5000))
    sizes = [p.shape for p in net.parameters()]
    lens = [p.numel() for p in net.parameters()]

    def step_ga():
        optim.zero_grad()
        x = next(ds).cuda()
        loss = causalLLMLoss(net(x), x, V)
        loss.backward()
        if world > 1:
            dist.barrier()
        tmp = []
        for p in net.parameters():
            tmp.append(torch.zeros_like(p).view(-1) if p.grad is None else p.grad.view(-1))
            p.grad = None
        flat = torch.cat(tmp)                                   # (the reference copies this to the CPU for gloo)
        if world > 1:
            dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        for p, g, s in zip(net.parameters(), torch.split(flat, lens), sizes):
            p.grad = g.view(s) / world
        optim.step()
        torch.cuda.empty_cache()

    def step_wa():
        optim.zero_grad()
        x = next(ds).cuda()
        causalLLMLoss(net(x), x, V).backward()
        optim.step()
        if world > 1:
            dist.barrier()
        flat = torch.cat([p.data.view(-1) for p in net.parameters()])
        if world > 1:
            dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        with torch.no_grad():
            for p, w, s in zip(net.parameters(), torch.split(flat, lens), sizes):
                p.copy_(w.view(s) / world)                      # written back (intro_DP_WA.py:67 assigns to the loop variable)

    ms = _timed(step_ga if mode == "ga" else step_wa, 3, 10)
    return {"model": "LLaMA 6x288, vocab 32000, batch 3 x 256 per rank, Adam", "ms_per_step": round(ms, 3),
            "tokens_per_s": round(world * B * T / ms * 1e3, 1)}


def fedavg_mnist(world: int, rank: int) -> dict:
    import numpy as np
    from torch.utils.data import DataLoader, TensorDataset
    from ddl25spring_b200.core.data import SyntheticMNIST, split
    from ddl25spring_b200.models.mnist_cnn import MnistCnn
    ds = SyntheticMNIST(seed=0)
    subsets = split(ds.train, 100, True, 10)
    K = 8 * world
    torch.manual_seed(10)
    glob = MnistCnn().cuda()
    worker = MnistCnn().cuda()
    xs = ds.train.images.cuda()
    ys = ds.train.targets.cuda()
    xt, yt = ds.test.images.cuda(), ds.test.targets.cuda()
    rng = np.random.default_rng(10)

    def one_round():
        chosen = rng.choice(100, K, replace=False)
        weights = [p.detach().clone() for p in glob.parameters()]
        acc = [torch.zeros_like(p) for p in weights]
        n_tot = sum(len(subsets[int(c)]) for c in chosen)
        for pos, c in enumerate(chosen):
            if pos % world != rank:
                continue
            idx = torch.as_tensor(np.asarray(subsets[int(c)].indices), device="cuda")
            with torch.no_grad():
                for p, w in zip(worker.parameters(), weights):
                    p.copy_(w)
            opt = torch.optim.SGD(worker.parameters(), lr=0.01)
            worker.train()
            g = torch.Generator(device="cuda").manual_seed(int(c))
            perm = idx[torch.randperm(len(idx), device="cuda", generator=g)]
            for lo in range(0, len(perm), 100):
                b = perm[lo:lo + 100]
                opt.zero_grad()
                F.nll_loss(worker(xs[b]), ys[b]).backward()
                opt.step()
            with torch.no_grad():
                for a, p in zip(acc, worker.parameters()):
                    a.add_(p, alpha=len(idx) / n_tot)
        flat = torch.cat([a.view(-1) for a in acc])
        if world > 1:
            dist.all_reduce(flat)
        with torch.no_grad():
            off = 0
            for p in glob.parameters():
                p.copy_(flat[off:off + p.numel()].view_as(p))
                off += p.numel()
            glob.eval()
            sl = slice(rank, None, world)
            correct = (glob(xt[sl]).argmax(1) == yt[sl]).sum()
            if world > 1:
                dist.all_reduce(correct)
        return float(correct) / len(yt)

    ms = _timed(one_round, 2, 5)
    return {"config": "FedAvg MnistCnn, 8 clients per GPU per round, B=100, E=1, + 10 000-sample test (stock autograd, data resident)",
            "ms_per_round": round(ms, 2), "samples_per_s": round(K * 600 / ms * 1e3, 1)}


def robust(world: int, rank: int) -> dict:
    D, G_ = 11_181_696, 8
    K = G_ * world
    f = max(1, int(0.2 * K))
    g = torch.Generator(device="cuda").manual_seed(rank)
    base = 0.05 * torch.randn(D, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
    rows = base[None, :] + 0.01 * torch.randn(G_, D, device="cuda", generator=g)
    allrows = torch.empty(world, G_, D, device="cuda")
    t = int(-(-0.2 * K // 1))

    def gather():
        if world > 1:
            dist.all_gather_into_tensor(allrows.view(-1), rows.view(-1))
            return allrows.view(K, D)
        return rows

    def krum():
        A = gather()
        Ad = A - base[None, :]
        dd = torch.cdist(Ad, Ad).pow(2)
        dd.fill_diagonal_(float("inf"))
        sc = torch.sort(dd, dim=1).values[:, :K - f - 2].sum(1)
        return A[torch.argsort(sc)[:K - 2 * f]].mean(0)
    out = {"K": K, "row_floats": D}
    out["mean_ms"] = round(_timed(lambda: gather().mean(0), 1, 3), 3)
    out["krum_ms"] = round(_timed(krum, 1, 3), 3)
    out["median_ms"] = round(_timed(lambda: torch.median(gather(), dim=0).values, 1, 2), 3)
    out["trimmed_mean_ms"] = round(_timed(lambda: torch.sort(gather(), dim=0).values[t:K - t].mean(0), 1, 2), 3)
    return out


def vfl(world: int, rank: int) -> dict:
    Bsz, emb = 4096, 64
    torch.manual_seed(1000 + rank)
    bottom = torch.nn.Sequential(torch.nn.Linear(32, emb), torch.nn.ReLU(), torch.nn.Linear(emb, emb), torch.nn.ReLU()).cuda()
    torch.manual_seed(0)
    top = torch.nn.Sequential(torch.nn.Linear(emb * world, 128), torch.nn.LeakyReLU(), torch.nn.Linear(128, 256), torch.nn.LeakyReLU(),
                              torch.nn.Linear(256, 2)).cuda()
    opt = torch.optim.AdamW(list(bottom.parameters()) + list(top.parameters()))
    x = torch.randn(Bsz, 32, device="cuda")
    y = torch.randint(0, 2, (Bsz,), device="cuda")

    class Gather(torch.autograd.Function):
        @staticmethod
        def forward(ctx, e):
            parts = [torch.empty_like(e) for _ in range(world)]
            if world > 1:
                dist.all_gather(parts, e.contiguous())
            else:
                parts = [e]
            return torch.cat(parts, 1)

        @staticmethod
        def backward(ctx, g):
            return g[:, rank * emb:(rank + 1) * emb].contiguous()

    def step():
        opt.zero_grad()
        F.cross_entropy(top(Gather.apply(bottom(x))), y).backward()
        opt.step()
    ms = _timed(step, 3, 20)
    return {"config": f"one party per GPU, B={Bsz}, {emb}-wide embeddings, top 128-256-2, eager", "ms_per_step": round(ms, 4),
            "samples_per_s": round(Bsz / ms * 1e3, 1)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="dp_ga,dp_wa,fedavg,robust,vfl")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    out = {"impl": "reference algorithms on torch.distributed nccl + ATen", "world": world}
    for name in a.only.split(","):
        try:
            if name == "dp_ga":
                out[name] = dp_llama(world, rank, "ga")
            elif name == "dp_wa":
                out[name] = dp_llama(world, rank, "wa")
            elif name == "fedavg":
                out[name] = fedavg_mnist(world, rank)
            elif name == "robust":
                out[name] = robust(world, rank)
            elif name == "vfl":
                out[name] = vfl(world, rank)
        except Exception as e:      # one section failing must not hide the others
            out[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
        torch.cuda.empty_cache()
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
