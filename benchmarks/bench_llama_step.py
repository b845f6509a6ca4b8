#!/usr/bin/env python
"""The DP / PP labs' LLaMA training step (lab/tutorial_1b: dmodel 288, 6 heads, 6 layers, ctx 256, vocab 32 000, batch 3, Adam)
on one B200: plain PyTorch modules vs this framework's kernels (tcgen05 linears, RMSNorm, flash-style causal attention, fused
causal-LM loss, fused Adam), each eager and replayed from a CUDA graph (parallel/dp.py graph=True).  Device-timed."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from ddl25spring_b200.models.llama import LLama, causalLLMLoss
from ddl25spring_b200.ops.nn import kernel_causal_lm_loss, use_kernels
from ddl25spring_b200.parallel.dp import DataParallel
from ddl25spring_b200.utils.timing import L2Flusher


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=3)
    ap.add_argument("--steps", type=int, default=20)
    a = ap.parse_args()
    dev = torch.device("cuda")
    fl = L2Flusher(dev)
    V = 32_000
    tok = torch.randint(0, V, (a.batch, 256), device=dev)
    out = {"batch": a.batch, "seq": 256, "model": "LLaMA 6x288 (vocab 32000)", "ms_per_step": {}, "tokens_per_s": {}}
    for arm in ("torch_eager", "torch_graph", "kernels_eager", "kernels_graph"):
        torch.manual_seed(0)
        net = LLama(None, V, 288, 6, dev, 6, 256)
        kernels = arm.startswith("kernels")
        if kernels:
            use_kernels(net)
        loss_fn = (lambda m, t: kernel_causal_lm_loss(m(t), t)) if kernels else (lambda m, t: causalLLMLoss(m(t), t, V))
        dp = DataParallel(net, optimizer="adam", lr=8e-4, graph=arm.endswith("graph"))
        for _ in range(5):
            loss = dp.train_step(loss_fn, tok)
        torch.cuda.synchronize()
        tot = 0.0
        for _ in range(a.steps):
            fl.flush()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); loss = dp.train_step(loss_fn, tok); e1.record()
            torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        ms = tot / a.steps
        out["ms_per_step"][arm] = round(ms, 3)
        out["tokens_per_s"][arm] = round(a.batch * 256 / ms * 1e3, 1)
        out.setdefault("final_loss", {})[arm] = round(float(loss), 4)
        del dp, net
        torch.cuda.empty_cache()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
