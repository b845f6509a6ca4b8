#!/usr/bin/env python
"""One DCGAN iteration (64x64 images, batch B): discriminator step on real + fake, then generator step — the local update of
the federated DCGAN (gen/fed_dcgan.py) — on ATen/cuDNN (models/dcgan.py) vs the NHWC tcgen05 stack (models/dcgan_nhwc.py),
eager and replayed from a CUDA graph.  Gradients only (the fused Adam step is identical in both arms).  Device-timed."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from ddl25spring_b200.models import dcgan, dcgan_nhwc
from ddl25spring_b200.utils.graphs import GraphedStep
from ddl25spring_b200.utils.timing import L2Flusher


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    dev = torch.device("cuda")
    fl = L2Flusher(dev)
    torch.manual_seed(0)
    real = torch.rand(a.batch, 3, 64, 64, device=dev) * 2 - 1
    z = torch.randn(a.batch, 100, device=dev)
    ones, zeros = torch.ones(a.batch, device=dev), torch.zeros(a.batch, device=dev)
    out = {"batch": a.batch, "ms_per_iter": {}}
    for arm in ("aten", "nhwc"):
        G, D = (dcgan.Generator(), dcgan.Discriminator()) if arm == "aten" else (dcgan_nhwc.GeneratorNHWC(), dcgan_nhwc.DiscriminatorNHWC())
        G, D = G.to(dev), D.to(dev)

        def it(real_, z_):
            for p in list(G.parameters()) + list(D.parameters()):
                p.grad = None
            fake = G(z_)
            d_loss = F.binary_cross_entropy_with_logits(D(real_), ones) + F.binary_cross_entropy_with_logits(D(fake.detach()), zeros)
            d_loss.backward()
            g_loss = F.binary_cross_entropy_with_logits(D(fake), ones)
            g_loss.backward()
            return d_loss.detach() + g_loss.detach()

        for mode in ("eager", "graph"):
            fn = it if mode == "eager" else GraphedStep(it, (real, z))
            for _ in range(3):
                fn(real, z)
            torch.cuda.synchronize()
            tot = 0.0
            for _ in range(a.iters):
                fl.flush()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); fn(real, z); e1.record()
                torch.cuda.synchronize()
                tot += e0.elapsed_time(e1)
            out["ms_per_iter"][f"{arm}_{mode}"] = round(tot / a.iters, 3)
    m = out["ms_per_iter"]
    out["speedup_graph"] = round(m["aten_graph"] / m["nhwc_graph"], 3)
    out["images_per_s_nhwc_graph"] = round(a.batch / m["nhwc_graph"] * 1e3, 1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
