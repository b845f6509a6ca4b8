#!/usr/bin/env python
"""ResNet-18 training step (forward + backward, batch B of 3x32x32) on one B200:
ATen/cuDNN model (models/resnet.py, TF32 convs = torch default) vs the channels-last tcgen05 tap-GEMM model
(models/resnet_nhwc.py).  CUDA-event timing, L2 flushed between iterations.  Also times single conv layers.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.nn.functional as F

from ddl25spring_b200 import ops
from ddl25spring_b200.models.resnet import ResNet18
from ddl25spring_b200.models.resnet_nhwc import ResNet18NHWC, from_torch_layout
from ddl25spring_b200.ops import conv as K
from ddl25spring_b200.utils.graphs import GraphedStep
from ddl25spring_b200.utils.timing import L2Flusher


def time_fn(fn, flusher, iters=20, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flusher.flush()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    return tot / iters


def time_graphed(fn, flusher, iters=20, rep=1):
    """GPU time of one ``fn()`` when its launches are replayed from a CUDA graph (no host launch gaps); ``rep`` > 1
    captures that many back-to-back calls and divides (steady-state, L2-warm cost of a small op inside a step)."""
    g = GraphedStep(lambda: [fn() for _ in range(rep)][-1], ())
    return time_fn(g, flusher, iters=iters, warmup=3) / rep


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--layers", action="store_true")
    args = ap.parse_args()
    dev = torch.device("cuda")
    fl = L2Flusher(dev)
    torch.manual_seed(0)
    ref = ResNet18(10).to(dev)
    net = ResNet18NHWC(10).to(dev)
    net.load_state_dict(from_torch_layout(ref.state_dict()))
    x = torch.randn(args.batch, 3, 32, 32, device=dev)
    t = torch.randint(0, 10, (args.batch,), device=dev)

    def step(m):
        def f():
            for p in m.parameters():
                p.grad = None
            F.cross_entropy(m(x), t).backward()
        return f

    out = {"batch": args.batch}
    out["aten_cudnn_tf32_ms"] = time_fn(step(ref), fl)
    ops.reset_launch_count()
    step(net)()
    out["kernel_launches_per_step"] = ops.launch_count()
    out["tcgen05_nhwc_ms"] = time_fn(step(net), fl)
    out["speedup"] = out["aten_cudnn_tf32_ms"] / out["tcgen05_nhwc_ms"]
    out["aten_cudnn_tf32_graph_ms"] = time_graphed(step(ref), fl)
    out["tcgen05_nhwc_graph_ms"] = time_graphed(step(net), fl)
    out["speedup_graph"] = out["aten_cudnn_tf32_graph_ms"] / out["tcgen05_nhwc_graph_ms"]
    out["samples_per_s_graph"] = args.batch / out["tcgen05_nhwc_graph_ms"] * 1e3
    with torch.no_grad():
        ref.eval(); net.eval()
        out["aten_eval_ms"] = time_fn(lambda: ref(x), fl)
        out["tcgen05_eval_ms"] = time_fn(lambda: net(x), fl)
    print(json.dumps(out))

    if args.layers:
        for (N, H, Cin, Cout, k, s, p) in [(args.batch, 8, 64, 64, 3, 1, 1), (args.batch, 8, 64, 128, 3, 2, 1), (args.batch, 4, 128, 128, 3, 1, 1),
                                          (args.batch, 2, 256, 256, 3, 1, 1), (args.batch, 1, 512, 512, 3, 1, 1), (64, 32, 64, 64, 3, 1, 1),
                                          (64, 32, 128, 128, 3, 1, 1)]:
            xh = torch.randn(N, H, H, Cin, device=dev)
            w = torch.randn(Cout, k, k, Cin, device=dev)
            with torch.no_grad():
                xh, w = xh.detach(), w.detach()
                xc, wc = xh.permute(0, 3, 1, 2).contiguous(), w.permute(0, 3, 1, 2).contiguous()
                xcl, wcl = xh.permute(0, 3, 1, 2), wc.contiguous(memory_format=torch.channels_last)
                y = K.conv2d_nhwc(xh, w, s, p)
                gy = torch.randn_like(y)
                gyc, gycl = gy.permute(0, 3, 1, 2).contiguous(), gy.permute(0, 3, 1, 2)
                flops = 2.0 * y.numel() * Cin * k * k
                r = {"shape": [N, H, H, Cin, Cout, k, s]}

                def ours_fwd():
                    ph = K._phase_split(xh) if s == 2 else None
                    return K._fprop(xh, ph, w, s, p), ph

                def ours_bwd(ph):
                    return K._dgrad(gy, w, s, p, H, H), K._wgrad(gy, xh if s == 1 else None, ph, tuple(w.shape), s, p)

                _, ph0 = ours_fwd()
                cbw = lambda g, x_, w_: torch.ops.aten.convolution_backward(g, x_, w_, None, [s, s], [p, p], [1, 1], False, [0, 0], 1,
                                                                            [True, True, False])
                r["fwd_ms"] = time_graphed(ours_fwd, fl, rep=8)
                r["fwd_cudnn_ms"] = time_graphed(lambda: F.conv2d(xc, wc, None, s, p), fl, rep=8)
                r["fwd_cudnn_cl_ms"] = time_graphed(lambda: F.conv2d(xcl, wcl, None, s, p), fl, rep=8)
                r["bwd_ms"] = time_graphed(lambda: ours_bwd(ph0), fl, rep=8)
                r["bwd_cudnn_ms"] = time_graphed(lambda: cbw(gyc, xc, wc), fl, rep=8)
                r["bwd_cudnn_cl_ms"] = time_graphed(lambda: cbw(gycl, xcl, wcl), fl, rep=8)
            r["fwd_tflops"] = flops / r["fwd_ms"] / 1e9
            print(json.dumps(r))
        # BatchNorm(+ReLU) forward / backward on [B*8*8, 64] and [B, 512] (layer1 / layer4 shapes)
        for (R_, C_) in [(args.batch * 64, 64), (args.batch * 16, 128), (args.batch, 512)]:
            xb = torch.randn(R_, C_, device=dev)
            bn = torch.nn.BatchNorm2d(C_).to(dev)
            with torch.no_grad():
                yb = K._BnFn.apply(xb, bn.weight, bn.bias, None, bn.running_mean, bn.running_var, bn.num_batches_tracked, True, 0.1, 1e-5, True)
            gb = torch.randn_like(yb)
            mean = torch.zeros(C_, device=dev); rstd = torch.ones(C_, device=dev)
            ws = K._workspace(dev)
            from ddl25spring_b200.ops import _lib
            L = _lib.lib()

            def bn_f():
                return K._BnFn.forward(_Ctx(), xb, bn.weight, bn.bias, None, bn.running_mean, bn.running_var, None, True, 0.1, 1e-5, True)

            def bn_b():
                dx = torch.empty_like(xb); dg = torch.empty_like(mean); db = torch.empty_like(mean)
                _lib.check(L.ddl_bn_nhwc_bwd(_lib.p(xb), _lib.p(yb), _lib.p(gb), _lib.p(dx), None, _lib.p(mean), _lib.p(rstd), _lib.p(bn.weight),
                                             _lib.p(dg), _lib.p(db), _lib.p(ws["bn_partial"]), _lib.p(ws["bn_counters"]), _lib.i64(R_), int(C_), 1,
                                             _lib.f32(0.0), _lib.stream()))
                return dx

            xn = xb.view(-1, 1, 1, C_).permute(0, 3, 1, 2)
            with torch.no_grad():
                print(json.dumps({"bn_shape": [R_, C_], "fwd_ms": time_graphed(bn_f, fl, rep=8), "bwd_ms": time_graphed(bn_b, fl, rep=8),
                                  "fwd_aten_ms": time_graphed(lambda: F.relu(F.batch_norm(xb, None, None, bn.weight, bn.bias, True)), fl, rep=8)}))


class _Ctx:
    def save_for_backward(self, *a):
        pass


if __name__ == "__main__":
    main()
