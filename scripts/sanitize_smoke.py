"""Tiny forward + backward through every kernel added for the NHWC / LLaMA paths (odd sizes, partial tiles): a quick on-GPU smoke;
compute-sanitizer is closed on this pool, so out-of-bounds checking relies on the oracle comparisons in tests/ (small and odd shapes)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl25spring_b200 import ops
from ddl25spring_b200.ops import conv as K
from ddl25spring_b200.ops.nn import kernel_causal_attention, kernel_rope_qkv, kernel_swiglu
from ddl25spring_b200.models.llama import rope_tables

torch.manual_seed(0)
dev = "cuda"
for (N, H, Cin, Cout, k, s, p) in [(2, 8, 32, 64, 3, 1, 1), (2, 8, 32, 32, 3, 2, 1), (3, 6, 32, 32, 1, 2, 0), (2, 4, 64, 32, 4, 2, 1)]:
    x = torch.randn(N, H, H, Cin, device=dev, requires_grad=True)
    w = torch.randn(Cout, k, k, Cin, device=dev, requires_grad=True)
    y = K.conv2d_nhwc(x, w, s, p)
    y.square().sum().backward()
xt = torch.randn(2, 4, 4, 64, device=dev, requires_grad=True)
wt = torch.randn(64, 4, 4, 32, device=dev, requires_grad=True)
K.conv_transpose2d_nhwc(xt, wt, None, 2, 1, act="tanh").sum().backward()
bn = torch.nn.BatchNorm2d(64).to(dev)
xb = torch.randn(2, 8, 8, 64, device=dev, requires_grad=True)
K.batch_norm_nhwc(xb, bn, relu=True, residual=torch.randn_like(xb)).sum().backward()
xm = torch.randn(2, 8, 8, 64, device=dev, requires_grad=True)
K.global_avg_pool_nhwc(K.max_pool3x3s2_nhwc(xm)).sum().backward()
img = torch.randn(2, 3, 32, 32, device=dev)
ws = torch.zeros(64, 7, 7, 4, device=dev); ws[..., :3].normal_(); ws.requires_grad_()
K.stem_conv7x7s2(img, ws).sum().backward()
q, kk, v = (torch.randn(1, 2, 70, 48, device=dev, requires_grad=True) for _ in range(3))
kernel_causal_attention(q, kk, v).sum().backward()
cos, sin = rope_tables(48, 128, device=dev)
qkv = torch.randn(1, 70, 3 * 2 * 48, device=dev, requires_grad=True)
a, b, c = kernel_rope_qkv(qkv, cos, sin, 2)
(a.sum() + b.sum() + c.sum()).backward()
g, u = (torch.randn(4, 64, device=dev, requires_grad=True) for _ in range(2))
kernel_swiglu(g, u).sum().backward()
ops.u8_normalize(torch.randint(0, 256, (16, 784), dtype=torch.uint8, device=dev), 0.1307, 0.3081)
torch.cuda.synchronize()
print("sanitize smoke ok")
