"""Small single-GPU drivers for `ncu --set full` captures of the round-2 kernels.

  python scripts/profile_r2_kernels.py gram8 | gram64 | median64 | linear | attention | persist
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl25spring_b200 import ops
from ddl25spring_b200.ops import gemm as G
from ddl25spring_b200.ops import nn as knn

which = sys.argv[1] if len(sys.argv) > 1 else "gram8"
D = 11_194_496
torch.manual_seed(0)
if which in ("gram8", "gram64"):
    K = 8 if which == "gram8" else 64
    c = 0.05 * torch.randn(D, device="cuda")
    U = c[None, :] + 0.01 * torch.randn(K, D, device="cuda")
    for _ in range(4):
        g = G.gram_tcgen05(U, center=c) if K == 8 else G.gram_tcgen05(U)
    torch.cuda.synchronize()
    print(which, float(g[0, 0]))
elif which == "median64":
    U = 0.01 * torch.randn(64, D, device="cuda")
    for _ in range(4):
        m = ops.coordinate_median(U)
    torch.cuda.synchronize()
    print(which, float(m[0]))
elif which == "linear":
    x = torch.randn(768, 288, device="cuda")
    w = torch.randn(864, 288, device="cuda")
    for _ in range(4):
        y = G.gemm(x, w, precise=True)
    torch.cuda.synchronize()
    print(which, float(y[0, 0]))
elif which == "attention":
    q, k, v = (torch.randn(3, 6, 256, 48, device="cuda", requires_grad=True) for _ in range(3))
    for _ in range(3):
        o = knn.kernel_causal_attention(q, k, v)
        o.sum().backward()
    torch.cuda.synchronize()
    print(which, float(o[0, 0, 0, 0]))
elif which == "persist":
    # the LLaMA lab's vocabulary projection: 768 x 32 000 x 288, 3xTF32 -> 1 500 tiles on the persistent kernel
    x = torch.randn(768, 288, device="cuda")
    w = torch.randn(32000, 288, device="cuda")
    for _ in range(4):
        y = G.gemm(x, w, precise=True)
    torch.cuda.synchronize()
    print(which, float(y[0, 0]))
