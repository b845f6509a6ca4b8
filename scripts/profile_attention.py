"""Causal attention forward + backward at the lab shape (B*H = 18, T = 256, D = 48): target for `ncu --set full -k regex:attn_`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl25spring_b200.ops.nn import kernel_causal_attention
q, k, v = (torch.randn(3, 6, 256, 48, device="cuda", requires_grad=True) for _ in range(3))
for _ in range(3):
    o = kernel_causal_attention(q, k, v)
    o.sum().backward()
torch.cuda.synchronize()
print("ok", float(o.abs().mean()))
