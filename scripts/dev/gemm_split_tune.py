"""Split-K choice for the LLaMA lab's GEMM shapes, timed the way the training step runs them: dependent calls inside one
CUDA graph (programmatic dependent launch, separate splitk_reduce kernel).  Prints us per call for each n_split."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from ddl25spring_b200.ops import gemm as G
G.AUTO_SPLIT = False

def t(fn, n=40, reps=8):
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for _ in range(2): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(n): fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (n * reps) * 1e3

T, D, F, V = 768, 288, 768, 32000
shapes = []
for name, n_out, k_in in [("qkv", 3 * D, D), ("wo", D, D), ("gate_up", 2 * F, D), ("down", D, F), ("head", V, D)]:
    shapes.append((name + ".fwd", T, n_out, k_in, False, False))       # y = x w^T
    shapes.append((name + ".dgrad", T, k_in, n_out, False, True))      # dx = dy w
    shapes.append((name + ".wgrad", n_out, k_in, T, True, True))       # dw = dy^T x
splits = [1, 2, 3, 4, 6, 8, 12]
print("shape (M, N, K)".ljust(40) + "".join(f"x{s:<7d}" for s in splits) + " auto")
for name, M, N, K, a_mn, b_mn in shapes:
    a = torch.randn((K, M) if a_mn else (M, K), device="cuda")
    b = torch.randn((K, N) if b_mn else (N, K), device="cuda")
    out = torch.empty(M, N, device="cuda")
    row = []
    for ns in splits:
        if ns > max(1, (K // 32) // 2): row.append(float("nan")); continue
        row.append(t(lambda: G.gemm(a, b, a_mn=a_mn, b_mn=b_mn, precise=True, n_split=ns, out=out)))
    G.AUTO_SPLIT = True
    auto = t(lambda: G.gemm(a, b, a_mn=a_mn, b_mn=b_mn, precise=True, out=out))
    G.AUTO_SPLIT = False
    print(f"{name:14s} ({M:5d},{N:5d},{K:5d})".ljust(40) + "".join(f"{v:<8.1f}" for v in row) + f" {auto:.1f}", flush=True)
