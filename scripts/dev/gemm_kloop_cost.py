"""Fixed cost vs per-k-block cost of the tcgen05 GEMM (TF32 and 3xTF32) at the LLaMA lab's shapes: K sweep, one split."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from ddl25spring_b200.ops import gemm as G
G.AUTO_SPLIT = False

def t(fn, n=50, reps=10):
    """Device time per call: n dependent calls captured into one CUDA graph (no host launch cost), replayed reps times."""
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for _ in range(3): fn()
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

for (M, N) in [(768, 864), (768, 288), (768, 1536)]:
    for K in [32, 288, 1152, 4608]:
        a, b = torch.randn(M, K, device="cuda"), torch.randn(N, K, device="cuda")
        out = torch.empty(M, N, device="cuda")
        row = []
        for precise in (False, True):
            for ns in (1, 3):
                if ns > 1 and K < 288: row.append(float("nan")); continue
                row.append(t(lambda: G.gemm(a, b, precise=precise, n_split=ns, out=out)))
        print(f"M={M} N={N} K={K:5d}  tf32 x1 {row[0]:7.1f} us  x3 {row[1]:7.1f} us | 3xtf32 x1 {row[2]:7.1f} us  x3 {row[3]:7.1f} us", flush=True)
