"""Run the descriptor row-shift probe (csrc/dbg_probe.cu) and print which base_offset rule works."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl25spring_b200.ops import _lib

L = _lib.lib()
torch.manual_seed(0)
A = torch.randn(256, 32, device="cuda")
B = torch.randn(64, 32, device="cuda")
for mode in (0, 1, 2):
    oks = []
    for shift in range(0, 64):
        out = torch.zeros(128, 64, device="cuda")
        _lib.check(L.ddl_dbg_probe_shift(_lib.p(A), _lib.p(B), _lib.p(out), shift, mode, _lib.stream()), "probe")
        torch.cuda.synchronize()
        exp = A[shift:shift + 128].double() @ B.double().T
        err = float((out.double() - exp).abs().max() / exp.abs().max())
        oks.append(err < 5e-3)
    print(f"mode {mode}: ok shifts = {[i for i, o in enumerate(oks) if o]}")

A2 = torch.randn(128, 32, device="cuda")
B2 = torch.randn(32, 64, device="cuda")
oks = []
for shift in range(0, 90):
    out = torch.zeros(128, 64, device="cuda")
    _lib.check(L.ddl_dbg_probe_mn(_lib.p(A2), _lib.p(B2), _lib.p(out), shift, _lib.stream()), "probe_mn")
    torch.cuda.synchronize()
    exp = torch.zeros(128, 64, dtype=torch.float64, device="cuda")
    ok = True
    for j in range(4):
        if shift + 32 + j > 128:
            continue
        e = A2[shift + j: shift + j + 32].double().T @ B2.double()      # [32 c, 64 n]
        got = out[j * 32:(j + 1) * 32].double()
        ok &= float((got - e).abs().max() / e.abs().max()) < 5e-3
    oks.append(ok)
print(f"MN-major overlapped atoms: ok shifts = {[i for i, o in enumerate(oks) if o]}")
