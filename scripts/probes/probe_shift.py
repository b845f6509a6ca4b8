"""Run the descriptor row-shift probe (scripts/probes/dbg_probe.cu) and print which base_offset rule works.

The probe kernels are NOT part of libddl25_kernels.so: this script compiles them into their own shared object next to
itself (linking the product library for the tensor-map encoder and error plumbing) and loads that."""
import ctypes, os, subprocess, sys
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import torch
from ddl25spring_b200 import build as _build
from ddl25spring_b200.ops import _lib

_lib.lib()                                             # product library first (RTLD_GLOBAL symbols the probe links against)
SO = os.path.join(HERE, "libddl25_probes.so")
if not os.path.exists(SO) or os.path.getmtime(SO) < os.path.getmtime(os.path.join(HERE, "dbg_probe.cu")):
    subprocess.run([_build.NVCC, *_build.ARCH_FLAGS, *_build.CFLAGS, "-shared", os.path.join(HERE, "dbg_probe.cu"), "-o", SO,
                    "-L", str(_build.LIBDIR), "-lddl25_kernels", f"-Xlinker=-rpath={_build.LIBDIR}", "-cudart", "static"], check=True)
L = ctypes.CDLL(SO)
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
