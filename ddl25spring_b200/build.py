"""In-tree build of the sm_100a kernel library.

``python -m ddl25spring_b200.build`` compiles every ``csrc/*.cu`` with
``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`` into
``ddl25spring_b200/lib/libddl25_kernels.so`` (git-ignored, travels to the GPU box with the
repo snapshot).  Objects are cached by source mtime so incremental rebuilds take seconds.
nvcc cross-compiles without a GPU, so this is also the CPU-side "does it build" check.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
LIBDIR = PKG / "lib"
OBJDIR = PKG / "lib" / "obj"
LIB = LIBDIR / "libddl25_kernels.so"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
CFLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
          "-I", str(ROOT / "include"), "-I", str(CSRC)]


def _deps_mtime() -> float:
    hdrs = list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + list((ROOT / "include").glob("*.h"))
    return max((h.stat().st_mtime for h in hdrs), default=0.0)


def _compile(src: Path, hdr_mtime: float, verbose: bool) -> Path:
    obj = OBJDIR / (src.stem + ".o")
    if obj.exists() and obj.stat().st_mtime > max(src.stat().st_mtime, hdr_mtime):
        return obj
    cmd = [NVCC, *ARCH_FLAGS, *CFLAGS, "-Xptxas", "-v" if verbose else "-warn-spills", "-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    if verbose or "spill" in r.stderr.lower():
        sys.stderr.write(r.stderr)
    return obj


def build(verbose: bool = False, force: bool = False) -> Path:
    OBJDIR.mkdir(parents=True, exist_ok=True)
    srcs = sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cpp"))
    if not srcs:
        raise RuntimeError("no CUDA sources found")
    if force:
        for o in OBJDIR.glob("*.o"):
            o.unlink()
    stems = {s.stem for s in srcs}
    stale = [o for o in OBJDIR.glob("*.o") if o.stem not in stems]      # objects of sources that were removed / moved out
    for o in stale:
        o.unlink()
    if stale and LIB.exists():
        LIB.unlink()
    hdr_mtime = _deps_mtime()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, hdr_mtime, verbose), srcs))
    newest = max(o.stat().st_mtime for o in objs)
    if not LIB.exists() or LIB.stat().st_mtime < newest:
        cmd = [NVCC, *ARCH_FLAGS, "-shared", "-o", str(LIB), *map(str, objs), "-cudart", "static", "-ldl", "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    p = build(verbose="-v" in sys.argv, force="-f" in sys.argv)
    print(p)
