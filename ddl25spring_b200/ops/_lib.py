"""ctypes binding to ``libddl25_kernels.so`` (the hand-written sm_100a kernels).

The library is built in-tree by ``ddl25spring_b200.build`` and loaded lazily.  On a machine
with a GPU a missing or unloadable library is a hard error – the CUDA ops never fall back to
PyTorch silently.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import torch

_LIB = None
LIB_PATH = Path(__file__).resolve().parent.parent / "lib" / "libddl25_kernels.so"


class KernelError(RuntimeError):
    pass


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        if not LIB_PATH.exists():
            from .. import build as _build
            _build.build()
        _LIB = C.CDLL(str(LIB_PATH))
        _LIB.ddl_last_error.restype = C.c_char_p
        _LIB.ddl_arch.restype = C.c_char_p
        _LIB.ddl_launch_count.restype = C.c_int64
        _LIB.ddl_bn_nhwc_ws_floats.restype = C.c_int64
    return _LIB


def available() -> bool:
    try:
        lib()
        return True
    except Exception:
        return False


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise KernelError(f"{what} failed (rc={rc}): {lib().ddl_last_error().decode()}")


def launch_count() -> int:
    return int(lib().ddl_launch_count())


def reset_launch_count() -> None:
    lib().ddl_reset_launch_count()


def p(t: torch.Tensor | None) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def i64(x) -> C.c_int64:
    return C.c_int64(int(x))


def f32(x) -> C.c_float:
    return C.c_float(float(x))


def host_floats(xs) -> C.Array:
    xs = [float(x) for x in xs]
    return (C.c_float * len(xs))(*xs)


def host_ptrs(ptrs) -> C.Array:
    ptrs = [int(x) for x in ptrs]
    return (C.c_void_p * len(ptrs))(*ptrs)


def req_f32(t: torch.Tensor, name: str = "tensor") -> torch.Tensor:
    if t.dtype != torch.float32 or not t.is_cuda:
        raise KernelError(f"{name}: expected a CUDA float32 tensor, got {t.dtype} on {t.device}")
    return t
