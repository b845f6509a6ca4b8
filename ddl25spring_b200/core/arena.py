"""Flat parameter arenas.

Every model in the framework keeps its parameters in ONE contiguous fp32 buffer with the
module's parameters re-bound as views into it.  That turns every aggregator (FedAvg,
Krum, median, trimmed mean), every data-parallel all-reduce and every optimizer step into
a 1-D problem over a single pointer, which is what the sm_100a kernels in ``csrc/``
consume.  It is the device-resident replacement for the reference's per-round
``cat([p.grad.view(-1) ...])`` (lab/tutorial_1b/DP/gradient_aggr/intro_DP_GA.py:55-65) and
for its list-of-CPU-tensors "messages" (lab/tutorial_1a/hfl_complete.py:277,332,356).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Iterable, Iterator

import torch
from torch import nn

# Every tensor starts on a 16-byte boundary (float4 / TMA friendly); the arena length is
# padded to a multiple of ARENA_PAD floats so vector kernels never need a scalar tail.
TENSOR_ALIGN = 4
ARENA_PAD = 128


@dataclass(frozen=True)
class TensorSpec:
    name: str
    shape: tuple[int, ...]
    offset: int
    numel: int
    is_buffer: bool = False


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def build_specs(named: Iterable[tuple[str, torch.Tensor]], *, buffers: bool = False,
                start: int = 0) -> tuple[list[TensorSpec], int]:
    specs: list[TensorSpec] = []
    off = start
    for name, t in named:
        off = _round_up(off, TENSOR_ALIGN)
        specs.append(TensorSpec(name, tuple(t.shape), off, t.numel(), buffers))
        off += t.numel()
    return specs, off


class FlatArena:
    """Contiguous fp32 storage for a module's parameters (and, optionally, its float buffers).

    ``arena.flat``  – 1-D parameter buffer, ``arena.grad`` – matching gradient buffer
    (allocated lazily by :meth:`bind_grads`).  ``arena.views()`` yields per-tensor views in
    ``module.parameters()`` order, so code written against the reference's
    "list of tensors" message format keeps working.
    """

    def __init__(self, specs: list[TensorSpec], size: int, *, device=None, dtype=torch.float32,
                 flat: torch.Tensor | None = None):
        self.specs = specs
        self.size = _round_up(max(size, 1), ARENA_PAD)
        self.n_param_elems = sum(s.numel for s in specs)
        if flat is None:
            flat = torch.zeros(self.size, device=device, dtype=dtype)
        else:
            assert flat.numel() == self.size and flat.dim() == 1
        self.flat = flat
        self.grad: torch.Tensor | None = None
        self.module: nn.Module | None = None

    # ------------------------------------------------------------------ construction
    @classmethod
    def like_module(cls, module: nn.Module, *, include_buffers: bool = False, device=None) -> "FlatArena":
        specs, end = build_specs(module.named_parameters())
        if include_buffers:
            bufs = [(n, b) for n, b in module.named_buffers() if b.dtype.is_floating_point]
            bspecs, end = build_specs(bufs, buffers=True, start=end)
            specs = specs + bspecs
        if device is None:
            try:
                device = next(module.parameters()).device
            except StopIteration:
                device = torch.device("cpu")
        return cls(specs, end, device=device)

    @classmethod
    def from_module(cls, module: nn.Module, *, include_buffers: bool = False, with_grad: bool = False,
                    device=None) -> "FlatArena":
        """Move ``module``'s parameters into a new arena and re-bind them as views."""
        arena = cls.like_module(module, include_buffers=include_buffers, device=device)
        arena.bind(module)
        if with_grad:
            arena.bind_grads()
        return arena

    def bind(self, module: nn.Module) -> None:
        params = dict(module.named_parameters())
        bufs = dict(module.named_buffers())
        with torch.no_grad():
            for s, v in zip(self.specs, self.views()):
                src = bufs[s.name] if s.is_buffer else params[s.name]
                v.copy_(src.detach().to(v.device, v.dtype))
                src.data = v  # parameter / buffer now aliases the arena
        self.module = module

    def rebase(self, storage: torch.Tensor) -> None:
        """Move the arena onto caller-provided storage (e.g. a symmetric / peer-mapped buffer) and re-bind the module's
        parameters and buffers as views into it.  ``storage``: 1-D, same dtype / device, at least ``size`` elements."""
        assert storage.dim() == 1 and storage.numel() >= self.size and storage.dtype == self.flat.dtype
        new = storage[:self.size]
        with torch.no_grad():
            new.copy_(self.flat)
        self.flat = new
        if self.module is not None:
            params = dict(self.module.named_parameters())
            bufs = dict(self.module.named_buffers())
            for s, v in zip(self.specs, self.views()):
                (bufs[s.name] if s.is_buffer else params[s.name]).data = v

    def bind_grads(self) -> torch.Tensor:
        """Allocate the gradient arena and alias every ``param.grad`` into it."""
        if self.grad is None:
            self.grad = torch.zeros_like(self.flat)
        if self.module is not None:
            params = dict(self.module.named_parameters())
            for s in self.specs:
                if s.is_buffer:
                    continue
                p = params[s.name]
                p.grad = self.grad[s.offset:s.offset + s.numel].view(s.shape)
                # ops.nn: a weight-gradient GEMM may accumulate straight into this slot (C = A B + C in the epilogue) instead
                # of handing a temporary to autograd's add kernel; post-accumulate-grad hooks are still run
                p._ddl_direct_grad = True
        return self.grad

    # ------------------------------------------------------------------ access
    def view_of(self, spec: TensorSpec, flat: torch.Tensor | None = None) -> torch.Tensor:
        flat = self.flat if flat is None else flat
        return flat[..., spec.offset:spec.offset + spec.numel].view(*flat.shape[:-1], *spec.shape)

    def views(self, flat: torch.Tensor | None = None, *, params_only: bool = False) -> Iterator[torch.Tensor]:
        for s in self.specs:
            if params_only and s.is_buffer:
                continue
            yield self.view_of(s, flat)

    def to_list(self, flat: torch.Tensor | None = None, *, clone: bool = True, device=None) -> list[torch.Tensor]:
        """The reference's message format: a list of per-parameter tensors (hfl_complete.py:332)."""
        out = []
        for v in self.views(flat, params_only=True):
            v = v.detach()
            if device is not None:
                v = v.to(device)
            out.append(v.clone() if clone else v)
        return out

    def load_list(self, tensors: Iterable[torch.Tensor], flat: torch.Tensor | None = None) -> None:
        with torch.no_grad():
            for v, t in zip(self.views(flat, params_only=True), tensors):
                v.copy_(t)

    def flatten_list(self, tensors: Iterable[torch.Tensor]) -> torch.Tensor:
        out = torch.zeros(self.size, device=self.flat.device, dtype=self.flat.dtype)
        self.load_list(tensors, out)
        return out

    # ------------------------------------------------------------------ state_dict interop
    def state_dict(self, flat: torch.Tensor | None = None) -> dict[str, torch.Tensor]:
        return {s.name: self.view_of(s, flat).detach().clone() for s in self.specs}

    def load_state_dict(self, sd: dict[str, torch.Tensor], flat: torch.Tensor | None = None, strict: bool = True) -> None:
        with torch.no_grad():
            for s in self.specs:
                if s.name not in sd:
                    if strict:
                        raise KeyError(f"missing key {s.name!r} in state_dict")
                    continue
                self.view_of(s, flat).copy_(sd[s.name])

    def new_stack(self, n: int) -> torch.Tensor:
        """``[n, size]`` buffer with the same layout per row – the packed-clients arena."""
        return torch.zeros(n, self.size, device=self.flat.device, dtype=self.flat.dtype)

    def __repr__(self) -> str:  # pragma: no cover
        return f"FlatArena({len(self.specs)} tensors, {self.n_param_elems} elems, padded {self.size}, {self.flat.device})"
