"""Fused optimizers over a flat parameter arena with a ``torch.optim``-like face.

``ArenaOptimizer(module, "adamw")`` moves the module's parameters into one contiguous fp32 arena (core/arena.py) with the
gradients pre-bound as views of a matching gradient arena, so ``zero_grad()`` is one memset and ``step()`` ONE kernel launch
over the whole model (K15: ``ddl_sgd_step`` / ``ddl_adam_step``) instead of torch.optim's per-tensor (or multi-tensor) loop.
Replaces the reference's ``torch.optim.SGD`` / ``Adam`` / ``AdamW`` objects (lab/tutorial_1a/hfl_complete.py:196,319;
lab/tutorial_2b/vfl.py:50; lab/tutorial_2a/generative-modeling.py:154; lab/tutorial_2b/exercise_3.py:191;
lab/tutorial_2a/centralized.py:46) with the same update rules and defaults.

:func:`make_optimizer` picks it for CUDA modules and the stock ``torch.optim`` class on CPU (the lab plumbing path).
"""
from __future__ import annotations

import torch
from torch import nn

from . import adam_step, sgd_step
from ..core.arena import FlatArena

_DEFAULT_WD = {"sgd": 0.0, "adam": 0.0, "adamw": 0.01}


class ArenaOptimizer:
    def __init__(self, module: nn.Module, kind: str = "adamw", lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float | None = None, momentum: float = 0.0, capturable: bool = False):
        assert kind in ("sgd", "adam", "adamw")
        # capturable: the Adam step counter lives in a device int (bumped by a one-element kernel), so zero_grad() + step() are
        # plain launches with no per-call host value and a whole training step can be captured into a CUDA graph
        self.capturable = capturable
        self.step_t = None
        self.module, self.kind, self.lr, self.betas, self.eps, self.momentum = module, kind, lr, betas, eps, momentum
        self.wd = _DEFAULT_WD[kind] if weight_decay is None else weight_decay
        self.arena: FlatArena | None = None
        self.m = self.v = None
        self.steps = 0

    def _bind(self) -> FlatArena:
        """(Re)bind lazily: ``module.to(device)`` after construction replaces ``param.data`` and would orphan an early arena."""
        a = self.arena
        first = next(iter(self.module.parameters()))
        if a is None or a.flat.device != first.device or first.data_ptr() != a.view_of(a.specs[0]).data_ptr():
            a = self.arena = FlatArena.from_module(self.module, with_grad=True)
            if self.kind != "sgd" or self.momentum:
                self.m = torch.zeros_like(a.flat)
            if self.kind != "sgd":
                self.v = torch.zeros_like(a.flat)
            if self.capturable and a.flat.is_cuda:
                self.step_t = torch.zeros((), dtype=torch.int32, device=a.flat.device)
        return a

    def _rebind_grads(self, a: FlatArena) -> None:
        params = dict(self.module.named_parameters())
        for sp in a.specs:
            p = params[sp.name]
            view = a.grad[sp.offset:sp.offset + sp.numel].view(sp.shape)
            if p.grad is None:
                view.zero_()
                p.grad = view
            elif p.grad.data_ptr() != view.data_ptr():
                view.copy_(p.grad)
                p.grad = view

    def zero_grad(self, set_to_none: bool = False) -> None:
        a = self._bind()
        a.grad.zero_()
        self._rebind_grads(a)

    @torch.no_grad()
    def step(self) -> None:
        a = self._bind()
        self._rebind_grads(a)
        self.steps += 1
        if self.kind == "sgd":
            sgd_step(a.flat, a.grad, self.lr, self.wd, self.momentum, self.m)
        else:
            step = self.steps
            if self.step_t is not None:
                self.step_t.add_(1)
                step = self.step_t
            adam_step(a.flat, a.grad, self.m, self.v, step, self.lr, self.betas[0], self.betas[1], self.eps, self.wd,
                      decoupled=self.kind == "adamw")

    def state_dict(self) -> dict:
        return {"kind": self.kind, "steps": self.steps, "m": self.m, "v": self.v}


def make_optimizer(module: nn.Module, kind: str = "adamw", **kw):
    """Fused arena optimizer for CUDA modules, the stock torch.optim class otherwise (same defaults either way)."""
    try:
        dev = next(module.parameters()).device
    except StopIteration:
        dev = torch.device("cpu")
    if dev.type == "cuda":
        return ArenaOptimizer(module, kind, **kw)
    cls = {"sgd": torch.optim.SGD, "adam": torch.optim.Adam, "adamw": torch.optim.AdamW}[kind]
    if kind == "sgd":
        kw.setdefault("lr", 1e-3)
    return cls(module.parameters(), **kw)
