"""CUDA-graph capture of a whole training (or inference) step.

A ResNet-18 step at the lab's batch sizes is ~260 kernel launches of a few microseconds each: issued one by one from
Python the GPU idles two thirds of the time (measured on B200: 6.0 ms wall per step for 2.0 ms of kernel time).  The
kernels of this package take all their arguments by value — tensor maps included — so a step is a fixed launch sequence
over fixed addresses and can be recorded once and replayed with a single ``cudaGraphLaunch``.

``GraphedStep(fn, example_inputs)`` warms ``fn`` up on a side stream, captures one call (forward, backward and the
optimizer step if ``fn`` does them) into a ``torch.cuda.CUDAGraph`` with static input / output tensors, and replays it
on every call.  On CPU (or with ``enabled=False``) it just calls ``fn``.
"""
from __future__ import annotations

import gc
from typing import Callable, Sequence

import torch


class GraphedStep:
    def __init__(self, fn: Callable, example_inputs: Sequence[torch.Tensor], warmup: int = 3, enabled: bool = True):
        self.fn = fn
        self.enabled = bool(enabled) and all(t.is_cuda for t in example_inputs) and torch.cuda.is_available()
        self.graph = None
        self.static_in = None
        self.static_out = None
        if not self.enabled:
            return
        self.static_in = [t.clone() for t in example_inputs]
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):          # lazy initialisation (workspaces, cuDNN plans, smem attributes) happens here
                fn(*self.static_in)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        # A dead reference cycle that still owns an earlier CUDAGraph (a previous model's step) must not be collected WHILE
        # this capture runs: destroying a graph is not a capturable call and invalidates the capture.  torch.cuda.graph no
        # longer collects on entry, so collect now and keep the cyclic collector off until the capture has ended.
        gc.collect()
        gc_was_on = gc.isenabled()
        gc.disable()
        try:
            with torch.cuda.graph(self.graph):
                self.static_out = fn(*self.static_in)
        finally:
            if gc_was_on:
                gc.enable()

    def __call__(self, *inputs):
        if not self.enabled:
            return self.fn(*inputs)
        for dst, src in zip(self.static_in, inputs):
            if dst.data_ptr() != src.data_ptr():
                dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.static_out


def graphed_train_step(model: torch.nn.Module, loss_fn: Callable, optimizer, example_x: torch.Tensor, example_y: torch.Tensor,
                       enabled: bool = True) -> GraphedStep:
    """One optimisation step ``loss = loss_fn(model(x), y); loss.backward(); optimizer.step()`` as a replayable graph.

    ``optimizer`` may be ``None`` (gradients only), a ``torch.optim`` optimizer constructed with ``capturable=True`` where
    it has that flag, or any object with ``zero_grad`` / ``step`` whose step is plain kernel launches
    (``ddl25spring_b200.fl.client.FlatSGD``-style arena optimizers)."""

    def step(x, y):
        if optimizer is not None:
            optimizer.zero_grad(set_to_none=True)
        else:
            for p in model.parameters():
                p.grad = None
        loss = loss_fn(model(x), y)
        loss.backward()
        if optimizer is not None:
            optimizer.step()
        return loss.detach()

    return GraphedStep(step, (example_x, example_y), enabled=enabled)
