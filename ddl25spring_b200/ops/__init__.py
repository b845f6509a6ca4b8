"""Operator front-end.

Each op takes torch tensors.  CUDA tensors run the hand-written sm_100a kernel from
``libddl25_kernels.so`` (no fallback: a missing library raises); CPU tensors run the plain
PyTorch statement of the same math in :mod:`.reference`, which is also the numerics oracle
for the GPU tests.  This is device placement, not a multi-backend layer: there is exactly
one GPU implementation and it targets B200 only.
"""
from __future__ import annotations

import torch

from . import reference as ref
from . import _lib
from ._lib import KernelError, launch_count, reset_launch_count  # noqa: F401


def _rows(updates: torch.Tensor):
    assert updates.dim() == 2 and updates.stride(1) == 1
    return updates.shape[0], updates.shape[1], updates.stride(0)


# --------------------------------------------------------------------------- aggregation
def weighted_sum(updates: torch.Tensor, weights, out: torch.Tensor | None = None, *, base: torch.Tensor | None = None,
                 alpha: float = 0.0, beta: float = 1.0) -> torch.Tensor:
    """``out = alpha*base + beta * sum_k weights[k] * updates[k]`` (K1; FedAvg / FedSGD server)."""
    K, D, ld = _rows(updates)
    if out is None:
        out = torch.empty(D, device=updates.device, dtype=updates.dtype)
    dev_w = None
    if torch.is_tensor(weights) and weights.is_cuda:
        dev_w, w = weights.contiguous().float(), None      # stays on the GPU: no host sync
    else:
        w = [float(x) for x in (weights.tolist() if torch.is_tensor(weights) else weights)]
    if not updates.is_cuda:
        res = ref.weighted_sum(updates, torch.tensor(w, dtype=updates.dtype)) * beta
        if base is not None:
            res = res + alpha * base
        out.copy_(res)
        return out
    _lib.req_f32(updates, "updates")
    L = _lib.lib()
    _lib.check(L.ddl_weighted_sum(_lib.p(updates), _lib.i64(ld), K, None if w is None else _lib.host_floats(w),
                                  _lib.p(dev_w), _lib.p(base), _lib.f32(alpha), _lib.f32(beta), _lib.p(out),
                                  _lib.i64(D), _lib.stream()),
               "ddl_weighted_sum")
    return out


def coordinate_median(updates: torch.Tensor, lower: bool = False, out: torch.Tensor | None = None) -> torch.Tensor:
    K, D, ld = _rows(updates)
    if not updates.is_cuda:
        res = ref.coordinate_median(updates, lower)
        return res if out is None else out.copy_(res)
    if out is None:
        out = torch.empty(D, device=updates.device, dtype=updates.dtype)
    L = _lib.lib()
    _lib.check(L.ddl_coordinate_median(_lib.p(_lib.req_f32(updates)), _lib.i64(ld), K, int(lower), _lib.p(out),
                                       _lib.i64(D), _lib.stream()), "ddl_coordinate_median")
    return out


def trimmed_mean(updates: torch.Tensor, trim: int, out: torch.Tensor | None = None) -> torch.Tensor:
    K, D, ld = _rows(updates)
    if not updates.is_cuda:
        res = ref.trimmed_mean(updates, trim)
        return res if out is None else out.copy_(res)
    if out is None:
        out = torch.empty(D, device=updates.device, dtype=updates.dtype)
    L = _lib.lib()
    _lib.check(L.ddl_trimmed_mean(_lib.p(_lib.req_f32(updates)), _lib.i64(ld), K, int(trim), _lib.p(out),
                                  _lib.i64(D), _lib.stream()), "ddl_trimmed_mean")
    return out


def gram(updates: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """``G = U U^T`` (K3).  CUDA: split-K tcgen05 (tf32) kernel; CPU: matmul."""
    if not updates.is_cuda:
        res = ref.gram(updates)
        return res if out is None else out.copy_(res)
    from .gemm import gram_tcgen05
    return gram_tcgen05(updates, out)


def krum_select(updates: torch.Tensor, n_byzantine: int, multi: int = 1, gram_matrix: torch.Tensor | None = None):
    """Return ``(selected_indices[multi], selection_weights[K], scores[K])`` (K3)."""
    K = updates.shape[0]
    if not updates.is_cuda:
        d = None
        if gram_matrix is not None:
            d = ref.pairwise_sq_dists_from_gram(gram_matrix.double())
        scores = ref.krum_scores(updates, n_byzantine, d)
        sel = torch.argsort(scores, stable=True)[:multi]
        w = torch.zeros(K, dtype=updates.dtype)
        w[sel] = 1.0 / multi
        return sel, w, scores.to(updates.dtype)
    G = gram(updates) if gram_matrix is None else gram_matrix
    scores = torch.empty(K, device=updates.device, dtype=torch.float32)
    sel = torch.empty(multi, device=updates.device, dtype=torch.int32)
    selw = torch.empty(K, device=updates.device, dtype=torch.float32)
    L = _lib.lib()
    _lib.check(L.ddl_krum_select(_lib.p(G), int(G.stride(0)), K, int(n_byzantine), int(multi), _lib.p(scores),
                                 _lib.p(sel), _lib.p(selw), _lib.stream()), "ddl_krum_select")
    return sel.long(), selw, scores


def attack_perturb(update: torch.Tensor, global_flat: torch.Tensor | None, scale: float = 1.0, noise: float = 0.0,
                   seed: int = 0, out: torch.Tensor | None = None) -> torch.Tensor:
    """Sign-flip / scaling attack ``g - scale*(u - g)`` (+ optional uniform noise) (K6)."""
    if out is None:
        out = torch.empty_like(update)
    if not update.is_cuda:
        g = torch.zeros_like(update) if global_flat is None else global_flat
        res = ref.sign_flip(update, g, scale)
        if noise:
            gen = torch.Generator().manual_seed(seed)
            res = res + noise * (2 * torch.rand(update.shape, generator=gen) - 1)
        return out.copy_(res)
    L = _lib.lib()
    _lib.check(L.ddl_attack_perturb(_lib.p(_lib.req_f32(update)), _lib.p(global_flat), _lib.p(out), _lib.f32(scale),
                                    _lib.f32(noise), int(seed) & 0xFFFFFFFF, _lib.i64(update.numel()), _lib.stream()),
               "ddl_attack_perturb")
    return out


# --------------------------------------------------------------------------- optimizers
def sgd_step(params: torch.Tensor, grads: torch.Tensor, lr: float, weight_decay: float = 0.0, momentum: float = 0.0,
             momentum_buf: torch.Tensor | None = None, grad_scale: float = 1.0) -> torch.Tensor:
    if not params.is_cuda:
        g = grads * grad_scale if grad_scale != 1.0 else grads
        return ref.sgd_step(params, g, lr, weight_decay, momentum, momentum_buf)
    L = _lib.lib()
    _lib.check(L.ddl_sgd_step(_lib.p(_lib.req_f32(params)), _lib.p(grads), _lib.p(momentum_buf), _lib.i64(params.numel()),
                              _lib.f32(lr), _lib.f32(weight_decay), _lib.f32(momentum), _lib.f32(grad_scale),
                              _lib.stream()), "ddl_sgd_step")
    return params


def adam_step(params, grads, exp_avg, exp_avg_sq, step: int, lr: float = 1e-3, beta1: float = 0.9,
              beta2: float = 0.999, eps: float = 1e-8, weight_decay: float = 0.0, decoupled: bool = False,
              grad_scale: float = 1.0):
    if not params.is_cuda:
        g = grads * grad_scale if grad_scale != 1.0 else grads
        return ref.adam_step(params, g, exp_avg, exp_avg_sq, int(step), lr, beta1, beta2, eps, weight_decay, decoupled)
    L = _lib.lib()
    if torch.is_tensor(step):          # device-resident int32 counter: the step can live inside a CUDA graph
        assert step.is_cuda and step.dtype == torch.int32
        _lib.check(L.ddl_adam_step_dev(_lib.p(_lib.req_f32(params)), _lib.p(grads), _lib.p(exp_avg), _lib.p(exp_avg_sq),
                                       _lib.i64(params.numel()), _lib.p(step), _lib.f32(lr), _lib.f32(beta1), _lib.f32(beta2),
                                       _lib.f32(eps), _lib.f32(weight_decay), int(decoupled), _lib.f32(grad_scale),
                                       _lib.stream()), "ddl_adam_step_dev")
        return params
    _lib.check(L.ddl_adam_step(_lib.p(_lib.req_f32(params)), _lib.p(grads), _lib.p(exp_avg), _lib.p(exp_avg_sq),
                               _lib.i64(params.numel()), int(step), _lib.f32(lr), _lib.f32(beta1), _lib.f32(beta2),
                               _lib.f32(eps), _lib.f32(weight_decay), int(decoupled), _lib.f32(grad_scale),
                               _lib.stream()), "ddl_adam_step")
    return params


# --------------------------------------------------------------------------- losses / eval
def softmax_xent(logits: torch.Tensor, target: torch.Tensor, *, reduction: str = "mean", want_grad: bool = True,
                 ignore_index: int = -100):
    """Cross-entropy value and d(loss)/d(logits) in one pass (K13).  ``target`` int64 [N] or float [N, C]."""
    N, Cc = logits.shape
    soft = target.is_floating_point()
    scale = 1.0 / N if reduction == "mean" else 1.0
    if not logits.is_cuda:
        loss, grad = (ref.softmax_ce_soft if soft else ref.logsoftmax_nll)(logits, target)
        if reduction != "mean":
            loss, grad = loss * N, grad * N
        return loss, (grad if want_grad else None)
    _lib.req_f32(logits, "logits")
    assert logits.stride(1) == 1
    loss = torch.zeros((), device=logits.device, dtype=torch.float32)
    grad = torch.empty_like(logits) if want_grad else None
    L = _lib.lib()
    _lib.check(L.ddl_softmax_xent(_lib.p(logits), _lib.i64(logits.stride(0)),
                                  _lib.p(None if soft else target), _lib.i64(1 if soft else target.stride(0)),
                                  _lib.p(target.contiguous() if soft else None), _lib.p(grad), _lib.p(loss), _lib.p(None),
                                  _lib.i64(N), int(Cc), _lib.f32(scale), _lib.f32(scale), _lib.i64(ignore_index),
                                  _lib.stream()), "ddl_softmax_xent")
    return loss, grad


def eval_top1(logits: torch.Tensor, target: torch.Tensor, count: torch.Tensor | None = None) -> torch.Tensor:
    """``count += #correct`` (K17); returns the int32 device counter."""
    if not logits.is_cuda:
        c = ref.eval_top1(logits, target).to(torch.int32)
        return c if count is None else count.add_(c)
    if count is None:
        count = torch.zeros((), device=logits.device, dtype=torch.int32)
    assert logits.stride(1) == 1
    target = target.contiguous()
    L = _lib.lib()
    _lib.check(L.ddl_eval_top1(_lib.p(_lib.req_f32(logits)), _lib.i64(logits.stride(0)), _lib.p(target),
                               _lib.i64(logits.shape[0]), int(logits.shape[1]), _lib.p(count), _lib.stream()),
               "ddl_eval_top1")
    return count


def vae_loss(x_recon, x, mu, logvar, want_grad: bool = False):
    """MSE(sum)+KL (K14); with ``want_grad`` also returns (d/dx_recon, d/dmu, d/dlogvar)."""
    if not x.is_cuda:
        loss = ref.vae_reparam_loss(x_recon, x, mu, logvar)
        if not want_grad:
            return loss
        return loss, (2 * (x_recon - x), mu.clone(), 0.5 * (logvar.exp() - 1))
    loss = torch.zeros((), device=x.device, dtype=torch.float32)
    gs = (torch.empty_like(x_recon), torch.empty_like(mu), torch.empty_like(logvar)) if want_grad else (None,) * 3
    L = _lib.lib()
    _lib.check(L.ddl_vae_loss(_lib.p(x_recon.contiguous()), _lib.p(x.contiguous()), _lib.i64(x.numel()),
                              _lib.p(mu.contiguous()), _lib.p(logvar.contiguous()), _lib.i64(mu.numel()), _lib.p(loss),
                              _lib.p(gs[0]), _lib.p(gs[1]), _lib.p(gs[2]), _lib.f32(1.0), _lib.stream()), "ddl_vae_loss")
    return (loss, gs) if want_grad else loss


def u8_normalize(u8: torch.Tensor, mean: float, std: float, out: torch.Tensor | None = None) -> torch.Tensor:
    """``(u8 / 255 - mean) / std`` as fp32, bit-identical on CPU and GPU (the input transform of hfl_complete.py:19-24)."""
    if out is None:
        out = torch.empty(u8.shape, dtype=torch.float32, device=u8.device)
    if not u8.is_cuda:
        out.copy_(u8.float().div_(255.0).sub_(mean).div_(std))
        return out
    assert u8.dtype == torch.uint8 and u8.is_contiguous() and out.is_contiguous() and u8.numel() % 4 == 0
    _lib.check(_lib.lib().ddl_u8_normalize(_lib.p(u8), _lib.p(out), _lib.i64(u8.numel()), _lib.f32(mean), _lib.f32(std), _lib.stream()),
               "ddl_u8_normalize")
    return out
