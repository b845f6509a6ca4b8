"""Autograd bridges from ``torch.nn`` modules to the hand-written kernels.

``KernelLinear`` / ``kernel_linear`` run forward, input-gradient and weight-gradient of a Linear layer on the tcgen05
GEMM (K10: one kernel, K-major and MN-major operands, no transposes); ``kernel_cross_entropy`` is the fused
softmax-xent (K13: value + gradient in one pass, incl. the shifted causal-LM form); ``kernel_rms_norm`` the RMSNorm
pair.  On CPU tensors every function falls back to the plain PyTorch statement of the same math (the oracle).
``use_kernels(model)`` swaps the ``nn.Linear`` modules of a model (e.g. the LLaMA of models/llama.py) in place.
"""
from __future__ import annotations

import os

import torch
import torch.nn.functional as F
from torch import nn

from . import _lib
from .gemm import gemm


def _ceil32(n: int) -> int:
    return (n + 31) // 32 * 32


def pad2d(src: torch.Tensor, rows_dst: int, cols_dst: int, fill_col: int = -1, fill_vec: torch.Tensor | None = None,
          fill_val: float = 0.0) -> torch.Tensor:
    """Zero-padded copy of a 2-D fp32 CUDA tensor (row-contiguous), optionally with one filled column (csrc/elementwise.cu)."""
    assert src.dim() == 2 and src.stride(1) == 1
    dst = torch.empty(rows_dst, cols_dst, device=src.device, dtype=torch.float32)
    _lib.check(_lib.lib().ddl_pad2d(_lib.p(src), _lib.i64(src.stride(0)), int(src.shape[0]), int(src.shape[1]), _lib.p(dst),
                                    _lib.i64(cols_dst), int(rows_dst), int(cols_dst), int(fill_col), _lib.p(fill_vec), _lib.f32(fill_val),
                                    _lib.stream()), "ddl_pad2d")
    return dst


def col_sum(x: torch.Tensor) -> torch.Tensor:
    """``x.sum(0)`` of a 2-D fp32 CUDA tensor (bias gradients), deterministic."""
    assert x.dim() == 2 and x.stride(1) == 1
    out = torch.empty(x.shape[1], device=x.device, dtype=torch.float32)
    _lib.check(_lib.lib().ddl_col_sum(_lib.p(x), _lib.i64(x.stride(0)), int(x.shape[0]), int(x.shape[1]), _lib.p(out), _lib.f32(1.0),
                                      _lib.f32(0.0), _lib.stream()), "ddl_col_sum")
    return out


# Input and weight gradient of a linear layer only share inputs: with CONCURRENT_WGRAD the weight-gradient GEMM runs on the
# side stream of the current workspace scope (a parallel branch inside a captured step), as ops/conv.py does for convolutions.
CONCURRENT_WGRAD = os.environ.get("DDL_CONCURRENT_WGRAD", "1") == "1"


def _fork_pays() -> bool:
    """A side-stream branch only helps where the GPU, not the host, sets the pace: while a step is being captured into a CUDA
    graph.  Issued eagerly, the fork / join costs more host time than the overlap returns (LLaMA step, eager: 8.2 -> 8.9 ms)."""
    if not CONCURRENT_WGRAD:
        return False
    from .conv import _workspace
    _workspace(torch.device("cuda", torch.cuda.current_device()))      # exists before any capture (warm-up steps come through here)
    return torch.cuda.is_current_stream_capturing()


def _side_and_main(side_fn, main_fn):
    from .conv import _workspace
    main, side = torch.cuda.current_stream(), _workspace(torch.device("cuda", torch.cuda.current_device()))["side"]
    side.wait_stream(main)
    with torch.cuda.stream(side):
        a = side_fn()
    b = main_fn()
    main.wait_stream(side)
    return a, b


def _direct_grad_slot(w: torch.Tensor):
    """The gradient slot of a parameter bound to a flat gradient arena (``core.arena.FlatArena.bind_grads`` marks them), when a
    backward kernel may accumulate straight into it: the weight-gradient GEMM then runs as ``C = A B + C`` on the slot - no
    temporary and no autograd add kernel per parameter per step (45 launches and two 37 MB read-modify-write passes of the
    LLaMA lab step).  The backward then returns no gradient for the parameter; autograd still runs the parameter's
    post-accumulate-grad hooks (DataParallel's bucket launches) - checked by tests/test_nn_bridge_gpu.py.
    ``None`` -> return the gradient to autograd as usual (no slot yet, or a double-backward graph)."""
    if not getattr(w, "_ddl_direct_grad", False) or torch.is_grad_enabled():
        return None
    g = w.grad
    if g is None or g.dtype != torch.float32 or not g.is_contiguous() or g.shape != w.shape or g.data_ptr() % 16:
        return None
    return g


class _LinearFn(torch.autograd.Function):
    """nn.Linear on the tcgen05 GEMM, forward + both gradients, for ANY layer width.

    Widths that are multiples of 32 (LLaMA, the ResNet head's input) use the operands as they are.  Ragged widths (the tabular
    models: 30 -> 64 -> ... -> 2, party slabs of 6 .. 26 columns) are handled by operand preparation, one pad2d launch each:
    X and W are zero-padded to multiples of 32 along the reduction, and the bias is FOLDED INTO THE GEMM - X gets a column of
    ones, W the bias vector in the same column - so the forward needs no bias epilogue and the weight-gradient GEMM
    ``dY^T [X | 1]`` delivers ``db`` in that column for free.  No ATen / cuBLAS GEMM on either path."""

    @staticmethod
    def forward(ctx, x, w, b, precise):
        x2 = x.reshape(-1, x.shape[-1])
        if x2.stride(1) != 1:
            x2 = x2.contiguous()
        N, K = w.shape
        M = x2.shape[0]
        ctx.has_bias, ctx.precise, ctx.shape = b is not None, precise, x.shape
        ctx.aligned = K % 32 == 0 and N % 32 == 0 and w.stride(1) == 1 and (w.stride(0) * 4) % 16 == 0 and (x2.stride(0) * 4) % 16 == 0
        if ctx.aligned:
            y = gemm(x2, w, bias=b, precise=precise)
            ctx.save_for_backward(x2, w)
        else:
            Kp, Np = _ceil32(K + (1 if b is not None else 0)), _ceil32(N)
            xp = pad2d(x2, M, Kp, K if b is not None else -1, None, 1.0)
            wp = pad2d(w if w.stride(1) == 1 else w.contiguous(), Np, Kp, K if b is not None else -1, b, 0.0)
            y = gemm(xp, wp[:N], precise=precise)                   # [M, N] contiguous; rows N.. of wp only serve the backward
            ctx.save_for_backward(xp, wp)
        ctx.dims = (M, N, K)
        return y.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, gy):
        M, N, K = ctx.dims
        g2 = gy.reshape(-1, N)
        gx = gw = gb = None
        if ctx.aligned:
            x2, w = ctx.saved_tensors
            g2 = g2.contiguous()
            def wgrad():
                slot = _direct_grad_slot(w)
                if slot is not None:
                    gemm(g2, x2, a_mn=True, b_mn=True, precise=ctx.precise, out=slot, beta=1.0)
                    return None
                return gemm(g2, x2, a_mn=True, b_mn=True, precise=ctx.precise)

            if ctx.needs_input_grad[0] and ctx.needs_input_grad[1] and _fork_pays():
                gw, gx = _side_and_main(wgrad, lambda: gemm(g2, w, b_mn=True, precise=ctx.precise).view(ctx.shape))
            else:
                if ctx.needs_input_grad[0]:
                    gx = gemm(g2, w, b_mn=True, precise=ctx.precise).view(ctx.shape)
                if ctx.needs_input_grad[1]:
                    gw = wgrad()
            if ctx.has_bias and ctx.needs_input_grad[2]:
                gb = col_sum(g2)
            return gx, gw, gb, None
        xp, wp = ctx.saved_tensors
        Np, Kp = wp.shape
        g2p = pad2d(g2 if g2.stride(1) == 1 else g2.contiguous(), M, Np)
        if ctx.needs_input_grad[0]:
            gx = gemm(g2p, wp, b_mn=True, precise=ctx.precise)[:, :K].reshape(ctx.shape)
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            dwp = gemm(g2p, xp, a_mn=True, b_mn=True, precise=ctx.precise)          # [Np, Kp]; column K holds dY^T 1 = db
            if ctx.needs_input_grad[1]:
                gw = dwp[:N, :K]
            if ctx.has_bias and ctx.needs_input_grad[2]:
                gb = dwp[:N, K]
        return gx, gw, gb, None


def kernel_linear(x, w, b=None, precise: bool = True):
    if not x.is_cuda or x.dtype != torch.float32 or w.dtype != torch.float32 or x.numel() == 0:
        return F.linear(x, w, b)
    return _LinearFn.apply(x, w, b, precise)


class KernelLinear(nn.Linear):
    precise = True

    def forward(self, x):
        return kernel_linear(x, self.weight, self.bias, self.precise)


class _XentFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target, ignore_index, n_valid):
        from . import softmax_xent
        if n_valid is None:
            n_valid = (target != ignore_index).sum().clamp_min(1)
        loss, grad = softmax_xent(logits, target, reduction="sum", want_grad=True, ignore_index=ignore_index)
        ctx.save_for_backward(grad)
        ctx.n_valid = n_valid
        return loss / n_valid

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * (g / ctx.n_valid), None, None, None


def kernel_cross_entropy(logits: torch.Tensor, target: torch.Tensor, ignore_index: int = -100, n_valid: int | None = None):
    """mean cross-entropy over the rows whose target != ignore_index (``F.cross_entropy`` semantics).  ``n_valid``: the number
    of such rows when the caller knows it (saves the device-side count)."""
    if not logits.is_cuda or logits.dtype != torch.float32:
        return F.cross_entropy(logits, target, ignore_index=ignore_index)
    return _XentFn.apply(logits.contiguous(), target.contiguous(), ignore_index, n_valid)


def kernel_classification_loss(logits: torch.Tensor, target: torch.Tensor):
    """Mean cross-entropy of a classifier batch in which every row counts (no ignore_index): the fused kernel alone - no
    device-side count of valid rows.  The ``loss_fn`` of the ResNet-18 DP / federated runs."""
    return kernel_cross_entropy(logits, target, n_valid=target.numel())


class _SoftXentFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        from . import softmax_xent
        loss, grad = softmax_xent(logits, target, reduction="mean", want_grad=True)
        ctx.save_for_backward(grad)
        return loss

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * g, None


def kernel_soft_cross_entropy(logits: torch.Tensor, target: torch.Tensor):
    """Mean cross-entropy against float (one-hot / soft) targets - ``nn.CrossEntropyLoss()(logits, probs)`` as used by
    lab/tutorial_2b/vfl.py:51,79 - value and gradient from one launch of the fused kernel."""
    if not logits.is_cuda or logits.dtype != torch.float32:
        return F.cross_entropy(logits, target)
    return _SoftXentFn.apply(logits.contiguous(), target.contiguous().float())


class _BceConstFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        x = logits.contiguous().view(-1)
        n = x.numel()
        loss = torch.zeros((), device=x.device, dtype=torch.float32)
        grad = torch.empty_like(x)
        _lib.check(_lib.lib().ddl_bce_logits_const(_lib.p(x), _lib.i64(n), _lib.f32(target), _lib.f32(1.0 / max(n, 1)), _lib.p(loss),
                                                   _lib.p(grad), _lib.stream()), "ddl_bce_logits_const")
        ctx.save_for_backward(grad)
        ctx.shape = logits.shape
        return loss

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return (grad * g).view(ctx.shape), None


def kernel_bce_logits(logits: torch.Tensor, target: float):
    """``F.binary_cross_entropy_with_logits(logits, full_like(logits, target))`` (mean) - value and gradient from one launch;
    the DCGAN discriminator / generator losses (real = 1.0, fake = 0.0)."""
    if not logits.is_cuda or logits.dtype != torch.float32:
        return F.binary_cross_entropy_with_logits(logits, torch.full_like(logits, float(target)))
    return _BceConstFn.apply(logits, float(target))


class KernelSoftCrossEntropy(nn.Module):
    """Drop-in for ``nn.CrossEntropyLoss()`` with float targets (module form of :func:`kernel_soft_cross_entropy`)."""

    def forward(self, logits, target):
        return kernel_soft_cross_entropy(logits, target)


def kernel_causal_lm_loss(logits: torch.Tensor, tokens: torch.Tensor):
    """Shifted next-token loss (lab/tutorial_1b/primer/intro.py:29) through the fused kernel.  The logits are used in place
    ([B*T, V], no sliced copy): row (b, t) is scored against token t + 1 and the last position of every sequence is ignored."""
    B, T, V = logits.shape
    if not logits.is_cuda:
        return kernel_cross_entropy(logits[:, :-1].reshape(-1, V), tokens.to(logits.device)[:, 1:].reshape(-1))
    tgt = torch.full((B, T), -100, dtype=torch.int64, device=logits.device)
    tgt[:, :-1] = tokens.to(logits.device)[:, 1:]
    return kernel_cross_entropy(logits.reshape(B * T, V), tgt.view(-1), n_valid=B * (T - 1))


class _RMSNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, eps):
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        y = torch.empty_like(x2)
        rstd = torch.empty(x2.shape[0], device=x.device, dtype=torch.float32)
        _lib.check(_lib.lib().ddl_rmsnorm_fwd(_lib.p(x2), _lib.p(w), _lib.p(y), _lib.p(rstd), _lib.i64(x2.shape[0]), int(x2.shape[1]),
                                              _lib.f32(eps), _lib.stream()), "ddl_rmsnorm_fwd")
        ctx.save_for_backward(x2, w, rstd)
        ctx.shape = x.shape
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, gy):
        x2, w, rstd = ctx.saved_tensors
        g2 = gy.reshape(-1, gy.shape[-1]).contiguous()
        dx, dw = torch.empty_like(x2), torch.zeros_like(w)
        _lib.check(_lib.lib().ddl_rmsnorm_bwd(_lib.p(x2), _lib.p(w), _lib.p(g2), _lib.p(rstd), _lib.p(dx), _lib.p(dw),
                                              _lib.i64(x2.shape[0]), int(x2.shape[1]), _lib.stream()), "ddl_rmsnorm_bwd")
        return dx.view(ctx.shape), dw, None


def kernel_rms_norm(x, w, eps: float = 1e-5):
    if not x.is_cuda or x.dtype != torch.float32:
        return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps) * w
    return _RMSNormFn.apply(x, w.contiguous(), eps)


class _BatchNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, run_mean, run_var, training, momentum, eps):
        xc = x.contiguous()
        N, Cc = xc.shape[0], xc.shape[1]
        S = xc.numel() // (N * Cc)
        y = torch.empty_like(xc)
        mean = torch.empty(Cc, device=x.device, dtype=torch.float32)
        rstd = torch.empty_like(mean)
        _lib.check(_lib.lib().ddl_batchnorm_fwd(_lib.p(xc), _lib.p(gamma), _lib.p(beta), _lib.p(y), _lib.p(mean), _lib.p(rstd),
                                                _lib.p(run_mean), _lib.p(run_var), int(N), int(Cc), int(S), _lib.f32(eps),
                                                _lib.f32(momentum), int(training), _lib.stream()), "ddl_batchnorm_fwd")
        ctx.save_for_backward(xc, gamma, mean, rstd)
        ctx.dims = (N, Cc, S)
        ctx.training = training
        return y

    @staticmethod
    def backward(ctx, gy):
        xc, gamma, mean, rstd = ctx.saved_tensors
        N, Cc, S = ctx.dims
        assert ctx.training, "backward through eval-mode BatchNorm is not on a kernel path"
        gyc = gy.contiguous()
        dx = torch.empty_like(xc)
        dg, db = torch.empty_like(mean), torch.empty_like(mean)
        _lib.check(_lib.lib().ddl_batchnorm_bwd(_lib.p(xc), _lib.p(gyc), _lib.p(gamma), _lib.p(mean), _lib.p(rstd), _lib.p(dx), _lib.p(dg),
                                                _lib.p(db), int(N), int(Cc), int(S), _lib.stream()), "ddl_batchnorm_bwd")
        return dx, dg, db, None, None, None, None, None


def kernel_batch_norm(x, bn: nn.modules.batchnorm._BatchNorm):
    """BatchNorm1d/2d forward (+ backward) on the K11 kernels; falls back to torch off-GPU or in eval-with-grad."""
    if not x.is_cuda or x.dtype != torch.float32 or bn.weight is None or not bn.track_running_stats or bn.momentum is None \
            or (not bn.training and torch.is_grad_enabled() and x.requires_grad):
        return F.batch_norm(x, bn.running_mean, bn.running_var, bn.weight, bn.bias, bn.training, bn.momentum or 0.0, bn.eps)
    if bn.training and bn.num_batches_tracked is not None:
        bn.num_batches_tracked.add_(1)
    return _BatchNormFn.apply(x, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.training, bn.momentum, bn.eps)


class _CausalAttnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, scale):
        B, H, T, D = q.shape
        qc, kc, vc = q.contiguous(), k.contiguous(), v.contiguous()
        o = torch.empty_like(qc)
        lse = torch.empty(B * H, T, device=q.device, dtype=torch.float32)
        _lib.check(_lib.lib().ddl_causal_attention_fwd(_lib.p(qc), _lib.p(kc), _lib.p(vc), _lib.p(o), _lib.p(lse), int(B * H), int(T),
                                                       int(D), _lib.f32(scale), _lib.stream()), "ddl_causal_attention_fwd")
        ctx.save_for_backward(qc, kc, vc, o, lse)
        ctx.scale = scale
        return o

    @staticmethod
    def backward(ctx, go):
        q, k, v, o, lse = ctx.saved_tensors
        B, H, T, D = q.shape
        goc = go.contiguous()
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        delta = torch.empty_like(lse)
        _lib.check(_lib.lib().ddl_causal_attention_bwd(_lib.p(q), _lib.p(k), _lib.p(v), _lib.p(o), _lib.p(goc), _lib.p(lse), _lib.p(dq),
                                                       _lib.p(dk), _lib.p(dv), _lib.p(delta), int(B * H), int(T), int(D),
                                                       _lib.f32(ctx.scale), _lib.stream()), "ddl_causal_attention_bwd")
        return dq, dk, dv, None


def _padded_heads(t: torch.Tensor, HP: int):
    """[B, H, T, hd] -> ([G, T, HP] zero-padded, contiguous).  Free when ``t`` is the [..., :hd] view of a padded buffer written
    by the fused QKV split + RoPE kernel; otherwise one pad2d launch."""
    B, H, T, hd = t.shape
    if hd == HP and t.is_contiguous():
        return t.view(B * H, T, HP)
    if t.stride() == (H * T * HP, T * HP, HP, 1):
        return t.as_strided((B * H, T, HP), (T * HP, HP, 1))
    tc = t.contiguous().view(B * H * T, hd)
    return pad2d(tc, B * H * T, HP).view(B * H, T, HP)


class _CausalAttnGemmFn(torch.autograd.Function):
    """Causal attention with every product on the tcgen05 GEMM (csrc/attention_tc.cu explains the kernel sequence)."""

    @staticmethod
    def forward(ctx, q, k, v, scale):
        B, H, T, hd = q.shape
        HP = _ceil32(hd)
        qp, kp, vp = _padded_heads(q, HP), _padded_heads(k, HP), _padded_heads(v, HP)
        L = _lib.lib()
        S = gemm(qp, kp, alpha=scale, precise=True)                                   # [G, T, T] = scale Q K^T
        _lib.check(L.ddl_causal_softmax(_lib.p(S), _lib.i64(B * H * T), int(T), _lib.stream()), "ddl_causal_softmax")
        op = gemm(S, vp, b_mn=True, precise=True)                                     # [G, T, HP] = P V
        ctx.save_for_backward(qp, kp, vp, S, op)
        ctx.dims, ctx.scale = (B, H, T, hd, HP), scale
        return op.view(B, H, T, HP)[..., :hd]

    @staticmethod
    def backward(ctx, go):
        qp, kp, vp, P, op = ctx.saved_tensors
        B, H, T, hd, HP = ctx.dims
        G = B * H
        L = _lib.lib()
        if go.stride(3) != 1:
            go = go.contiguous()
        dop = torch.empty(G, T, HP, device=go.device, dtype=torch.float32)
        delta = torch.empty(G * T, device=go.device, dtype=torch.float32)
        _lib.check(L.ddl_attn_bwd_prep(_lib.p(go), _lib.i64(go.stride(0)), _lib.i64(go.stride(1)), _lib.i64(go.stride(2)), _lib.p(op),
                                       _lib.p(dop), _lib.p(delta), int(B), int(H), int(T), int(hd), int(HP), _lib.stream()),
                   "ddl_attn_bwd_prep")
        if _fork_pays():
            # dV only needs P and dO, dK only dS and Q: both leave the critical path prep -> dP -> dS -> dQ and run on the side
            # stream of the workspace scope (parallel branches of a captured step); outputs are allocated before the fork
            from .conv import _workspace
            main, side = torch.cuda.current_stream(), _workspace(go.device)["side"]
            dv, dk = torch.empty_like(dop), torch.empty_like(dop)
            side.wait_stream(main)
            with torch.cuda.stream(side):
                gemm(P, dop, a_mn=True, b_mn=True, precise=True, out=dv)              # P^T dO
            dS = gemm(dop, vp, precise=True)                                          # dP = dO V^T
            _lib.check(L.ddl_attn_ds(_lib.p(P), _lib.p(dS), _lib.p(delta), _lib.i64(G * T), int(T), _lib.f32(ctx.scale), _lib.stream()),
                       "ddl_attn_ds")
            side.wait_stream(main)
            with torch.cuda.stream(side):
                gemm(dS, qp, a_mn=True, b_mn=True, precise=True, out=dk)              # dS^T Q
            dq = gemm(dS, kp, b_mn=True, precise=True)                                # dS K
            main.wait_stream(side)
        else:
            dS = gemm(dop, vp, precise=True)                                          # dP = dO V^T
            _lib.check(L.ddl_attn_ds(_lib.p(P), _lib.p(dS), _lib.p(delta), _lib.i64(G * T), int(T), _lib.f32(ctx.scale), _lib.stream()),
                       "ddl_attn_ds")
            dv = gemm(P, dop, a_mn=True, b_mn=True, precise=True)                     # P^T dO
            dq = gemm(dS, kp, b_mn=True, precise=True)                                # dS K
            dk = gemm(dS, qp, a_mn=True, b_mn=True, precise=True)                     # dS^T Q
        view = lambda t: t.view(B, H, T, HP)[..., :hd]
        return view(dq), view(dk), view(dv), None


def kernel_causal_attention(q, k, v):
    """softmax(q k^T / sqrt(D) + causal mask) v for [B, H, T, D] tensors (K16).  T a multiple of 32: grouped tcgen05 GEMMs +
    row-wise softmax kernels; other T (D in 16/32/48/64): the CUDA-core flash kernels."""
    D, T = q.shape[-1], q.shape[-2]
    if not q.is_cuda or q.dtype != torch.float32:
        return F.scaled_dot_product_attention(q, k, v, is_causal=True)
    if T % 32 == 0 and D % 2 == 0:
        return _CausalAttnGemmFn.apply(q, k, v, 1.0 / (D ** 0.5))
    if D not in (16, 32, 48, 64):
        return F.scaled_dot_product_attention(q, k, v, is_causal=True)
    return _CausalAttnFn.apply(q, k, v, 1.0 / (D ** 0.5))


def kernel_causal_attention_simt(q, k, v):
    """The CUDA-core flash kernels (attention.cu), kept selectable for comparisons."""
    D = q.shape[-1]
    return _CausalAttnFn.apply(q, k, v, 1.0 / (D ** 0.5))


class _RopeQkvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, cos, sin, H):
        B, T, C3 = qkv.shape
        hd = C3 // (3 * H)
        HP = _ceil32(hd)                    # q, k, v are written with the head dim zero-padded to a multiple of 32: legal
        qkv = qkv.contiguous()              # MN-major operands for the attention GEMMs; callers see [..., :hd] views
        q = torch.empty((B, H, T, HP), device=qkv.device, dtype=torch.float32)
        k, v = torch.empty_like(q), torch.empty_like(q)
        _lib.check(_lib.lib().ddl_rope_qkv(_lib.p(qkv), _lib.p(cos), _lib.p(sin), _lib.p(q), _lib.p(k), _lib.p(v), int(B), int(T), int(H),
                                           int(hd), int(HP), 0, _lib.stream()), "ddl_rope_qkv")
        ctx.save_for_backward(cos, sin)
        ctx.dims = (B, T, H, hd, HP)
        return q[..., :hd], k[..., :hd], v[..., :hd]

    @staticmethod
    def backward(ctx, gq, gk, gv):
        cos, sin = ctx.saved_tensors
        B, T, H, hd, HP = ctx.dims
        g = torch.empty((B, T, 3 * H * hd), device=gq.device, dtype=torch.float32)

        def pitched(t):                     # gradients arrive either as [..., :hd] views of HP-pitched buffers, or dense
            if t.stride() == (H * T * HP, T * HP, HP, 1):
                return t, HP
            return t.contiguous(), hd
        (gq, pq), (gk, pk), (gv, pv) = pitched(gq), pitched(gk), pitched(gv)
        if not (pq == pk == pv):
            gq, gk, gv, pq = gq.contiguous(), gk.contiguous(), gv.contiguous(), hd
        _lib.check(_lib.lib().ddl_rope_qkv(_lib.p(g), _lib.p(cos), _lib.p(sin), _lib.p(gq), _lib.p(gk), _lib.p(gv), int(B), int(T), int(H),
                                           int(hd), int(pq), 1, _lib.stream()), "ddl_rope_qkv")
        return g, None, None, None


def kernel_rope_qkv(qkv, cos, sin, n_heads: int):
    """``qkv`` [B, T, 3*H*hd] -> rotary-embedded q, k and v, each [B, H, T, hd] (one kernel each way); the returned tensors
    are views of buffers whose head dim is zero-padded to a multiple of 32."""
    return _RopeQkvFn.apply(qkv, cos.contiguous(), sin.contiguous(), n_heads)


class _SwiGluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, g, u):
        g, u = g.contiguous(), u.contiguous()
        h = torch.empty_like(g)
        _lib.check(_lib.lib().ddl_swiglu_fwd(_lib.p(g), _lib.p(u), _lib.p(h), _lib.i64(g.numel()), _lib.stream()), "ddl_swiglu_fwd")
        ctx.save_for_backward(g, u)
        return h

    @staticmethod
    def backward(ctx, dh):
        g, u = ctx.saved_tensors
        dg, du = torch.empty_like(g), torch.empty_like(u)
        _lib.check(_lib.lib().ddl_swiglu_bwd(_lib.p(g), _lib.p(u), _lib.p(dh.contiguous()), _lib.p(dg), _lib.p(du), _lib.i64(g.numel()),
                                             _lib.stream()), "ddl_swiglu_bwd")
        return dg, du


def kernel_swiglu(g, u):
    """``silu(g) * u`` (the gated-MLP activation) in one kernel each way."""
    if not g.is_cuda or g.dtype != torch.float32 or g.numel() % 4:
        return F.silu(g) * u
    return _SwiGluFn.apply(g, u)


def _stacked(wg: torch.Tensor, wu: torch.Tensor) -> bool:
    """True when ``wu`` starts where ``wg`` ends in the same storage, i.e. [wg; wu] is one contiguous [2F, D] matrix (adjacent
    views of a flat arena, or the pair laid out by :func:`stack_gate_up`)."""
    return (wg.shape == wu.shape and wg.is_contiguous() and wu.is_contiguous()
            and wg.untyped_storage().data_ptr() == wu.untyped_storage().data_ptr()
            and wu.storage_offset() == wg.storage_offset() + wg.numel())


def stack_gate_up(mlp) -> None:
    """Lay ``mlp.w_gate.weight`` and ``mlp.w_up.weight`` out back to back in one buffer (values kept, Parameters kept: only
    their ``.data`` moves).  core.arena.FlatArena packs parameters in registration order, which keeps the pair adjacent."""
    wg, wu = mlp.w_gate.weight, mlp.w_up.weight
    if _stacked(wg.data, wu.data) or wg.shape != wu.shape:
        return
    buf = torch.cat([wg.data.reshape(-1), wu.data.reshape(-1)])
    n = wg.numel()
    wg.data, wu.data = buf[:n].view_as(wg), buf[n:].view_as(wu)


class _GateUpSwiGluFn(torch.autograd.Function):
    """``silu(x Wg^T) * (x Wu^T)`` with the two projections as ONE GEMM against the stacked weight [Wg; Wu] (N = 2F), a SwiGLU
    kernel on the stacked [M, 2F] result, and a backward of ONE dgrad GEMM (K = 2F) + ONE wgrad GEMM instead of two each plus
    the add of the two input gradients.  Replaces (reference): simplellm's gated MLP inside every LLaMA block
    (lab/tutorial_1b/primer/intro.py:17 via ``LLama``)."""

    @staticmethod
    def forward(ctx, x, wg, wu, precise):
        Fh, D = wg.shape
        x2 = x.reshape(-1, D)
        if x2.stride(1) != 1 or x2.stride(0) != D:
            x2 = x2.contiguous()
        M = x2.shape[0]
        w2 = torch.as_strided(wg, (2 * Fh, D), (D, 1))
        gu = gemm(x2, w2, precise=precise)
        h = torch.empty(M, Fh, device=x.device, dtype=torch.float32)
        _lib.check(_lib.lib().ddl_swiglu_gu_fwd(_lib.p(gu), _lib.p(h), _lib.i64(M), int(Fh), _lib.stream()), "ddl_swiglu_gu_fwd")
        ctx.save_for_backward(x2, wg, gu)
        ctx.wu = wu                      # only its gradient slot is needed in the backward
        ctx.precise, ctx.shape = precise, x.shape
        return h.view(*x.shape[:-1], Fh)

    @staticmethod
    def backward(ctx, dh):
        x2, wg, gu = ctx.saved_tensors
        Fh, D = wg.shape
        M = x2.shape[0]
        dh2 = dh.reshape(M, Fh)
        if not dh2.is_contiguous():
            dh2 = dh2.contiguous()
        dgu = torch.empty_like(gu)
        _lib.check(_lib.lib().ddl_swiglu_gu_bwd(_lib.p(gu), _lib.p(dh2), _lib.p(dgu), _lib.i64(M), int(Fh), _lib.stream()), "ddl_swiglu_gu_bwd")
        w2 = torch.as_strided(wg, (2 * Fh, D), (D, 1))
        gx = gwg = gwu = None

        def wgrad():
            wu = ctx.wu
            sg, su = _direct_grad_slot(wg), _direct_grad_slot(wu)
            if sg is not None and su is not None and _stacked(sg, su):       # the two slots are back to back like the weights
                gemm(dgu, x2, a_mn=True, b_mn=True, precise=ctx.precise, out=torch.as_strided(sg, (2 * Fh, D), (D, 1)), beta=1.0)
                return None, None
            gw2 = gemm(dgu, x2, a_mn=True, b_mn=True, precise=ctx.precise)           # [2F, D]
            return gw2[:Fh], gw2[Fh:]

        need_w = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
        if ctx.needs_input_grad[0] and need_w and _fork_pays():
            (gwg, gwu), gx = _side_and_main(wgrad, lambda: gemm(dgu, w2, b_mn=True, precise=ctx.precise).view(ctx.shape))
        else:
            if ctx.needs_input_grad[0]:
                gx = gemm(dgu, w2, b_mn=True, precise=ctx.precise).view(ctx.shape)
            if need_w:
                gwg, gwu = wgrad()
        return gx, gwg, gwu, None


def kernel_gate_up_swiglu(x, wg, wu, precise: bool = True):
    """The gated MLP's first half.  Stacked weights (see :func:`stack_gate_up`) take the one-GEMM path; anything else the
    two-GEMM path through :func:`kernel_linear` and :func:`kernel_swiglu`."""
    Fh, D = wg.shape
    if (x.is_cuda and x.dtype == torch.float32 and wg.dtype == torch.float32 and x.numel() and D % 32 == 0 and Fh % 32 == 0
            and _stacked(wg, wu)):
        return _GateUpSwiGluFn.apply(x, wg, wu, precise)
    return kernel_swiglu(kernel_linear(x, wg, None, precise), kernel_linear(x, wu, None, precise))


class _EmbeddingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tokens, weight, padding_idx):
        tok = tokens.contiguous().view(-1)
        V, d = weight.shape
        out = torch.empty(tok.numel(), d, device=weight.device, dtype=torch.float32)
        _lib.check(_lib.lib().ddl_embedding_fwd(_lib.p(weight), _lib.p(tok), _lib.p(out), _lib.i64(tok.numel()), int(d), _lib.i64(V),
                                                _lib.stream()), "ddl_embedding_fwd")
        ctx.save_for_backward(tok, weight)
        ctx.shape, ctx.pad = (V, d), -1 if padding_idx is None else int(padding_idx)
        return out.view(*tokens.shape, d)

    @staticmethod
    def backward(ctx, g):
        tok, weight = ctx.saved_tensors
        V, d = ctx.shape
        g2 = g.reshape(-1, d)
        if g2.stride(1) != 1 or g2.stride(0) != d:
            g2 = g2.contiguous()
        slot = _direct_grad_slot(weight)         # the scatter-add goes straight into the arena slot: no V x d temporary
        dW = slot if slot is not None else torch.zeros(V, d, device=g.device, dtype=torch.float32)
        _lib.check(_lib.lib().ddl_embedding_bwd(_lib.p(g2), _lib.p(tok), _lib.p(dW), _lib.i64(tok.numel()), int(d), _lib.i64(V),
                                                _lib.i64(ctx.pad), _lib.stream()), "ddl_embedding_bwd")
        if slot is not None:
            return None, None, None
        return None, dW, None


def kernel_embedding(tokens: torch.Tensor, emb: nn.Embedding):
    """``nn.Embedding`` forward (row gather) and backward (scatter-add into the weight gradient) on kernels."""
    w = emb.weight
    if not w.is_cuda or w.dtype != torch.float32 or w.shape[1] % 4 or tokens.dtype != torch.int64 or emb.max_norm is not None \
            or not w.is_contiguous():
        return F.embedding(tokens, w, emb.padding_idx)
    return _EmbeddingFn.apply(tokens.to(w.device), w, emb.padding_idx)


def use_kernels(model: nn.Module, precise: bool = True) -> nn.Module:
    """Route every ``nn.Linear`` (and RMSNorm of models/llama.py) of ``model`` through the kernels, in place."""
    from ..models.llama import Attention, GatedMLP, RMSNorm
    for mod in model.modules():
        if type(mod) is nn.Linear:
            mod.__class__ = KernelLinear
            mod.precise = precise
        elif isinstance(mod, RMSNorm):
            mod.forward = (lambda m: (lambda x: kernel_rms_norm(x, m.weight, m.eps)))(mod)
        elif isinstance(mod, (nn.BatchNorm1d, nn.BatchNorm2d)):
            mod.forward = (lambda m: (lambda x: kernel_batch_norm(x, m)))(mod)
        elif type(mod) is nn.Embedding:
            mod.forward = (lambda m: (lambda tokens: kernel_embedding(tokens, m)))(mod)
        elif isinstance(mod, Attention):
            mod.attn_fn = kernel_causal_attention
            mod.qkv_fn = kernel_rope_qkv
        elif isinstance(mod, GatedMLP):
            mod.act_fn = kernel_swiglu
            if mod.w_gate.bias is None and mod.w_up.bias is None:
                stack_gate_up(mod)
                mod.gate_up_fn = (lambda pr: (lambda x, wg, wu: kernel_gate_up_swiglu(x, wg, wu, pr)))(precise)
    return model
