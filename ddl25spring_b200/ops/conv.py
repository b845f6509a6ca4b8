"""Channels-last convolution stack on the tcgen05 tap-GEMM kernels (K9 general, K11/K12 NHWC forms).

Activations are NHWC fp32 ``[N, H, W, C]``; convolution weights are OHWI ``[Cout, kh, kw, Cin]``.  A convolution is
described to the kernel as a list of *taps* ``(map, dh, dw, weight tap)`` (csrc/conv_tcgen05.cu); this module derives
the tap lists for forward, input gradient and weight gradient of stride-1 / stride-2 convolutions and of stride-2
transposed convolutions, and wraps them in autograd Functions:

===============================  ===============================================================
``conv2d_nhwc``                   nn.Conv2d(k, stride 1|2, padding)            forward / dgrad / wgrad
``conv_transpose2d_nhwc``         nn.ConvTranspose2d(k, stride 1|2, padding)   (the conv's dgrad as a forward op)
``batch_norm_nhwc``               BatchNorm2d (+ ReLU, + residual add) fused
``max_pool3x3s2_nhwc``            F.max_pool2d(x, 3, 2, 1)
``global_avg_pool_nhwc``          adaptive_avg_pool2d(x, 1)
``stem_conv7x7s2``                the 7x7/2 stem of ResNet-18 on the NCHW network input (im2col + GEMM)
===============================  ===============================================================

On CPU tensors every function evaluates the same math with plain PyTorch ops (the oracle used by the tests).
Reference call sites replaced: nn.Conv2d at lab/tutorial_1a/hfl_complete.py:51,53 (general shapes) and the ResNet-18 /
DCGAN layers of the north-star configs.
"""
from __future__ import annotations

import ctypes as C
import functools

import torch
import torch.nn.functional as F

import contextlib

from . import _lib
from .gemm import gemm

ACT = {None: 0, "none": 0, "relu": 1, "leaky_relu": 2, "tanh": 3}
_WS: dict = {}
_SCOPE = {"id": 0, "concurrent": 0}      # a plain global, NOT thread-local: backward passes run on autograd's worker thread and must see the scope


@contextlib.contextmanager
def workspace_scope(scope_id: int):
    """Give the convolution / BatchNorm calls issued inside the block their OWN workspace (split-K scratch, side stream,
    BatchNorm partials).  Work that runs concurrently on different CUDA streams - e.g. the federated clients a server trains
    side by side (fl/server.py ``concurrent_clients``) - must not share scratch; a CUDA graph captured inside a scope keeps
    that scope's buffers."""
    prev = _SCOPE["id"]
    _SCOPE["id"] = int(scope_id)
    _SCOPE["concurrent"] += 1            # kernels that wait for sibling CTAs (cooperative split-K fix-up) are not used here:
    try:                                 # two such grids sharing the SMs could each hold what the other still needs
        yield
    finally:
        _SCOPE["id"] = prev
        _SCOPE["concurrent"] -= 1


def _workspace(device) -> dict:
    key = (device.type, device.index, _SCOPE["id"])
    ws = _WS.get(key)
    if ws is None:
        ws = {"scratch": torch.empty(16 << 20, device=device, dtype=torch.float32),      # split-K partial tiles (64 MB)
              "scratch_side": torch.empty(8 << 20, device=device, dtype=torch.float32),  # ... of the weight-gradient branch
              "side": torch.cuda.Stream(device=device),
              "bn_partial": torch.empty(1 << 19, device=device, dtype=torch.float32),
              "bn_counters": torch.zeros(64, device=device, dtype=torch.int32)}
        _WS[key] = ws
    return ws


def _ints(xs):
    return (C.c_int * len(xs))(*[int(v) for v in xs])


# ------------------------------------------------------------------------------------------------ tap lists
def fwd_taps(kh: int, kw: int, stride: int, pad: int):
    """Taps of the forward convolution.  stride 1: map 0 is x itself; stride 2: maps 0..3 are the phases 2a+b."""
    taps = []
    for r in range(kh):
        for s in range(kw):
            qh, qw = r - pad, s - pad
            if stride == 1:
                taps.append((0, qh, qw, r * kw + s))
            else:
                a, b = qh % 2, qw % 2
                taps.append((2 * a + b, (qh - a) // 2, (qw - b) // 2, r * kw + s))
    return taps


def dgrad_taps(kh: int, kw: int, stride: int, pad: int):
    """Per output phase (one entry for stride 1, four for stride 2): taps reading dY."""
    if stride == 1:
        return [[(0, pad - r, pad - s, r * kw + s) for r in range(kh) for s in range(kw)]]
    phases = []
    for a in range(2):
        for b in range(2):
            taps = [(0, (a + pad - r) // 2, (b + pad - s) // 2, r * kw + s)
                    for r in range(kh) for s in range(kw) if (a + pad - r) % 2 == 0 and (b + pad - s) % 2 == 0]
            phases.append(taps)
    return phases


@functools.lru_cache(maxsize=None)
def _tap_arrays(taps: tuple):
    """ctypes int arrays (map, dh, dw, weight tap) of a tap list - built once per distinct list (host time matters in eager mode)."""
    return tuple(_ints([t[i] for t in taps]) for i in range(4))


@functools.lru_cache(maxsize=None)
def _fwd_taps_c(kh, kw, stride, pad):
    return tuple(fwd_taps(kh, kw, stride, pad))


@functools.lru_cache(maxsize=None)
def _dgrad_taps_c(kh, kw, stride, pad):
    """(all taps concatenated, ctypes phase_begin or None) of the input gradient."""
    phases = dgrad_taps(kh, kw, stride, pad)
    if stride == 1:
        return tuple(phases[0]), None
    taps, begin = [], [0]
    for t in phases:
        taps += t
        begin.append(len(taps))
    return tuple(taps), _ints(begin)


def _tap_gemm(maps, N, Ha, Wa, Ca, w, w_mn, taps, out, Hh, Ww, o_off, o_sn, o_sh, o_sw, bias=None, alpha=1.0, beta=0.0, act=None,
              slope=0.2, phase_begin=None, phase_off=None):
    Cout, T, Cin = w.shape[0], w.shape[1] * w.shape[2], w.shape[3]
    ws = _workspace(out.device)
    ptrs = _lib.host_ptrs([m.data_ptr() for m in maps])
    n_phases = len(phase_off) if phase_off else 1
    a0, a1, a2, a3 = _tap_arrays(taps)
    _lib.check(_lib.lib().ddl_conv_tap_gemm(
        ptrs, len(maps), int(N), int(Ha), int(Wa), int(Ca), _lib.p(w), int(Cout), int(T), int(Cin), int(w_mn),
        a0, a1, a2, a3, len(taps),
        n_phases, phase_begin if phase_off else None, (C.c_int64 * n_phases)(*phase_off) if phase_off else None,
        _lib.p(out), int(Hh), int(Ww), _lib.i64(o_off), _lib.i64(o_sn), _lib.i64(o_sh), _lib.i64(o_sw), _lib.p(bias),
        _lib.f32(alpha), _lib.f32(beta), int(ACT[act]), _lib.f32(slope), _lib.p(ws["scratch"]), _lib.i64(ws["scratch"].numel()),
        _lib.stream()), "ddl_conv_tap_gemm")


def _phase_split(x):
    N, H, W, Cc = x.shape
    ph = torch.empty((4, N, H // 2, W // 2, Cc), device=x.device, dtype=torch.float32)
    _lib.check(_lib.lib().ddl_phase_split(_lib.p(x), _lib.p(ph), int(N), int(H), int(W), int(Cc), _lib.stream()), "ddl_phase_split")
    return ph


def _out_size(H, k, stride, pad):
    return (H + 2 * pad - k) // stride + 1


def _fprop(x, ph, w, stride, pad, bias=None, act=None, slope=0.2):
    """x [N,H,W,Cin] (stride 1) or its phases ph [4,N,H/2,W/2,Cin] (stride 2) -> y [N,Ho,Wo,Cout]."""
    Cout, kh, kw, Cin = w.shape
    if stride == 1:
        N, H, W, _ = x.shape
        maps, Ha, Wa = [x], H, W
    else:
        N, Ha, Wa = ph.shape[1], ph.shape[2], ph.shape[3]
        H, W = 2 * Ha, 2 * Wa
        maps = [ph[i] for i in range(4)]
    Ho, Wo = _out_size(H, kh, stride, pad), _out_size(W, kw, stride, pad)
    y = torch.empty((N, Ho, Wo, Cout), device=w.device, dtype=torch.float32)
    _tap_gemm(maps, N, Ha, Wa, Cin, w, 0, _fwd_taps_c(kh, kw, stride, pad), y, Ho, Wo, 0, Ho * Wo * Cout, Wo * Cout, Cout, bias=bias,
              act=act, slope=slope)
    return y


def _dgrad(dy, w, stride, pad, H, W, bias=None, act=None, slope=0.2):
    """dy [N,Ho,Wo,Cout] -> dx [N,H,W,Cin] (also the forward of a transposed convolution, with bias / activation)."""
    Cout, kh, kw, Cin = w.shape
    N, Ho, Wo, _ = dy.shape
    taps, begin = _dgrad_taps_c(kh, kw, stride, pad)
    dx = torch.empty((N, H, W, Cin), device=dy.device, dtype=torch.float32)
    if stride == 1:
        _tap_gemm([dy], N, Ho, Wo, Cout, w, 1, taps, dx, H, W, 0, H * W * Cin, W * Cin, Cin, bias=bias, act=act, slope=slope)
        return dx
    assert H % 2 == 0 and W % 2 == 0
    # the four output phases (h%2, w%2) in ONE launch (grid.z = phase); a phase without taps is written as act(bias)
    offs = [((idx >> 1) * W + (idx & 1)) * Cin for idx in range(4)]
    _tap_gemm([dy], N, Ho, Wo, Cout, w, 1, taps, dx, H // 2, W // 2, 0, H * W * Cin, 2 * W * Cin, 2 * Cin, bias=bias, act=act, slope=slope,
              phase_begin=begin, phase_off=offs)
    return dx


def _wgrad(dy, x, ph, wshape, stride, pad, out=None, scratch="scratch"):
    Cout, kh, kw, Cin = wshape
    N, Ho, Wo, _ = dy.shape
    if stride == 1:
        maps, Hx, Wx = [x], x.shape[1], x.shape[2]
    else:
        maps, Hx, Wx = [ph[i] for i in range(4)], ph.shape[2], ph.shape[3]
    taps = _fwd_taps_c(kh, kw, stride, pad)
    a0, a1, a2, a3 = _tap_arrays(taps)
    dw = out if out is not None else torch.empty(wshape, device=dy.device, dtype=torch.float32)
    ws = _workspace(dy.device)
    _lib.check(_lib.lib().ddl_conv_wgrad(
        _lib.p(dy), int(N), int(Ho), int(Wo), int(Cout), _lib.host_ptrs([m.data_ptr() for m in maps]), len(maps), int(Hx), int(Wx),
        int(Cin), a0, a1, a2, a3,
        len(taps), _lib.p(dw), int(kh * kw), _lib.f32(1.0), _lib.f32(0.0), _lib.p(ws[scratch]), _lib.i64(ws[scratch].numel()),
        _lib.stream()), "ddl_conv_wgrad")
    return dw


def _dgrad_and_wgrad(gy, w, x, ph, stride, pad, H, W):
    """Input and weight gradient of one convolution on two streams: they only share inputs, so the weight-gradient
    kernels run on a side stream (a parallel branch when the step is captured into a CUDA graph) and leave the
    backward critical path; both streams are joined before returning."""
    ws = _workspace(gy.device)
    main, side = torch.cuda.current_stream(), ws["side"]
    gw = torch.empty(tuple(w.shape), device=gy.device, dtype=torch.float32)
    side.wait_stream(main)
    with torch.cuda.stream(side):
        _wgrad(gy, x, ph, tuple(w.shape), stride, pad, out=gw, scratch="scratch_side")
    gx = _dgrad(gy, w, stride, pad, H, W)
    main.wait_stream(side)
    return gx, gw


# ------------------------------------------------------------------------------------------------ conv2d
def _torch_conv_nhwc(x, w, stride, pad, bias=None):
    y = F.conv2d(x.permute(0, 3, 1, 2), w.permute(0, 3, 1, 2), bias, stride, pad)
    return y.permute(0, 2, 3, 1).contiguous()


class _ConvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, stride, pad, ph=None, act=None, slope=0.2):
        x = x.contiguous()
        if stride == 2 and ph is None:
            ph = _phase_split(x)
        y = _fprop(x, ph, w, stride, pad, act=act, slope=slope)
        ctx.save_for_backward(x if stride == 1 else ph, w, y if act is not None else None)
        ctx.cfg = (stride, pad, x.shape[1], x.shape[2], act, slope)
        return y

    @staticmethod
    def backward(ctx, gy):
        saved, w, y = ctx.saved_tensors
        stride, pad, H, W, act, slope = ctx.cfg
        gy = _act_backward(gy.contiguous(), y, act, slope)
        gx = gw = None
        xs, phs = (saved, None) if stride == 1 else (None, saved)
        if ctx.needs_input_grad[0] and ctx.needs_input_grad[1]:
            gx, gw = _dgrad_and_wgrad(gy, w, xs, phs, stride, pad, H, W)
        elif ctx.needs_input_grad[0]:
            gx = _dgrad(gy, w, stride, pad, H, W)
        elif ctx.needs_input_grad[1]:
            gw = _wgrad(gy, xs, phs, tuple(w.shape), stride, pad)
        return gx, gw, None, None, None, None, None


def _act_backward(gy, y, act, slope):
    """Gradient through the epilogue activation, from the saved OUTPUT y."""
    if act == "relu":
        return gy * (y > 0)
    if act == "leaky_relu":
        return torch.where(y > 0, gy, gy * slope)
    if act == "tanh":
        return gy * (1 - y * y)
    return gy


def _kernel_ok(x, w, stride):
    cin, cout = w.shape[3], w.shape[0]
    ok = lambda c: c % 4 == 0 and (c < 32 or c % 32 == 0)
    return (x.is_cuda and x.dtype == torch.float32 and w.is_contiguous() and ok(cin) and ok(cout) and stride in (1, 2)
            and w.shape[1] * w.shape[2] <= 16 and (stride == 1 or (x.shape[1] % 2 == 0 and x.shape[2] % 2 == 0)))


def conv2d_nhwc(x: torch.Tensor, w: torch.Tensor, stride: int = 1, padding: int = 0, phases: torch.Tensor | None = None,
                act: str | None = None, slope: float = 0.2) -> torch.Tensor:
    """``x`` [N,H,W,Cin], ``w`` [Cout,kh,kw,Cin] -> [N,Ho,Wo,Cout] (TF32 tensor-core math, fp32 accumulate).

    ``phases``: the result of ``stride2_phases(x)`` when several stride-2 convolutions read the same input."""
    if not _kernel_ok(x, w, stride):
        y = _torch_conv_nhwc(x, w, stride, padding)
        return {"relu": F.relu, "leaky_relu": lambda t: F.leaky_relu(t, slope), "tanh": torch.tanh, None: lambda t: t}[act](y)
    return _ConvFn.apply(x, w, stride, padding, phases, act, slope)


def stride2_phases(x: torch.Tensor) -> torch.Tensor | None:
    """Even/odd sub-images ``[4, N, H/2, W/2, C]`` of ``x`` (input layout of the stride-2 kernels); None off the kernel path."""
    if not (x.is_cuda and x.dtype == torch.float32 and x.shape[1] % 2 == 0 and x.shape[2] % 2 == 0 and x.shape[3] % 4 == 0):
        return None
    with torch.no_grad():
        return _phase_split(x.contiguous())


# ------------------------------------------------------------------------------------------------ conv_transpose2d
class _ConvTFn(torch.autograd.Function):
    """y = conv_transpose(x, w): x [N,Hi,Wi,Ci], w stored as the OHWI weight of the conv it transposes: [Ci, kh, kw, Co]."""

    @staticmethod
    def forward(ctx, x, w, bias, stride, pad, act, slope):
        x = x.contiguous()
        Ci, kh, kw, Co = w.shape
        N, Hi, Wi, _ = x.shape
        H, W = (Hi - 1) * stride - 2 * pad + kh, (Wi - 1) * stride - 2 * pad + kw
        y = _dgrad(x, w, stride, pad, H, W, bias=bias, act=act, slope=slope)
        ctx.save_for_backward(x, w, y if act is not None else None)
        ctx.cfg = (stride, pad, act, slope, bias is not None)
        return y

    @staticmethod
    def backward(ctx, gy):
        x, w, y = ctx.saved_tensors
        stride, pad, act, slope, has_bias = ctx.cfg
        gy = _act_backward(gy.contiguous(), y, act, slope)
        ph = _phase_split(gy) if stride == 2 else None
        gx = gw = gb = None
        if ctx.needs_input_grad[0]:
            gx = _fprop(gy, ph, w, stride, pad)
        if ctx.needs_input_grad[1]:
            gw = _wgrad(x, gy if stride == 1 else None, ph, tuple(w.shape), stride, pad)
        if has_bias and ctx.needs_input_grad[2]:
            gb = gy.sum((0, 1, 2))
        return gx, gw, gb, None, None, None, None


def conv_transpose2d_nhwc(x, w, bias=None, stride: int = 2, padding: int = 1, act: str | None = None, slope: float = 0.2):
    """``x`` [N,Hi,Wi,Ci], ``w`` [Ci,kh,kw,Co] -> act(conv_transpose2d(x) + bias) [N,H,W,Co]."""
    if not _kernel_ok(x, w, 1) or stride not in (1, 2):
        wt = w.permute(0, 3, 1, 2)                                            # torch layout [Ci, Co, kh, kw]
        y = F.conv_transpose2d(x.permute(0, 3, 1, 2), wt, bias, stride, padding).permute(0, 2, 3, 1).contiguous()
        if act == "relu":
            y = F.relu(y)
        elif act == "leaky_relu":
            y = F.leaky_relu(y, slope)
        elif act == "tanh":
            y = torch.tanh(y)
        return y
    return _ConvTFn.apply(x, w, bias, stride, padding, act, slope)


# ------------------------------------------------------------------------------------------------ BatchNorm (+ReLU +residual)
class _BnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, res, run_mean, run_var, n_tracked, training, momentum, eps, relu, slope=0.0):
        x = x.contiguous()
        Cc = x.shape[-1]
        R = x.numel() // Cc
        ws = _workspace(x.device)
        L = _lib.lib()
        if training:
            mean = torch.empty(Cc, device=x.device, dtype=torch.float32)
            rstd = torch.empty_like(mean)
            assert L.ddl_bn_nhwc_ws_floats(_lib.i64(R), int(Cc)) <= ws["bn_partial"].numel()
            _lib.check(L.ddl_bn_nhwc_stats(_lib.p(x), _lib.i64(R), int(Cc), _lib.p(ws["bn_partial"]), _lib.p(ws["bn_counters"]), _lib.p(mean),
                                           _lib.p(rstd), _lib.p(run_mean), _lib.p(run_var), _lib.p(n_tracked), _lib.f32(eps), _lib.f32(momentum),
                                           _lib.stream()), "ddl_bn_nhwc_stats")
        else:
            mean, rstd = run_mean, torch.rsqrt(run_var + eps)
        y = torch.empty_like(x)
        resc = res.contiguous() if res is not None else None
        _lib.check(L.ddl_bn_nhwc_apply(_lib.p(x), _lib.p(resc), _lib.p(y), _lib.p(mean), _lib.p(rstd), _lib.p(gamma), _lib.p(beta),
                                       _lib.i64(R), int(Cc), int(relu), _lib.f32(slope), _lib.stream()), "ddl_bn_nhwc_apply")
        ctx.save_for_backward(x, y, gamma, mean, rstd)
        ctx.cfg = (R, Cc, relu, res is not None, training, slope)
        return y

    @staticmethod
    def backward(ctx, gy):
        x, y, gamma, mean, rstd = ctx.saved_tensors
        R, Cc, relu, has_res, training, slope = ctx.cfg
        assert training, "backward through eval-mode BatchNorm is not on a kernel path"
        gy = gy.contiguous()
        ws = _workspace(x.device)
        dx = torch.empty_like(x)
        dres = torch.empty_like(x) if has_res else None
        dg, db = torch.empty_like(mean), torch.empty_like(mean)
        _lib.check(_lib.lib().ddl_bn_nhwc_bwd(_lib.p(x), _lib.p(y), _lib.p(gy), _lib.p(dx), _lib.p(dres), _lib.p(mean), _lib.p(rstd),
                                              _lib.p(gamma), _lib.p(dg), _lib.p(db), _lib.p(ws["bn_partial"]), _lib.p(ws["bn_counters"]),
                                              _lib.i64(R), int(Cc), int(relu), _lib.f32(slope), _lib.stream()), "ddl_bn_nhwc_bwd")
        return dx, dg, db, dres, None, None, None, None, None, None, None, None


def batch_norm_nhwc(x, bn: torch.nn.BatchNorm2d, relu: bool = False, residual: torch.Tensor | None = None, leaky: float | None = None):
    """``act(BatchNorm(x) + residual?)`` for channels-last ``x`` [..., C]; statistics over all leading dims.
    ``relu=True``: ReLU; ``leaky=s``: LeakyReLU(s)."""
    Cc = x.shape[-1]
    kernel = x.is_cuda and x.dtype == torch.float32 and (Cc % 128 == 0 or Cc in (4, 8, 16, 32, 64)) and bn.track_running_stats and bn.momentum is not None \
        and (bn.training or not (torch.is_grad_enabled() and x.requires_grad))
    if not kernel:
        Cc = x.shape[-1]
        y = F.batch_norm(x.reshape(-1, Cc), bn.running_mean, bn.running_var, bn.weight, bn.bias, bn.training, bn.momentum or 0.0,
                         bn.eps).view(x.shape)
        if residual is not None:
            y = y + residual
        if leaky is not None:
            return F.leaky_relu(y, leaky)
        return F.relu(y) if relu else y
    return _BnFn.apply(x, bn.weight, bn.bias, residual, bn.running_mean, bn.running_var, bn.num_batches_tracked, bn.training, bn.momentum,
                       bn.eps, 2 if leaky is not None else int(relu), float(leaky or 0.0))


# ------------------------------------------------------------------------------------------------ pools
class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = x.contiguous()
        N, H, W, Cc = x.shape
        Ho, Wo = (H + 1) // 2, (W + 1) // 2
        y = torch.empty((N, Ho, Wo, Cc), device=x.device, dtype=torch.float32)
        code = torch.empty((N, Ho, Wo, Cc), device=x.device, dtype=torch.uint8)
        _lib.check(_lib.lib().ddl_maxpool3x3s2_fwd(_lib.p(x), _lib.p(y), _lib.p(code), int(N), int(H), int(W), int(Cc), _lib.stream()),
                   "ddl_maxpool3x3s2_fwd")
        ctx.save_for_backward(code)
        ctx.shape = (N, H, W, Cc)
        return y

    @staticmethod
    def backward(ctx, gy):
        (code,) = ctx.saved_tensors
        N, H, W, Cc = ctx.shape
        dx = torch.empty(ctx.shape, device=gy.device, dtype=torch.float32)
        _lib.check(_lib.lib().ddl_maxpool3x3s2_bwd(_lib.p(gy.contiguous()), _lib.p(code), _lib.p(dx), int(N), int(H), int(W), int(Cc),
                                                   _lib.stream()), "ddl_maxpool3x3s2_bwd")
        return dx


def max_pool3x3s2_nhwc(x):
    if not x.is_cuda or x.dtype != torch.float32:
        return F.max_pool2d(x.permute(0, 3, 1, 2), 3, 2, 1).permute(0, 2, 3, 1).contiguous()
    return _MaxPoolFn.apply(x)


class _AvgPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = x.contiguous()
        N, H, W, Cc = x.shape
        y = torch.empty((N, Cc), device=x.device, dtype=torch.float32)
        _lib.check(_lib.lib().ddl_avgpool_nhwc_fwd(_lib.p(x), _lib.p(y), int(N), int(H * W), int(Cc), _lib.stream()), "ddl_avgpool_nhwc_fwd")
        ctx.shape = (N, H, W, Cc)
        return y

    @staticmethod
    def backward(ctx, gy):
        N, H, W, Cc = ctx.shape
        dx = torch.empty(ctx.shape, device=gy.device, dtype=torch.float32)
        _lib.check(_lib.lib().ddl_avgpool_nhwc_bwd(_lib.p(gy.contiguous()), _lib.p(dx), int(N), int(H * W), int(Cc), _lib.stream()),
                   "ddl_avgpool_nhwc_bwd")
        return dx


def global_avg_pool_nhwc(x):
    if not x.is_cuda or x.dtype != torch.float32:
        return x.mean((1, 2))
    return _AvgPoolFn.apply(x)


# ------------------------------------------------------------------------------------------------ 7x7/2 stem
def _im2col_stem_torch(x):
    """[N, Cin<=4, H, W] -> [N*Ho*Wo, 196] with column order (r, s, ci4)."""
    N, Cin, H, W = x.shape
    xp = F.pad(x, (0, 0, 0, 0, 0, 4 - Cin))
    cols = F.unfold(xp, 7, 1, 3, 2)                                    # [N, 4*49, L], rows ordered (ci, r, s)
    L = cols.shape[-1]
    return cols.view(N, 4, 49, L).permute(0, 3, 2, 1).reshape(N * L, 196)


class _StemFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w):
        x = x.contiguous()
        N, Cin, H, W = x.shape
        Ho, Wo = (H + 1) // 2, (W + 1) // 2
        col = torch.empty((N * Ho * Wo, 224), device=x.device, dtype=torch.float32)     # 196 patch values + 28 zeros per row
        _lib.check(_lib.lib().ddl_im2col_stem7(_lib.p(x), _lib.p(col), int(N), int(Cin), int(H), int(W), _lib.stream()), "ddl_im2col_stem7")
        w2 = w.view(w.shape[0], 196)
        y = gemm(col[:, :196], w2)
        ctx.save_for_backward(col)
        ctx.wshape = tuple(w.shape)
        return y.view(N, Ho, Wo, w.shape[0])

    @staticmethod
    def backward(ctx, gy):
        (col,) = ctx.saved_tensors
        Cout = ctx.wshape[0]
        g2 = gy.reshape(-1, Cout).contiguous()
        # dW [Cout, 224] = dY^T col, both operands MN-major, K = pixels (split-K across the SMs)
        gw = gemm(g2, col, a_mn=True, b_mn=True, n_split=max(1, min(64, g2.shape[0] // 1024)))
        return None, gw[:, :196].contiguous().view(ctx.wshape)


def stem_conv7x7s2(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """``x`` NCHW [N, Cin<=4, H, W], ``w`` [Cout, 7, 7, 4] -> NHWC [N, Ho, Wo, Cout] (conv 7x7, stride 2, padding 3, no bias)."""
    if not x.is_cuda or x.dtype != torch.float32:
        N, _, H, W = x.shape
        col = _im2col_stem_torch(x)
        return (col @ w.reshape(w.shape[0], 196).t()).view(N, (H + 1) // 2, (W + 1) // 2, w.shape[0])
    return _StemFn.apply(x, w)


# ------------------------------------------------------------------------------------------------ host oracle of the tap algebra
def _shift2d(t, dh, dw, Hh, Ww):
    """out[n, i, j, :] = t[n, i+dh, j+dw, :] (zero outside t) for i < Hh, j < Ww."""
    N, Ha, Wa, Cc = t.shape
    out = t.new_zeros((N, Hh, Ww, Cc))
    i0, i1 = max(0, -dh), min(Hh, Ha - dh)
    j0, j1 = max(0, -dw), min(Ww, Wa - dw)
    if i1 > i0 and j1 > j0:
        out[:, i0:i1, j0:j1] = t[:, i0 + dh:i1 + dh, j0 + dw:j1 + dw]
    return out


def tap_conv_reference(x, w, stride, pad):
    """Forward convolution evaluated literally as the kernel's tap list (plain PyTorch; used by the CPU tests)."""
    Cout, kh, kw, Cin = w.shape
    N, H, W, _ = x.shape
    Ho, Wo = _out_size(H, kh, stride, pad), _out_size(W, kw, stride, pad)
    maps = [x] if stride == 1 else [x[:, a::2, b::2] for a in range(2) for b in range(2)]
    y = x.new_zeros((N, Ho, Wo, Cout))
    wt = w.reshape(Cout, kh * kw, Cin)
    for m, dh, dw, t in fwd_taps(kh, kw, stride, pad):
        y += _shift2d(maps[m], dh, dw, Ho, Wo) @ wt[:, t].t()
    return y


def tap_dgrad_reference(dy, w, stride, pad, H, W):
    Cout, kh, kw, Cin = w.shape
    N = dy.shape[0]
    wt = w.reshape(Cout, kh * kw, Cin)
    dx = dy.new_zeros((N, H, W, Cin))
    for idx, taps in enumerate(dgrad_taps(kh, kw, stride, pad)):
        a, b = (idx >> 1, idx & 1) if stride == 2 else (0, 0)
        Hh, Ww = (H // 2, W // 2) if stride == 2 else (H, W)
        acc = dy.new_zeros((N, Hh, Ww, Cin))
        for _, dh, dw, t in taps:
            acc += _shift2d(dy, dh, dw, Hh, Ww) @ wt[:, t]
        if stride == 2:
            dx[:, a::2, b::2] = acc
        else:
            dx = acc
    return dx
