"""Packed MnistCnn engine: G simulated clients trained in lock-step on one GPU.

This is the B200-native replacement for the reference's sequential per-client loop
(lab/tutorial_1a/hfl_complete.py:365-375 calling ``WeightClient.update`` -> ``train_epoch``
:71-80): the clients dealt to this GPU hold their weights as rows of one ``[G, P]`` arena and
every step of local SGD is ONE sequence of grouped kernels over all of them (csrc/mnist_*.cu,
csrc/gemm_tcgen05.cu).  No autograd, no ATen compute, no per-client Python loop.

Per step (all clients at once):

=====================  =====================================================================
conv1 + ReLU            CUDA cores, gathers the minibatch through the client's permutation
conv2 + ReLU + pool     tcgen05 implicit GEMM (TF32), fused bias/ReLU/max-pool/dropout epilogue
  + dropout1
fc1 + bias + ReLU       tcgen05 grouped GEMM, 3xTF32 (fp32-grade), split-K with ordered reduce
head                    dropout2, fc2, log-softmax, NLL, backward, SGD on fc2 / fc1.bias
fc1 dgrad / wgrad       tcgen05 GEMMs on MN-major operands; wgrad epilogue applies SGD in place
pool/ReLU backward      scatter through the saved arg-max mask to the 26-pitch gradient grid
conv2 dgrad / wgrad     tcgen05 implicit GEMMs (shifted window), ordered partial reduction
conv1 wgrad             CUDA cores
small-tensor SGD        conv1 / conv2 weights + biases, re-packs conv2 weights for the next step
=====================  =====================================================================
"""
from __future__ import annotations

import ctypes as C

import torch

from ..ops import _lib
from ..ops.gemm import gemm

P = 1_200_000
OFF = {"conv1.weight": 0, "conv1.bias": 288, "conv2.weight": 320, "conv2.bias": 18752, "fc1.weight": 18816,
       "fc1.bias": 1198464, "fc2.weight": 1198592, "fc2.bias": 1199872}
PIX1, FLAT, FC1, NCLS = 676, 9216, 128, 10
MAX_G = 16


def _ints(xs):
    return (C.c_int * len(xs))(*[int(x) for x in xs])


class PackedMnistEngine:
    """Owns the activation / gradient buffers for ``G`` clients x ``B`` rows and runs grouped steps."""

    def __init__(self, G: int, B: int, device="cuda", p_drop1: float = 0.25, p_drop2: float = 0.5, train: bool = True):
        assert 1 <= G <= MAX_G
        self.G, self.B, self.device = G, B, torch.device(device)
        self.p1, self.p2 = p_drop1, p_drop2
        rows = G * B
        f32 = dict(device=self.device, dtype=torch.float32)
        self.a1 = torch.zeros(rows * PIX1, 32, **f32)
        self.pooled = torch.zeros(G, B, FLAT, **f32)
        self.h1 = torch.zeros(G, B, FC1, **f32)
        self.wt_fwd = torch.zeros(G * 9 * 64, 32, **f32)
        self.wt_dgrad = torch.zeros(G * 9 * 64, 32, **f32)
        self.n_split_fc1 = 18 if train else 9    # K = 9216: 512-long chains, 8 clients x 18 = 144 CTAs
        if train:
            self.pmask = torch.zeros(rows, FLAT, device=self.device, dtype=torch.uint8)
            self.dz1 = torch.zeros(G, B, FC1, **f32)
            self.dpool = torch.zeros(G, B, FLAT, **f32)
            self.dy2 = torch.zeros(rows * PIX1, 64, **f32)
            self.da1 = torch.zeros(rows * PIX1, 32, **f32)
            self.loss = torch.zeros(G, **f32)
        else:
            self.pmask = None
        self.L = _lib.lib()

    # ------------------------------------------------------------------ forward pieces
    def prep_weights(self, W: torch.Tensor) -> None:
        _lib.check(self.L.ddl_mnist_prep_conv2(_lib.p(W), _lib.p(self.wt_fwd), _lib.p(self.wt_dgrad), self.G,
                                               _lib.stream()), "ddl_mnist_prep_conv2")

    def forward_features(self, W, X, x_stride, perm, perm_stride, sample_off, bsz, train: bool, seed: int) -> None:
        """conv1 -> conv2/pool(/dropout1) -> fc1+ReLU; results in ``self.pooled`` / ``self.h1``."""
        L, G, B = self.L, self.G, self.B
        cb = _ints(bsz)
        _lib.check(L.ddl_mnist_conv1_fwd(_lib.p(W), _lib.p(X), _lib.i64(x_stride), _lib.p(perm), _lib.i64(perm_stride),
                                         int(sample_off), _lib.p(self.a1), G, B, cb, _lib.stream()), "conv1_fwd")
        _lib.check(L.ddl_mnist_conv2_pool_fwd(_lib.p(W), _lib.p(self.wt_fwd), _lib.p(self.a1), _lib.p(self.pooled),
                                              _lib.p(self.pmask), G, B, cb, int(train), _lib.f32(self.p1),
                                              int(seed) & 0xFFFFFFFF, _lib.stream()), "conv2_pool_fwd")
        w1 = W.view(G, P)[:, OFF["fc1.weight"]:OFF["fc1.weight"] + FC1 * FLAT].unflatten(1, (FC1, FLAT))
        b1 = W.view(G, P)[:, OFF["fc1.bias"]:OFF["fc1.bias"] + FC1]
        gemm(self.pooled, w1, bias=b1, act="relu", precise=True, n_split=self.n_split_fc1, out=self.h1)

    # ------------------------------------------------------------------ evaluation
    @torch.no_grad()
    def eval_logits(self, W: torch.Tensor, X: torch.Tensor, Y: torch.Tensor | None = None, correct: torch.Tensor | None = None,
                    want_logits: bool = True):
        """Forward ``X`` [n, 1, 28, 28] through the model ``W`` ([P] flat) in chunks of ``G*B`` rows
        (engine must have been built with G == 1)."""
        assert self.G == 1
        n = X.shape[0]
        X = X.reshape(n, 784)
        logits = torch.empty(n, NCLS, device=self.device) if want_logits else None
        self.prep_weights(W)
        for lo in range(0, n, self.B):
            m = min(self.B, n - lo)
            self.forward_features(W, X, 0, None, 0, lo, [m], False, 0)
            _lib.check(self.L.ddl_mnist_head_eval(_lib.p(W), _lib.p(self.h1), _lib.p(Y), int(lo), int(m),
                                                  _lib.p(logits[lo:] if want_logits else None), _lib.p(correct),
                                                  _lib.stream()), "head_eval")
        return logits


# ====================================================================== training (appended methods)
def _alloc_train_scratch(self) -> None:
    if getattr(self, "wg_part", None) is None:
        f32 = dict(device=self.device, dtype=torch.float32)
        self.n_cta = int(self.L.ddl_mnist_wgrad_ctas(self.G))
        self.wg_part = torch.zeros(self.G * self.n_cta * 3 * 128 * 64, **f32)
        self.c2b_part = torch.zeros(self.G * self.B, 64, **f32)
        self.c1_part = torch.zeros(self.G * self.B, 320, **f32)
        self.head_scratch = torch.zeros(int(self.L.ddl_mnist_head_scratch_floats(self.G)), **f32)


def train_step(self, W, X, x_stride, Y, y_stride, perm, perm_stride, sample_off, bsz, lr: float, seed: int) -> None:
    """One lock-step SGD step of all G clients (forward, backward and in-place update of ``W``)."""
    _alloc_train_scratch(self)
    L, G, B, st = self.L, self.G, self.B, _lib.stream
    cb = _ints(bsz)
    Wv = W.view(G, P)
    w1 = Wv[:, OFF["fc1.weight"]:OFF["fc1.weight"] + FC1 * FLAT].unflatten(1, (FC1, FLAT))
    self.forward_features(W, X, x_stride, perm, perm_stride, sample_off, bsz, True, seed)
    _lib.check(L.ddl_mnist_head_train(_lib.p(W), _lib.p(self.h1), _lib.p(Y), _lib.i64(y_stride), _lib.p(perm),
                                      _lib.i64(perm_stride), int(sample_off), _lib.p(self.dz1), _lib.p(self.loss), _lib.p(self.head_scratch), G, B, cb,
                                      _lib.f32(self.p2), int(seed) & 0xFFFFFFFF, _lib.f32(lr), st()), "head_train")
    # fc1 backward: dgrad first (reads the old weights), then wgrad with the SGD step fused in its epilogue
    gemm(self.dz1, w1, b_mn=True, precise=True, out=self.dpool)
    gemm(self.dz1, self.pooled, a_mn=True, b_mn=True, precise=True, alpha=-lr, beta=1.0, out=w1)
    _lib.check(L.ddl_mnist_pool_bwd(_lib.p(self.dpool), _lib.p(self.pmask), _lib.p(self.dy2), _lib.p(self.c2b_part), G, B,
                                    cb, _lib.f32(self.p1), st()), "pool_bwd")
    _lib.check(L.ddl_mnist_conv2_dgrad(_lib.p(self.wt_dgrad), _lib.p(self.dy2), _lib.p(self.a1), _lib.p(self.da1), G, B, cb,
                                       st()), "conv2_dgrad")
    _lib.check(L.ddl_mnist_conv2_wgrad(_lib.p(self.a1), _lib.p(self.dy2), _lib.p(self.wg_part), G, B, cb, st()),
               "conv2_wgrad")
    _lib.check(L.ddl_mnist_conv1_wgrad(_lib.p(X), _lib.i64(x_stride), _lib.p(perm), _lib.i64(perm_stride), int(sample_off),
                                       _lib.p(self.da1), _lib.p(self.c1_part), G, B, cb, st()), "conv1_wgrad")
    _lib.check(L.ddl_mnist_small_update(_lib.p(W), _lib.p(self.wg_part), self.n_cta, _lib.p(self.c2b_part),
                                        _lib.p(self.c1_part), _lib.p(self.wt_fwd), _lib.p(self.wt_dgrad), G, B, cb,
                                        _lib.f32(lr), st()), "small_update")


def local_train(self, W, X, Y, perm, counts, batch_size: int, epochs: int, lr: float, seed: int) -> int:
    """``epochs`` local epochs of minibatch SGD for all clients at once.

    ``W`` [G, P] weights (updated in place), ``X`` [G, n_max, 784], ``Y`` [G, n_max] int64,
    ``perm`` [G, epochs * n_max] int32 sample orders (epoch e at column offset e * n_max),
    ``counts`` host list of per-client sample counts.  Returns the number of steps run."""
    n_max = X.shape[1]
    assert batch_size <= self.B
    self.prep_weights(W)
    steps = 0
    n_steps = -(-max(counts) // batch_size)
    for e in range(epochs):
        for s in range(n_steps):
            off = s * batch_size
            bsz = [max(0, min(batch_size, int(c) - off)) for c in counts]
            train_step(self, W, X, n_max * 784, Y, n_max, perm, perm.shape[1], e * n_max + off, bsz, lr,
                       (seed * 2654435761 + e * 7919 + s * 104729) & 0xFFFFFFFF)
            steps += 1
    return steps


PackedMnistEngine.train_step = train_step
PackedMnistEngine.local_train = local_train
