"""Plain PyTorch restatements of every op the CUDA library implements.

These are (a) the numerics oracles the GPU tests compare each sm_100a kernel against and
(b) the implementation used for CPU tensors (lab plumbing over gloo, BASELINE.json config 1).
They are written for clarity, not speed.  Robust-aggregation rules follow the published
algorithms: Krum / Multi-Krum (Blanchard et al. 2017), coordinate-wise median
and trimmed mean (Yin et al. 2018); the reference tree contains no code for them.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------- aggregation
def weighted_sum(updates: torch.Tensor, weights: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """``out[d] = sum_k weights[k] * updates[k, d]`` — FedAvg's aggregator with
    ``weights[k] = n_k / sum n`` (lab/tutorial_1a/hfl_complete.py:370-378)."""
    res = (updates * weights.to(updates.dtype)[:, None]).sum(0)
    if out is not None:
        out.copy_(res)
        return out
    return res


def weighted_sum_sgd(updates, weights, params, lr: float) -> torch.Tensor:
    """FedSGD server: ``params -= lr * sum_k w_k g_k`` (hfl_complete.py:291-306)."""
    params.sub_(weighted_sum(updates, weights), alpha=lr)
    return params


def coordinate_median(updates: torch.Tensor, lower: bool = False) -> torch.Tensor:
    """Per-coordinate median over the K rows.  For even K the default is the mean of the two
    middle values (numpy convention); ``lower=True`` is ``torch.median``'s lower median."""
    K = updates.shape[0]
    s = torch.sort(updates, dim=0).values
    if K % 2 == 1 or lower:
        return s[(K - 1) // 2].clone()
    return 0.5 * (s[K // 2 - 1] + s[K // 2])


def trimmed_mean(updates: torch.Tensor, trim: int) -> torch.Tensor:
    """Per coordinate, drop the ``trim`` largest and ``trim`` smallest values, mean the rest."""
    K = updates.shape[0]
    assert 0 <= 2 * trim < K
    s = torch.sort(updates, dim=0).values
    return s[trim:K - trim].mean(0)


def gram(updates: torch.Tensor) -> torch.Tensor:
    """``G = U U^T`` in the input dtype (fp64 in the oracle tests)."""
    return updates @ updates.T


def pairwise_sq_dists_from_gram(G: torch.Tensor) -> torch.Tensor:
    d = torch.diagonal(G)
    return (d[:, None] + d[None, :] - 2 * G).clamp_min(0)


def krum_scores(updates: torch.Tensor, n_byzantine: int, dists: torch.Tensor | None = None) -> torch.Tensor:
    """score_i = sum of the ``K - f - 2`` smallest squared distances from i to the others."""
    K = updates.shape[0]
    if dists is None:
        dists = torch.cdist(updates.double(), updates.double()).pow(2)
    m = K - n_byzantine - 2
    assert m >= 1, "Krum needs K >= f + 3"
    d = dists.clone().double()
    d.fill_diagonal_(float("inf"))
    return torch.sort(d, dim=1).values[:, :m].sum(1)


def krum_select(updates: torch.Tensor, n_byzantine: int, multi: int = 1, dists=None) -> torch.Tensor:
    """Indices of the ``multi`` best-scoring updates (multi=1: Krum; >1: Multi-Krum)."""
    scores = krum_scores(updates, n_byzantine, dists)
    return torch.argsort(scores, stable=True)[:multi]


def krum(updates: torch.Tensor, n_byzantine: int, multi: int = 1) -> torch.Tensor:
    sel = krum_select(updates, n_byzantine, multi)
    return updates[sel].mean(0)


# --------------------------------------------------------------------------- attacks
def sign_flip(update: torch.Tensor, global_flat: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
    """Malicious weights ``w_g - s (w_k - w_g)``; for a gradient message pass ``global_flat=0``."""
    return global_flat - scale * (update - global_flat)


def gaussian_noise_attack(update, sigma: float, generator=None):
    return update + sigma * torch.randn(update.shape, generator=generator, device=update.device, dtype=update.dtype)


# --------------------------------------------------------------------------- optimizers (flat)
def sgd_step(p, g, lr: float, weight_decay: float = 0.0, momentum: float = 0.0, buf=None):
    if weight_decay:
        g = g + weight_decay * p
    if momentum:
        buf.mul_(momentum).add_(g)
        g = buf
    p.sub_(g, alpha=lr)
    return p


def adam_step(p, g, m, v, step: int, lr: float = 1e-3, beta1: float = 0.9, beta2: float = 0.999,
              eps: float = 1e-8, weight_decay: float = 0.0, decoupled: bool = False):
    """torch.optim.Adam (decoupled=False) / AdamW (decoupled=True) single-tensor update."""
    if weight_decay and decoupled:
        p.mul_(1 - lr * weight_decay)
    elif weight_decay:
        g = g + weight_decay * p
    m.mul_(beta1).add_(g, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    bc1, bc2 = 1 - beta1 ** step, 1 - beta2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)
    return p


# --------------------------------------------------------------------------- losses / eval
def logsoftmax_nll(logits: torch.Tensor, target: torch.Tensor):
    """mean NLL of log_softmax(logits) and d(loss)/d(logits) (hfl_complete.py:62,78)."""
    logp = F.log_softmax(logits.float(), dim=1)
    loss = F.nll_loss(logp, target)
    grad = (logp.exp() - F.one_hot(target, logits.shape[1]).float()) / logits.shape[0]
    return loss, grad


def softmax_ce_soft(logits: torch.Tensor, target_probs: torch.Tensor):
    """CrossEntropyLoss with float (one-hot / soft) targets (lab/tutorial_2b/vfl.py:59,79)."""
    logp = F.log_softmax(logits.float(), dim=1)
    loss = -(target_probs * logp).sum(1).mean()
    grad = (logp.exp() * target_probs.sum(1, keepdim=True) - target_probs) / logits.shape[0]
    return loss, grad


def causal_lm_loss(logits: torch.Tensor, tokens: torch.Tensor):
    V = logits.shape[-1]
    return F.cross_entropy(logits[:, :-1].reshape(-1, V).float(), tokens[:, 1:].reshape(-1))


def vae_reparam_loss(x_recon, x, mu, logvar):
    return F.mse_loss(x_recon, x, reduction="sum") - 0.5 * torch.sum(1 + logvar - mu.pow(2) - logvar.exp())


def eval_top1(logits: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """Number of correct top-1 predictions (hfl_complete.py:172-183)."""
    return (logits.argmax(1) == target).sum()


# --------------------------------------------------------------------------- dense layers
def linear_act(x, w, b=None, act: str = "none", slope: float = 0.01):
    y = F.linear(x, w, b)
    if act == "relu":
        return F.relu(y)
    if act == "leaky_relu":
        return F.leaky_relu(y, slope)
    return y


def gather_linear(parts: list[torch.Tensor], w, b=None, act: str = "none"):
    """VFL top-model entry: ``act(cat(parts, 1) @ w.T + b)`` (vfl.py:36-37)."""
    return linear_act(torch.cat(parts, dim=1), w, b, act)


# --------------------------------------------------------------------------- MnistCnn pieces
def mnist_forward(flat_params: torch.Tensor, x: torch.Tensor, layout, train_masks=None):
    """Functional MnistCnn forward over a flat parameter vector (eval mode unless masks given)."""
    p = {}
    off = 0
    for name, shape in layout:
        off = (off + 3) // 4 * 4
        n = math.prod(shape)
        p[name] = flat_params[off:off + n].view(shape)
        off += n
    h = F.relu(F.conv2d(x, p["conv1.weight"], p["conv1.bias"]))
    h = F.relu(F.conv2d(h, p["conv2.weight"], p["conv2.bias"]))
    h = F.max_pool2d(h, 2)
    if train_masks is not None:
        h = h * train_masks[0]
    h = F.relu(F.linear(h.flatten(1), p["fc1.weight"], p["fc1.bias"]))
    if train_masks is not None:
        h = h * train_masks[1]
    return F.linear(h, p["fc2.weight"], p["fc2.bias"])
