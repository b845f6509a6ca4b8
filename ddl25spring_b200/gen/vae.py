"""Tabular VAE training and the synthetic-data utility evaluation (lab/tutorial_2a).

* :func:`train_evaluator`        – the 49-epoch full-batch AdamW loop of centralized.py:31-70, keeping a
                                   *cloned* best state dict (the reference keeps an alias).
* :func:`synthetic_data_utility` – generative-modeling.py:133-211: train the VAE on ``[X_train, y]``, sample a
                                   synthetic set of the same size, train one evaluator on real and one on
                                   synthetic data, test both on the real test set ("train on synthetic, test on real").
The VAE loss (MSE(sum) + KL) runs through the fused kernel on CUDA tensors (K14) with an autograd bridge.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn, optim

from .. import ops
from ..core.checkpoint import snapshot
from ..ops.nn import kernel_cross_entropy, use_kernels
from ..ops.optim import make_optimizer
from ..models.tabular import Autoencoder, HeartDiseaseNN, vae_loss


class _FusedVAELoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x_recon, x, mu, logvar):
        loss, grads = ops.vae_loss(x_recon.contiguous(), x.contiguous(), mu.contiguous(), logvar.contiguous(), want_grad=True)
        ctx.save_for_backward(*grads)
        return loss

    @staticmethod
    def backward(ctx, g):
        gx, gm, gl = ctx.saved_tensors
        return g * gx, None, g * gm, g * gl


def fused_vae_loss(x_recon, x, mu, logvar):
    """``customLoss`` with value and gradients from one kernel launch on GPU; plain torch on CPU."""
    if x.is_cuda:
        return _FusedVAELoss.apply(x_recon, x, mu, logvar)
    return vae_loss(x_recon, x, mu, logvar)


def train_evaluator(x_train, y_train, x_test, y_test, epochs: int = 49, model: nn.Module | None = None, log=None):
    """Returns ``(model with best-on-test weights, best test accuracy, history)``."""
    net = model or HeartDiseaseNN(x_train.shape[1]).to(x_train.device)
    if x_train.is_cuda:            # B200 path: every Linear on the tcgen05 GEMM (forward + both gradients), fused cross-entropy,
        use_kernels(net)           # one fused AdamW launch over the flat parameter arena
    opt = make_optimizer(net, "adamw")
    best_acc, best = -1.0, snapshot(net)
    hist = []
    for epoch in range(1, epochs + 1):
        net.train()
        opt.zero_grad()
        out = net(x_train)
        loss = kernel_cross_entropy(out, y_train)
        loss.backward()
        opt.step()
        net.eval()
        with torch.no_grad():
            acc = float((net(x_test).argmax(1) == y_test).float().mean())
        hist.append((float(loss.detach()), acc))
        if log:
            log(epoch, hist[-1])
        if acc > best_acc:
            best_acc, best = acc, snapshot(net)          # a real copy
    net.load_state_dict(best)
    return net, best_acc, hist


def train_tabular_vae(real_data: torch.Tensor, H: int = 48, H2: int = 32, latent_dim: int = 16, epochs: int = 200,
                      batch_size: int = 64, lr: float = 1e-3, zero_grad_per_step: bool = True, log=None) -> Autoencoder:
    model = Autoencoder(real_data.shape[1], H, H2, latent_dim).to(real_data.device)
    if real_data.is_cuda:          # Linear + BatchNorm1d on kernels, fused VAE loss, fused Adam over the arena
        use_kernels(model)
    opt = make_optimizer(model, "adam", lr=lr)
    model.train()
    model.train_with_settings(epochs, batch_size, real_data, opt, fused_vae_loss, zero_grad_per_step=zero_grad_per_step, log=log)
    return model


def synthetic_data_utility(x_train, y_train, x_test, y_test, vae_epochs: int = 200, eval_epochs: int = 49, seed: int = 42):
    """Returns ``{"real": acc, "synthetic": acc}`` on the real test set."""
    torch.manual_seed(seed)
    real = torch.cat([x_train, y_train.float().view(-1, 1)], dim=1)
    vae = train_tabular_vae(real, epochs=vae_epochs)
    vae.train()
    with torch.no_grad():
        _, mu, logvar = vae(real)
    synth = torch.from_numpy(vae.sample(len(real), logvar, mu)).to(real.device)
    sx, sy = synth[:, :-1].float(), synth[:, -1].long().clamp(0, 1)
    _, acc_real, _ = train_evaluator(x_train, y_train, x_test, y_test, eval_epochs)
    _, acc_syn, _ = train_evaluator(sx, sy, x_test, y_test, eval_epochs)
    return {"real": acc_real, "synthetic": acc_syn}
