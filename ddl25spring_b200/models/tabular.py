"""Tabular models of tutorials 2a/2b.

* :class:`HeartDiseaseNN`  – evaluator MLP 30-64-128-256-2 (lab/tutorial_2a/centralized.py:13-28)
* :class:`Autoencoder`     – tabular VAE (lab/tutorial_2a/generative-modeling.py:13-118)
* :class:`VAELoss`         – MSE(sum)+KL (generative-modeling.py:121-130)
* :class:`BottomModel` / :class:`TopModel` – split-NN parts (lab/tutorial_2b/vfl.py:11-40)
* :class:`ClientEncoder` / :class:`ClientDecoder` / :class:`ServerVAE` / :class:`VFLVAE`
  – vertical generative FL (lab/tutorial_2b/exercise_3.py:10-147)

Attribute names follow the reference so ``state_dict`` keys are interchangeable.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F
from torch import nn


class HeartDiseaseNN(nn.Module):
    """43 842 parameters."""

    def __init__(self, in_features: int = 30, n_classes: int = 2, p_drop: float = 0.1):
        super().__init__()
        widths = [in_features, 64, 128, 256]
        self.fc1, self.fc2, self.fc3 = (nn.Linear(a, b) for a, b in zip(widths[:-1], widths[1:]))
        self.fc4 = nn.Linear(widths[-1], n_classes)
        self.act = nn.LeakyReLU()
        self.dropout = nn.Dropout(p_drop)

    def forward(self, x):
        for fc in (self.fc1, self.fc2, self.fc3):
            x = self.act(fc(x))
        return self.fc4(self.dropout(x))


def _lin_bn(mod: nn.Module, lin_name: str, bn_name: str, a: int, b: int) -> None:
    setattr(mod, lin_name, nn.Linear(a, b))
    setattr(mod, bn_name, nn.BatchNorm1d(b))


class Autoencoder(nn.Module):
    """Tabular VAE: 3x(Linear+BN+ReLU) encoder -> fc1+BN -> mu / logvar; mirrored decoder whose
    last BatchNorm has no activation.  D_in=31,H=48,H2=32,latent=16 -> 10 781 parameters."""

    ENC = (("linear1", "lin_bn1"), ("linear2", "lin_bn2"), ("linear3", "lin_bn3"), ("fc1", "bn1"))
    DEC = (("fc3", "fc_bn3"), ("fc4", "fc_bn4"), ("linear4", "lin_bn4"), ("linear5", "lin_bn5"),
           ("linear6", "lin_bn6"))

    def __init__(self, D_in: int, H: int = 50, H2: int = 12, latent_dim: int = 3):
        super().__init__()
        enc_w = [D_in, H, H2, H2, latent_dim]
        for (l, b), i, o in zip(self.ENC, enc_w[:-1], enc_w[1:]):
            _lin_bn(self, l, b, i, o)
        self.fc21 = nn.Linear(latent_dim, latent_dim)   # mu
        self.fc22 = nn.Linear(latent_dim, latent_dim)   # log-variance
        dec_w = [latent_dim, latent_dim, H2, H2, H, D_in]
        for (l, b), i, o in zip(self.DEC, dec_w[:-1], dec_w[1:]):
            _lin_bn(self, l, b, i, o)
        self.relu = nn.ReLU()
        self.optimizer = None
        self.criterion = None

    def encode(self, x):
        for l, b in self.ENC:
            x = F.relu(getattr(self, b)(getattr(self, l)(x)))
        return self.fc21(x), self.fc22(x)

    def reparameterize(self, mu, logvar):
        if not self.training:
            return mu
        return mu + torch.randn_like(mu) * torch.exp(0.5 * logvar)

    def decode(self, z):
        for i, (l, b) in enumerate(self.DEC):
            z = getattr(self, b)(getattr(self, l)(z))
            if i + 1 < len(self.DEC):
                z = F.relu(z)
        return z

    def forward(self, x):
        mu, logvar = self.encode(x)
        return self.decode(self.reparameterize(mu, logvar)), mu, logvar

    def train_with_settings(self, epochs, batch_sz, real_data, optimizer, loss_fn, *,
                            zero_grad_per_step: bool = True, log=None):
        """Minibatch training loop (generative-modeling.py:81-102).

        The reference zeroes gradients once per *epoch* so they accumulate across minibatches
        (SURVEY §9); ``zero_grad_per_step=False`` reproduces that, the default fixes it.
        """
        self.optimizer, self.criterion = optimizer, loss_fn
        n = len(real_data)
        history = []
        for epoch in range(epochs):
            optimizer.zero_grad()
            total = 0.0
            nb = 0
            for lo in range(0, n, batch_sz):
                mb = real_data[lo:lo + batch_sz]
                if zero_grad_per_step:
                    optimizer.zero_grad()
                out, mu, logvar = self(mb)
                loss = loss_fn(out, mb, mu, logvar)
                loss.backward()
                optimizer.step()
                total += float(loss.detach())
                nb += 1
            history.append(total / max(nb, 1))
            if log is not None:
                log(epoch, history[-1])
        return history

    @torch.no_grad()
    def sample(self, nr_samples: int, logvar: torch.Tensor, mu: torch.Tensor, *, generator=None,
               round_last: bool = True) -> np.ndarray:
        """Draw z ~ N(mean_batch(mu), mean_batch(sigma)) and decode; the last column (the label)
        is clipped to [0,1] and rounded (generative-modeling.py:104-118)."""
        was_training = self.training
        self.eval()
        m, s = mu.mean(0), torch.exp(logvar / 2).mean(0)
        z = m + s * torch.randn(nr_samples, m.numel(), generator=generator, device=m.device)
        pred = self.decode(z).cpu().numpy()
        self.train(was_training)
        if round_last:
            pred[:, -1] = np.round(np.clip(pred[:, -1], 0, 1))
        return pred


def vae_loss(x_recon, x, mu, logvar):
    """MSE(sum) + KL,  KL = -1/2 sum(1 + logvar - mu^2 - exp(logvar))."""
    return F.mse_loss(x_recon, x, reduction="sum") - 0.5 * torch.sum(1 + logvar - mu.pow(2) - logvar.exp())


class VAELoss(nn.Module):
    def forward(self, x_recon, x, mu, logvar):
        return vae_loss(x_recon, x, mu, logvar)


customLoss = VAELoss  # the reference's name


# --------------------------------------------------------------------------- split NN (VFL)
class BottomModel(nn.Module):
    """Party model: Linear(in->out) ReLU Linear(out->out) ReLU dropout(.1); out = 2*in in the labs."""

    def __init__(self, in_feat: int, out_feat: int, p_drop: float = 0.1):
        super().__init__()
        self.local_out_dim = out_feat
        self.fc1 = nn.Linear(in_feat, out_feat)
        self.fc2 = nn.Linear(out_feat, out_feat)
        self.act = nn.ReLU()
        self.dropout = nn.Dropout(p_drop)

    def forward(self, x):
        return self.dropout(self.act(self.fc2(self.act(self.fc1(x)))))


class TopModel(nn.Module):
    """Server model over the concatenated party embeddings: cat -> 128 -> 256 -> n_outs.

    The reference applies LeakyReLU *and dropout* to the logits (vfl.py:39-40);
    ``act_on_logits=True`` keeps that shape, ``False`` returns plain logits.
    """

    def __init__(self, local_models_or_dims, n_outs: int = 2, act_on_logits: bool = True, p_drop: float = 0.1):
        super().__init__()
        dims = [m if isinstance(m, int) else m.local_out_dim for m in local_models_or_dims]
        self.party_dims = dims
        self.in_size = sum(dims)
        self.fc1 = nn.Linear(self.in_size, 128)
        self.fc2 = nn.Linear(128, 256)
        self.fc3 = nn.Linear(256, n_outs)
        self.act = nn.LeakyReLU()
        self.dropout = nn.Dropout(p_drop)
        self.act_on_logits = act_on_logits

    def forward_from_concat(self, h):
        h = self.act(self.fc2(self.act(self.fc1(h))))
        h = self.fc3(h)
        return self.dropout(self.act(h)) if self.act_on_logits else h

    def forward(self, party_outs):
        return self.forward_from_concat(torch.cat(list(party_outs), dim=1))


# --------------------------------------------------------------------------- VFL-VAE
class ClientEncoder(nn.Module):
    """in -> 48 -> 32 -> 32 -> latent, each Linear+BN+ReLU (exercise_3.py:10-31)."""
    STACK = (("linear1", "bn1"), ("linear2", "bn2"), ("linear3", "bn3"), ("fc", "bn_fc"))

    def __init__(self, input_dim: int, latent_dim: int):
        super().__init__()
        w = [input_dim, 48, 32, 32, latent_dim]
        for (l, b), i, o in zip(self.STACK, w[:-1], w[1:]):
            _lin_bn(self, l, b, i, o)
        self.relu = nn.ReLU()

    def forward(self, x):
        for l, b in self.STACK:
            x = F.relu(getattr(self, b)(getattr(self, l)(x)))
        return x


class ClientDecoder(nn.Module):
    """latent -> latent -> 32 -> 48 -> out; last BN without activation (exercise_3.py:33-54)."""
    STACK = (("linear1", "bn1"), ("linear2", "bn2"), ("linear3", "bn3"), ("linear4", "bn4"))

    def __init__(self, latent_dim: int, output_dim: int):
        super().__init__()
        w = [latent_dim, latent_dim, 32, 48, output_dim]
        for (l, b), i, o in zip(self.STACK, w[:-1], w[1:]):
            _lin_bn(self, l, b, i, o)
        self.relu = nn.ReLU()

    def forward(self, z):
        for i, (l, b) in enumerate(self.STACK):
            z = getattr(self, b)(getattr(self, l)(z))
            if i + 1 < len(self.STACK):
                z = F.relu(z)
        return z


class ServerVAE(Autoencoder):
    """VAE over the concatenated party latents (exercise_3.py:56-113); same stack as
    :class:`Autoencoder` with defaults H=48, H2=32, latent=16."""

    def __init__(self, D_in: int, H: int = 48, H2: int = 32, latent_dim: int = 16):
        super().__init__(D_in, H, H2, latent_dim)


class VFLVAE(nn.Module):
    """Per-party encoders -> concat (logical all-gather) -> server VAE -> column slices
    (logical scatter) -> per-party decoders (exercise_3.py:115-138)."""

    def __init__(self, client_encoders, server_vae, client_decoders, client_latent_dim: int):
        super().__init__()
        self.client_encoders = nn.ModuleList(client_encoders)
        self.server_vae = server_vae
        self.client_decoders = nn.ModuleList(client_decoders)
        self.client_latent_dim = client_latent_dim

    def forward(self, x_clients):
        latents = [enc(x) for enc, x in zip(self.client_encoders, x_clients)]
        concat_latent = torch.cat(latents, dim=1)
        recon_concat, mu, logvar = self.server_vae(concat_latent)
        parts = recon_concat.split(self.client_latent_dim, dim=1)
        recon_clients = [dec(p) for dec, p in zip(self.client_decoders, parts)]
        return recon_clients, mu, logvar, concat_latent, recon_concat


def combined_loss(x_clients, recon_clients, concat_latent, recon_concat, mu, logvar):
    """sum_party MSE(sum) + latent MSE(sum) + KL (exercise_3.py:140-147)."""
    loss = sum(F.mse_loss(r, o, reduction="sum") for o, r in zip(x_clients, recon_clients))
    loss = loss + F.mse_loss(recon_concat, concat_latent, reduction="sum")
    return loss - 0.5 * torch.sum(1 + logvar - mu.pow(2) - logvar.exp())
