"""Vertical federated learning (split learning): party bottom models feeding a server top model.

Parity with lab/tutorial_2b:

* :class:`VFLNetwork` — ``train_with_settings(epochs, batch_sz, n_cli, cli_features, x, y)``, ``forward``,
  ``test(x, y) -> (accuracy, loss)`` (vfl.py:43-102).  Two reference quirks are options, not defaults
  (SURVEY §9): ``freeze_bottoms=True`` reproduces the unregistered-bottom-models bug (vfl.py:48-50: only
  the top model trains), ``zero_grad_per_step=False`` the zero-grad-once-per-epoch accumulation (vfl.py:62).
* partitioners — contiguous raw-column split (vfl.py:105-141), seeded random feature permutations
  (exercise_1.py:107-163), balanced split for 2/4/6/8 parties (exercise_2.py:106-179).
* :func:`train_vfl_vae` — the vertical VAE loop of exercise_3.py:191-203.
* :class:`DistributedVFL` — parties on different ranks/GPUs.  The party -> server crossing
  (``torch.cat(party_outs, 1)`` at vfl.py:36) becomes an all-gather of embeddings, its backward the
  matching scatter of embedding gradients.  On B200 the gather is fused into the top model's first GEMM
  (:mod:`parallel.symm_impl`, K7); with gloo/NCCL it is ``all_gather`` + ``cat``.
"""
from __future__ import annotations

from typing import Callable, Sequence

import numpy as np
import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch import nn, optim

from ..core.data import HEART_RAW_COLUMNS, HeartTable
from ..models.tabular import BottomModel, TopModel, VFLVAE, combined_loss


# ------------------------------------------------------------------------------------------ partitioners
def contiguous_raw_partition(table: HeartTable, n_parties: int = 4) -> list[list[str]]:
    """Split the RAW columns ``(n-1) x floor(13/n)`` + remainder, then expand categorical columns to their
    one-hot names (vfl.py:116-141): 4 parties -> 7/4/6/13 encoded features."""
    raw = list(HEART_RAW_COLUMNS)
    per = (n_parties - 1) * [len(raw) // n_parties]
    per.append(len(raw) - sum(per))
    out, start = [], 0
    for k in per:
        names: list[str] = []
        for c in raw[start:start + k]:
            names += table.raw_to_encoded[c]
        out.append(names)
        start += k
    return out


def balanced_counts(n_features: int, n_parties: int) -> list[int]:
    """``base + (1 if i < extra)`` (exercise_2.py:129-132)."""
    base, extra = divmod(n_features, n_parties)
    return [base + (1 if i < extra else 0) for i in range(n_parties)]


def balanced_partition(columns: Sequence[str], n_parties: int) -> list[list[str]]:
    out, start = [], 0
    for k in balanced_counts(len(columns), n_parties):
        out.append(list(columns[start:start + k]))
        start += k
    return out


def random_permutation_partition(columns: Sequence[str], n_parties: int, seed: int) -> list[list[str]]:
    """Seeded permutation of the encoded columns, then ``(n-1) x floor`` + remainder (exercise_1.py:117-129:
    30 columns / 4 parties -> 7/7/7/9)."""
    rng = np.random.default_rng(seed)
    cols = [columns[i] for i in rng.permutation(len(columns))]
    per = (n_parties - 1) * [len(cols) // n_parties]
    per.append(len(cols) - sum(per))
    out, start = [], 0
    for k in per:
        out.append(cols[start:start + k])
        start += k
    return out


def _select(x, feats, columns=None) -> torch.Tensor:
    """Feature slice of ``x`` for one party: DataFrame + names, tensor + names (needs ``columns``) or tensor + indices."""
    if hasattr(x, "loc") and hasattr(x, "columns"):            # pandas DataFrame
        return torch.tensor(x[list(feats)].to_numpy(dtype="float32"))
    if len(feats) and isinstance(feats[0], str):
        pos = {c: i for i, c in enumerate(columns)}
        feats = [pos[f] for f in feats]
    return x[:, list(feats)].float()


def _targets(y) -> torch.Tensor:
    if hasattr(y, "to_numpy"):
        y = torch.tensor(y.to_numpy(dtype="float32"))
    if y.dim() == 1:                                           # class indices -> float one-hot (vfl.py:59,79)
        y = F.one_hot(y.long(), int(y.max()) + 1).float()
    return y.float()


# ------------------------------------------------------------------------------------------ VFLNetwork
class VFLNetwork(nn.Module):
    def __init__(self, local_models: Sequence[BottomModel], n_outs: int = 2, *, freeze_bottoms: bool = False,
                 act_on_logits: bool = True, columns: Sequence[str] | None = None):
        super().__init__()
        self.num_cli = None
        self.cli_features = None
        self.columns = list(columns) if columns is not None else None
        self.freeze_bottoms = freeze_bottoms
        if freeze_bottoms:
            self.bottom_models = list(local_models)            # plain list: NOT registered, like the reference
        else:
            self.bottom_models = nn.ModuleList(local_models)
        self.top_model = TopModel(local_models, n_outs, act_on_logits=act_on_logits)
        self.optimizer = optim.AdamW(self.parameters())
        self.criterion = nn.CrossEntropyLoss()
        self._on_kernels = False

    def _kernel_path(self, device) -> None:
        """On a GPU: every Linear of the bottoms and the top model runs on the tcgen05 GEMM (forward, dgrad, wgrad with the
        bias folded in), the soft-target cross-entropy is the fused kernel and AdamW is ONE launch over the flat arena of the
        registered parameters (the frozen-bottoms quirk is kept: unregistered bottoms are not in the arena)."""
        if self._on_kernels or torch.device(device).type != "cuda":
            return
        from ..ops.nn import KernelSoftCrossEntropy, use_kernels
        from ..ops.optim import ArenaOptimizer
        use_kernels(self)
        for m in self.bottom_models:
            use_kernels(m)
        self.optimizer = ArenaOptimizer(self, "adamw")
        self.criterion = KernelSoftCrossEntropy()
        self._on_kernels = True

    def forward(self, x: Sequence[torch.Tensor]) -> torch.Tensor:
        local_outs = [m(xi) for m, xi in zip(self.bottom_models, x)]
        return self.top_model(local_outs)                      # cat(dim=1) = the party -> server gather

    def train_with_settings(self, epochs: int, batch_sz: int, n_cli: int, cli_features, x, y, *,
                            zero_grad_per_step: bool = True, log_loss: Callable[[int, float], None] | None = None,
                            verbose: bool = False) -> list[float]:
        self.num_cli, self.cli_features = n_cli, cli_features
        dev = next(self.top_model.parameters()).device
        self._kernel_path(dev)
        x_train = [_select(x, feats, self.columns).to(dev) for feats in cli_features]
        y_train = _targets(y).to(dev)
        n = len(y_train)
        history = []
        for epoch in range(epochs):
            self.optimizer.zero_grad()
            total_loss, correct, nb = 0.0, 0, 0
            for lo in range(0, n, batch_sz):
                xb = [xi[lo:lo + batch_sz] for xi in x_train]
                yb = y_train[lo:lo + batch_sz]
                if zero_grad_per_step:
                    self.optimizer.zero_grad()
                outs = self.forward(xb)
                correct += int((outs.argmax(1) == yb.argmax(1)).sum())
                loss = self.criterion(outs, yb)
                loss.backward()
                self.optimizer.step()
                total_loss += float(loss.detach())
                nb += 1
            history.append(total_loss / nb)
            if log_loss is not None:
                log_loss(epoch, history[-1])
            if verbose:
                print(f"Epoch: {epoch} Train accuracy: {correct * 100 / n:.2f}% Loss: {history[-1]:.3f}")
        return history

    @torch.no_grad()
    def test(self, x, y):
        was = self.training
        dev = next(self.top_model.parameters()).device
        x_test = [_select(x, feats, self.columns).to(dev) for feats in self.cli_features]
        y_test = _targets(y).to(dev)
        outs = self.forward(x_test)
        acc = (outs.argmax(1) == y_test.argmax(1)).float().mean()
        loss = self.criterion(outs, y_test)
        self.train(was)
        return acc, loss


def build_vfl(table: HeartTable, client_feature_names: Sequence[Sequence[str]], outs_per_feature: int = 2, **kw) -> VFLNetwork:
    bottoms = [BottomModel(len(f), outs_per_feature * len(f)) for f in client_feature_names]
    return VFLNetwork(bottoms, 2, columns=table.columns, **kw)


# ------------------------------------------------------------------------------------------ VFL-VAE
def train_vfl_vae(model: VFLVAE, x_clients: Sequence[torch.Tensor], epochs: int = 1000, lr: float = 1e-3,
                  log: Callable[[int, float], None] | None = None) -> list[float]:
    """Full-batch Adam loop of exercise_3.py:191-203; returns the per-epoch combined loss."""
    if next(model.parameters()).is_cuda:       # Linear / BatchNorm1d on kernels, fused Adam over the arena
        from ..ops.nn import use_kernels
        use_kernels(model)
    from ..ops.optim import make_optimizer
    opt = make_optimizer(model, "adam", lr=lr)
    losses = []
    for epoch in range(epochs):
        model.train()
        opt.zero_grad()
        recon, mu, logvar, concat_latent, recon_concat = model(x_clients)
        loss = combined_loss(x_clients, recon, concat_latent, recon_concat, mu, logvar)
        loss.backward()
        opt.step()
        losses.append(float(loss.detach()))
        if log is not None:
            log(epoch, losses[-1])
    return losses


# ------------------------------------------------------------------------------------------ distributed
class _GatherCat(torch.autograd.Function):
    """Forward: all-gather party embeddings and concatenate along features.  Backward: every rank keeps the
    slice of the gradient that belongs to its own embedding (the scatter of vfl.py:81)."""

    @staticmethod
    def forward(ctx, emb: torch.Tensor, dims: tuple, rank: int, group):
        ctx.rank, ctx.dims = rank, dims
        dmax = max(dims)                                     # collectives need equal shapes: pad, gather, trim
        padded = emb if emb.shape[1] == dmax else F.pad(emb, (0, dmax - emb.shape[1]))
        parts = [torch.empty(emb.shape[0], dmax, dtype=emb.dtype, device=emb.device) for _ in dims]
        dist.all_gather(parts, padded.contiguous(), group=group)
        return torch.cat([p[:, :d] for p, d in zip(parts, dims)], dim=1)

    @staticmethod
    def backward(ctx, grad):
        lo = sum(ctx.dims[:ctx.rank])
        return grad[:, lo:lo + ctx.dims[ctx.rank]].contiguous(), None, None, None


class DistributedVFL(nn.Module):
    """One party per rank.  Each rank owns its bottom model (private features never leave the rank); the top
    model is replicated and fed the gathered embeddings, so its gradients are identical on every rank and
    need no all-reduce, and each rank back-propagates only its own embedding's gradient slice.

    ``gather`` may be replaced by a fused implementation with the signature
    ``gather_linear(emb, dims, rank, weight, bias) -> pre-activation of top.fc1`` (K7).
    """

    def __init__(self, in_features: int, party_dims: Sequence[int], n_outs: int = 2, *, group=None, act_on_logits: bool = True,
                 gather_linear=None, seed: int = 0):
        super().__init__()
        self.group = group
        self.rank = dist.get_rank(group)
        self.dims = tuple(int(d) for d in party_dims)
        torch.manual_seed(seed + 1000 + self.rank)
        self.bottom = BottomModel(in_features, self.dims[self.rank])
        torch.manual_seed(seed)                                  # identical top model everywhere
        self.top_model = TopModel(list(self.dims), n_outs, act_on_logits=act_on_logits)
        self.gather_linear = gather_linear

    def forward(self, x_local: torch.Tensor) -> torch.Tensor:
        emb = self.bottom(x_local)
        if self.gather_linear is not None:
            t = self.top_model
            h = t.act(self.gather_linear(emb, self.dims, self.rank, t.fc1.weight, t.fc1.bias))
            h = t.fc3(t.act(t.fc2(h)))
            return t.dropout(t.act(h)) if t.act_on_logits else h
        return self.top_model.forward_from_concat(_GatherCat.apply(emb, self.dims, self.rank, self.group))


# ------------------------------------------------------------------------------------------ distributed VFL-VAE
class _SliceScatter(torch.autograd.Function):
    """The server -> party crossing of the vertical VAE (exercise_3.py:132-137).  Forward: party ``rank`` takes its column
    slice of the (replicated) reconstructed latent.  Backward: the gradient of the replicated tensor is the CONCATENATION of
    every party's slice gradient, so the slices are all-gathered - afterwards every replica of the server VAE back-propagates
    the same, complete gradient and the replicas stay identical without any parameter all-reduce."""

    @staticmethod
    def forward(ctx, recon_concat: torch.Tensor, rank: int, width: int, n_parties: int, group):
        ctx.rank, ctx.width, ctx.n, ctx.group = rank, width, n_parties, group
        return recon_concat[:, rank * width:(rank + 1) * width].contiguous()

    @staticmethod
    def backward(ctx, g):
        parts = [torch.empty_like(g) for _ in range(ctx.n)]
        dist.all_gather(parts, g.contiguous(), group=ctx.group)
        return torch.cat(parts, dim=1), None, None, None, None


class DistributedVFLVAE(nn.Module):
    """The vertical VAE of lab/tutorial_2b/exercise_3.py:115-147 with one party per rank (GPU).

    Each rank owns its party's encoder and decoder - private features and reconstructions never leave it - and a replica of
    the server VAE.  Two crossings per forward: the party latents are all-gathered into ``concat_latent`` (its backward keeps
    the own slice: :class:`_GatherCat`), and the reconstructed latent is sliced per party (its backward all-gathers the slice
    gradients: :class:`_SliceScatter`).  The server replicas see identical inputs and identical noise (same seed, same RNG
    consumption), compute identical gradients and therefore need no all-reduce.  The gather is NOT fused into the server's
    first GEMM here: the loss needs ``concat_latent`` itself (the latent reconstruction term), so it has to exist anyway.
    ``loss()`` returns this rank's share (own reconstruction + the replicated server terms) whose backward is exact;
    ``total_loss`` all-reduces the party terms for reporting."""

    def __init__(self, in_features: int, n_parties: int, client_latent_dim: int = 8, *, group=None, seed: int = 0):
        super().__init__()
        from ..models.tabular import ClientDecoder, ClientEncoder, ServerVAE
        self.group, self.n, self.L = group, n_parties, client_latent_dim
        self.rank = dist.get_rank(group)
        torch.manual_seed(seed + 1000 + self.rank)
        self.encoder = ClientEncoder(in_features, client_latent_dim)
        self.decoder = ClientDecoder(client_latent_dim, in_features)
        torch.manual_seed(seed)                                  # identical server replica (and noise stream) everywhere
        self.server_vae = ServerVAE(n_parties * client_latent_dim)

    def forward(self, x_local: torch.Tensor):
        latent = self.encoder(x_local)
        concat = _GatherCat.apply(latent, (self.L,) * self.n, self.rank, self.group)
        recon_concat, mu, logvar = self.server_vae(concat)
        mine = _SliceScatter.apply(recon_concat, self.rank, self.L, self.n, self.group)
        return self.decoder(mine), mu, logvar, concat, recon_concat

    def loss(self, x_local, out):
        recon, mu, logvar, concat, recon_concat = out
        party = F.mse_loss(recon, x_local, reduction="sum")
        server = F.mse_loss(recon_concat, concat, reduction="sum") - 0.5 * torch.sum(1 + logvar - mu.pow(2) - logvar.exp())
        return party + server, party.detach(), server.detach()


def train_distributed_vfl_vae(model: DistributedVFLVAE, x_local: torch.Tensor, epochs: int = 1000, lr: float = 1e-3, log=None) -> list[float]:
    """Full-batch Adam loop of exercise_3.py:191-203 on every rank; returns the per-epoch TOTAL combined loss."""
    if next(model.parameters()).is_cuda:
        from ..ops.nn import use_kernels
        use_kernels(model)
    from ..ops.optim import make_optimizer
    opt = make_optimizer(model, "adam", lr=lr)
    losses = []
    for epoch in range(epochs):
        model.train()
        opt.zero_grad()
        loss, party, server = model.loss(x_local, model(x_local))
        loss.backward()
        opt.step()
        tot = party.clone()
        dist.all_reduce(tot, group=model.group)
        losses.append(float(tot + server))
        if log is not None:
            log(epoch, losses[-1])
    return losses
