"""Federated DCGAN (BASELINE.json config 5, north star): FedAvg of generator AND discriminator.

Every client holds 64x64 images, runs ``local_steps`` of the standard DCGAN update (D on real + fake,
then G; BCE-with-logits), and sends both networks' flat arenas — parameters AND BatchNorm running
statistics (``include_buffers=True``) — which the server averages with the sample-count weights through
the same K1 aggregation kernel as every other FedAvg in this framework.  With ``torch.distributed``
initialised the clients are dealt to the ranks and the partial sums are all-reduced.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist
import torch.nn.functional as F

from .. import ops
from ..ops.nn import kernel_bce_logits
from ..core.arena import FlatArena
from ..core.seeds import clients_per_round
from ..models.dcgan import Discriminator, Generator


class DCGANPair:
    def __init__(self, device, nz=100, ngf=64, ndf=64, kernel_convs: bool = False):
        if kernel_convs:      # channels-last tcgen05 (transposed) convolutions + fused NHWC BatchNorm (models/dcgan_nhwc.py)
            from ..models.dcgan_nhwc import DiscriminatorNHWC, GeneratorNHWC
            self.G, self.D = GeneratorNHWC(nz, ngf).to(device), DiscriminatorNHWC(ndf).to(device)
        else:
            self.G, self.D = Generator(nz, ngf).to(device), Discriminator(ndf).to(device)
        self.ga = FlatArena.from_module(self.G, include_buffers=True, with_grad=True)
        self.da = FlatArena.from_module(self.D, include_buffers=True, with_grad=True)
        self.nz, self.device = nz, device

    def n_param_elems(self):
        return (sum(s.numel for s in self.ga.specs if not s.is_buffer), sum(s.numel for s in self.da.specs if not s.is_buffer))


def dcgan_local_steps(pair: DCGANPair, images: torch.Tensor, steps: int, batch: int, lr: float, gen: torch.Generator,
                      m_g, v_g, m_d, v_d, step0: int, beta1: float = 0.5):
    """``steps`` DCGAN iterations with fused Adam over the flat arenas; returns (d_loss, g_loss) of the last step."""
    G, D, dev = pair.G, pair.D, pair.device
    G.train(); D.train()
    n = len(images)
    d_loss = g_loss = torch.zeros((), device=dev)
    pg, pd = pair.n_param_elems()
    for s in range(steps):
        idx = torch.randint(0, n, (batch,), generator=gen)
        real = images[idx].to(dev)
        z = torch.randn(batch, pair.nz, generator=gen).to(dev)
        # --- D step
        pair.da.grad.zero_()
        fake = G(z)
        d_loss = kernel_bce_logits(D(real), 1.0) + kernel_bce_logits(D(fake.detach()), 0.0)
        d_loss.backward()
        ops.adam_step(pair.da.flat[:_pad4(pd)], pair.da.grad[:_pad4(pd)], m_d, v_d, step0 + s + 1, lr, beta1, 0.999)
        # --- G step
        pair.ga.grad.zero_()
        g_loss = kernel_bce_logits(D(fake), 1.0)
        g_loss.backward()
        ops.adam_step(pair.ga.flat[:_pad4(pg)], pair.ga.grad[:_pad4(pg)], m_g, v_g, step0 + s + 1, lr, beta1, 0.999)
    return float(d_loss.detach()), float(g_loss.detach())


class GraphedDCGANStep:
    """One full DCGAN iteration (D step on real + fake, fused Adam on D, G step through the updated D, fused Adam on G) captured
    into ONE CUDA graph over the pair's static arenas.  The Adam moments and the bias-correction step live in static device
    buffers (``ddl_adam_step_dev`` reads the step from memory), so the same graph serves every client: its moments are copied
    in before and out after its local steps."""

    def __init__(self, pair: DCGANPair, batch: int, lr: float, beta1: float = 0.5):
        from ..utils.graphs import GraphedStep
        dev = pair.device
        pg, pd = pair.n_param_elems()
        self.pg4, self.pd4 = _pad4(pg), _pad4(pd)
        z = lambda n: torch.zeros(n, device=dev)
        self.m_g, self.v_g, self.m_d, self.v_d = z(self.pg4), z(self.pg4), z(self.pd4), z(self.pd4)
        self.step_t = torch.zeros((), dtype=torch.int32, device=dev)
        G, D = pair.G, pair.D

        def it(real, zz):
            ones, zeros = torch.ones(batch, device=dev), torch.zeros(batch, device=dev)
            self.step_t.add_(1)
            pair.da.grad.zero_()
            fake = G(zz)
            d_loss = kernel_bce_logits(D(real), 1.0) + kernel_bce_logits(D(fake.detach()), 0.0)
            d_loss.backward()
            ops.adam_step(pair.da.flat[:self.pd4], pair.da.grad[:self.pd4], self.m_d, self.v_d, self.step_t, lr, beta1, 0.999)
            pair.ga.grad.zero_()
            g_loss = kernel_bce_logits(D(fake), 1.0)
            g_loss.backward()
            ops.adam_step(pair.ga.flat[:self.pg4], pair.ga.grad[:self.pg4], self.m_g, self.v_g, self.step_t, lr, beta1, 0.999)
            return torch.stack([d_loss.detach(), g_loss.detach()])

        saved = [t.detach().clone() for t in (pair.ga.flat, pair.da.flat)]
        buffers = [(b, b.clone()) for m in (G, D) for b in m.buffers()]
        ex_real = torch.zeros(batch, 3, 64, 64, device=dev)
        ex_z = torch.zeros(batch, pair.nz, device=dev)
        G.train(); D.train()
        self.graph = GraphedStep(it, (ex_real, ex_z))
        with torch.no_grad():                       # undo the warm-up / capture iterations
            pair.ga.flat.copy_(saved[0]); pair.da.flat.copy_(saved[1])
            for b, v in buffers:
                b.copy_(v)
            for t in (self.m_g, self.v_g, self.m_d, self.v_d):
                t.zero_()
            self.step_t.zero_()

    def load(self, m_g, v_g, m_d, v_d, step0: int) -> None:
        self.m_g.copy_(m_g); self.v_g.copy_(v_g); self.m_d.copy_(m_d); self.v_d.copy_(v_d)
        self.step_t.fill_(step0)

    def store(self, m_g, v_g, m_d, v_d) -> None:
        m_g.copy_(self.m_g); v_g.copy_(self.v_g); m_d.copy_(self.m_d); v_d.copy_(self.v_d)

    def __call__(self, real, z):
        return self.graph(real, z)


def dcgan_local_steps_graphed(pair: DCGANPair, gstep: GraphedDCGANStep, images: torch.Tensor, steps: int, batch: int,
                              gen: torch.Generator, m_g, v_g, m_d, v_d, step0: int):
    """Same iterations and the same random draws as :func:`dcgan_local_steps`, each replayed from the captured graph."""
    dev = pair.device
    pair.G.train(); pair.D.train()
    gstep.load(m_g, v_g, m_d, v_d, step0)
    losses = None
    for _ in range(steps):
        idx = torch.randint(0, len(images), (batch,), generator=gen)
        real = images[idx].to(dev)
        z = torch.randn(batch, pair.nz, generator=gen).to(dev)
        losses = gstep(real, z)
    gstep.store(m_g, v_g, m_d, v_d)
    d_loss, g_loss = losses.tolist()
    return d_loss, g_loss


def _pad4(n: int) -> int:
    return (n + 3) // 4 * 4


class FederatedDCGAN:
    def __init__(self, client_images: list[torch.Tensor], client_fraction: float = 1.0, local_steps: int = 2, batch_size: int = 16,
                 lr: float = 2e-4, seed: int = 0, device=None, ngf: int = 64, ndf: int = 64, kernel_convs: bool | None = None,
                 graph: bool | None = None, collective=None):
        self.device = torch.device(device) if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.clients = client_images
        self.counts = [len(c) for c in client_images]
        self.k = clients_per_round(len(client_images), client_fraction)
        self.local_steps, self.batch, self.lr, self.seed = local_steps, batch_size, lr, seed
        self.rng = np.random.default_rng(seed)
        torch.manual_seed(seed)
        if kernel_convs is None:
            kernel_convs = self.device.type == "cuda"
        self.kernel_convs = kernel_convs
        # one CUDA graph per iteration (GraphedDCGANStep): default on the kernel path, where launch overhead dominates
        self.graph = (kernel_convs if graph is None else graph) and self.device.type == "cuda"
        self._gstep = None
        self.glob = DCGANPair(self.device, ngf=ngf, ndf=ndf, kernel_convs=kernel_convs)
        self.work = DCGANPair(self.device, ngf=ngf, ndf=ndf, kernel_convs=kernel_convs)
        self.rank, self.world = (dist.get_rank(), dist.get_world_size()) if dist.is_initialized() else (0, 1)
        pg, pd = self.glob.n_param_elems()
        z = lambda n: torch.zeros(_pad4(n), device=self.device)
        # Multi-GPU aggregation.  collective = a parallel.symm_impl.SymmetricArena: each rank pre-reduces ITS clients' G and D
        # (one [G | D] row, HBM speed), then ONE fused peer-memory kernel per rank sums the W partial rows shard by shard and
        # stores the result into every rank's replica - which IS the global G / D arena (rebased onto the symmetric buffer),
        # so the new global model needs no further copy.  None: torch.distributed all_reduce (the baseline).
        self.collective = collective if (collective is not None and self.world > 1) else None
        if self.collective is not None:
            a = self.collective
            self._pg, self._pd = self.glob.ga.size, self.glob.da.size
            n = self._pg + self._pd
            self._part = a.alloc("dcgan_partial", n * 4)
            res = a.alloc("dcgan_global", n * 4)
            self._part_ptrs, self._res_ptrs = a.peer_ptrs("dcgan_partial"), a.peer_ptrs("dcgan_global")
            self.glob.ga.rebase(res[:self._pg])
            self.glob.da.rebase(res[self._pg:n])
            self._n_sym = n
            a.barrier()
        self._adam = {}                      # per-client Adam moments persist across rounds
        self._zeros = z
        self.history: list[tuple[float, float]] = []

    def _moments(self, c):
        if c not in self._adam:
            pg, pd = self.glob.n_param_elems()
            self._adam[c] = [self._zeros(pg), self._zeros(pg), self._zeros(pd), self._zeros(pd), 0]
        return self._adam[c]

    def run(self, nr_rounds: int):
        for r in range(nr_rounds):
            chosen = self.rng.choice(len(self.clients), self.k, replace=False)
            w = np.asarray([self.counts[int(c)] for c in chosen], dtype=np.float64)
            w = w / w.sum()
            mine = [(p, int(c)) for p, c in enumerate(chosen) if p % self.world == self.rank]
            UG = torch.zeros(max(1, len(mine)), self.glob.ga.size, device=self.device)
            UD = torch.zeros(max(1, len(mine)), self.glob.da.size, device=self.device)
            losses = []
            for i, (pos, c) in enumerate(mine):
                with torch.no_grad():
                    self.work.ga.flat.copy_(self.glob.ga.flat)
                    self.work.da.flat.copy_(self.glob.da.flat)
                mom = self._moments(c)
                gen = torch.Generator().manual_seed(self.seed + c + 1 + r * self.k)
                if self.graph and len(self.clients[c]) >= 1:
                    if self._gstep is None:
                        self._gstep = GraphedDCGANStep(self.work, self.batch, self.lr)
                        with torch.no_grad():        # capture ran on the work pair: reload the global model
                            self.work.ga.flat.copy_(self.glob.ga.flat)
                            self.work.da.flat.copy_(self.glob.da.flat)
                    losses.append(dcgan_local_steps_graphed(self.work, self._gstep, self.clients[c], self.local_steps, self.batch, gen,
                                                            mom[0], mom[1], mom[2], mom[3], mom[4]))
                else:
                    losses.append(dcgan_local_steps(self.work, self.clients[c], self.local_steps, self.batch, self.lr, gen,
                                                    mom[0], mom[1], mom[2], mom[3], mom[4]))
                mom[4] += self.local_steps
                UG[i].copy_(self.work.ga.flat.detach())
                UD[i].copy_(self.work.da.flat.detach())
            lw = [float(w[pos]) for pos, _ in mine] or [0.0]
            with torch.no_grad():
                if self.collective is not None:
                    from ..parallel.symm_impl import shard_bounds
                    a = self.collective
                    if mine:
                        ops.weighted_sum(UG, lw, self._part[:self._pg])
                        ops.weighted_sum(UD, lw, self._part[self._pg:self._n_sym])
                    else:
                        self._part.zero_()
                    d0, d1 = shard_bounds(self._n_sym, a.world, a.rank)
                    a.reduce(0, self._part_ptrs, [1.0] * a.world, self._res_ptrs, d0, d1)     # writes the global G | D everywhere
                else:
                    aggG = ops.weighted_sum(UG, lw)
                    aggD = ops.weighted_sum(UD, lw)
                    if self.world > 1:
                        dist.all_reduce(aggG); dist.all_reduce(aggD)
                    self.glob.ga.flat.copy_(aggG)
                    self.glob.da.flat.copy_(aggD)
            if losses:
                self.history.append(tuple(np.mean(losses, axis=0)))
        if self.collective is not None:
            torch.cuda.synchronize()
            self.collective.check()
        return self.history

    @torch.no_grad()
    def sample(self, n: int, seed: int = 0) -> torch.Tensor:
        self.glob.G.eval()
        z = torch.randn(n, self.glob.nz, generator=torch.Generator().manual_seed(seed)).to(self.device)
        return self.glob.G(z)
