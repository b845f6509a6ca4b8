"""Vertical FL: partitioners, VFLNetwork (incl. reference quirks as options), VFL-VAE, distributed parties."""
import pytest
import torch

from ddl25spring_b200.core.data import SyntheticHeart
from ddl25spring_b200.models.tabular import BottomModel, ClientDecoder, ClientEncoder, ServerVAE, VFLVAE
from ddl25spring_b200.parallel import vfl
from tests.dist_utils import run_distributed


@pytest.fixture(scope="module")
def table():
    return SyntheticHeart().encoded()


def test_heart_schema(table):
    assert table.x.shape == (1025, 30) and len(table.columns) == 30
    assert float(table.x.min()) >= 0 and float(table.x.max()) <= 1


def test_partitioners(table):
    parts = vfl.contiguous_raw_partition(table, 4)
    assert [len(p) for p in parts] == [7, 4, 6, 13]              # SURVEY C33 [measured]
    assert sum(parts, []) == table.columns
    assert [len(p) for p in vfl.random_permutation_partition(table.columns, 4, 0)] == [7, 7, 7, 9]
    a, b = vfl.random_permutation_partition(table.columns, 4, 1), vfl.random_permutation_partition(table.columns, 4, 2)
    assert a != b and sorted(sum(a, [])) == sorted(table.columns)
    assert vfl.balanced_counts(30, 8) == [4, 4, 4, 4, 4, 4, 3, 3]
    for n in (2, 4, 6, 8):
        assert sum(len(p) for p in vfl.balanced_partition(table.columns, n)) == 30


def test_vfl_network_trains_and_quirk_flags(table):
    torch.manual_seed(42)
    parts = vfl.contiguous_raw_partition(table, 4)
    xtr, ytr, xte, yte = table.x[:821], table.y[:821], table.x[821:], table.y[821:]
    net = vfl.build_vfl(table, parts)
    n_all = sum(p.numel() for p in net.parameters())
    assert n_all == 43086                                         # top 41 346 + bottoms 1 740
    logged = []
    hist = net.train_with_settings(15, 64, 4, parts, xtr, ytr, log_loss=lambda e, l: logged.append(l))
    assert hist[-1] < hist[0] and logged == hist
    acc, loss = net.test(xte, yte)
    assert 0.5 < float(acc) <= 1.0
    frozen = vfl.build_vfl(table, parts, freeze_bottoms=True)
    assert sum(p.numel() for p in frozen.parameters()) == 41346   # the reference's optimizer sees only the top model
    before = [p.detach().clone() for m in frozen.bottom_models for p in m.parameters()]
    frozen.train_with_settings(2, 64, 4, parts, xtr, ytr, zero_grad_per_step=False)
    after = [p.detach() for m in frozen.bottom_models for p in m.parameters()]
    assert all(torch.equal(a, b) for a, b in zip(before, after))


def test_vfl_vae_loss_decreases(table):
    torch.manual_seed(0)
    data = torch.cat([table.x, table.y[:, None].float()], 1)
    data = (data - data.mean(0)) / data.std(0).clamp_min(1e-6)
    counts = vfl.balanced_counts(31, 4)
    xs = list(data.split(counts, dim=1))
    model = VFLVAE([ClientEncoder(c, 8) for c in counts], ServerVAE(32), [ClientDecoder(8, c) for c in counts], 8)
    losses = vfl.train_vfl_vae(model, xs, epochs=40)
    assert losses[-1] < 0.7 * losses[0]


def _dvfl_worker(rank, world, steps):
    tab = SyntheticHeart().encoded()
    parts = vfl.balanced_partition(tab.columns, world)
    idx = tab.column_index(parts[rank])
    x_local, y = tab.x[:256, idx], tab.y[:256]
    dims = [2 * len(p) for p in parts]
    net = vfl.DistributedVFL(len(idx), dims, act_on_logits=False, seed=3)
    net.bottom.dropout.p = 0.0
    opt = torch.optim.SGD(net.parameters(), lr=0.1)
    losses = []
    for _ in range(steps):
        opt.zero_grad()
        loss = torch.nn.functional.cross_entropy(net(x_local), y)
        loss.backward()
        opt.step()
        losses.append(float(loss))
    return losses, {k: v.clone() for k, v in net.top_model.state_dict().items()}, {k: v.clone() for k, v in net.bottom.state_dict().items()}


@pytest.mark.timeout(150)
def test_distributed_vfl_equals_single_process():
    world, steps = 4, 3
    outs = run_distributed(_dvfl_worker, world, (steps,), timeout=120)
    # single-process restatement with the same per-party / top initialisation
    tab = SyntheticHeart().encoded()
    parts = vfl.balanced_partition(tab.columns, world)
    dims = [2 * len(p) for p in parts]
    bottoms = []
    for r in range(world):
        torch.manual_seed(3 + 1000 + r)
        b = BottomModel(len(parts[r]), dims[r]); b.dropout.p = 0.0
        bottoms.append(b)
    torch.manual_seed(3)
    from ddl25spring_b200.models.tabular import TopModel
    top = TopModel(dims, 2, act_on_logits=False)
    params = [p for b in bottoms for p in b.parameters()] + list(top.parameters())
    opt = torch.optim.SGD(params, lr=0.1)
    xs = [tab.x[:256, tab.column_index(p)] for p in parts]
    y = tab.y[:256]
    ref_losses = []
    for _ in range(steps):
        opt.zero_grad()
        loss = torch.nn.functional.cross_entropy(top([b(x) for b, x in zip(bottoms, xs)]), y)
        loss.backward(); opt.step(); ref_losses.append(float(loss))
    for r, (losses, top_sd, bot_sd) in enumerate(outs):
        assert losses == pytest.approx(ref_losses, rel=1e-5)
        for k, v in top.state_dict().items():
            torch.testing.assert_close(top_sd[k], v, rtol=1e-5, atol=1e-6)
        for k, v in bottoms[r].state_dict().items():
            torch.testing.assert_close(bot_sd[k], v, rtol=1e-5, atol=1e-6)


def _dvae_worker(rank, world, epochs):
    xs = list(torch.randn(200, 28, generator=torch.Generator().manual_seed(11)).split(7, dim=1))   # well-conditioned columns
    model = vfl.DistributedVFLVAE(7, world, 8, seed=5)
    torch.manual_seed(99)                                    # the reparameterisation noise stream, identical on every rank
    losses = vfl.train_distributed_vfl_vae(model, xs[rank], epochs=epochs)
    return losses, {k: v.clone() for k, v in model.state_dict().items()}


@pytest.mark.timeout(200)
def test_distributed_vfl_vae_equals_single_process():
    """One party per rank (gather of latents up, slice-scatter of reconstructed latents down, exercise_3.py:128-137) gives the
    same losses and the same weights as the single-process VFLVAE built from the same initialisation."""
    world, epochs = 4, 3
    outs = run_distributed(_dvae_worker, world, (epochs,), timeout=150)
    xs = list(torch.randn(200, 28, generator=torch.Generator().manual_seed(11)).split(7, dim=1))
    torch.set_num_threads(1)                                 # same reduction order as the single-threaded worker ranks
    encs, decs = [], []
    for r in range(world):
        torch.manual_seed(5 + 1000 + r)
        encs.append(ClientEncoder(7, 8)); decs.append(ClientDecoder(8, 7))
    torch.manual_seed(5)
    ref = VFLVAE(encs, ServerVAE(32), decs, 8)
    torch.manual_seed(99)
    ref_losses = vfl.train_vfl_vae(ref, xs, epochs=epochs)
    for r, (losses, sd) in enumerate(outs):
        assert losses == pytest.approx(ref_losses, rel=1e-4)
        for k, v in ref.server_vae.state_dict().items():
            torch.testing.assert_close(sd["server_vae." + k], v, rtol=1e-4, atol=1e-5)
        for k, v in ref.client_encoders[r].state_dict().items():
            torch.testing.assert_close(sd["encoder." + k], v, rtol=1e-4, atol=1e-5)
        for k, v in ref.client_decoders[r].state_dict().items():
            torch.testing.assert_close(sd["decoder." + k], v, rtol=1e-4, atol=1e-5)
