"""Data-parallel gradient bucketing / overlap hooks and weight averaging over gloo."""
import pytest
import torch
import torch.nn.functional as F

from ddl25spring_b200.core.arena import FlatArena
from ddl25spring_b200.parallel.dp import DataParallel, WeightAveraging, make_buckets
from tests.dist_utils import run_distributed


def _mlp():
    torch.manual_seed(0)
    return torch.nn.Sequential(torch.nn.Linear(20, 64), torch.nn.ReLU(), torch.nn.Linear(64, 64), torch.nn.ReLU(),
                               torch.nn.Linear(64, 5))


def _data(rank, n=16):
    g = torch.Generator().manual_seed(100 + rank)
    return torch.randn(n, 20, generator=g), torch.randint(0, 5, (n,), generator=g)


def test_buckets_cover_arena_in_reverse_order():
    m = _mlp()
    arena = FlatArena.from_module(m, with_grad=True)
    buckets, owner = make_buckets(arena, 4096 * 4)
    assert len(buckets) > 1
    assert buckets[0].hi >= buckets[-1].hi and buckets[-1].lo == 0
    covered = sorted((b.lo, b.hi) for b in buckets)
    for (l0, h0), (l1, h1) in zip(covered, covered[1:]):
        assert h0 == l1
    assert set(owner) == {n for n, _ in m.named_parameters()}
    assert sum(b.n_params for b in buckets) == len(list(m.parameters()))


def _dp_worker(rank, world, opt):
    m = _mlp()
    dp = DataParallel(m, bucket_mb=0.004, optimizer=opt, lr=0.05)
    x, y = _data(rank)
    for _ in range(3):
        dp.train_step(lambda mod, a, b: F.cross_entropy(mod(a), b), x, y)
    return dp.arena.flat.clone(), len(dp.buckets)


@pytest.mark.timeout(120)
@pytest.mark.parametrize("opt", ["sgd", "adam"])
def test_dp_matches_single_process_on_the_union_batch(opt):
    world = 3
    outs = run_distributed(_dp_worker, world, (opt,), timeout=100)
    for flat, nb in outs[1:]:
        assert torch.equal(flat, outs[0][0]), "replicas must stay bit-identical"
    assert outs[0][1] > 1
    m = _mlp()
    arena = FlatArena.from_module(m, with_grad=True)
    o = (torch.optim.SGD if opt == "sgd" else torch.optim.Adam)(m.parameters(), lr=0.05)
    xs, ys = zip(*[_data(r) for r in range(world)])
    X, Y = torch.cat(xs), torch.cat(ys)
    for _ in range(3):
        o.zero_grad(set_to_none=False)
        F.cross_entropy(m(X), Y).backward()
        o.step()
    torch.testing.assert_close(outs[0][0], arena.flat, rtol=2e-4, atol=2e-6)


def _wa_worker(rank, world):
    m = _mlp()
    wa = WeightAveraging(m, lr=0.05)
    x, y = _data(rank)
    for _ in range(2):
        F.cross_entropy(m(x), y).backward()
        wa.step()
    return wa.arena.flat.clone()


@pytest.mark.timeout(120)
def test_weight_averaging_writes_the_average_back():
    outs = run_distributed(_wa_worker, 3, timeout=100)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[1], outs[2])
    # one step of plain SGD + averaging == SGD on the mean gradient, so after 2 steps it equals DP-SGD
    m = _mlp()
    arena = FlatArena.from_module(m, with_grad=True)
    xs, ys = zip(*[_data(r) for r in range(3)])
    o = torch.optim.SGD(m.parameters(), lr=0.05)
    for _ in range(2):
        o.zero_grad(set_to_none=False)
        F.cross_entropy(m(torch.cat(xs)), torch.cat(ys)).backward()
        o.step()
    torch.testing.assert_close(outs[0], arena.flat, rtol=2e-4, atol=2e-6)
