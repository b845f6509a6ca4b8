"""Byzantine attacks and robust aggregation rules against hand-computable cases and fp64 oracles."""
import numpy as np
import pytest
import torch

from ddl25spring_b200 import ops
from ddl25spring_b200.core.data import SyntheticMNIST, split
from ddl25spring_b200.fl.aggregators import CoordinateMedian, Krum, TrimmedMean, WeightedMean, make_aggregator
from ddl25spring_b200.fl.attacks import NoiseAttack, ScaledUpdate, SignFlip, choose_malicious
from ddl25spring_b200.fl.server import FedAvgServer
from ddl25spring_b200.models.mnist_cnn import MnistCnn
from ddl25spring_b200.ops import reference as ref


def test_hand_cases():
    u = torch.tensor([[1.0, 10.0], [2.0, 20.0], [3.0, 30.0], [100.0, -100.0]])
    w = torch.full((4,), 0.25)
    assert torch.allclose(WeightedMean()(u, w), torch.tensor([26.5, -10.0]))
    assert torch.allclose(CoordinateMedian()(u, w), torch.tensor([2.5, 15.0]))
    assert torch.allclose(CoordinateMedian(lower=True)(u, w), u.median(0).values)
    assert torch.allclose(TrimmedMean(trim=1)(u, w), torch.tensor([2.5, 15.0]))
    k = Krum(n_byzantine=0)                                              # m = K - f - 2 = 2 nearest neighbours
    out = k(u, w)
    assert int(k.last_selected[0]) == 1 and torch.allclose(out, u[1])       # middle honest point wins
    mk = Krum(n_byzantine=0, multi=3)
    assert torch.allclose(mk(u, w), u[:3].mean(0))
    assert TrimmedMean(beta=0.2).trim_count(64) == 13 and TrimmedMean(beta=0.2).trim_count(3) == 1


@pytest.mark.parametrize("K,f", [(10, 2), (64, 12)])
def test_krum_matches_cdist_oracle(K, f):
    g = torch.Generator().manual_seed(K)
    u = torch.randn(K, 500, generator=g)
    u[:f] *= -4
    scores = ref.krum_scores(u, f)
    d = torch.cdist(u.double(), u.double()).pow(2)
    m = K - f - 2
    exp = torch.stack([torch.sort(torch.cat([d[i, :i], d[i, i + 1:]])).values[:m].sum() for i in range(K)])
    torch.testing.assert_close(scores, exp)
    sel, w, _ = ops.krum_select(u, f, 1)
    assert int(sel[0]) >= f and float(w.sum()) == pytest.approx(1.0)
    G = ref.gram(u.double())
    torch.testing.assert_close(ref.pairwise_sq_dists_from_gram(G), d, rtol=1e-9, atol=1e-9)


def test_attacks():
    g, u = torch.zeros(8), torch.ones(8)
    assert torch.allclose(SignFlip(2.0)(u, g, client=0, nr_round=0), -2 * u)
    assert torch.allclose(SignFlip(1.0)(u, None, client=0, nr_round=0), -u)          # gradient message
    assert torch.allclose(ScaledUpdate(10.0)(u, g, client=0, nr_round=0), 10 * u)
    n = NoiseAttack(0.5, seed=1)(u, g, client=3, nr_round=2)
    assert (n - u).abs().max() <= 0.5 + 1e-6 and not torch.equal(n, u)
    m = choose_malicious(64, 0.2, 0)
    assert m.sum() == 13 and m.dtype == bool
    assert make_aggregator("multi-krum", n_byzantine=2, multi=3).multi == 3


def test_robust_fedavg_survives_sign_flip_where_mean_does_not():
    ds = SyntheticMNIST(n_train=1500, n_test=300, seed=5, noise=0.7)
    subsets = split(ds.train, 10, True, 10)
    mal = np.zeros(10, dtype=bool)
    mal[[1, 4]] = True                                                            # 20 % attackers
    accs = {}
    for name, agg in [("mean", WeightedMean()), ("median", CoordinateMedian()), ("krum", Krum(2, multi=4)),
                      ("trimmed", TrimmedMean(beta=0.2))]:
        srv = FedAvgServer(0.05, 50, subsets, 1.0, 1, 10, model_fn=lambda: MnistCnn(0.0, 0.0), test_set=ds.test, device="cpu",
                           aggregator=agg, attack=SignFlip(4.0), malicious=mal)
        accs[name] = srv.run(3).test_accuracy[-1]
    assert accs["median"] > accs["mean"] + 10 and accs["krum"] > accs["mean"] + 10 and accs["trimmed"] > accs["mean"] + 10, accs
