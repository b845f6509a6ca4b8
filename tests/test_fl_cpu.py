"""FL plumbing on CPU (BASELINE.json config 1; SURVEY §4 'Unit (CPU, no GPU)')."""
import copy

import numpy as np
import pytest
import torch
import torch.nn.functional as F
from torch.utils.data import DataLoader, Subset

from ddl25spring_b200.core.data import SyntheticMNIST, split, split_indices
from ddl25spring_b200.core.results import RunResult, message_count
from ddl25spring_b200.core.seeds import client_round_seed, clients_per_round, epoch_permutations
from ddl25spring_b200.fl.server import CentralizedServer, FedAvgServer, FedSgdGradientServer, FedSgdWeightServer
from ddl25spring_b200.models.mnist_cnn import MnistCnn


@pytest.fixture(scope="module")
def mnist():
    return SyntheticMNIST(n_train=1200, n_test=400, seed=3, noise=0.7)


def nodrop():
    return MnistCnn(p_drop1=0.0, p_drop2=0.0)


def test_split_iid_is_a_partition(mnist):
    parts = split_indices(mnist.train.targets.numpy(), 10, True, 0)
    allidx = np.sort(np.concatenate(parts))
    assert np.array_equal(allidx, np.arange(1200))
    assert max(map(len, parts)) - min(map(len, parts)) <= 1


def test_split_non_iid_two_shards_per_client(mnist):
    t = mnist.train.targets.numpy()
    parts = split_indices(t, 20, False, 1)
    assert np.array_equal(np.sort(np.concatenate(parts)), np.arange(1200))
    # each client holds 2 label-sorted shards -> at most 4 distinct labels, usually 2
    assert max(len(np.unique(t[p])) for p in parts) <= 4
    assert np.mean([len(np.unique(t[p])) for p in parts]) < 3.5


def test_seed_and_message_formulas():
    assert client_round_seed(10, 4, 3, 10) == 10 + 4 + 1 + 30
    assert clients_per_round(100, 0.1) == 10 and clients_per_round(100, 0.001) == 1
    assert [message_count(r, 10) for r in range(3)] == [20, 40, 60]


def test_epoch_permutations_match_dataloader():
    ds = torch.utils.data.TensorDataset(torch.arange(97))
    g = torch.Generator()
    dl = DataLoader(ds, batch_size=16, shuffle=True, generator=g)
    g.manual_seed(77)
    got = [torch.cat([b[0] for b in dl]) for _ in range(3)]
    for a, b in zip(got, epoch_permutations(97, 77, 3)):
        assert torch.equal(a, b)


def test_run_result_dataframe():
    r = RunResult("FedAvg", 10, 0.1, -1, 1, 0.01, 10)
    r.wall_time += [0.1, 0.2]
    r.message_count += [2, 4]
    r.test_accuracy += [10.0, 20.0]
    df = r.as_df()
    assert list(df["Round"]) == [1, 2] and "Wall time" not in df.columns
    assert df["B"].iloc[0] == "\N{INFINITY}" and "\N{GREEK SMALL LETTER ETA}" in df.columns


def test_fedsgd_weight_equals_fedsgd_gradient(mnist):
    """The reference's only correctness check (lab/homework-1.ipynb cells 13-18)."""
    subsets = split(mnist.train, 10, True, 10)
    kw = dict(model_fn=nodrop, test_set=mnist.test, device="cpu")
    a = FedSgdGradientServer(0.05, subsets, 0.5, 10, **kw).run(3)
    b = FedSgdWeightServer(0.05, subsets, 0.5, 10, **kw).run(3)
    assert a.test_accuracy == b.test_accuracy
    assert a.message_count == b.message_count == [10, 20, 30]


def _reference_fedavg(mnist, subsets, lr, B, C, E, seed, rounds):
    """Independent restatement of the reference algorithm with a real DataLoader and torch.optim.SGD."""
    torch.manual_seed(seed)
    glob = nodrop()
    rng = np.random.default_rng(seed)
    K = max(1, round(C * len(subsets)))
    accs = []
    for r in range(rounds):
        weights = [p.detach().clone() for p in glob.parameters()]
        chosen = rng.choice(len(subsets), K, replace=False)
        tot = sum(len(subsets[i]) for i in chosen)
        acc = [torch.zeros_like(w) for w in weights]
        for c in chosen:
            m = nodrop()
            with torch.no_grad():
                for p, w in zip(m.parameters(), weights):
                    p.copy_(w)
            g = torch.Generator()
            dl = DataLoader(subsets[int(c)], batch_size=B, shuffle=True, generator=g)
            g.manual_seed(seed + int(c) + 1 + r * K)
            opt = torch.optim.SGD(m.parameters(), lr=lr)
            m.train()
            for _ in range(E):
                for x, y in dl:
                    opt.zero_grad()
                    F.nll_loss(m(x), y).backward()
                    opt.step()
            for a, p in zip(acc, m.parameters()):
                a += len(subsets[int(c)]) / tot * p.detach()
        with torch.no_grad():
            for p, a in zip(glob.parameters(), acc):
                p.copy_(a)
        glob.eval()
        with torch.no_grad():
            pred = glob(mnist.test.images).argmax(1)
        accs.append(100.0 * int((pred == mnist.test.targets).sum()) / len(mnist.test))
    return accs, [p.detach() for p in glob.parameters()]


def test_fedavg_matches_reference_algorithm(mnist):
    subsets = split(mnist.train, 6, True, 10)
    srv = FedAvgServer(0.05, 50, subsets, 0.5, 2, 10, model_fn=nodrop, test_set=mnist.test, device="cpu")
    res = srv.run(2)
    accs, params = _reference_fedavg(mnist, subsets, 0.05, 50, 0.5, 2, 10, 2)
    for p, q in zip(srv.model.parameters(), params):
        torch.testing.assert_close(p.detach(), q, rtol=1e-4, atol=1e-5)
    assert res.test_accuracy == pytest.approx(accs, abs=0.51)
    assert res.message_count == [6, 12]


def test_fedavg_learns_and_list_api(mnist):
    subsets = split(mnist.train, 10, True, 10)
    srv = FedAvgServer(0.05, 50, subsets, 0.3, 1, 10, test_set=mnist.test, device="cpu")
    before = srv.test()
    res = srv.run(3)
    assert res.test_accuracy[-1] > before + 20
    # message-level API parity: list of tensors in, list of tensors out
    w = [p.detach().clone() for p in srv.model.parameters()]
    out = srv.clients[0].update(w, seed=1)
    assert [o.shape for o in out] == [x.shape for x in w] and all(o.device.type == "cpu" for o in out)


def test_centralized_runs(mnist):
    srv = CentralizedServer(0.05, 100, 10, train_set=mnist.train, test_set=mnist.test, device="cpu")
    res = srv.run(2)
    assert res.message_count == [0, 0] and res.test_accuracy[-1] > 30


def test_killed_run_resumes_to_the_same_result(mnist, tmp_path):
    """Round-level checkpoint / resume: a run interrupted after round 2 and continued by a NEW server object
    from the checkpoint file gives the same RunResult and the same final weights as the uninterrupted run - with dropout
    on, so torch's RNG state has to be part of the checkpoint, and with a JSONL logger attached."""
    import json

    from ddl25spring_b200.core.checkpoint import RoundCheckpointer
    from ddl25spring_b200.utils.trace import JsonlLogger, TRACER

    subsets = split(mnist.train, 8, True, 10)

    def make():
        return FedAvgServer(0.05, 50, subsets, 0.5, 1, 10, test_set=mnist.test, device="cpu")

    full = make()
    ref = full.run(4)
    ck = RoundCheckpointer(str(tmp_path / "ck"), every=1)
    log = str(tmp_path / "rounds.jsonl")
    first = make()
    part = first.run(2, checkpointer=ck, logger=JsonlLogger(log, rank=0))        # "killed" after round 2
    assert part.test_accuracy == ref.test_accuracy[:2]
    del first
    second = make()                                                             # fresh process state: new model, new RNGs
    res = second.run(4, checkpointer=ck, logger=JsonlLogger(log, rank=0))
    assert res.test_accuracy == ref.test_accuracy and res.message_count == ref.message_count
    assert torch.equal(second.arena.flat, full.arena.flat)
    recs = [json.loads(l) for l in open(log)]
    assert [r["round"] for r in recs] == [1, 2, 3, 4] and recs[-1]["test_accuracy"] == ref.test_accuracy[-1]
    rep = TRACER.report()
    assert rep["server/aggregate"]["calls"] >= 4 and rep["clients/local_update"]["calls"] >= 4
