"""Generative workloads on CPU: tabular VAE utility eval, federated DCGAN, checkpoints, tracing."""
import os

import numpy as np
import pytest
import torch

from ddl25spring_b200.core import checkpoint
from ddl25spring_b200.core.data import SyntheticHeart, train_test_split_tensor
from ddl25spring_b200.gen.fed_dcgan import FederatedDCGAN
from ddl25spring_b200.gen.vae import synthetic_data_utility, train_evaluator
from ddl25spring_b200.models.mnist_cnn import MnistCnn
from ddl25spring_b200.utils.trace import TRACER, Watchdog, span


def test_evaluator_and_synthetic_utility():
    tab = SyntheticHeart().encoded()
    xtr, ytr, xte, yte = train_test_split_tensor(tab.x, tab.y)
    _, acc, hist = train_evaluator(xtr, ytr, xte, yte, epochs=30)
    assert acc > 0.6 and len(hist) == 30
    res = synthetic_data_utility(xtr, ytr, xte, yte, vae_epochs=15, eval_epochs=20)
    assert set(res) == {"real", "synthetic"} and res["real"] > 0.6 and 0.3 <= res["synthetic"] <= 1.0


def test_federated_dcgan_round_averages_g_and_d():
    torch.manual_seed(0)
    clients = [torch.tanh(torch.randn(8, 3, 64, 64)) for _ in range(3)]
    fed = FederatedDCGAN(clients, client_fraction=1.0, local_steps=1, batch_size=4, device="cpu", ngf=8, ndf=8)
    g0 = fed.glob.ga.flat.clone()
    hist = fed.run(1)
    assert len(hist) == 1 and all(np.isfinite(hist[0]))
    assert not torch.equal(g0, fed.glob.ga.flat)
    assert fed.sample(2).shape == (2, 3, 64, 64)
    # BN running stats are part of the averaged message
    assert any(s.is_buffer for s in fed.glob.ga.specs)


def test_checkpoint_roundtrip_and_resume(tmp_path):
    m = MnistCnn()
    p = os.path.join(tmp_path, "m.pt")
    checkpoint.save(m, p)
    m2 = MnistCnn()
    checkpoint.load(m2, p)
    for a, b in zip(m.parameters(), m2.parameters()):
        assert torch.equal(a, b)
    snap = checkpoint.snapshot(m)
    with torch.no_grad():
        m.fc2.bias.add_(1.0)
    assert not torch.equal(snap["fc2.bias"], m.fc2.bias)          # a copy, not an alias
    from ddl25spring_b200.core.arena import FlatArena
    from ddl25spring_b200.core.results import RunResult
    arena = FlatArena.from_module(m2)
    ck = checkpoint.RoundCheckpointer(str(tmp_path), every=1)
    rng = np.random.default_rng(5)
    rng.random(3)
    res = RunResult("FedAvg", 10, 0.1, 50, 1, 0.01, 5, wall_time=[0.1], message_count=[2], test_accuracy=[50.0])
    ck.maybe_save(0, arena, rng, res)
    expect_next = rng.random()
    rng2 = np.random.default_rng(0)
    arena2 = FlatArena.from_module(MnistCnn())
    start, res2 = ck.resume(arena2, rng2)
    assert start == 1 and res2.test_accuracy == [50.0] and rng2.random() == expect_next
    assert torch.equal(arena.flat, arena2.flat)


def test_tracer_and_watchdog():
    with span("unit"):
        pass
    assert TRACER.report()["unit"]["calls"] >= 1
    fired = []
    wd = Watchdog(0.2, on_timeout=lambda: fired.append(1)).start()
    import time
    time.sleep(0.6)
    wd.stop()
    assert fired
