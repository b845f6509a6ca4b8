"""Smoke tests of the example drivers (tiny sizes, CPU)."""
import subprocess
import sys
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_hfl_experiment_driver():
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import hfl_experiments
    res = hfl_experiments.main(["--nr_clients", "6", "--client_fraction", "0.5", "--nr_rounds", "2", "--batch_size", "50",
                                "--engine", "generic", "--aggregator", "median", "--attack", "sign_flip",
                                "--malicious_fraction", "0.2"], n_train=600, n_test=200)
    assert len(res.test_accuracy) == 2 and res.message_count == [6, 12]


def test_llm_primer_loss_decreases():
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import llm_primer
    losses = llm_primer.main(iters=12, dmodel=48, num_heads=4, n_layers=2, seq_l=32, batch_size=3, vocab=256, device="cpu", log=None)
    assert losses[-1] < losses[0]


@pytest.mark.timeout(300)
def test_dp_and_pp_launch_scripts():
    env = dict(os.environ, PORT="29611")
    r = subprocess.run(["bash", "scripts/run_dp.sh", "2", "--mode", "wa", "--iters", "2", "--dmodel", "48", "--vocab", "128"], cwd=ROOT,
                       capture_output=True, text=True, timeout=200, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    env["PORT"] = "29612"
    r = subprocess.run(["bash", "scripts/run_pp.sh", "4", "--schedule", "1f1b", "--pipelines", "2", "--micro", "2", "--iters", "2",
                        "--dmodel", "48", "--vocab", "128"], cwd=ROOT, capture_output=True, text=True, timeout=200, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "elapsed" in r.stdout


def test_robust_resnet_fl_example_runs_on_cpu():
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import robust_resnet_fl
    res = robust_resnet_fl.main(["--aggregator", "median", "--rounds", "1", "--clients", "3", "--samples", "16", "--batch", "8",
                                 "--malicious_fraction", "0.34"])
    assert len(res.test_accuracy) == 1 and res.test_accuracy[0] == res.test_accuracy[0]
