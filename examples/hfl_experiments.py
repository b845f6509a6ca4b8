"""Horizontal-FL experiment grid (lab/homework-1.ipynb cells 6-36): FedSGD vs FedAvg over N, C, E, IID / non-IID.

  python examples/hfl_experiments.py --nr_rounds 10 [--engine packed|generic] [--aggregator krum --attack sign_flip ...]
      one configuration (defaults are the notebook's: N=100, lr=0.01, C=0.1, E=1, B=100, 10 rounds, IID, seed 10 - cell 6)
  python examples/hfl_experiments.py --sweep [--out results/hfl] [--nr_rounds 10] [--quick]
      the homework's sweeps in one go: N in {10, 50, 100} (cell 24), C in {0.01, 0.1, 0.2} (cell 26), E in {1, 2, 4} (cell 32),
      IID vs non-IID (cell 34) for FedSGD and FedAvg; writes ONE DataFrame (sweep.csv) and the accuracy-vs-round line plots
      of cells 24-36 as SVG files (sweep_N.svg, sweep_C.svg, sweep_E.svg, sweep_iid.svg)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl25spring_b200.core.config import FLConfig
from ddl25spring_b200.core.data import SyntheticMNIST, split
from ddl25spring_b200.fl.aggregators import make_aggregator
from ddl25spring_b200.fl.attacks import choose_malicious, make_attack
from ddl25spring_b200.fl.server import CentralizedServer, FedAvgServer, FedSgdGradientServer, FedSgdWeightServer


def build_server(cfg: FLConfig, ds):
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    kw = {}
    if cfg.aggregator != "mean":
        k = max(1, round(cfg.client_fraction * cfg.nr_clients))
        extra = dict(n_byzantine=max(1, int(cfg.malicious_fraction * k))) if "krum" in cfg.aggregator else {}
        kw["aggregator"] = make_aggregator(cfg.aggregator, **extra)
    if cfg.attack != "none":
        kw["attack"] = make_attack(cfg.attack)
        kw["malicious"] = choose_malicious(cfg.nr_clients, cfg.malicious_fraction, cfg.seed)
    if cfg.algorithm == "centralized":
        return CentralizedServer(cfg.lr, cfg.batch_size, cfg.seed, train_set=ds.train, test_set=ds.test, device=dev)
    subsets = split(ds.train, cfg.nr_clients, cfg.iid, cfg.seed)
    if cfg.algorithm == "fedsgd":
        return FedSgdGradientServer(cfg.lr, subsets, cfg.client_fraction, cfg.seed, test_set=ds.test, device=dev, **kw)
    if cfg.algorithm == "fedsgd_weight":
        return FedSgdWeightServer(cfg.lr, subsets, cfg.client_fraction, cfg.seed, test_set=ds.test, device=dev, **kw)
    use_packed = cfg.engine == "packed" or (cfg.engine == "auto" and dev == "cuda")
    if use_packed:
        from ddl25spring_b200.fl.packed_server import PackedFedAvgServer
        return PackedFedAvgServer(cfg.lr, cfg.batch_size, subsets, cfg.client_fraction, cfg.nr_local_epochs, cfg.seed,
                                  test_set=ds.test, device=dev, **kw)
    return FedAvgServer(cfg.lr, cfg.batch_size, subsets, cfg.client_fraction, cfg.nr_local_epochs, cfg.seed, test_set=ds.test,
                        device=dev, **kw)


def sweep(argv=None, n_train=60_000, n_test=10_000):
    """N x C x E / IID sweeps of lab/homework-1.ipynb cells 24-36; returns the concatenated DataFrame."""
    import argparse
    from dataclasses import replace

    import pandas as pd

    from ddl25spring_b200.utils.plot import plot_runs
    ap = argparse.ArgumentParser()
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--out", default="results/hfl")
    ap.add_argument("--nr_rounds", type=int, default=10)
    ap.add_argument("--quick", action="store_true", help="small grid / few clients (smoke test)")
    a = ap.parse_args(argv)
    os.makedirs(a.out, exist_ok=True)
    ds = SyntheticMNIST(n_train=n_train, n_test=n_test, seed=0)
    base = FLConfig(nr_rounds=a.nr_rounds)
    Ns, Cs, Es = ((10, 50, 100), (0.01, 0.1, 0.2), (1, 2, 4)) if not a.quick else ((4, 8), (0.25, 0.5), (1, 2))
    if a.quick:
        base = replace(base, nr_clients=8, client_fraction=0.5, batch_size=50)
    grid = [("N", replace(base, algorithm=alg, nr_clients=n)) for alg in ("fedsgd", "fedavg") for n in Ns]
    grid += [("C", replace(base, algorithm=alg, client_fraction=c)) for alg in ("fedsgd", "fedavg") for c in Cs]
    grid += [("E", replace(base, algorithm="fedavg", nr_local_epochs=e)) for e in Es]
    grid += [("iid", replace(base, algorithm=alg, iid=iid)) for alg in ("fedsgd", "fedavg") for iid in (True, False)]
    frames = []
    for sweep_name, cfg in grid:
        res = build_server(cfg, ds).run(cfg.nr_rounds)
        df = res.as_df(perf=True)
        df["Sweep"], df["Iid"] = sweep_name, cfg.iid
        frames.append(df)
        print(f"[{sweep_name}] {cfg.algorithm} N={cfg.nr_clients} C={cfg.client_fraction} E={cfg.nr_local_epochs} iid={cfg.iid}: "
              f"final accuracy {res.test_accuracy[-1]:.2f} %", flush=True)
    out = pd.concat(frames, ignore_index=True)
    out.to_csv(os.path.join(a.out, "sweep.csv"), index=False)
    for name, hue in (("N", ("Algorithm", "N")), ("C", ("Algorithm", "C")), ("E", ("Algorithm", "E")), ("iid", ("Algorithm", "Iid"))):
        plot_runs(out[out["Sweep"] == name], os.path.join(a.out, f"sweep_{name}.svg"), hue=hue,
                  title=f"test accuracy per round, sweep over {name}")
    return out


def main(argv=None, n_train=60_000, n_test=10_000):
    if "--sweep" in (argv if argv is not None else sys.argv[1:]):
        return sweep(argv, n_train, n_test)
    cfg = FLConfig.from_args(argv)
    ds = SyntheticMNIST(n_train=n_train, n_test=n_test, seed=0)
    res = build_server(cfg, ds).run(cfg.nr_rounds)
    print(res.as_df(perf=True).to_string(index=False))
    return res


if __name__ == "__main__":
    main()
