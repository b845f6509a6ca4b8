"""Horizontal-FL experiment grid (lab/homework-1.ipynb cells 6-36): FedSGD vs FedAvg over N, C, E, IID / non-IID.

  python examples/hfl_experiments.py --nr_rounds 10 [--engine packed|generic] [--aggregator krum --attack sign_flip ...]
Defaults are the notebook's: N=100, lr=0.01, C=0.1, E=1, B=100, 10 rounds, IID, seed 10 (cell 6)."""
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


def main(argv=None, n_train=60_000, n_test=10_000):
    cfg = FLConfig.from_args(argv)
    ds = SyntheticMNIST(n_train=n_train, n_test=n_test, seed=0)
    res = build_server(cfg, ds).run(cfg.nr_rounds)
    print(res.as_df(perf=True).to_string(index=False))
    return res


if __name__ == "__main__":
    main()
