"""Worker of section 10 of horizontal_federated_learning.ipynb: spawned ranks import their entry function from a module on
disk (functions defined in a notebook cell cannot be pickled into a spawned process)."""


def fl_rank(rank, world, n_train, n_test):
    import torch

    from ddl25spring_b200.core.data import SyntheticMNIST, split
    from ddl25spring_b200.fl.server import FedAvgServer
    d = SyntheticMNIST(n_train=n_train, n_test=n_test, seed=0)
    device = f"cuda:{rank}" if torch.cuda.is_available() and torch.cuda.device_count() >= world else "cpu"
    srv = FedAvgServer(0.05, 100, split(d.train, 8, True, 10), 0.5, 1, 10, test_set=d.test, device=device)
    print(f"rank {rank}/{world} on {device}", flush=True)
    return srv.run(2).test_accuracy
