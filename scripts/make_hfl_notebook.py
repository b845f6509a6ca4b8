#!/usr/bin/env python
"""Write notebooks/horizontal_federated_learning.ipynb (the guided tutorial; counterpart of
lab/tutorial_1a/horizontal-federated-learning.ipynb cells 5-37) from the cell list below, stored clean (no outputs)."""
import json
import os

CELLS = []


def md(text):
    CELLS.append({"cell_type": "markdown", "metadata": {}, "source": text.strip("\n").splitlines(keepends=True)})


def code(text):
    CELLS.append({"cell_type": "code", "metadata": {}, "execution_count": None, "outputs": [],
                  "source": text.strip("\n").splitlines(keepends=True)})


md("""
# Horizontal federated learning on one (or eight) B200

The tutorial of `lab/tutorial_1a/horizontal-federated-learning.ipynb`, on `ddl25spring_b200`: the same client-local-train /
server-aggregate hooks, FedSGD and FedAvg, IID and non-IID splits, and the homework-1 sweeps - plus Byzantine clients and robust
aggregation.  On a GPU the FedAvg server trains all sampled clients in lock-step on hand-written sm_100a kernels; on a CPU the
same classes run the plain PyTorch statement of the same math, so every cell below runs anywhere.

Sections: 1 data and splits - 2 the model and one local epoch - 3 the two hooks (write your own client) - 4 centralised baseline -
5 FedSGD, twice - 6 FedAvg - 7 accuracy per round, plotted - 8 the homework sweeps - 9 attacks and defenses - 10 several ranks
from inside the notebook - 11 checkpoints and resume.
""")
code("""
import sys, os
sys.path.insert(0, os.path.abspath('..')); sys.path.insert(0, os.getcwd())
import tempfile
OUT = tempfile.mkdtemp()                                       # plots and per-rank logs of this session
import numpy as np
import pandas as pd
import torch
from ddl25spring_b200.core.data import SyntheticMNIST, split
from ddl25spring_b200.fl.server import CentralizedServer, FedSgdGradientServer, FedSgdWeightServer, FedAvgServer
ds = SyntheticMNIST(n_train=6000, n_test=1000, seed=0)          # MNIST-shaped synthetic data (no network on the box)
dev = 'cuda' if torch.cuda.is_available() else 'cpu'
LR, B, SEED, ROUNDS = 0.05, 100, 10, 3
dev
""")
md("""
## 1. Data and client splits

`split(dataset, nr_clients, iid, seed)` is the reference's `split` (`hfl_complete.py:91-104`): IID = a random permutation cut
into equal parts; non-IID = McMahan's pathological split - sort by label, cut into `2 * nr_clients` shards, two shards per client,
so most clients see only two digits.  The label histogram per client makes the difference visible.
""")
code("""
def label_table(subsets, n_show=5):
    rows = []
    for i, s in enumerate(subsets[:n_show]):
        y = ds.train.targets[np.asarray(s.indices)]
        rows.append(pd.Series(np.bincount(y.numpy(), minlength=10), name=f'client {i}'))
    return pd.DataFrame(rows)

iid_subsets, noniid_subsets = split(ds.train, 10, True, SEED), split(ds.train, 10, False, SEED)
assert sorted(i for s in iid_subsets for i in s.indices) == list(range(len(ds.train)))      # a partition of the training set
label_table(iid_subsets), label_table(noniid_subsets)
""")
md("""
## 2. The model and one local epoch

`MnistCnn` is the lab's network (conv 1-32, conv 32-64, max-pool, dropout, fc 9216-128, fc 128-10, log-softmax; 1 199 882
parameters).  One local epoch is the lab's `train_epoch` (`hfl_complete.py:71-80`): minibatches in a seeded order, NLL loss, SGD.
""")
code("""
from ddl25spring_b200.models.mnist_cnn import MnistCnn
from ddl25spring_b200.core.seeds import epoch_permutations
import torch.nn.functional as F

def train_epoch(model, x, y, optimizer, batch_size, seed):
    model.train()
    perm = epoch_permutations(len(y), seed, 1)[0]              # the DataLoader(shuffle=True, generator=g) order for this seed
    for lo in range(0, len(y), batch_size):
        idx = perm[lo:lo + batch_size]
        optimizer.zero_grad()
        loss = F.nll_loss(model(x[idx]), y[idx])
        loss.backward()
        optimizer.step()
    return float(loss)

torch.manual_seed(SEED)
m = MnistCnn()
print(sum(p.numel() for p in m.parameters()), 'parameters')
xs, ys = ds.train.images[:600], ds.train.targets[:600]
[round(train_epoch(m, xs, ys, torch.optim.SGD(m.parameters(), lr=LR), B, SEED + e), 3) for e in range(2)]
""")
md("""
## 3. The two hooks

Everything federated is two hooks:

* **client-local-train** - `Client.update(weights, seed) -> message` (`hfl_complete.py:145-155`).  A message is a list of tensors
  in `model.parameters()` order: gradients for FedSGD, weights for FedAvg.  On the device the same call exists as
  `update_flat(global_flat, seed, out=row)`, which writes the message straight into the client's row of the server's
  `[clients, parameters]` matrix.
* **server-aggregate** - `Server.run(nr_rounds) -> RunResult` samples clients, collects their messages, applies an
  `Aggregator` (`fl/aggregators.py`) and evaluates with `Server.test()`.

Exercise of the lab: write the FedAvg client yourself.  `MyWeightClient` below does what `WeightClient` does, using the
`train_epoch` of section 2, and is dropped into the stock server.
""")
code("""
from ddl25spring_b200.fl.client import Client, WeightClient

class MyWeightClient(Client):
    def __init__(self, data, lr, batch_size, nr_epochs, **kw):
        super().__init__(data, batch_size, **kw)
        self.lr, self.nr_epochs = lr, nr_epochs

    def update_flat(self, global_flat, seed, out=None):
        slot = self.slot                                   # borrowed model + flat parameter arena
        with torch.no_grad():
            slot.flat.copy_(global_flat)                   # server -> client
        opt = torch.optim.SGD(slot.model.parameters(), lr=self.lr)
        for e, perm in enumerate(epoch_permutations(self.n, seed, self.nr_epochs)):
            slot.model.train()
            for lo in range(0, self.n, self.batch_size):
                idx = perm[lo:lo + self.batch_size]
                opt.zero_grad()
                F.nll_loss(slot.model(self.x[idx]), self.y[idx]).backward()
                opt.step()
        msg = slot.flat.detach()
        return msg.clone() if out is None else out.copy_(msg)

nodrop = lambda: MnistCnn(p_drop1=0.0, p_drop2=0.0)
srv_mine = FedAvgServer(LR, B, iid_subsets, 0.3, 1, SEED, model_fn=nodrop, test_set=ds.test, device='cpu')
srv_mine.clients = [MyWeightClient(s, LR, B, 1, model_fn=nodrop, device='cpu', slot=srv_mine.worker_slot) for s in iid_subsets]
srv_stock = FedAvgServer(LR, B, iid_subsets, 0.3, 1, SEED, model_fn=nodrop, test_set=ds.test, device='cpu')
a, b = srv_mine.run(2), srv_stock.run(2)
assert a.test_accuracy == b.test_accuracy                  # the hand-written client is the stock client
a.as_df()
""")
md("## 4. Centralised baseline (`hfl_complete.py:193-216`): plain minibatch SGD over the whole set, one round = one epoch")
code("""
central = CentralizedServer(LR, B, SEED, train_set=ds.train, test_set=ds.test, device=dev).run(2)
central.as_df()
""")
md("""
## 5. FedSGD, twice (homework A1)

`FedSgdGradientServer` averages client **gradients** and takes one server SGD step; `FedSgdWeightServer` routes the same
gradients through the `.grad` arena and an optimizer step.  The homework shows their accuracies are identical round for round
(`lab/homework-1.ipynb` cells 13-18); so do the two frames below.
""")
code("""
sgd_g = FedSgdGradientServer(LR, iid_subsets, 0.5, SEED, test_set=ds.test, device=dev).run(ROUNDS)
sgd_w = FedSgdWeightServer(LR, iid_subsets, 0.5, SEED, test_set=ds.test, device=dev).run(ROUNDS)
diff = sgd_g.as_df()['Test accuracy'] - sgd_w.as_df()['Test accuracy']
assert diff.abs().max() == 0.0
sgd_g.as_df(), diff
""")
md("""
## 6. FedAvg: clients train locally for E epochs and send weights (`hfl_complete.py:336-390`)

`Message count` is the lab's `2 * (round + 1) * clients_per_round`; `wall_time` keeps the lab's *simulated-parallel* meaning,
`device_time_ms` (shown with `perf=True`) is the measured device time of the round.
""")
code("""
fedavg = FedAvgServer(LR, B, iid_subsets, 0.5, 1, SEED, test_set=ds.test, device=dev).run(ROUNDS)
fedavg.as_df(perf=True)
""")
md("## 7. Accuracy per round, plotted (the seaborn line plot of the lab's last cell, written as SVG - no plotting stack needed)")
code("""
from ddl25spring_b200.utils.plot import plot_runs
runs = pd.concat([central.as_df(), sgd_g.as_df(), fedavg.as_df()], ignore_index=True)
(plot_runs(runs, os.path.join(OUT, 'accuracy_per_round.svg'), hue=('Algorithm',), title='test accuracy per round'))
""")
md("""
## 8. The homework sweeps: clients per round C, local epochs E, IID vs non-IID (`lab/homework-1.ipynb` cells 24-36)

`examples/hfl_experiments.py --sweep` runs the full N x C x E / IID grid and writes one DataFrame plus the plots; here a small
slice of it inline.
""")
code("""
rows = []
for iid in (True, False):
    for E in (1, 2):
        srv = FedAvgServer(LR, B, split(ds.train, 10, iid, SEED), 0.5, E, SEED, test_set=ds.test, device=dev)
        df = srv.run(ROUNDS).as_df(); df['Iid'] = iid
        rows.append(df)
grid = pd.concat(rows, ignore_index=True)
(plot_runs(grid, os.path.join(OUT, 'sweep_E_iid.svg'), hue=('E', 'Iid'), title='FedAvg: local epochs and data heterogeneity'))
""")
md("""
## 9. Attacks and defenses, and the B200 path

20 % of the clients flip the sign of their update (`fl/attacks.py`); the server defends with Krum, the coordinate-wise median or
a trimmed mean (`fl/aggregators.py`; on a GPU each rule is one or two hand-written kernels).  `PackedFedAvgServer` has the same
constructor and hooks as `FedAvgServer` but trains all sampled clients of a GPU concurrently with grouped kernels.
""")
code("""
from ddl25spring_b200.fl.packed_server import PackedFedAvgServer      # GPU only: all sampled clients in lock-step kernels
from ddl25spring_b200.fl.aggregators import make_aggregator
from ddl25spring_b200.fl.attacks import SignFlip, choose_malicious
Srv = PackedFedAvgServer if dev == 'cuda' else FedAvgServer          # same constructor, same hooks
mal = choose_malicious(10, 0.2, 0)
rows = []
for name, kw in (('mean', {}), ('krum', dict(n_byzantine=1)), ('median', {}), ('trimmed_mean', {})):
    srv = Srv(LR, B, iid_subsets, 0.5, 1, SEED, test_set=ds.test, device=dev, aggregator=make_aggregator(name, **kw),
              attack=SignFlip(4.0), malicious=mal)
    df = srv.run(ROUNDS).as_df(); df['Defense'] = name
    rows.append(df)
(plot_runs(pd.concat(rows, ignore_index=True), os.path.join(OUT, 'defenses.svg'), hue=('Defense',), title='20 % sign-flip attackers'))
""")
md("""
## 10. Several ranks from inside the notebook

The lab forks its DP ranks from a notebook cell and redirects their output to `out{rank}.txt`
(`intro_DP_GA_notebook.ipynb` cells 0-1).  `utils.launch.spawn` does that with a timeout and error propagation: one rank per GPU
over NCCL when the machine has the GPUs, gloo ranks on the CPU otherwise.  Here two ranks run federated rounds with the sampled
clients dealt round-robin to the ranks.
""")
code("""
from ddl25spring_b200.utils.launch import spawn

def fl_rank(rank, world, n_train, n_test):
    import torch
    from ddl25spring_b200.core.data import SyntheticMNIST, split
    from ddl25spring_b200.fl.server import FedAvgServer
    d = SyntheticMNIST(n_train=n_train, n_test=n_test, seed=0)
    device = f'cuda:{rank}' if torch.cuda.is_available() and torch.cuda.device_count() >= world else 'cpu'
    srv = FedAvgServer(0.05, 100, split(d.train, 8, True, 10), 0.5, 1, 10, test_set=d.test, device=device)
    print(f'rank {rank}/{world} on {device}', flush=True)
    return srv.run(2).test_accuracy

if __name__ in ('__main__', '__notebook__'):
    import fl_rank_job                                      # spawned processes import the worker from a module on disk
    accs = spawn(fl_rank_job.fl_rank, 2, (len(ds.train), len(ds.test)), timeout=300, log_dir=OUT)
    assert accs[0] == accs[1]                               # every rank ends the round with the same global model
    print(accs[0], open(os.path.join(OUT, 'out1.txt')).read().strip())
""")
md("""
## 11. Checkpoints and resume

`run(nr_rounds, checkpointer=RoundCheckpointer(dir))` writes one atomic file per round (model `state_dict`, next round, RNG
states, the `RunResult` so far) and continues from it when it exists: kill after round 2, resume, same result.
""")
code("""
import shutil, tempfile
from ddl25spring_b200.core.checkpoint import RoundCheckpointer
from ddl25spring_b200.utils.trace import JsonlLogger, TRACER
make = lambda: FedAvgServer(LR, B, iid_subsets, 0.5, 1, SEED, test_set=ds.test, device='cpu')
ref = make().run(4)
d = tempfile.mkdtemp()
ck = RoundCheckpointer(d)
make().run(2, checkpointer=ck)                                       # "killed" after two rounds
resumed = make().run(4, checkpointer=ck, logger=JsonlLogger(os.path.join(d, 'rounds.jsonl')))
assert resumed.test_accuracy == ref.test_accuracy
shutil.rmtree(d)
pd.DataFrame(TRACER.report()).T                                      # where the host / device time of the rounds went
""")

nb = {"cells": CELLS, "metadata": {"kernelspec": {"display_name": "Python 3", "language": "python", "name": "python3"},
                                   "language_info": {"name": "python"}}, "nbformat": 4, "nbformat_minor": 5}
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "notebooks", "horizontal_federated_learning.ipynb")
with open(out, "w") as f:
    json.dump(nb, f, indent=1)
    f.write("\n")
print("wrote", os.path.normpath(out))
