#!/usr/bin/env python
"""Write notebooks/data_and_pipeline_parallel.ipynb (counterpart of lab/tutorial_1b/DP/gradient_aggr/intro_DP_GA_notebook.ipynb,
extended to weight aggregation and the pipeline labs) from the cell list below, stored clean (no outputs)."""
import json
import os

CELLS = []


def md(text):
    CELLS.append({"cell_type": "markdown", "metadata": {}, "source": text.strip("\n").splitlines(keepends=True)})


def code(text):
    CELLS.append({"cell_type": "code", "metadata": {}, "execution_count": None, "outputs": [],
                  "source": text.strip("\n").splitlines(keepends=True)})


md("""
# Data and pipeline parallelism, launched from a notebook

The reference starts its data-parallel ranks from a notebook with `multiprocessing` and per-rank `out{rank}.txt` files
(`lab/tutorial_1b/DP/gradient_aggr/intro_DP_GA_notebook.ipynb`), and its pipeline labs from bash loops.  Here every experiment
is one call of `ddl25spring_b200.utils.launch.spawn(fn, world, args)`: `world` processes rendezvous on 127.0.0.1 (NCCL with one
GPU per rank when the machine has enough GPUs, gloo on CPU otherwise), per-rank results come back as Python values, a failing
rank raises here with its traceback, and a deadlock is a time-out instead of a hung kernel.

Sections: 1 the model and one process - 2 gradient aggregation - 3 weight aggregation - 4 pipeline schedules as data -
5 a 1F1B pipeline on three ranks - 6 two pipelines x two stages (the hybrid the reference's own attempt deadlocks on).
""")
code("""
import sys, os
sys.path.insert(0, os.path.abspath('..')); sys.path.insert(0, os.getcwd())
import tempfile
OUT = tempfile.mkdtemp()                                       # per-rank logs and plots of this session
import torch
from ddl25spring_b200.utils.launch import spawn
from ddl25spring_b200.utils.plot import line_plot
import parallel_rank_jobs as jobs                              # the rank programs live in a module (spawned processes import it)
jobs.VOCAB, jobs.DMODEL, jobs.LAYERS, jobs.CTX
""")
md("""
## 1. The model, on one process

A small LLaMA (`models/llama.py`: RMSNorm, rotary attention, gated MLP - the architecture `simplellm` provides to the labs) on a
synthetic token stream.  The labs' size is 6 layers x 288 (`primer/intro.py:13-17`); the notebook uses a toy size so that CPU
ranks finish in seconds.
""")
code("""
from ddl25spring_b200.core.data import SyntheticTokens
from ddl25spring_b200.models.llama import LLama, causalLLMLoss
torch.manual_seed(0)
net = LLama(None, jobs.VOCAB, jobs.DMODEL, jobs.HEADS, 'cpu', jobs.LAYERS, jobs.CTX)
opt = torch.optim.Adam(net.parameters(), lr=3e-3)
ds = iter(SyntheticTokens(jobs.VOCAB, 2, jobs.CTX))
single = []
for _ in range(4):
    x = next(ds)
    loss = causalLLMLoss(net(x), x, jobs.VOCAB)
    opt.zero_grad(); loss.backward(); opt.step()
    single.append(float(loss.detach()))
single
""")
md("""
## 2. Data parallel: gradient aggregation

Every rank runs forward and backward on its own slice of the data; the gradients are summed over ranks and divided by the
world size before the optimizer step (`intro_DP_GA.py:53-67`).  `DataParallel` keeps parameters and gradients in flat arenas,
reduces them bucket by bucket from autograd hooks while the backward pass is still running and applies one fused optimizer
kernel; on GPUs the reduce, the optimizer step and the parameter broadcast are ONE peer-memory kernel per bucket inside the
captured step.  Replicas that start equal must stay equal: the parameter checksums of the ranks are compared below.
""")
code("""
res = spawn(jobs.dp_rank, 2, args=('grad', 4), log_dir=OUT, timeout=300)
losses = [r[0] for r in res]; sums = [r[1] for r in res]
assert abs(sums[0] - sums[1]) < 1e-6 * max(1.0, abs(sums[0])), sums      # the replicas are still identical
print(open(os.path.join(OUT, 'out0.txt')).read())                        # the reference's out{rank}.txt
losses
""")
md("""
## 3. Data parallel: weight aggregation

Each rank takes its own optimizer step and the WEIGHTS are averaged afterwards (`intro_DP_WA.py:52-67`).  The reference computes
the average and assigns it to the loop variable (`:67`), so its replicas silently drift apart; `WeightAveraging` writes the
average back.
""")
code("""
res = spawn(jobs.dp_rank, 2, args=('weight', 4), log_dir=OUT, timeout=300)
assert abs(res[0][1] - res[1][1]) < 1e-6 * max(1.0, abs(res[0][1]))
[r[0] for r in res]
""")
md("""
## 4. Pipeline schedules are data

A schedule is a list of `F`/`B` operations per stage (`parallel/pp.py`).  `check_schedule` plays all stages against each other
with blocking-receive semantics and reports the number of time steps - or raises on a deadlock - before any process is started;
`p2p_program` lists the sends and receives a stage will issue.  GPipe runs all forwards, then all backwards; 1F1B starts
backwards as soon as the first micro-batch reaches the last stage, which bounds the activations a stage holds.
""")
code("""
from ddl25spring_b200.parallel.pp import SCHEDULES, check_schedule, p2p_program
for name in ('gpipe', '1f1b'):
    print(name, 'time steps for 3 stages x 3 micro-batches:', check_schedule(name, 3, 3))
    for s in range(3):
        print('  stage', s, ' '.join(f'{op.kind}{op.mb}' for op in SCHEDULES[name](3, s, 3)))
p2p_program('1f1b', 3, 1, 3)[:6]
""")
code("""
# a schedule that cannot work - the last stage wants its backward before its forward - is caught by the simulator, not by a hang
from ddl25spring_b200.parallel.pp import Op
SCHEDULES['broken'] = lambda n_stages, stage, n_micro: ([Op('B', 0), Op('F', 0)] if stage == n_stages - 1 else [Op('F', 0), Op('B', 0)])
try:
    check_schedule('broken', 3, 1)
    caught = None
except RuntimeError as e:
    caught = str(e)
del SCHEDULES['broken']
assert caught is not None
caught
""")
md("""
## 5. A 1F1B pipeline on three ranks

The model is cut into three stages (`LLama.split_stages`); activations travel forward and gradients backward as point-to-point
messages (`dist.isend` / `irecv`, one NCCL communicator per direction on GPUs).  The gradients equal those of the unsplit model
(tests/test_pp_cpu.py); the last stage returns the mean micro-batch loss per iteration.
""")
code("""
res = spawn(jobs.pp_rank, 3, args=('1f1b', 3, 1, 4), log_dir=OUT, timeout=300)
pp_losses = res[-1]
svg = line_plot({'1 process': (list(range(len(single))), single), '3-stage 1F1B': (list(range(len(pp_losses))), pp_losses)},
                os.path.join(OUT, 'loss.svg'), title='training loss', xlabel='iteration', ylabel='loss')
assert res[0] is None and res[1] is None and len(pp_losses) == 4
pp_losses
""")
md("""
## 6. Two pipelines x two stages

The hybrid of the homework (`intro_PP_1F1B_MP.py`): ranks 0-1 form one pipeline, ranks 2-3 a second one on different data, and
after every iteration each stage averages its gradients with the same stage of the other pipeline.  The reference's attempt
never completes an iteration (`lab/Abgabe/outputs/out_MP0.txt`); here the mesh creates every group on every rank in the same
order and the per-stage reduction is ONE collective over the stage's flat gradient arena.
""")
code("""
res = spawn(jobs.pp_rank, 4, args=('1f1b', 2, 2, 3), log_dir=OUT, timeout=300)
hyb = [r for r in res if r is not None]
assert len(hyb) == 2 and all(len(h) == 3 for h in hyb)        # both pipelines finished all iterations
hyb
""")

nb = {"cells": CELLS, "metadata": {"kernelspec": {"display_name": "Python 3", "language": "python", "name": "python3"},
                                   "language_info": {"name": "python"}}, "nbformat": 4, "nbformat_minor": 5}
path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "notebooks", "data_and_pipeline_parallel.ipynb")
with open(path, "w") as f:
    json.dump(nb, f, indent=1)
    f.write("\n")
print("wrote", path, len(CELLS), "cells")
