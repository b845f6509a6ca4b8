"""ddl25spring_b200 — a B200-native (sm_100a) federated / distributed training framework.

Capability parity target: the DDL25Spring labs (horizontal FL, DP, PP, vertical FL,
federated generative models, Byzantine attacks + robust aggregation).  See SURVEY.md
for the component inventory and DESIGN.md for how each one maps onto this package.

Layout
------
core/      flat parameter arenas, seeds, RunResult, synthetic data, split(), checkpoints
models/    MnistCnn, heart MLP, tabular VAE, VFL bottom/top, VFL-VAE, ResNet-18, DCGAN, LLaMA
ops/       ctypes bindings to the hand-written sm_100a kernels (+ torch fp32 oracles)
fl/        Client / Server hooks, FedSGD / FedAvg / Centralized, attacks, robust aggregators
parallel/  DP (bucketed gradient all-reduce, weight averaging), PP schedules, PPxDP mesh,
           symmetric peer memory, vertical-FL party parallelism
gen/       VAE / VFL-VAE / federated DCGAN training loops and the synthetic-data utility eval
utils/     timing, logging, clocks sampling, launch helpers
csrc/      CUDA sources (built in-tree into libddl25_kernels.so by `build.py`)
"""

__version__ = "0.1.0"

from .core.results import RunResult  # noqa: F401
from .core.arena import FlatArena  # noqa: F401
