"""Synthetic datasets of the labs' shapes, and the client sharding function.

There is no network in the build or benchmark environment, so every workload runs on
synthetic data that has the *shape and statistics* of the lab data and enough class
structure that "accuracy goes up when you train" is a meaningful test:

============================  ==========================================  =========================
reference data                 cite                                        synthetic stand-in
============================  ==========================================  =========================
MNIST 60k/10k, 1x28x28         lab/tutorial_1a/hfl_complete.py:19-31       :class:`SyntheticMNIST`
CIFAR-10-shaped 3x32x32        BASELINE.json config 2/4 (north star)       :class:`SyntheticCIFAR`
heart.csv 1025 x (13+target)   lab/tutorial_2a/centralized.py:31-44        :class:`SyntheticHeart`
TinyStories token stream       lab/tutorial_1b/primer/intro.py:20          :class:`SyntheticTokens`
64x64 RGB images for DCGAN     BASELINE.json config 5 (north star)         :class:`SyntheticImages64`
============================  ==========================================  =========================
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import numpy.random as npr
import torch
from torch.utils.data import Dataset

MNIST_MEAN, MNIST_STD = 0.1307, 0.3081


# --------------------------------------------------------------------------- image sets
class TensorImageSet(Dataset):
    """An in-memory ``(images, targets)`` dataset; ``images`` are already normalised fp32."""

    def __init__(self, images: torch.Tensor, targets: torch.Tensor):
        assert len(images) == len(targets)
        self.images = images
        self.targets = targets

    def __len__(self) -> int:
        return len(self.targets)

    def __getitem__(self, i):
        return self.images[i], int(self.targets[i])

    def pin(self) -> "TensorImageSet":
        if torch.cuda.is_available() and not self.images.is_pinned():
            self.images = self.images.pin_memory()
            self.targets = self.targets.pin_memory()
        return self


def _class_templates(rng: npr.Generator, n_classes: int, shape: tuple[int, ...], smooth: int = 3) -> np.ndarray:
    """Smooth random per-class prototypes (low-pass filtered noise)."""
    t = rng.standard_normal((n_classes, *shape)).astype(np.float32)
    for _ in range(smooth):  # cheap separable box blur along the two spatial axes
        t = (t + np.roll(t, 1, -1) + np.roll(t, -1, -1)) / 3.0
        t = (t + np.roll(t, 1, -2) + np.roll(t, -1, -2)) / 3.0
    t /= t.std(axis=tuple(range(1, t.ndim)), keepdims=True) + 1e-6
    return t


def _make_images(n: int, n_classes: int, shape: tuple[int, ...], seed: int, noise: float,
                 templates: np.ndarray | None = None):
    rng = npr.default_rng(seed)
    if templates is None:
        templates = _class_templates(rng, n_classes, shape)
    y = rng.integers(0, n_classes, size=n)
    x = templates[y] + noise * rng.standard_normal((n, *shape)).astype(np.float32)
    return torch.from_numpy(x.astype(np.float32)), torch.from_numpy(y.astype(np.int64)), templates


class SyntheticMNIST:
    """MNIST-shaped train/test pair.

    The canonical form is uint8 pixels + uint8 labels, exactly like the real IDX files, so the
    same synthetic set can be (a) normalised once into fp32 ``(x/255 - 0.1307)/0.3081`` for this
    framework (``.train`` / ``.test``; hfl_complete.py:19-24) and (b) written to disk in IDX
    format for the unmodified reference + torchvision to load (``write_idx``; bench.py
    ``--impl reference``)."""

    def __init__(self, n_train: int = 60_000, n_test: int = 10_000, seed: int = 0, noise: float = 1.0):
        rng = npr.default_rng(seed)
        tpl = _class_templates(rng, 10, (1, 28, 28))

        def make(n, sd):
            r = npr.default_rng(sd)
            y = r.integers(0, 10, size=n)
            x = 110.0 + 60.0 * tpl[y] + 60.0 * noise * r.standard_normal((n, 1, 28, 28)).astype(np.float32)
            return np.clip(np.rint(x), 0, 255).astype(np.uint8), y.astype(np.uint8)

        self.train_u8, self.train_labels = make(n_train, seed + 101)
        self.test_u8, self.test_labels = make(n_test, seed + 202)
        self.train = TensorImageSet(self.normalize(self.train_u8), torch.from_numpy(self.train_labels.astype(np.int64)))
        self.test = TensorImageSet(self.normalize(self.test_u8), torch.from_numpy(self.test_labels.astype(np.int64)))

    @staticmethod
    def normalize(u8: np.ndarray) -> torch.Tensor:
        x = torch.from_numpy(u8).float().div_(255.0)
        return x.sub_(MNIST_MEAN).div_(MNIST_STD)

    def write_idx(self, raw_dir) -> None:
        """Write the four MNIST IDX files torchvision looks for under ``<root>/MNIST/raw``."""
        import os
        import struct

        os.makedirs(raw_dir, exist_ok=True)

        def images(name, a):
            with open(os.path.join(raw_dir, name), "wb") as f:
                f.write(struct.pack(">IIII", 2051, a.shape[0], 28, 28))
                f.write(a.reshape(a.shape[0], 28, 28).tobytes())

        def labels(name, a):
            with open(os.path.join(raw_dir, name), "wb") as f:
                f.write(struct.pack(">II", 2049, a.shape[0]))
                f.write(a.tobytes())

        images("train-images-idx3-ubyte", self.train_u8)
        labels("train-labels-idx1-ubyte", self.train_labels)
        images("t10k-images-idx3-ubyte", self.test_u8)
        labels("t10k-labels-idx1-ubyte", self.test_labels)


class SyntheticCIFAR:
    """CIFAR-10-shaped train/test pair: 3x32x32 fp32, 10 classes."""

    def __init__(self, n_train: int = 50_000, n_test: int = 10_000, seed: int = 0, noise: float = 1.0):
        xtr, ytr, tpl = _make_images(n_train, 10, (3, 32, 32), seed, noise)
        xte, yte, _ = _make_images(n_test, 10, (3, 32, 32), seed + 1, noise, tpl)
        self.train = TensorImageSet(xtr, ytr)
        self.test = TensorImageSet(xte, yte)


class SyntheticImages64(Dataset):
    """``n`` RGB 64x64 images in [-1, 1] (tanh range) for the DCGAN workloads."""

    def __init__(self, n: int = 8192, seed: int = 0):
        rng = npr.default_rng(seed)
        tpl = _class_templates(rng, 16, (3, 64, 64), smooth=6)
        y = rng.integers(0, 16, size=n)
        x = np.tanh(tpl[y] + 0.3 * rng.standard_normal((n, 3, 64, 64)).astype(np.float32))
        self.images = torch.from_numpy(x.astype(np.float32))

    def __len__(self):
        return len(self.images)

    def __getitem__(self, i):
        return self.images[i]


# --------------------------------------------------------------------------- client sharding
def split_indices(targets, nr_clients: int, iid: bool, seed: int) -> list[np.ndarray]:
    """Client index sets.  Same algorithm as ``split`` in hfl_complete.py:91-104.

    IID: ``array_split(permutation(N), n)``.  Non-IID (McMahan pathological): sort by label,
    cut into ``2n`` shards, deal two random shards to every client.
    """
    rng = npr.default_rng(seed)
    n = len(targets)
    if iid:
        return [np.asarray(s, dtype=np.int64) for s in np.array_split(rng.permutation(n), nr_clients)]
    sorted_indices = np.argsort(np.asarray(targets), kind="stable")
    shards = np.array_split(sorted_indices, 2 * nr_clients)
    order = rng.permutation(len(shards))
    return [np.concatenate([shards[i] for i in pair]).astype(np.int64)
            for pair in order.reshape(nr_clients, 2)]


def split(dataset, nr_clients: int, iid: bool, seed: int):
    """``split(nr_clients, iid, seed) -> list[Subset]`` with the dataset made explicit."""
    from torch.utils.data import Subset

    targets = dataset.targets if hasattr(dataset, "targets") else [t for _, t in dataset]
    return [Subset(dataset, idx.tolist()) for idx in split_indices(np.asarray(targets), nr_clients, iid, seed)]


# --------------------------------------------------------------------------- heart (tabular)
HEART_NUMERIC = ["age", "trestbps", "chol", "thalach", "oldpeak"]
# categorical column -> cardinality, as in heart.csv (8 categorical columns -> 25 one-hot
# columns, 30 encoded features in total: lab/tutorial_2a/centralized.py:33-34)
HEART_CATEGORICAL = {"sex": 2, "cp": 4, "fbs": 2, "restecg": 3, "exang": 2, "slope": 3, "ca": 5, "thal": 4}
HEART_RAW_COLUMNS = ["age", "sex", "cp", "trestbps", "chol", "fbs", "restecg", "thalach", "exang",
                     "oldpeak", "slope", "ca", "thal"]


@dataclass
class HeartTable:
    """One-hot encoded, min-max scaled heart table: ``x`` [N, 30] fp32, ``y`` [N] int64."""
    x: torch.Tensor
    y: torch.Tensor
    columns: list[str]          # encoded column names, raw-column order preserved
    raw_to_encoded: dict[str, list[str]]

    def column_index(self, names: list[str]) -> list[int]:
        pos = {c: i for i, c in enumerate(self.columns)}
        return [pos[n] for n in names]


class SyntheticHeart:
    """Heart-disease-shaped table: 1025 rows, 5 numeric + 8 categorical columns, binary target.

    ``load_csv(path)`` accepts the real ``heart.csv`` schema (same columns + ``target``) when a
    user has it; the synthetic generator draws categorical levels and numeric features whose
    distribution depends on the label so the evaluator MLP can learn something.
    """

    def __init__(self, n: int = 1025, seed: int = 42):
        rng = npr.default_rng(seed)
        y = rng.integers(0, 2, size=n)
        raw: dict[str, np.ndarray] = {}
        for j, c in enumerate(HEART_NUMERIC):
            raw[c] = rng.normal(loc=0.6 * (y - 0.5) * (1 if j % 2 == 0 else -1), scale=1.0, size=n)
        for j, (c, k) in enumerate(HEART_CATEGORICAL.items()):
            p0 = rng.dirichlet(np.ones(k))
            p1 = rng.dirichlet(np.ones(k))
            lv = np.where(y == 1, rng.choice(k, size=n, p=p1), rng.choice(k, size=n, p=p0))
            lv[:k] = np.arange(k)  # every level appears at least once
            raw[c] = lv
        self.raw = raw
        self.target = y.astype(np.int64)

    @staticmethod
    def load_csv(path: str) -> "SyntheticHeart":
        import pandas as pd

        df = pd.read_csv(path)
        obj = SyntheticHeart.__new__(SyntheticHeart)
        obj.raw = {c: df[c].to_numpy() for c in HEART_RAW_COLUMNS}
        obj.target = df["target"].to_numpy().astype(np.int64)
        return obj

    def encoded(self) -> HeartTable:
        cols: list[str] = []
        mats: list[np.ndarray] = []
        raw_to_enc: dict[str, list[str]] = {}
        for c in HEART_RAW_COLUMNS:
            v = self.raw[c]
            if c in HEART_CATEGORICAL:
                levels = np.unique(v)
                names = [f"{c}_{int(l)}" for l in levels]
                mats.append((v[:, None] == levels[None, :]).astype(np.float32))
            else:
                names = [c]
                mats.append(v.astype(np.float32)[:, None])
            raw_to_enc[c] = names
            cols += names
        x = np.concatenate(mats, axis=1)
        lo, hi = x.min(0, keepdims=True), x.max(0, keepdims=True)
        x = (x - lo) / np.maximum(hi - lo, 1e-12)  # MinMaxScaler (centralized.py:40-42)
        return HeartTable(torch.from_numpy(x), torch.from_numpy(self.target), cols, raw_to_enc)


def train_test_split_tensor(x: torch.Tensor, y: torch.Tensor, test_frac: float = 0.2, seed: int = 42):
    g = torch.Generator().manual_seed(seed)
    perm = torch.randperm(len(x), generator=g)
    n_test = int(round(len(x) * test_frac))
    te, tr = perm[:n_test], perm[n_test:]
    return x[tr], y[tr], x[te], y[te]


# --------------------------------------------------------------------------- tokens
class SyntheticTokens:
    """Deterministic token stream standing in for ``TinyStories(tokenizer, batch_size, seq_l, skip)``.

    Iterating yields ``LongTensor[batch_size, seq_l]``.  ``skip`` offsets the stream so DP ranks
    see different data (intro_DP_GA.py:29) and two iterators with the same ``skip`` stay in
    lock-step (the PP scripts rely on that: intro_PP_1F1B_MB.py:32,41).  Tokens follow a
    first-order Markov chain so a causal LM can drive the loss well below ``ln(vocab)``.
    """

    def __init__(self, vocab_size: int = 32_000, batch_size: int = 3, seq_l: int = 256, skip: int = 0,
                 seed: int = 0, branching: int = 4, pad_id: int = 0):
        self.vocab_size, self.batch_size, self.seq_l, self.skip = vocab_size, batch_size, seq_l, skip
        self.pad_id = pad_id
        g = torch.Generator().manual_seed(seed)
        # successor table: each token can be followed by one of `branching` tokens
        self.succ = torch.randint(1, vocab_size, (vocab_size, branching), generator=g)
        self.seed = seed
        self.branching = branching

    def batch(self, index: int) -> torch.Tensor:
        g = torch.Generator().manual_seed(self.seed * 1_000_003 + index + 17)
        cur = torch.randint(1, self.vocab_size, (self.batch_size,), generator=g)
        choices = torch.randint(0, self.branching, (self.batch_size, self.seq_l), generator=g)
        out = torch.empty(self.batch_size, self.seq_l, dtype=torch.long)
        for t in range(self.seq_l):
            out[:, t] = cur
            cur = self.succ[cur, choices[:, t]]
        return out

    def __iter__(self):
        i = self.skip
        while True:
            yield self.batch(i)
            i += 1


class SimpleTokenizer:
    """The two attributes the labs read from ``SPTokenizer`` (primer/intro.py:15-18)."""

    def __init__(self, vocab_size: int = 32_000, pad_id: int = 0):
        self.vocab_size = vocab_size
        self.pad_id = pad_id
