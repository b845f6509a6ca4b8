"""Server-side aggregation rules — the *server-aggregate* hook.

An :class:`Aggregator` maps the stacked client messages ``updates[K, D]`` (one flat arena row
per sampled client) plus the FedAvg weights ``n_k / sum n`` to one flat vector.  The reference
has exactly one rule, the sample-count-weighted mean (lab/tutorial_1a/hfl_complete.py:291-299,
370-378); Krum / Multi-Krum, coordinate-wise median and trimmed mean are the robust defenses
the north star adds (BASELINE.json).  On CUDA tensors each rule is one or two
launches of the sm_100a kernels in ``csrc/aggregate.cu`` / ``csrc/gemm_tcgen05.cu``.

BatchNorm buffers (ResNet-18 / DCGAN): arenas built with ``include_buffers=True`` carry the
running statistics in the same flat vector.  The weighted mean averages them like parameters;
the robust rules apply to the *whole* vector too (a Byzantine client can poison running
statistics just as well as weights), except Krum which scores on the parameter prefix only
(``score_dims``) so that slowly varying statistics do not dominate the distances.
"""
from __future__ import annotations

from abc import ABC, abstractmethod

import torch

from .. import ops


class Aggregator(ABC):
    name = "aggregator"

    @abstractmethod
    def __call__(self, updates: torch.Tensor, weights: torch.Tensor, *, out: torch.Tensor | None = None,
                 global_flat: torch.Tensor | None = None) -> torch.Tensor:
        """``updates`` [K, D] fp32, ``weights`` [K] (host or device) -> flat [D]."""

    def __repr__(self):
        return f"{type(self).__name__}()"


class WeightedMean(Aggregator):
    """FedAvg / FedSGD rule: ``sum_k w_k u_k`` (K1)."""
    name = "mean"

    def __call__(self, updates, weights, *, out=None, global_flat=None):
        return ops.weighted_sum(updates, weights, out)


class CoordinateMedian(Aggregator):
    """Coordinate-wise median (K4).  ``lower=True`` reproduces ``torch.median`` for even K."""
    name = "median"

    def __init__(self, lower: bool = False):
        self.lower = lower

    def __call__(self, updates, weights, *, out=None, global_flat=None):
        return ops.coordinate_median(updates, self.lower, out)


class TrimmedMean(Aggregator):
    """Coordinate-wise trimmed mean (K5): drop ``ceil(beta*K)`` values at each end."""
    name = "trimmed_mean"

    def __init__(self, beta: float = 0.2, trim: int | None = None):
        self.beta, self.trim = beta, trim

    def trim_count(self, K: int) -> int:
        t = self.trim if self.trim is not None else int(-(-self.beta * K // 1))
        return max(0, min(t, (K - 1) // 2))

    def __call__(self, updates, weights, *, out=None, global_flat=None):
        return ops.trimmed_mean(updates, self.trim_count(updates.shape[0]), out)


class Krum(Aggregator):
    """Krum (``multi=1``) / Multi-Krum (``multi=m``) (K3).

    Scores are computed from the Gram matrix of update *deltas* ``u_k - global`` (smaller norms
    than raw weights, so the TF32 Gram loses less to cancellation); the selected rows are then
    averaged with the K1 kernel using the selection weights.
    """
    name = "krum"

    def __init__(self, n_byzantine: int, multi: int = 1, precise: bool = False, score_dims: int | None = None,
                 fused_gather: bool = False):
        self.f, self.multi, self.precise, self.score_dims = n_byzantine, multi, precise, score_dims
        # multi-GPU only: one kernel does the all-gather of the updates AND the Gram GEMM (TMA k-blocks from peer HBM)
        self.fused_gather = fused_gather
        self.last_selected: torch.Tensor | None = None
        self.last_scores: torch.Tensor | None = None

    def __call__(self, updates, weights, *, out=None, global_flat=None):
        U = updates if self.score_dims is None else updates[:, :self.score_dims]
        if U.is_cuda:
            # the global model is subtracted from the staged tiles inside the Gram kernel (3xTF32): no K x D delta matrix
            from ..ops.gemm import gram_tcgen05
            center = None
            if global_flat is not None:
                center = global_flat if self.score_dims is None else global_flat[:self.score_dims]
            if center is not None and (U.shape[1] % 4 or U.stride(1) != 1 or center.data_ptr() % 16):
                U, center = (U - center[None, :]).contiguous(), None           # ragged score_dims: rare, not the hot path
            G = gram_tcgen05(U, precise=self.precise, center=center)
            sel, selw, scores = ops.krum_select(U, self.f, self.multi, gram_matrix=G)
        else:
            if global_flat is not None:
                U = U - (global_flat if self.score_dims is None else global_flat[:self.score_dims])[None, :]
            sel, selw, scores = ops.krum_select(U, self.f, self.multi)
        self.last_selected, self.last_scores = sel, scores
        return ops.weighted_sum(updates, selw, out)           # device weights; rows with weight 0 are not read

    def __repr__(self):
        return f"Krum(f={self.f}, multi={self.multi})"


def make_aggregator(name: str, **kw) -> Aggregator:
    name = name.lower().replace("-", "_")
    if name in ("mean", "fedavg", "weighted_mean"):
        return WeightedMean()
    if name == "median":
        return CoordinateMedian(**kw)
    if name in ("trimmed_mean", "trimmedmean"):
        return TrimmedMean(**kw)
    if name == "krum":
        return Krum(**kw)
    if name in ("multi_krum", "multikrum"):
        kw.setdefault("multi", 2)
        return Krum(**kw)
    raise ValueError(f"unknown aggregator {name!r}")
