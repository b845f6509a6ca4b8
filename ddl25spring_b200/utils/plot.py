"""Line plots without a plotting stack.

The labs draw accuracy-vs-round curves with seaborn (lab/tutorial_1a/horizontal-federated-learning.ipynb cell 37,
lab/homework-1.ipynb cells 24-36, lab/tutorial_2b/exercise_1.py:150-163).  Neither matplotlib nor seaborn is part of this
image, so :func:`line_plot` writes a self-contained SVG (axes, ticks, legend, one polyline per series) and returns the
markup as an :class:`SVGFigure`, which Jupyter renders inline.  With matplotlib installed the same call can also render
through it (``backend="matplotlib"``).
"""
from __future__ import annotations

from typing import Mapping, Sequence

class SVGFigure(str):
    """SVG markup that Jupyter renders inline when it is the value of a cell (``_repr_svg_``)."""

    def _repr_svg_(self):
        return str(self)


PALETTE = ["#1f77b4", "#ff7f0e", "#2ca02c", "#d62728", "#9467bd", "#8c564b", "#e377c2", "#7f7f7f", "#bcbd22", "#17becf"]


def _ticks(lo: float, hi: float, n: int = 5) -> list[float]:
    if hi <= lo:
        hi = lo + 1.0
    raw = (hi - lo) / n
    mag = 10 ** __import__("math").floor(__import__("math").log10(raw))
    step = min((m for m in (1, 2, 2.5, 5, 10) if m * mag >= raw), default=10) * mag
    t0 = __import__("math").ceil(lo / step) * step
    out = []
    while t0 <= hi + 1e-9:
        out.append(round(t0, 10))
        t0 += step
    return out


def line_plot(series: Mapping[str, tuple[Sequence[float], Sequence[float]]], path: str | None = None, *, title: str = "",
              xlabel: str = "", ylabel: str = "", width: int = 640, height: int = 400, backend: str = "svg") -> str:
    """``series``: label -> (xs, ys).  Writes ``path`` (if given) and returns the SVG text."""
    if backend == "matplotlib":
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        fig, ax = plt.subplots(figsize=(width / 100, height / 100))
        for k, (xs, ys) in series.items():
            ax.plot(xs, ys, marker="o", label=k)
        ax.set(title=title, xlabel=xlabel, ylabel=ylabel)
        ax.legend()
        if path:
            fig.savefig(path)
        plt.close(fig)
        return path or ""
    ml, mr, mt, mb = 60, 170, 36, 48
    xs_all = [x for xs, _ in series.values() for x in xs] or [0.0, 1.0]
    ys_all = [y for _, ys in series.values() for y in ys if y == y] or [0.0, 1.0]
    x0, x1, y0, y1 = min(xs_all), max(xs_all), min(ys_all), max(ys_all)
    if x1 == x0:
        x1 = x0 + 1
    pad = 0.05 * (y1 - y0 or 1.0)
    y0, y1 = y0 - pad, y1 + pad
    pw, ph = width - ml - mr, height - mt - mb

    def X(x):
        return ml + (x - x0) / (x1 - x0) * pw

    def Y(y):
        return mt + (1 - (y - y0) / (y1 - y0)) * ph

    o = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{width}" height="{height}" font-family="sans-serif" font-size="11">',
         f'<rect width="{width}" height="{height}" fill="white"/>',
         f'<text x="{ml + pw / 2}" y="18" text-anchor="middle" font-size="13">{title}</text>',
         f'<rect x="{ml}" y="{mt}" width="{pw}" height="{ph}" fill="none" stroke="#444"/>']
    for t in _ticks(x0, x1):
        o.append(f'<line x1="{X(t):.1f}" y1="{mt + ph}" x2="{X(t):.1f}" y2="{mt + ph + 4}" stroke="#444"/>'
                 f'<text x="{X(t):.1f}" y="{mt + ph + 16}" text-anchor="middle">{t:g}</text>')
    for t in _ticks(y0, y1):
        o.append(f'<line x1="{ml - 4}" y1="{Y(t):.1f}" x2="{ml + pw}" y2="{Y(t):.1f}" stroke="#ddd"/>'
                 f'<text x="{ml - 7}" y="{Y(t) + 4:.1f}" text-anchor="end">{t:g}</text>')
    o.append(f'<text x="{ml + pw / 2}" y="{height - 10}" text-anchor="middle">{xlabel}</text>')
    o.append(f'<text x="14" y="{mt + ph / 2}" text-anchor="middle" transform="rotate(-90 14 {mt + ph / 2})">{ylabel}</text>')
    for i, (k, (xs, ys)) in enumerate(series.items()):
        c = PALETTE[i % len(PALETTE)]
        pts = " ".join(f"{X(x):.1f},{Y(y):.1f}" for x, y in zip(xs, ys) if y == y)
        o.append(f'<polyline fill="none" stroke="{c}" stroke-width="1.8" points="{pts}"/>')
        o += [f'<circle cx="{X(x):.1f}" cy="{Y(y):.1f}" r="2.4" fill="{c}"/>' for x, y in zip(xs, ys) if y == y]
        ly = mt + 12 + 16 * i
        o.append(f'<line x1="{ml + pw + 10}" y1="{ly}" x2="{ml + pw + 30}" y2="{ly}" stroke="{c}" stroke-width="2"/>'
                 f'<text x="{ml + pw + 35}" y="{ly + 4}">{k}</text>')
    o.append("</svg>")
    svg = SVGFigure("\n".join(o))
    if path:
        with open(path, "w") as f:
            f.write(svg)
    return svg


def plot_runs(df, path: str | None = None, *, x: str = "Round", y: str = "Test accuracy", hue: Sequence[str] = ("Algorithm",),
              title: str = "") -> str:
    """Accuracy-vs-round curves of a concatenated ``RunResult.as_df()`` frame, one line per distinct ``hue`` combination
    (the seaborn ``lineplot(hue=...)`` of lab/homework-1.ipynb cell 32)."""
    series = {}
    for key, g in df.groupby(list(hue), sort=False):
        key = key if isinstance(key, tuple) else (key,)
        label = ", ".join(f"{h}={v}" for h, v in zip(hue, key))
        series[label] = (list(g[x]), list(g[y]))
    return line_plot(series, path, title=title, xlabel=x, ylabel=y)
