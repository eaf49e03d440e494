"""Dependency-free SVG line charts of speed-up / efficiency.

Counterpart of the reference's two PNG charts (`iamge1.png`, `iamge2.png`,
readme.md:102-108; generator not shipped): speed-up ``S = T_ref / T`` and
efficiency ``E = S / P`` per implementation.  Headless by construction — it
writes an SVG text file, no graphics stack is involved.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

_COLORS = ["#1f77b4", "#d62728", "#2ca02c", "#ff7f0e", "#9467bd", "#8c564b", "#17becf", "#7f7f7f"]


def _ticks(lo: float, hi: float, n: int = 5) -> List[float]:
    if hi <= lo:
        hi = lo + 1.0
    step = (hi - lo) / n
    return [lo + i * step for i in range(n + 1)]


def line_chart(series: Dict[str, Sequence[Tuple[float, float]]], title: str, xlabel: str, ylabel: str,
               width: int = 640, height: int = 400, logx: bool = False) -> str:
    """series: name -> [(x, y), ...].  Returns the SVG document as a string."""
    import math
    ml, mr, mt, mb = 60, 160, 40, 50
    pw, ph = width - ml - mr, height - mt - mb
    fx = (lambda v: math.log2(v)) if logx else (lambda v: v)
    xs = [fx(x) for pts in series.values() for x, _ in pts]
    ys = [y for pts in series.values() for _, y in pts]
    x0, x1 = min(xs), max(xs)
    y0, y1 = 0.0, max(ys) * 1.05 if ys else 1.0
    if x1 == x0:
        x1 = x0 + 1

    def px(x):
        return ml + (fx(x) - x0) / (x1 - x0) * pw

    def py(y):
        return mt + ph - (y - y0) / (y1 - y0) * ph

    out = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{width}" height="{height}" font-family="sans-serif" font-size="12">',
           f'<rect width="{width}" height="{height}" fill="white"/>',
           f'<text x="{width / 2}" y="20" text-anchor="middle" font-size="14">{title}</text>',
           f'<rect x="{ml}" y="{mt}" width="{pw}" height="{ph}" fill="none" stroke="black"/>']
    for t in _ticks(y0, y1):
        out.append(f'<line x1="{ml}" x2="{ml + pw}" y1="{py(t):.1f}" y2="{py(t):.1f}" stroke="#ddd"/>')
        out.append(f'<text x="{ml - 6}" y="{py(t) + 4:.1f}" text-anchor="end">{t:.3g}</text>')
    xt = sorted({x for pts in series.values() for x, _ in pts})
    for x in xt:
        out.append(f'<text x="{px(x):.1f}" y="{mt + ph + 16}" text-anchor="middle">{x:g}</text>')
    out.append(f'<text x="{ml + pw / 2}" y="{height - 10}" text-anchor="middle">{xlabel}</text>')
    out.append(f'<text x="16" y="{mt + ph / 2}" text-anchor="middle" transform="rotate(-90 16 {mt + ph / 2})">{ylabel}</text>')
    for i, (name, pts) in enumerate(series.items()):
        c = _COLORS[i % len(_COLORS)]
        path = " ".join(f"{'M' if k == 0 else 'L'}{px(x):.1f},{py(y):.1f}" for k, (x, y) in enumerate(pts))
        out.append(f'<path d="{path}" fill="none" stroke="{c}" stroke-width="2"/>')
        for x, y in pts:
            out.append(f'<circle cx="{px(x):.1f}" cy="{py(y):.1f}" r="3" fill="{c}"/>')
        out.append(f'<rect x="{ml + pw + 10}" y="{mt + 18 * i}" width="12" height="12" fill="{c}"/>')
        out.append(f'<text x="{ml + pw + 28}" y="{mt + 18 * i + 10}">{name}</text>')
    out.append("</svg>")
    return "\n".join(out)


def speedup_efficiency_charts(rows: Dict[str, Sequence[Tuple[int, float]]], t_ref: float, title: str):
    """rows: implementation -> [(P, seconds)].  Returns (speedup_svg, efficiency_svg)."""
    sp = {k: [(p, t_ref / t) for p, t in v] for k, v in rows.items()}
    ef = {k: [(p, t_ref / t / p) for p, t in v] for k, v in rows.items()}
    return (line_chart(sp, f"{title}: speed-up S = T_ref / T", "processes / threads / GPUs", "S", logx=True),
            line_chart(ef, f"{title}: efficiency E = S / P", "processes / threads / GPUs", "E", logx=True))
