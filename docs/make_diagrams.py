#!/usr/bin/env python
"""Draws the six data-flow diagrams of the tutorial (role of the reference's images/*.png,
README.md:87,123,153,187,226,259) as SVG, from a description of what every rank holds before
and after the call. Run: python docs/make_diagrams.py  -> images/<collective>.svg"""
from __future__ import annotations

import os

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "images")
N = 4
W, H, BW, BH, GAP = 640, 250, 120, 44, 30
COL = ["#4e79a7", "#f28e2b", "#59a14f", "#e15759"]


def box(x, y, text, fill, w=BW):
    return (f'<rect x="{x}" y="{y}" width="{w}" height="{BH}" rx="6" fill="{fill}" fill-opacity="0.25" '
            f'stroke="{fill}" stroke-width="2"/>'
            f'<text x="{x + w / 2}" y="{y + BH / 2 + 5}" text-anchor="middle" font-family="monospace" '
            f'font-size="14">{text}</text>')


def arrow(x1, y1, x2, y2, color):
    return (f'<line x1="{x1}" y1="{y1}" x2="{x2}" y2="{y2}" stroke="{color}" stroke-width="2" '
            f'marker-end="url(#a)"/>')


def diagram(name, title, before, after, edges):
    """before/after: text per rank; edges: (src rank, dst rank) pairs drawn top row -> bottom row."""
    xs = [GAP + r * (BW + GAP) for r in range(N)]
    top, bot = 60, 170
    s = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{W}" height="{H}" viewBox="0 0 {W} {H}">',
         '<defs><marker id="a" markerWidth="8" markerHeight="8" refX="7" refY="4" orient="auto">'
         '<path d="M0,0 L8,4 L0,8 z" fill="#444"/></marker></defs>',
         f'<text x="{W / 2}" y="28" text-anchor="middle" font-family="sans-serif" font-size="18" '
         f'font-weight="bold">{title}</text>']
    for r in range(N):
        s.append(f'<text x="{xs[r] + BW / 2}" y="{top - 8}" text-anchor="middle" font-family="sans-serif" '
                 f'font-size="12">rank {r}</text>')
        s.append(box(xs[r], top, before[r], COL[r]))
        s.append(box(xs[r], bot, after[r], COL[r]))
    for a, b in edges:
        s.append(arrow(xs[a] + BW / 2, top + BH, xs[b] + BW / 2, bot - 2, COL[a]))
    s.append('</svg>')
    with open(os.path.join(OUT, name + ".svg"), "w") as f:
        f.write("\n".join(s))


def main():
    os.makedirs(OUT, exist_ok=True)
    every = [(a, b) for a in range(N) for b in range(N)]
    diagram("reduce", "reduce(dst=0, op=SUM)", ["1"] * 4, ["4", "1", "1", "1"], [(r, 0) for r in range(N)])
    diagram("all_reduce", "all_reduce(op=SUM)", ["1"] * 4, ["4"] * 4, every)
    diagram("scatter", "scatter(src=0)", ["[1,2,3,4]", "-", "-", "-"], ["1", "2", "3", "4"],
            [(0, r) for r in range(N)])
    diagram("gather", "gather(dst=0)", ["0", "1", "2", "3"], ["[0,1,2,3]", "-", "-", "-"],
            [(r, 0) for r in range(N)])
    diagram("all_gather", "all_gather", ["0", "1", "2", "3"], ["[0,1,2,3]"] * 4, every)
    diagram("broadcast", "broadcast(src=0)", ["0", "?", "?", "?"], ["0"] * 4, [(0, r) for r in range(N)])
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
