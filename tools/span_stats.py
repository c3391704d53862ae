#!/usr/bin/env python
"""span_stats.py -- CPU-only: how many contiguous address spans does a cell's (p+1)^3 DoF gather need?

Input of the round-2 bulk-copy design for the vmult kernel: the 27 entity ranges of every cell
(entity table of the stand-in generator) are merged when the gap between them is <= --gap DoFs; prints the spans per
cell, the over-fetch and the number of distinct 128-byte lines, for comparison with the 45 segments / 37 lines per cell
of the per-(i,j)-column gather of k_laplace_apply.

usage: python tools/span_stats.py [--geometry hypercube] [--nref 5] [--degree 4] [--gap 0 2 8 32]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--geometry", default="hypercube")
    ap.add_argument("--nref", type=int, default=5)
    ap.add_argument("--degree", type=int, default=4)
    ap.add_argument("--gap", type=int, nargs="+", default=[0, 2, 8, 32])
    ap.add_argument("--max-cells", type=int, default=40000)
    args = ap.parse_args()
    import dealii_multigrid_b200 as mg

    d = mg.DoFMap(mg.Mesh(args.geometry, args.nref), args.degree)
    p = args.degree
    tab = d.eval_table().astype(np.int64)
    flag = (tab >> 31) & 1
    base = tab & 0x7FFFFFFF
    dims = np.array([(e % 3 == 1) + ((e // 3) % 3 == 1) + ((e // 9) == 1) for e in range(27)])
    length = (p - 1) ** dims                                                # DoFs of the entity (vertex 1, line p-1, ...)
    n = min(d.n_cells, args.max_cells)
    sel = np.linspace(0, d.n_cells - 1, n).astype(np.int64)                 # sample over the whole Morton range
    print(f"{args.geometry} L={args.nref} p={p}: {d.n_cells} cells, {(p + 1) ** 3} DoFs per cell, sampling {n} cells")
    for gap in args.gap:
        spans, fetched, lines, needed = [], [], [], []
        for c in sel:
            keep = flag[c] == 0                                             # Dirichlet entities are not read
            b, l = base[c][keep], length[keep]
            order = np.argsort(b, kind="stable")
            b, l = b[order], l[order]
            s0, s1, ns, nf = b[0], b[0] + l[0], 1, 0
            ln = set()
            for bb, ll in zip(b[1:], l[1:]):
                if bb <= s1 + gap:
                    s1 = max(s1, bb + ll)
                else:
                    nf += s1 - s0
                    ln.update(range(s0 * 8 // 128, (s1 * 8 - 1) // 128 + 1))
                    s0, s1, ns = bb, bb + ll, ns + 1
            nf += s1 - s0
            ln.update(range(s0 * 8 // 128, (s1 * 8 - 1) // 128 + 1))
            spans.append(ns); fetched.append(nf); lines.append(len(ln)); needed.append(int(l.sum()))
        spans, fetched, lines, needed = map(np.array, (spans, fetched, lines, needed))
        print(f"  gap<={gap:3d}: spans/cell mean {spans.mean():5.2f} max {spans.max():2d} | fetched/needed DoFs {fetched.sum() / needed.sum():5.3f}"
              f" | 128-byte lines/cell mean {lines.mean():5.2f} (FP64)")


if __name__ == "__main__":
    main()
