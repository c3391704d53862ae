#!/usr/bin/env python
"""span_table.py -- CPU-only: write the per-cell address spans of a mesh for tools/bench_bulk_cells.cu.

Every cell's 27 entity ranges (DoF units) are merged when the gap is <= --gap, rounded out to even DoF boundaries (16 bytes
for FP64, what cp.async.bulk needs) and written as uint32 triples (start_dof, n_dofs, staging_offset_dofs), MAX_SPANS per
cell (unused: n_dofs = 0).  File: int64 header (n_cells, max_spans, n_dofs_padded) + the table.

usage: python tools/span_table.py [--nref 6] [--degree 4] [--gap 0] [--out tools/spans_L6_p4.bin]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
MAX_SPANS = 20


def cell_spans(base, length, gap):
    """merge sorted entity ranges of one cell; returns list of (start, end) on even boundaries, merged again after rounding"""
    order = np.argsort(base, kind="stable")
    out = []
    for b, l in zip(base[order], length[order]):
        s, e = int(b) & ~1, (int(b + l) + 1) & ~1
        if out and s <= out[-1][1] + gap:
            out[-1][1] = max(out[-1][1], e)
        else:
            out.append([s, e])
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--geometry", default="hypercube")
    ap.add_argument("--nref", type=int, default=6)
    ap.add_argument("--degree", type=int, default=4)
    ap.add_argument("--gap", type=int, default=0)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import dealii_multigrid_b200 as mg

    d = mg.DoFMap(mg.Mesh(args.geometry, args.nref), args.degree)
    p = args.degree
    tab = d.eval_table().astype(np.int64)
    keep = ((tab >> 31) & 1) == 0
    base = tab & 0x7FFFFFFF
    dims = np.array([(e % 3 == 1) + ((e // 3) % 3 == 1) + ((e // 9) == 1) for e in range(27)])
    length = (p - 1) ** dims
    out = np.zeros((d.n_cells, MAX_SPANS, 3), dtype=np.uint32)
    n_spans = np.zeros(d.n_cells, dtype=np.int64)
    staged = np.zeros(d.n_cells, dtype=np.int64)
    for c in range(d.n_cells):
        sp = cell_spans(base[c][keep[c]], length[keep[c]], args.gap)
        if len(sp) > MAX_SPANS:
            raise RuntimeError("cell %d needs %d spans" % (c, len(sp)))
        off = 0
        for k, (s, e) in enumerate(sp):
            out[c, k] = (s, e - s, off)
            off += e - s
        n_spans[c], staged[c] = len(sp), off
    n_pad = (d.n_dofs_global + 1) & ~1
    path = args.out or os.path.join(os.path.dirname(os.path.abspath(__file__)), "spans_L%d_p%d.bin" % (args.nref, p))
    with open(path, "wb") as f:
        np.array([d.n_cells, MAX_SPANS, n_pad], dtype=np.int64).tofile(f)
        out.tofile(f)
    print(f"{path}: {d.n_cells} cells, spans/cell mean {n_spans.mean():.2f} max {n_spans.max()}, staged DoFs/cell mean {staged.mean():.1f} "
          f"max {staged.max()} (needed {(p + 1) ** 3}), {os.path.getsize(path) / 1e6:.1f} MB")


if __name__ == "__main__":
    main()
