#!/usr/bin/env python
"""time_vmult.py -- time the operator kernels in isolation on cuda:0 (CUDA events) for one mesh/degree.

usage: python tools/time_vmult.py [--geometry hypercube] [--nref 6] [--degree 4] [--number double] [--reps 20]
Prints one JSON line with the vmult time, GDoF/s and the 16 B/DoF (8 B/DoF for float) bandwidth figure.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--geometry", default="hypercube")
    ap.add_argument("--nref", type=int, default=6)
    ap.add_argument("--degree", type=int, default=4)
    ap.add_argument("--number", default="double")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--deform", default="none", choices=["none", "cube", "smooth"],
                    help="general-geometry kernel (mgb_operator_set_cell_vertices): undeformed or smoothly deformed vertices")
    args = ap.parse_args()
    import torch
    import dealii_multigrid_b200 as mg

    torch.cuda.set_device(0)
    mg.init(0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    mg.set_stream(stream.cuda_stream)
    nt = mg.DOUBLE if args.number == "double" else mg.FLOAT
    d = mg.DoFMap(mg.Mesh(args.geometry, args.nref), args.degree)
    op = mg.Operator(d, number_type=nt)
    if args.deform != "none":
        import numpy as np
        h = 2.0 / (1 << d.cell_level().astype(np.int64))
        corner = np.array([[(v >> a) & 1 for a in range(3)] for v in range(8)], dtype=float)
        xyz = (-1.0 + h[:, None] * d.cell_xyz())[:, None, :] + h[:, None, None] * corner[None, :, :]
        if args.deform == "smooth":
            x, y, z = xyz[..., 0].copy(), xyz[..., 1].copy(), xyz[..., 2].copy()
            xyz = np.stack([x + 0.08 * np.sin(np.pi * y) * np.cos(0.5 * np.pi * z), y + 0.06 * np.sin(np.pi * z) * x,
                            z + 0.07 * np.cos(np.pi * x) * np.sin(0.5 * np.pi * y)], axis=-1)
        op.set_cell_vertices(np.ascontiguousarray(xyz))
    src, dst = op.initialize_dof_vector(), op.initialize_dof_vector()
    src.set(1.0)
    for _ in range(3):
        op.vmult(dst, src)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    mg.profile_start(d.n_cells)
    e0.record()
    for _ in range(args.reps):
        op.vmult(dst, src)
    e1.record()
    torch.cuda.synchronize()
    k_ms, k_n = mg.profile_stop()
    ms = e0.elapsed_time(e1) / args.reps
    n = op.m()
    s = 8 if args.number == "double" else 4
    print(json.dumps({"geometry": args.geometry, "nref": args.nref, "degree": args.degree, "number": args.number, "deform": args.deform,
                      "n_dofs": n,
                      "n_cells": d.n_cells, "vmult_ms": ms, "cell_kernel_ms": k_ms / max(k_n, 1),
                      "vmult_gdofs": n / ms / 1e6, "cell_kernel_gdofs": n / (k_ms / max(k_n, 1)) / 1e6,
                      "cell_kernel_gbs_alg": 2 * s * n / (k_ms / max(k_n, 1)) / 1e6}))


if __name__ == "__main__":
    main()
