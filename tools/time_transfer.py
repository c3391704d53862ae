#!/usr/bin/env python
"""time_transfer.py -- time the two-level h-transfer kernels of one level pair on cuda:0 (CUDA events)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nref", type=int, default=6)
    ap.add_argument("--degree", type=int, default=4)
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    import torch
    import dealii_multigrid_b200 as mg

    torch.cuda.set_device(0)
    mg.init(0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    mg.set_stream(stream.cuda_stream)
    mf = mg.Mesh("hypercube", args.nref)
    mc = mf.coarsen_global()
    df, dc = mg.DoFMap(mf, args.degree), mg.DoFMap(mc, args.degree)
    tm = mg.TransferMap(mf, df, mc, dc)
    opf, opc = mg.Operator(df), mg.Operator(dc)
    T = mg.MGTwoLevelTransfer(df, dc, tm, opf, opc)
    vf, vc = opf.initialize_dof_vector(), opc.initialize_dof_vector()
    vf.set(1.0)
    vc.set(1.0)
    out = {"nref": args.nref, "degree": args.degree, "n_fine": opf.m(), "n_coarse": opc.m()}
    for name, fn, nbytes in (("prolongate_and_add", lambda: T.prolongate_and_add(vf, vc), 8 * (opc.m() + 2 * opf.m())),
                             ("restrict_and_add", lambda: T.restrict_and_add(vc, vf), 8 * (opf.m() + 2 * opc.m()))):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.reps
        out[name] = {"ms": ms, "gbs_alg": nbytes / ms / 1e6}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
