#!/usr/bin/env python
"""time_secondary.py -- the HBM-bound kernels beside the cell loop at the bench size (hypercube L, degree p, FP64) on cuda:0:
Chebyshev updates (k_cheb<T,0/1/2>), BLAS-1 (k_dot, k_sadd, k_add, k_cg_update through add_and_dot), the memset of dst, the
h-transfer pair.  CUDA events on the library's stream, inputs far larger than L2.  One JSON line per kernel:
algorithmic bytes, ms, GB/s, fraction of the measured copy peak."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nref", type=int, default=7)
    ap.add_argument("--degree", type=int, default=4)
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    import torch
    import dealii_multigrid_b200 as mg

    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except (OSError, KeyError, ValueError):
        peak = 6650.0
    torch.cuda.set_device(0)
    mg.init(0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    mg.set_stream(stream.cuda_stream)
    mf = mg.Mesh("hypercube", args.nref)
    mc = mf.coarsen_global()
    df, dc = mg.DoFMap(mf, args.degree), mg.DoFMap(mc, args.degree)
    opf, opc = mg.Operator(df), mg.Operator(dc)
    T = mg.MGTwoLevelTransfer(df, dc, mg.TransferMap(mf, df, mc, dc), opf, opc)
    n, nc, s = opf.m(), opc.m(), 8
    x, y, b = opf.initialize_dof_vector(), opf.initialize_dof_vector(), opf.initialize_dof_vector()
    xc = opc.initialize_dof_vector()
    x.set(1.0), y.set(0.5), b.set(0.25), xc.set(1.0)

    def timed(fn, reps=args.reps):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    def line(name, ms, nbytes, note=""):
        gbs = nbytes / ms / 1e6
        print(json.dumps({"kernel": name, "ms": round(ms, 4), "algorithmic_bytes": nbytes, "gbs": round(gbs, 1),
                          "frac_of_copy_peak": round(gbs / peak, 3), "note": note}), flush=True)

    print(json.dumps({"nref": args.nref, "degree": args.degree, "n_dofs": n, "vector_MB": n * s / 1e6, "copy_peak_gbs": peak}))
    line("memset (dst = 0 before a cell loop)", timed(lambda: y.set(0.0)), s * n, "1 stream")
    line("k_equ (y = a x)", timed(lambda: y.equ(2.0, x)), 2 * s * n, "2 streams")
    line("k_add (y += a x)", timed(lambda: y.add(1e-3, x)), 3 * s * n, "3 streams")
    line("k_sadd (y = s y + a x)", timed(lambda: y.sadd(0.5, 1e-3, x)), 3 * s * n, "3 streams")
    line("k_dot", timed(lambda: x.dot(y)), 2 * s * n, "2 streams + host read-back of the scalar")
    line("k_add_and_dot (y += a x; y.y)", timed(lambda: y.add_and_dot(1e-6, x)), 3 * s * n, "3 streams + read-back")
    # the smoother: subtract the cell loops (timed alone) from the whole call to get the vector updates
    t_vmult = timed(lambda: opf.vmult(y, x))
    line("cell loop incl. dst = 0 and identity rows (for reference)", t_vmult, 2 * s * n, "vmult")
    cheb = mg.PreconditionChebyshev(opf, 3, 20.0, 20)
    cheb.eigenvalues()
    t_pre = timed(lambda: cheb.vmult(y, b), max(3, args.reps // 4))
    t_post = timed(lambda: cheb.step(y, b), max(3, args.reps // 4))
    line("Chebyshev(3) vmult, zero guess: k_cheb<0> + 2 x (cell loop + k_cheb<1|2>)", t_pre, (3 + 5 + 6) * s * n + 2 * 2 * s * n,
         f"vector updates alone: {t_pre - 2 * t_vmult:.3f} ms for 14 streams = {(14 * s * n) / (t_pre - 2 * t_vmult) / 1e6:.0f} GB/s")
    line("Chebyshev(3) step: 3 x (cell loop + k_cheb<1|2|2>)", t_post, (5 + 6 + 6) * s * n + 3 * 2 * s * n,
         f"vector updates alone: {t_post - 3 * t_vmult:.3f} ms for 17 streams = {(17 * s * n) / (t_post - 3 * t_vmult) / 1e6:.0f} GB/s")
    line("k_transfer prolongate_and_add", timed(lambda: T.prolongate_and_add(x, xc)), s * (nc + 2 * n), "s (N_c + 2 N_f)")
    line("k_transfer restrict_and_add", timed(lambda: T.restrict_and_add(xc, x)), s * (n + 2 * nc), "s (N_f + 2 N_c)")


if __name__ == "__main__":
    main()
