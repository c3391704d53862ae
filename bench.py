#!/usr/bin/env python
"""bench.py -- V-cycle-preconditioned CG throughput (DoFs/s) of the B200 hot path, with the vmult roofline.

Contract: `python bench.py --gpus N --steps K --warmup W [--impl reference]` prints ONE JSON line on rank 0.
  * one "step" = one complete preconditioned-CG solve (SolverCG + PreconditionMG V-cycle, Chebyshev(3) smoothing,
    matrix-free coarse CG) of the reference's multigrid_throughput on the workload below
    (multigrid_throughput.cc:1238-1283; throughput = n_dofs * n_iterations / t_solve, :1282);
  * `value`: all inputs resident in HBM (Problem.solve_once);
  * `e2e`: the same solve through the host-buffer entry point of the C ABI (mgb_problem_solve_host: H2D of the
    right-hand side from pinned memory, solve, D2H of the solution), copies inside the timed region;
  * `roofline`: the dominant kernel (finest-level FP64 Laplace cell loop), timed live with CUDA-event pairs on the
    launching stream around every launch inside the timed region (mgb_profile_start/stop);
    algorithmic bytes = 16 B/DoF (read src once, write dst once);
  * `cpu_baseline` / `--impl reference`: the threaded C restatement of the reference's CPU algorithm (oracle/mgcpu.c:
    deal.II's 12-sweep cell kernel, Chebyshev, transfers, V-cycle, CG; OpenMP over coloured cells, thread count autotuned)
    timed on the host cores on a bounded sample (deal.II itself cannot be built here: SURVEY.md 8(c)).
Workload: BASELINE.json configs[1] -- 3-D hypercube Poisson, global coarsening, p=4, FP64 V-cycle, coarse CG.
N>1: one rank per GPU (torchrun), z-order partition of the same mesh (strong scaling), NCCL ghost exchange.
`--scaling weak` (N = 1 or 8 only: a refinement level multiplies the cell count by 8) keeps the DoFs per GPU fixed instead:
N = 8 then solves hypercube L = nref+1 (L=8: 1 076 890 625 DoFs, BASELINE.json configs[4]'s "~1B DoFs across 8 B200").
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "V-cycle-preconditioned CG DoFs/s (n_dofs * n_iterations / t_solve)"
# FP64 FMA peak measured on this pool's B200 with tools/bench_fp64.cu (MEASURED_PEAKS.json carries no FP64 figure)
FP64_PEAK_TFLOPS, FP64_PEAK_SOURCE = 34.2, "measured: tools/bench_fp64.cu, profiles/r2l_bench_fp64.txt (DFMA on all 148 SMs)"
UNIT = "DoFs/s"


def workload_params(n_ref, degree=4, number="double", coarse="cg"):
    return {"Type": "HMG-global", "GeometryType": "hypercube", "NRefGlobal": str(n_ref), "NRefLocal": "0", "Degree": str(degree),
            "Paraview": False, "Verbosity": False, "PartitionerName": "DefaultPolicy", "MinLevel": "0", "MinNCells": "0",
            "CoarseGridSolverType": coarse, "SmootherDegree": 3, "CoarseSolverNCycles": 1, "RelativeTolerance": 1e-4,
            "MGNumberType": number, "SimulationType": "Constant"}


def workload_name(n_ref, degree, number):
    n1 = degree * 2 ** n_ref + 1
    return (f"hypercube L={n_ref} p={degree} HMG-global ({n1}^3 = {n1 ** 3} DoFs), MGNumberType={number}, Chebyshev(3), "
            f"coarse CG, reltol 1e-4")


class ClockSampler:
    """nvidia-smi sampling during the timed region (B200_PROFILING.md 'clocks' line)"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])), mx.append(float(r[1])), pw.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "power_w_max": max(pw) if pw else None, "samples": len(sm)}


def measured_traffic(n_ref, degree, number, n_gpus):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture of this workload
    (profiles/vmult_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum), or None if it was not captured"""
    try:
        with open(os.path.join(ROOT, "profiles", "vmult_traffic.json")) as f:
            return json.load(f).get(f"hypercube_L{n_ref}_p{degree}_{number}_{n_gpus}gpu", {}).get("dram_bytes")
    except (OSError, ValueError):
        return None


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    except (OSError, KeyError, ValueError):
        return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------ CPU arm (oracle)
def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown CPU"


def cpu_solve_sample(n_ref, degree, steps, warmup, keep=False):
    """time the threaded C restatement of the reference's CPU algorithm (oracle/mgcpu.c, built on this machine with
    -march=native when a compiler is present) on the hypercube workload of NRefGlobal = n_ref, on as many host threads as
    speed the operator up (autotuned up to all hardware threads the process may use)"""
    import dealii_multigrid_b200 as mg  # host-side map generator only (no GPU call)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from util import mesh_sequence
    from oracle import reference as ref
    from oracle import cpu_solver
    from oracle.cpu_solver import CpuMultigrid

    t0 = time.perf_counter()
    build = cpu_solver.build_native()
    meshes = mesh_sequence("hypercube", n_ref)
    dofmaps = [mg.DoFMap(m, degree) for m in meshes]
    tmaps = [None] + [mg.TransferMap(meshes[l], dofmaps[l], meshes[l - 1], dofmaps[l - 1]) for l in range(1, len(meshes))]
    M = CpuMultigrid(dofmaps, tmaps, 3, 20.0, 20)
    threads = M.autotune_threads()
    b = ref.assemble_rhs_constant(M.maps[-1])
    n = M.maps[-1].n_dofs
    setup_s = time.perf_counter() - t0
    times, its, i = [], 0, 0
    while i < warmup + steps:
        t0 = time.perf_counter()
        _, its, _, _ = M.solve(b)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
        elif dt > 40.0:
            steps = 1  # a slow host: one timed solve keeps the whole run within a few minutes
        i += 1
    best = min(times)
    r = {"n_dofs": n, "n_iterations": its, "seconds": best, "seconds_all": times, "value": n * its / best, "cores": threads,
         "setup_s": setup_s, "build": build, "cpu": cpu_model(), "steps_run": len(times),
         "hw_threads": len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)}
    return (r, M, b) if keep else r


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (deal.II itself cannot be built here, so its
    threaded C restatement oracle/mgcpu.c) on the SAME workload as the GPU arm, on the host cores.  A solve of the 135M-DoF
    problem takes tens of seconds on the host, so the run is bounded to at most 1 warm-up + 2 timed solves; the line
    reports the counts that were actually run."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_ref = args.nref if args.cpu_nref is None else args.cpu_nref
    steps, warmup = min(max(1, args.steps), 2), min(max(0, args.warmup), 1)
    r = cpu_solve_sample(n_ref, args.degree, steps, warmup)
    steps = r["steps_run"]
    sample = (f"{workload_name(n_ref, args.degree, 'double')}: best of {steps} solve(s) after {warmup} warm-up, threaded C restatement "
              f"oracle/mgcpu.c ({r['build']}; cell kernel on 8 cells at once like deal.II's VectorizedArray) on {r['cores']} of {r['hw_threads']} hardware threads, {r['cpu']}; setup {r['setup_s']:.0f} s; "
              "a port, not the deal.II binary (which cannot be built here)")
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": warmup, "requested_steps": args.steps, "requested_warmup": args.warmup,
            "ms_per_step": r["seconds"] * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(n_ref, args.degree, "double"), "sample": sample,
                       "same_workload_as_gpu_arm": n_ref == args.nref and args.number == "double"},
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": sample},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "n_iterations": r["n_iterations"]}
    print(json.dumps(line))


def parity_check(mg, M, b_cpu, n_ref, degree, number):
    """bench-scale parity inside the driver-run artefact: the GPU path against the host restatement on the SAME mesh
    (NRefGlobal = n_ref): one vmult and one V-cycle of a seeded vector, and the CG iteration counts"""
    import numpy as np
    prob = mg.Problem(workload_params(n_ref, degree, number))
    fine = prob.n_levels - 1
    op = prob.fine_operator()
    n = M.maps[-1].n_dofs
    x = np.random.default_rng(7).uniform(-1.0, 1.0, n)
    src, dst = op.initialize_dof_vector(), op.initialize_dof_vector()
    src.from_numpy(x)
    op.vmult(dst, src)
    y_cpu = M.vmult(len(M.maps) - 1, x)
    e_vmult = float(np.linalg.norm(dst.to_numpy() - y_cpu) / np.linalg.norm(y_cpu))
    r = x.copy()
    r[M.maps[-1].constrained] = 0.0
    src.from_numpy(r)
    prob.vcycle(dst, src)
    z_cpu = M.vcycle(r)
    e_vcycle = float(np.linalg.norm(dst.to_numpy() - z_cpu) / np.linalg.norm(z_cpu))
    its_gpu = prob.solve_once()["n_iterations"]
    _, its_cpu, _, _ = M.solve(b_cpu)
    tol_v, tol_c = (1e-12, 1e-11) if number == "double" else (None, 1e-3)
    ok = its_gpu == its_cpu and e_vcycle < tol_c and (tol_v is None or e_vmult < tol_v)
    return {"nref": n_ref, "n_dofs": int(n), "level": fine, "vmult_rel_l2": e_vmult, "vcycle_rel_l2": e_vcycle, "its_gpu": int(its_gpu),
            "its_cpu": int(its_cpu), "tolerance": {"vmult": tol_v, "vcycle": tol_c}, "ok": bool(ok)}


# ------------------------------------------------------------------------------------------------ GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    import dealii_multigrid_b200 as mg

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the hot path has no CPU fallback)")
    torch.cuda.set_device(local)
    mg.init(local)
    # the library enqueues on this (non-default) stream, so torch's events below bracket its work and the V-cycle can be
    # replayed from a CUDA graph
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    mg.set_stream(stream.cuda_stream)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        uid = [mg.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        mg.comm_init(uid[0], rank, world)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    t_setup = time.perf_counter()
    prob = mg.Problem(workload_params(args.nref, args.degree, args.number), world, rank)
    fine = prob.n_levels - 1
    n_dofs = prob.level_n_dofs(fine)
    dm = prob.level_dofmap(fine)
    n_owned, n_cells_local = dm.n_owned, dm.n_cells
    t_setup = time.perf_counter() - t_setup

    # host buffers of the end-to-end arm (pinned)
    rhs_host = torch.empty(n_owned, dtype=torch.float64).pin_memory()
    sol_host = torch.empty(n_owned, dtype=torch.float64).pin_memory()
    rhs_host.copy_(torch.from_numpy(prob.rhs().to_numpy()))

    for _ in range(max(args.warmup, 1)):  # the first solve also runs the lazy Chebyshev eigenvalue estimation
        res = prob.solve_once()
    prob.solve_host(rhs_host.data_ptr(), sol_host.data_ptr())

    # ---- timed region 1: K solves, inputs resident in HBM
    sampler = ClockSampler(local) if rank == 0 else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    launches0 = mg.launch_count()
    ev0.record()
    its = 0
    for _ in range(args.steps):
        res = prob.solve_once()
        its += res["n_iterations"]
    ev1.record()
    barrier()
    launches = mg.launch_count() - launches0
    ms = max_over_ranks(ev0.elapsed_time(ev1))
    # ---- timed region 1b: the same K solves with a CUDA-event pair around every finest-level cell-loop launch (the
    # events force the eager launch path instead of the CUDA-graph replay, so this region is timed separately)
    barrier()
    mg.profile_start(n_cells_local)
    ev0.record()
    for _ in range(args.steps):
        prob.solve_once()
    ev1.record()
    barrier()
    k_ms, k_launches = mg.profile_stop()
    ms_prof = max_over_ranks(ev0.elapsed_time(ev1))
    # ---- timed region 2: the same K solves through host buffers (H2D rhs, D2H solution inside the timed region)
    barrier()
    ev0.record()
    its_e2e = 0
    for _ in range(args.steps):
        r2 = prob.solve_host(rhs_host.data_ptr(), sol_host.data_ptr())
        its_e2e += r2["n_iterations"]
    ev1.record()
    barrier()
    ms_e2e = max_over_ranks(ev0.elapsed_time(ev1))
    # ---- timed region 3: the same K host-buffer solves through the batch entry point (mgb_problem_solve_host_batch): every
    # step still uploads its right-hand side from pinned memory and downloads its solution inside the timed region, but the
    # copies of neighbouring steps run on copy streams behind the solves
    prob.solve_host_batch([rhs_host.data_ptr()] * 2, [sol_host.data_ptr()] * 2)   # creates the staging buffers (untimed)
    barrier()
    t_wall = time.perf_counter()
    ev0.record()
    rb = prob.solve_host_batch([rhs_host.data_ptr()] * args.steps, [sol_host.data_ptr()] * args.steps)
    ev1.record()
    barrier()
    t_wall = time.perf_counter() - t_wall
    its_batch = sum(r["n_iterations"] for r in rb)
    ms_batch = max_over_ranks(max(ev0.elapsed_time(ev1), t_wall * 1e3))   # copies run on other streams: the host clock bounds it too
    clocks = sampler.stop() if sampler else None

    value = n_dofs * its / (ms * 1e-3)
    e2e_single = n_dofs * its_e2e / (ms_e2e * 1e-3)
    e2e = n_dofs * its_batch / (ms_batch * 1e-3)
    peak, peak_src = measured_peaks()
    k_avg_s = (k_ms / max(k_launches, 1)) * 1e-3
    k_avg_s = max_over_ranks(k_avg_s)
    value_bytes = 8 if args.number == "double" else 4  # the finest level runs in MGNumberType inside the V-cycle
    alg_bytes = 2.0 * value_bytes * n_owned
    achieved = alg_bytes / k_avg_s / 1e9 if k_launches else 0.0
    n1 = args.degree + 1
    flops_ref = n_cells_local * (24.0 * n1 ** 4 + 9.0 * n1 ** 3)   # F_alg of the 12-sweep reference kernel
    # what k_laplace_apply actually executes per cell: 6 even-odd sweeps over n^2 lines (2 ne^2 + 2 no^2 - ne + no flop per line),
    # the diagonal scaling (1 add + 1 mul per node) and the scaling with h
    ne, no = (n1 + 1) // 2, n1 // 2
    flop_per_cell = 6 * n1 * n1 * (2 * ne * ne + 2 * no * no - ne + no) + 3 * n1 ** 3
    flops_actual = n_cells_local * float(flop_per_cell)
    if rank != 0:
        return
    cpu = parity = None
    if world == 1 and not args.no_cpu:
        cpu_nref = 6 if args.cpu_nref is None else args.cpu_nref
        c, M, b_cpu = cpu_solve_sample(cpu_nref, args.degree, 2, 1, keep=True)
        cpu = {"value": c["value"], "unit": UNIT, "cores": c["cores"], "kind": "port",
               "sample": f"{workload_name(cpu_nref, args.degree, 'double')} (a bounded sample: the GPU arm runs NRefGlobal={args.nref}; "
                         f"`--impl reference` times the full workload): best of 2 solves after 1 warm-up ({c['seconds']:.1f} s each), "
                         f"threaded C restatement oracle/mgcpu.c ({c['build']}) on {c['cores']} of {c['hw_threads']} hardware threads, {c['cpu']}"}
        parity = parity_check(mg, M, b_cpu, cpu_nref, args.degree, args.number)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 1),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f64" if args.number == "double" else "f64 outer CG / f32 V-cycle", "data": "synthetic",
        "config": {"workload": workload_name(args.nref, args.degree, args.number), "partition": f"z-order, {world} rank(s)",
                   "l2": "inputs larger than L2 (each fine-level vector %.0f MB, >10 vectors touched per iteration)" % (n_owned * 8 / 1e6),
                   "n_iterations_per_solve": its / args.steps, "setup_s": round(t_setup, 2), "n_dofs": int(n_dofs),
                   "n_dofs_per_gpu": int(n_dofs // world)},
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": 8 * n_owned, "d2h_bytes_per_step": 8 * n_owned,
                "ms_per_step": ms_batch / args.steps,
                "api": "mgb_problem_solve_host_batch: K host-buffer solves in one call, upload of step i+1 and download of step i-1 "
                       "behind solve i (max of the CUDA-event and the host clock around the call)",
                "one_call_per_step": {"value": e2e_single, "ms_per_step": ms_e2e / args.steps,
                                      "api": "mgb_problem_solve_host: H2D, solve, D2H strictly in sequence"}},
        "gpu_launches": launches,
        "roofline": {"kernel": f"k_laplace_apply<{args.degree},{args.number}> (finest level" + ("" if args.number == "double" else "; the outer CG's double-precision launches on the same mesh are timed with it, bytes counted for the float ones") + ")", "bound": "hbm", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": measured_traffic(args.nref, args.degree, args.number, world),
                     "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": alg_bytes, "launches_timed": k_launches, "avg_launch_ms": k_avg_s * 1e3,
                     "share_of_step": k_ms / ms_prof, "timed_in": "second pass of the same K solves with per-launch CUDA events "
                     "(eager launches, %.3f ms/step)" % (ms_prof / args.steps), "gdofs_per_s": n_owned / k_avg_s / 1e9,
                     "fp64": {"achieved_tflops": flops_actual / k_avg_s / 1e12, "peak_tflops": FP64_PEAK_TFLOPS,
                              "frac": flops_actual / k_avg_s / 1e12 / FP64_PEAK_TFLOPS, "flop_per_cell": flop_per_cell,
                              "peak_source": FP64_PEAK_SOURCE,
                              "reference_algorithm_equivalent_tflops": flops_ref / k_avg_s / 1e12,
                              "note": "achieved = flop the kernel executes (fast diagonalisation + even-odd); the equivalent figure "
                                      "counts the reference's 12-sweep kernel (24 n^4 + 9 n^3 flop per cell) at this speed"}},
        "cpu_baseline": cpu, "parity_check": parity, "clocks": clocks,
    }
    print(json.dumps(line))
    if parity is not None and not parity["ok"]:
        raise SystemExit("bench.py: parity_check failed: %s" % json.dumps(parity))
    if world > 1:
        dist.destroy_process_group()


def apply_scaling(args):
    """--scaling weak: one more global refinement per factor 8 in the rank count (fixed DoFs per GPU)"""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.scaling == "weak":
        extra = {1: 0, 8: 1, 64: 2}.get(world)
        if extra is None:
            raise SystemExit("bench.py --scaling weak: the hypercube hierarchy only offers weak-scaling points at 1, 8, 64 ranks")
        args.nref += extra


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--nref", type=int, default=7, help="NRefGlobal of the GPU workload (7: 513^3 = 135 005 697 DoFs)")
    ap.add_argument("--degree", type=int, default=4)
    ap.add_argument("--number", default="double", choices=["double", "float"], help="MGNumberType")
    ap.add_argument("--cpu-nref", type=int, default=None,
                    help="NRefGlobal of the CPU runs (default: 6 = 257^3 DoFs for the cpu_baseline sample of the GPU arm, --nref for --impl reference)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="strong: the same mesh over N ranks (default, what the driver's 1/2/4/8 runs use); weak: nref+1 at 8 ranks")
    args = ap.parse_args()
    apply_scaling(args)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
