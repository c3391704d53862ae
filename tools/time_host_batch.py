#!/usr/bin/env python
"""time_host_batch.py -- where the time of mgb_problem_solve_host_batch goes: device-resident solves, one call per step with host
buffers, and batches of several lengths (the fixed exposure is the first upload + the last download)."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dealii_multigrid_b200 as mg
import bench

nref = int(sys.argv[1]) if len(sys.argv) > 1 else 7
torch.cuda.set_device(0)
mg.init(0)
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
mg.set_stream(stream.cuda_stream)
prob = mg.Problem(bench.workload_params(nref))
n = prob.level_dofmap(prob.n_levels - 1).n_owned
rhs = torch.from_numpy(prob.rhs().to_numpy()).pin_memory()
sol = torch.empty(n, dtype=torch.float64).pin_memory()
for _ in range(3):
    prob.solve_once()
prob.solve_host_batch([rhs.data_ptr()] * 2, [sol.data_ptr()] * 2)


def wall(f, reps=1):
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(reps):
        f()
    torch.cuda.synchronize()
    return (time.perf_counter() - t) * 1e3 / reps


out = {"nref": nref, "n_owned": n, "solve_once_ms": wall(prob.solve_once, 5), "solve_host_ms": wall(lambda: prob.solve_host(rhs.data_ptr(), sol.data_ptr()), 5)}
dev = torch.empty(n, dtype=torch.float64, device="cuda")
out["h2d_ms"] = wall(lambda: dev.copy_(rhs, non_blocking=True), 5)
out["d2h_ms"] = wall(lambda: sol.copy_(dev, non_blocking=True), 5)
for k in (1, 2, 5, 10, 20):
    out[f"batch_{k}_ms_per_step"] = wall(lambda: prob.solve_host_batch([rhs.data_ptr()] * k, [sol.data_ptr()] * k)) / k
print(json.dumps(out))
