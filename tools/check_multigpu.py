#!/usr/bin/env python
"""check_multigpu.py -- parity of the N-rank path against the serial CPU oracle; run under torchrun, one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/check_multigpu.py

Every rank builds its part of the problem, applies the distributed vmult / V-cycle / CG solve (ghost exchange and dot
products over NCCL) and compares its owned slice with the oracle evaluated on the serial maps (small meshes only)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    os.environ.setdefault("MGB_MIN_CELLS_PER_RANK", "64")  # partition the small parity meshes too
    import torch
    import torch.distributed as dist
    import dealii_multigrid_b200 as mg
    from util import OracleHierarchy, OracleHierarchyLS, params, rel_l2, seeded

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    mg.init(local)
    stream = torch.cuda.Stream()  # non-default stream: after the first solve the V-cycle is replayed from a CUDA graph
    torch.cuda.set_stream(stream)
    mg.set_stream(stream.cuda_stream)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    uid = [mg.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    mg.comm_init(uid[0], rank, world)
    ok = True
    cases = [("hypercube", 3, 2, "HMG-global", "cg", "DefaultPolicy"), ("quadrant", 4, 4, "HMG-global", "cg", "DefaultPolicy"),
             ("annulus", 5, 4, "HPMG", "cg_with_chebyshev", "DefaultPolicy"), ("quadrant", 4, 3, "PMG", "cg", "DefaultPolicy"),
             ("quadrant", 4, 2, "HMG-local", "cg", "DefaultPolicy"), ("annulus", 6, 1, "HMG-local", "cg", "DefaultPolicy"),
             ("quadrant", 4, 4, "HPMG-local", "cg", "DefaultPolicy"),
             # PartitionerName policies (multigrid_throughput.cc:2066-2174): the result must not depend on the partition
             ("quadrant", 4, 4, "HMG-global", "cg", "FirstChildPolicy"), ("quadrant", 4, 2, "HMG-global", "cg", "CellWeightPolicy-2.5"),
             ("annulus", 5, 4, "HPMG", "cg", "FirstChildPolicy-2.0"), ("quadrant", 4, 2, "HMG-global", "cg", "MinimalGranularityPolicy-40"),
             ("quadrant", 4, 2, "HMG-global", "cg", "BalancedGranularityPartitionPolicy"),
             ("quadrant", 4, 2, "HMG-local", "cg", "CellWeightPolicy-2.0")]
    only = os.environ.get("MGB_CHECK_ONLY")  # e.g. "Policy": run the cases whose description contains the string
    for geometry, L, p, mg_type, coarse, policy in cases:
        if only and only not in f"{geometry} {mg_type} {policy}":
            continue
        prm = params(geometry, L, p, mg_type, coarse)
        prm["PartitionerName"] = policy
        prob = mg.Problem(prm, world, rank)
        if mg_type.endswith("local"):
            H = OracleHierarchyLS(geometry, L, p, mg_type, 3, coarse)
            H.maps, H.ops = [H.fine_maps], [H.fine_op]  # only the outer operator is compared level-wise below
        else:
            H = OracleHierarchy(geometry, L, p, mg_type, 3, coarse)
        d = prob.fine_dofmap()
        lo, hi = d.owned_begin, d.owned_begin + d.n_owned
        n = H.maps[-1].n_dofs
        x = seeded(n)
        op = prob.fine_operator()
        src, dst = op.initialize_dof_vector(), op.initialize_dof_vector()
        src.from_numpy(x[lo:hi])
        op.vmult(dst, src)
        e_vmult = rel_l2(dst.to_numpy(d.n_owned), H.ops[-1].vmult(x)[lo:hi])
        r = x.copy()
        r[H.maps[-1].constrained] = 0.0
        src.from_numpy(r[lo:hi])
        prob.vcycle(dst, src)
        e_vc = rel_l2(dst.to_numpy(d.n_owned), H.mg.vmult(r)[lo:hi])
        b = np.zeros(n)
        from oracle import reference as ref
        b = ref.assemble_rhs_constant(H.maps[-1])
        e_rhs = rel_l2(prob.rhs().to_numpy(d.n_owned), b[lo:hi])
        _, its, _ = H.solve(b.copy())
        res = prob.solve_once()
        x1 = prob.solution().to_numpy(d.n_owned)
        res2 = prob.solve_once()          # V-cycle captured (NCCL ghost exchange included) and replayed
        res3 = prob.solve_once()
        e_graph = rel_l2(prob.solution().to_numpy(d.n_owned), x1) if d.n_owned else 0.0
        src.from_numpy(r[lo:hi])
        prob.vcycle(dst, src)
        prob.vcycle(dst, src)
        e_vc = max(e_vc, rel_l2(dst.to_numpy(d.n_owned), H.mg.vmult(r)[lo:hi]))
        good = (e_vmult < 1e-12 and e_vc < 1e-11 and e_rhs < 1e-13 and res["n_iterations"] == its == res2["n_iterations"] == res3["n_iterations"]
                and e_graph < 1e-8)  # two solves differ by the order of the atomic adds, amplified by the CG iteration
        ok &= good
        print(f"[rank {rank}/{world}] {geometry} L={L} p={p} {mg_type} {policy}: owned={d.n_owned} ghost={d.n_ghost} vmult={e_vmult:.1e} "
              f"vcycle={e_vc:.1e} rhs={e_rhs:.1e} its={res['n_iterations']} (oracle {its}) replay={e_graph:.1e} {'OK' if good else 'FAIL'}", flush=True)
    t = torch.tensor([1.0 if ok else 0.0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    good = t.item() == 1.0
    if rank == 0:
        print("MULTIGPU PARITY", "OK" if good else "FAILED", flush=True)
    dist.barrier()
    torch.cuda.synchronize()
    # CUDA graphs that captured NCCL calls keep the communicator referenced (ncclCommDestroy would wait for them), and Python
    # may still hold problems through borrowed handles: leave without finalising, like bench.py
    sys.stdout.flush()
    if os.environ.get("MGB_CHECK_FINALIZE") == "1":  # exercise mgb_comm_finalize (releases the captured graphs first)
        mg.comm_finalize()
        print(f"[rank {rank}] comm_finalize returned", flush=True)
    os._exit(0 if good else 1)


if __name__ == "__main__":
    main()
