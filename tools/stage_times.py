#!/usr/bin/env python
"""stage_times.py -- per-stage times of one solve (the reference's result-table columns) on N ranks; rank 0 prints.
usage: [torchrun ...] tools/stage_times.py [--nref 7] [--degree 4]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)) + "/..")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nref", type=int, default=7)
    ap.add_argument("--degree", type=int, default=4)
    args = ap.parse_args()
    import torch
    import dealii_multigrid_b200 as mg
    from bench import workload_params

    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    mg.init(local)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    mg.set_stream(stream.cuda_stream)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        uid = [mg.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        mg.comm_init(uid[0], rank, world)
    prob = mg.Problem(workload_params(args.nref, args.degree), world, rank)
    for _ in range(3):
        r0 = prob.solve_once(False)
    r = prob.solve_once(True)
    if rank == 0:
        keys = ["time", "time_cg", "time_pre", "time_residuum", "time_res", "time_cs", "time_pro", "time_edge_pro", "time_post", "time_to_mg",
                "time_to_global", "n_iterations"]
        out = {k: r[k] for k in keys}
        out.update(ranks=world, nref=args.nref, time_graph_replay=r0["time"])
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(0)  # like bench.py: no communicator teardown while captured graphs are alive


if __name__ == "__main__":
    main()
