#!/usr/bin/env python
"""time_halo.py -- time update_ghost_values + compress(add) pairs of one level, NCCL send/recv against the NVLink peer-memory
kernels (device/halo.cu); run under torchrun, one rank per GPU.  Prints one line per mode on rank 0."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    os.environ["MGB_P2P"] = "1"
    ap = argparse.ArgumentParser()
    ap.add_argument("--nref", type=int, default=6)
    ap.add_argument("--degree", type=int, default=4)
    ap.add_argument("--reps", type=int, default=200)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import dealii_multigrid_b200 as mg

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    mg.init(local)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    mg.set_stream(stream.cuda_stream)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    uid = [mg.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    mg.comm_init(uid[0], rank, world)
    d = mg.DoFMap(mg.Mesh("hypercube", args.nref), args.degree, world, rank)
    for mode in ("nccl", "p2p"):
        op = mg.Operator(d)
        if mode == "p2p" and not mg.comm_enable_peer_exchange():
            print("peer exchange unavailable")
            break
        v = op.initialize_dof_vector()
        v.set(1.0)
        for _ in range(5):
            v.update_ghost_values()
            v.compress_add()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.reps):
            v.update_ghost_values()
            v.compress_add()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / args.reps], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        g = torch.tensor([float(d.n_ghost)], device="cuda")
        dist.all_reduce(g, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(json.dumps({"mode": mode, "ranks": world, "nref": args.nref, "degree": args.degree, "max_n_ghost": int(g.item()),
                              "us_per_update_plus_compress": 1e3 * t.item()}), flush=True)
    mg.comm_finalize()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
