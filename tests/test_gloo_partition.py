"""world_size-2 test of the N-rank host logic on CPU (gloo): every process builds its rank's DoF map and partitioner with the
product's generator, the ghost exchange of LinearAlgebra::distributed::Vector (update_ghost_values / compress(add))
is carried out over gloo exactly as the tables prescribe, the cell loop in between is the numpy oracle's.  The assembled result
must equal the serial operator -- this pins the partitioner tables (ghost ranges, import lists, their pairing between ranks)
that the GPU path hands to NCCL."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _exchange(part, send_chunks, recv_sizes, send_ranks, recv_ranks):
    """one halo exchange: post every recv, send every chunk, wait"""
    bufs = [torch.zeros(int(n), dtype=torch.float64) for n in recv_sizes]
    reqs = [dist.irecv(b, src=int(r)) for b, r in zip(bufs, recv_ranks)]
    reqs += [dist.isend(torch.from_numpy(np.ascontiguousarray(c)), dst=int(r)) for c, r in zip(send_chunks, send_ranks)]
    for q in reqs:
        q.wait()
    return [b.numpy() for b in bufs]


def _worker(rank, world, port, geometry, L, p, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import dealii_multigrid_b200 as mg
    from oracle import reference as ref
    from util import seeded
    mesh = mg.Mesh(geometry, L)
    d = mg.DoFMap(mesh, p, world, rank)
    maps = ref.LevelMaps.from_dofmap(d)
    P = d.partitioner()
    n_global = d.n_dofs_global
    x_global = seeded(n_global)
    x = np.zeros(d.n_owned + d.n_ghost)
    x[:d.n_owned] = x_global[d.owned_begin:d.owned_begin + d.n_owned]
    io, go = P["import_offsets"].astype(int), P["ghost_offsets"].astype(int)
    # update_ghost_values: owners send the entries listed in import_indices, ghosters receive into their ghost ranges
    send = [x[P["import_indices"][io[i]:io[i + 1]]] for i in range(len(P["import_ranks"]))]
    got = _exchange(P, send, [go[j + 1] - go[j] for j in range(len(P["ghost_ranks"]))], P["import_ranks"], P["ghost_ranks"])
    for j, g in enumerate(got):
        x[d.n_owned + go[j]:d.n_owned + go[j + 1]] = g
    np.testing.assert_array_equal(x[d.n_owned:], x_global[P["ghost_global"]])
    # cell loop on the local cells (no identity rows yet), then compress(add): ghost contributions go back to their owners
    y = ref.Operator(maps).cell_loop(x)
    send = [y[d.n_owned + go[j]:d.n_owned + go[j + 1]] for j in range(len(P["ghost_ranks"]))]
    got = _exchange(P, send, [io[i + 1] - io[i] for i in range(len(P["import_ranks"]))], P["ghost_ranks"], P["import_ranks"])
    for i, g in enumerate(got):
        np.add.at(y, P["import_indices"][io[i]:io[i + 1]], g)
    y = y[:d.n_owned]
    y[d.constrained()] = x[d.constrained()]
    # serial reference on rank 0's copy of the whole mesh
    serial = ref.Operator(ref.LevelMaps.from_dofmap(mg.DoFMap(mesh, p))).vmult(x_global)
    err = np.linalg.norm(y - serial[d.owned_begin:d.owned_begin + d.n_owned]) / np.linalg.norm(serial)
    # a distributed dot product (Allreduce of the owned parts) must equal the serial one
    t = torch.tensor([float(y @ y)], dtype=torch.float64)
    dist.all_reduce(t)
    out[rank] = (err, abs(t.item() - serial @ serial) / (serial @ serial), d.n_owned, d.n_ghost)
    dist.destroy_process_group()


@pytest.mark.parametrize("geometry,L,p", [("hypercube", 2, 2), ("quadrant", 3, 3)])
def test_two_rank_ghost_exchange_over_gloo(geometry, L, p):
    world = 2
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_worker, args=(world, _free_port(), geometry, L, p, out), nprocs=world, join=True)
        res = dict(out)
    assert set(res) == {0, 1}
    assert sum(r[2] for r in res.values()) > 0 and any(r[3] > 0 for r in res.values())  # somebody has ghosts
    for err, dot_err, _, _ in res.values():
        assert err < 1e-13 and dot_err < 1e-13
