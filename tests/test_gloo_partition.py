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


def _transfer_worker(rank, world, port, geometry, L, p, out):
    """the two-level transfer's internal ghosted fine vector (MGTwoLevelTransfer keeps its own partitioner: a rank's coarse
    cells may have children owned by the other rank): ghost update and compress(add) over gloo with the transfer map's tables"""
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import dealii_multigrid_b200 as mg
    from util import mesh_sequence, seeded
    seq = mesh_sequence(geometry, L)
    mf, mc = seq[-1], seq[-2]
    df, dc = mg.DoFMap(mf, p, world, rank), mg.DoFMap(mc, p, world, rank)
    tm = mg.TransferMap(mf, df, mc, dc)
    P = tm.partitioner()
    n_owned, n_ghost = P["n_owned"], P["n_ghost"]
    x_global = seeded(df.n_dofs_global)
    x = np.zeros(n_owned + n_ghost)
    x[:n_owned] = x_global[P["owned_begin"]:P["owned_begin"] + n_owned]
    io, go = P["import_offsets"].astype(int), P["ghost_offsets"].astype(int)
    send = [x[P["import_indices"][io[i]:io[i + 1]]] for i in range(len(P["import_ranks"]))]
    got = _exchange(P, send, [go[j + 1] - go[j] for j in range(len(P["ghost_ranks"]))], P["import_ranks"], P["ghost_ranks"])
    for j, g in enumerate(got):
        x[n_owned + go[j]:n_owned + go[j + 1]] = g
    ghost_ok = np.array_equal(x[n_owned:], x_global[P["ghost_global"]])
    # compress(add): every rank adds 1 to each entry its coarse cells' patches touch (owned or ghost); after the exchange an
    # owned entry holds the number of ranks touching it, and the global sum equals the sum of the ranks' touched counts
    touched = np.zeros(n_owned + n_ghost)
    own = tm.fine_own()
    esize = np.array([(p - 1) ** ((e % 3 == 1) + ((e // 3) % 3 == 1) + (e // 9 == 1)) for e in range(27)])
    for e in range(27):
        for k in range(esize[e]):
            touched[(own[:, e] & mg.INDEX_MASK) + k] = 1.0
    n_touched = float(touched.sum())
    send = [touched[n_owned + go[j]:n_owned + go[j + 1]] for j in range(len(P["ghost_ranks"]))]
    got = _exchange(P, send, [io[i + 1] - io[i] for i in range(len(P["import_ranks"]))], P["ghost_ranks"], P["import_ranks"])
    y = touched[:n_owned].copy()
    for i, g in enumerate(got):
        np.add.at(y, P["import_indices"][io[i]:io[i + 1]], g)
    t = torch.tensor([float(y.sum()), n_touched], dtype=torch.float64)
    dist.all_reduce(t)
    out[rank] = (bool(ghost_ok), float(t[0]), float(t[1]), float(y.max()), int(n_ghost))
    dist.destroy_process_group()


@pytest.mark.parametrize("geometry,L,p", [("hypercube", 2, 2), ("quadrant", 4, 2)])
def test_two_rank_transfer_partitioner_over_gloo(geometry, L, p):
    world = 2
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_transfer_worker, args=(world, _free_port(), geometry, L, p, out), nprocs=world, join=True)
        res = dict(out)
    assert set(res) == {0, 1}
    assert any(r[4] > 0 for r in res.values())                        # the transfer really needs fine values of the other rank
    for ghost_ok, total, touched, ymax, _ in res.values():
        assert ghost_ok and total == touched and ymax <= world
