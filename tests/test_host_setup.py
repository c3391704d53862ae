"""CPU tests of the stand-in setup (mesh, DoF numbering, constraints, transfer maps) and of the C ABI surface.
They verify the product's index maps through geometry with the oracle's independent formulations."""
import os
import re

import numpy as np
import pytest

import dealii_multigrid_b200 as mg
from dealii_multigrid_b200 import _lib
from oracle import fe as ofe
from oracle import reference as ref
from util import mesh_sequence

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_capi_exports_every_declared_symbol():
    """the shared library loads and exports exactly the entry points include/mgb200.h declares"""
    header = open(os.path.join(ROOT, "include", "mgb200.h")).read()
    declared = set(re.findall(r"\b(mgb_[a-z0-9_]+)\s*\(", header))
    lib = _lib.load()
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in mgb200.h but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)


def test_no_cpu_fallback_without_gpu():
    """compute entry points must fail loudly when no CUDA device is present (this container has none)"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(mg.MgbError, match="no CUDA device"):
        mg.init(0)
    with pytest.raises(mg.MgbError, match="no CUDA device"):
        mg.Vector(10)


@pytest.mark.parametrize("p", [1, 2, 3, 4, 5, 6])
def test_fe_1d_matches_oracle(p):
    a, b = mg.fe_1d(p), ofe.FE1D(p)
    np.testing.assert_allclose(a["nodes"], b.nodes, atol=1e-14)
    for k, ref_m in (("M", b.M), ("K", b.K), ("S0", b.S[0]), ("S1", b.S[1]), ("Ph", b.Ph)):
        np.testing.assert_allclose(a[k], ref_m, atol=2e-13, err_msg=k)
    assert abs(a["M"].sum() - 1.0) < 1e-14          # integral of 1
    np.testing.assert_allclose(a["K"].sum(axis=1), 0, atol=1e-12)  # constants in the kernel
    np.testing.assert_allclose(a["S0"].sum(axis=1), 1, atol=1e-13)
    if p > 1:
        np.testing.assert_allclose(mg.fe_p_embedding(p, max(p // 2, 1)), ofe.p_embedding(p, max(p // 2, 1)), atol=1e-13)


def _check_balance(cells):
    """2:1 balance across faces, edges and corners, by brute force"""
    leaves = {tuple(c) for c in cells.tolist()}
    for (l, x, y, z) in leaves:
        n = 1 << l
        for dx in (-1, 0, 1):
            for dy in (-1, 0, 1):
                for dz in (-1, 0, 1):
                    nx, ny, nz = x + dx, y + dy, z + dz
                    if (dx, dy, dz) == (0, 0, 0) or not (0 <= nx < n and 0 <= ny < n and 0 <= nz < n):
                        continue
                    for s in range(2, l + 1):  # an ancestor >= 2 levels coarser must not be a leaf
                        assert (l - s, nx >> s, ny >> s, nz >> s) not in leaves


@pytest.mark.parametrize("geometry,L,n_expected", [("hypercube", 3, 512), ("quadrant", 4, None), ("annulus", 6, None),
                                                    ("circle", 5, None), ("quadrant_flexible", 2, 64)])
def test_mesh_and_coarsening_sequence(geometry, L, n_expected):
    seq = mesh_sequence(geometry, L)
    fine = seq[-1]
    if n_expected:
        assert fine.n_cells == n_expected
    assert seq[0].n_cells == 1 and len(seq) == fine.n_global_levels
    vol = sum(8.0 ** (-int(c[0])) for c in fine.cells())
    assert abs(vol - 1.0) < 1e-12  # leaves tile the root cube
    for m in seq:
        _check_balance(m.cells())
    # every coarse cell is a fine leaf or the parent of 8 fine leaves
    for coarse, finer in zip(seq[:-1], seq[1:]):
        fl = {tuple(c) for c in finer.cells().tolist()}
        for (l, x, y, z) in coarse.cells().tolist():
            if (l, x, y, z) in fl:
                continue
            assert all((l + 1, 2 * x + (c & 1), 2 * y + ((c >> 1) & 1), 2 * z + ((c >> 2) & 1)) in fl for c in range(8))


def test_quadrant_matches_reference_rule():
    """create_quadrant (include/grid_generator.h:34-65): L=2 -> 8 cells, one of them refined again"""
    assert mg.Mesh("quadrant", 1).n_cells == 8
    assert mg.Mesh("quadrant", 2).n_cells == 15
    m = mg.Mesh("quadrant", 3)
    assert m.n_global_levels == 4


@pytest.mark.parametrize("L,p", [(2, 1), (3, 2), (2, 4), (1, 6)])
def test_hypercube_dof_count(L, p):
    d = mg.DoFMap(mg.Mesh("hypercube", L), p)
    assert d.n_dofs_global == (p * 2 ** L + 1) ** 3
    assert d.n_cells_hanging == 0
    # Dirichlet DoFs = boundary nodes
    assert d.n_constrained == (p * 2 ** L + 1) ** 3 - (p * 2 ** L - 1) ** 3


def _poly(xyz, p):
    x, y, z = xyz[..., 0], xyz[..., 1], xyz[..., 2]
    return 1.0 + 0.3 * x ** p - 0.7 * y ** p * x + 0.5 * z ** p * y ** max(p - 1, 0) + x * y * z


@pytest.mark.parametrize("geometry,L", [("hypercube", 2), ("quadrant", 3), ("quadrant", 4), ("annulus", 5), ("circle", 4)])
@pytest.mark.parametrize("p", [1, 2, 4])
def test_dofmap_reproduces_polynomials(geometry, L, p):
    """geometric pin of the index maps: with master DoFs set by nodal interpolation of a degree-p polynomial,
    gather through the evaluation table + hanging-node interpolation reproduces the polynomial at every cell's
    own nodes; own tables give every DoF one consistent support point."""
    d = mg.DoFMap(mg.Mesh(geometry, L), p)
    maps = ref.LevelMaps.from_dofmap(d)
    xyz_cells = maps.node_coordinates()
    # consistency of the numbering: every (cell, node) referencing a DoF sees the same coordinates
    xyz = maps.dof_coordinates()
    assert not np.isnan(xyz).any(), "a DoF is not referenced by any cell"
    np.testing.assert_allclose(xyz[maps.own_idx], xyz_cells, atol=1e-14)
    assert len(np.unique(maps.own_idx)) == d.n_dofs_global
    u = _poly(xyz, p)
    # slots hold real DoFs (Dirichlet flags ignored here: inhomogeneous values)
    for m in np.unique(maps.mask):
        cells = np.nonzero(maps.mask == m)[0]
        loc = u[maps.eval_idx[cells]]
        if m >> 3:
            loc = loc @ maps.C(m).T
        np.testing.assert_allclose(loc, _poly(xyz_cells[cells], p), atol=1e-11)
    # constrained list = Dirichlet (on the boundary) + hanging DoFs (own entity replaced by a coarse one)
    on_bnd = (np.abs(np.abs(xyz) - 1.0) < 1e-14).any(axis=1)
    hanging = np.zeros(d.n_dofs_global, dtype=bool)
    for m in np.unique(maps.mask):
        if m >> 3:
            cells = np.nonzero(maps.mask == m)[0]
            differs = maps.eval_idx[cells] != maps.own_idx[cells]
            hanging[maps.own_idx[cells][differs]] = True
    np.testing.assert_array_equal(np.nonzero(on_bnd | hanging)[0], maps.constrained)
    # flags: eval flag <=> slot DoF on the boundary; own flag <=> constrained
    slot_xyz_on_bnd = on_bnd[maps.eval_idx]
    np.testing.assert_array_equal(maps.eval_flag, slot_xyz_on_bnd)
    np.testing.assert_array_equal(maps.own_flag, (on_bnd | hanging)[maps.own_idx])
    # hanging DoFs are never masters
    assert not hanging[maps.eval_idx].any()


def test_full_index_map_matches_entity_table():
    d = mg.DoFMap(mg.Mesh("quadrant", 3), 3)
    idx, flag = ref.expand_table(d.eval_table(), 3)
    full = d.expand("eval")
    np.testing.assert_array_equal(full & mg.INDEX_MASK, idx)
    np.testing.assert_array_equal((full & mg.FLAG_BIT) != 0, flag)


@pytest.mark.parametrize("geometry,L,p", [("hypercube", 2, 2), ("quadrant", 4, 1), ("quadrant", 3, 4), ("annulus", 5, 2)])
def test_transfer_map_weights(geometry, L, p):
    seq = mesh_sequence(geometry, L)
    mf, mc = seq[-1], seq[-2]
    df, dc = mg.DoFMap(mf, p), mg.DoFMap(mc, p)
    tm = mg.TransferMap(mf, df, mc, dc)
    assert tm.n_children().astype(int).sum() == mf.n_cells
    fine, coarse = ref.LevelMaps.from_dofmap(df), ref.LevelMaps.from_dofmap(dc)
    w = ref.transfer_weights(fine, coarse, tm.fine_first(), tm.n_children())
    code = tm.weight_code().astype(float)
    ent, _ = ref.node_entity_tables(p)
    w_prod = np.where(code > 0, 1.0 / np.maximum(code, 1), 0.0)[:, ent]
    np.testing.assert_array_equal(w_prod, w[fine.own_idx])


POLICIES = [None, "CellWeightPolicy-2.5", "FirstChildPolicy", "FirstChildPolicy-2.0", "MinimalGranularityPolicy-40",
            "BalancedGranularityPartitionPolicy"]


def _ranges(policy, seq, n_ranks):
    """cell_begin per mesh of the coarse -> fine sequence (None: the default equal split inside DoFMap)"""
    return [None] * len(seq) if policy is None else list(mg.partition_sequence(policy, seq, n_ranks))


@pytest.mark.parametrize("policy", POLICIES)
@pytest.mark.parametrize("n_ranks", [2, 3, 8])
@pytest.mark.parametrize("geometry,L,p", [("hypercube", 2, 2), ("quadrant", 4, 3)])
def test_partitioned_dofmaps_are_consistent(geometry, L, p, n_ranks, policy):
    """rank-local maps describe the same global numbering as the serial map; ghost/import lists match pairwise -- for the
    default split and for every PartitionerName policy (ranks may then be empty, also in the middle of the rank order)"""
    seq = mesh_sequence(geometry, L)
    level = len(seq) - 1 if policy in (None, "CellWeightPolicy-2.5") else len(seq) - 2  # policies that act on coarse meshes
    mesh, cb = seq[level], _ranges(policy, seq, n_ranks)[level]
    serial = mg.DoFMap(mesh, p)
    parts = [mg.DoFMap(mesh, p, n_ranks, r, cell_begin=cb) for r in range(n_ranks)]
    if cb is not None:
        assert [d.cell_begin for d in parts] == list(cb[:-1])
    assert sum(d.n_owned for d in parts) == serial.n_dofs_global
    assert sum(d.n_cells for d in parts) == serial.n_cells
    s_eval = serial.eval_table()
    for d in parts:
        pt = d.partitioner()
        l2g = np.concatenate([np.arange(d.owned_begin, d.owned_begin + d.n_owned), pt["ghost_global"]])
        t = d.eval_table()
        g = l2g[t & mg.INDEX_MASK] | (t & mg.FLAG_BIT)
        np.testing.assert_array_equal(g, s_eval[d.cell_begin:d.cell_begin + d.n_cells])
        np.testing.assert_array_equal(l2g[d.constrained()], serial.constrained()[(serial.constrained() >= d.owned_begin) &
                                                                               (serial.constrained() < d.owned_begin + d.n_owned)])
    # what rank a imports from ... equals what rank b expects as ghosts, in the same order
    for a in parts:
        pa = a.partitioner()
        for i, b_rank in enumerate(pa["import_ranks"]):
            send = pa["import_indices"][pa["import_offsets"][i]:pa["import_offsets"][i + 1]] + a.owned_begin
            pb = parts[b_rank].partitioner()
            j = list(pb["ghost_ranks"]).index(a.rank)
            recv = pb["ghost_global"][pb["ghost_offsets"][j]:pb["ghost_offsets"][j + 1]]
            np.testing.assert_array_equal(send, recv)
    # ... and every ghost range has its import on the owner's side (no exchange partner is missed)
    for b in parts:
        for a_rank in b.partitioner()["ghost_ranks"]:
            assert b.rank in list(parts[a_rank].partitioner()["import_ranks"])


@pytest.mark.parametrize("p", [1, 2, 3, 4, 5, 6])
def test_fast_diagonalisation_of_the_1d_pencil(p):
    """the cell kernels evaluate A_e = h (Tm x Tm x Tm)^T diag(l_i+l_j+l_k) (Tm x Tm x Tm): K = Tm^T L Tm, M = Tm^T Tm,
    modes exactly symmetric / antisymmetric (even-odd sweeps), constant mode in the null space"""
    fe, (Tm, lam) = mg.fe_1d(p), mg.fe_1d_fdm(p)
    n, ne = p + 1, (p + 2) // 2
    assert np.abs(Tm.T @ np.diag(lam) @ Tm - fe["K"]).max() < 1e-13 * np.abs(fe["K"]).max()
    assert np.abs(Tm.T @ Tm - fe["M"]).max() < 1e-15
    assert np.array_equal(Tm[:ne], Tm[:ne, ::-1]) and np.array_equal(Tm[ne:], -Tm[ne:, ::-1])
    assert lam[0] == 0.0 and np.all(lam[1:] > 0)
    # 3-D cell matrix from the factorisation equals the Kronecker form
    if p <= 3:
        T3 = np.kron(Tm, np.kron(Tm, Tm))
        D = (lam[:, None, None] + lam[None, :, None] + lam[None, None, :]).ravel()
        M, K = fe["M"], fe["K"]
        A = np.kron(K, np.kron(M, M)) + np.kron(M, np.kron(K, M)) + np.kron(M, np.kron(M, K))
        assert np.abs(T3.T @ (D[:, None] * T3) - A).max() < 1e-13 * np.abs(A).max()


def _check_partitioned_transfer(mf, pf, mc, pc, cbf, cbc, n_ranks, ls=False):
    sf, sc = mg.DoFMap(mf, pf), mg.DoFMap(mc, pc)
    st = mg.TransferMap(mf, sf, mc, sc, local_smoothing=ls)
    s_own, s_first, s_nch, s_w = sf.own_table(), st.fine_first(), st.n_children(), st.weight_code()
    tms = []
    for r in range(n_ranks):
        df, dc = mg.DoFMap(mf, pf, n_ranks, r, cell_begin=cbf), mg.DoFMap(mc, pc, n_ranks, r, cell_begin=cbc)
        tm = mg.TransferMap(mf, df, mc, dc, local_smoothing=ls)
        tms.append((tm, df))
        assert tm.distributed and tm.n_coarse_cells == dc.n_cells
        pt = tm.partitioner()
        assert pt["owned_begin"] == df.owned_begin and pt["n_owned"] == df.n_owned
        l2g = np.concatenate([np.arange(df.owned_begin, df.owned_begin + df.n_owned), pt["ghost_global"]])
        own, first, nch, w = tm.fine_own(), tm.fine_first(), tm.n_children(), tm.weight_code()
        np.testing.assert_array_equal(nch, s_nch[dc.cell_begin:dc.cell_begin + dc.n_cells])
        esize = np.array([(pf - 1) ** ((e % 3 == 1) + ((e // 3) % 3 == 1) + (e // 9 == 1)) for e in range(27)])
        for c in range(dc.n_cells):
            for ch in range(nch[c]):
                slot, sfine = first[c] + ch, s_first[dc.cell_begin + c] + ch
                has = esize > 0
                np.testing.assert_array_equal(l2g[own[slot] & mg.INDEX_MASK][has], (s_own[sfine] & mg.INDEX_MASK)[has])
                np.testing.assert_array_equal((own[slot] & mg.FLAG_BIT)[has], (s_own[sfine] & mg.FLAG_BIT)[has])
                np.testing.assert_array_equal(w[slot], s_w[sfine])
    for tm, df in tms:
        pa = tm.partitioner()
        for i, b_rank in enumerate(pa["import_ranks"]):
            send = pa["import_indices"][pa["import_offsets"][i]:pa["import_offsets"][i + 1]] + df.owned_begin
            pb = tms[b_rank][0].partitioner()
            j = list(pb["ghost_ranks"]).index(df.rank)
            recv = pb["ghost_global"][pb["ghost_offsets"][j]:pb["ghost_offsets"][j + 1]]
            np.testing.assert_array_equal(send, recv)
        # every ghost group has a sender
        for j, o in enumerate(pa["ghost_ranks"]):
            assert df.rank in list(tms[o][0].partitioner()["import_ranks"])


@pytest.mark.parametrize("policy", [None, "FirstChildPolicy-2.0", "CellWeightPolicy-1.5"])
@pytest.mark.parametrize("n_ranks", [2, 3, 8])
@pytest.mark.parametrize("geometry,L,p,kind", [("hypercube", 2, 2, "h"), ("quadrant", 4, 3, "h"), ("quadrant", 3, 4, "p"), ("annulus", 5, 1, "h")])
def test_partitioned_transfer_maps_are_consistent(geometry, L, p, kind, n_ranks, policy):
    """several ranks: the coarse cells of a rank may have children on other ranks; the transfer's own fine-side table,
    weights and ghost/import lists must describe the serial transfer in the global numbering"""
    seq = mesh_sequence(geometry, L)
    cbs = _ranges(policy, seq, n_ranks)
    (mf, pf), (mc, pc) = ((seq[-1], p), (seq[-2], p)) if kind == "h" else ((seq[-1], p), (seq[-1], max(p // 2, 1)))
    cbf, cbc = (cbs[-1], cbs[-2]) if kind == "h" else (cbs[-1], cbs[-1])
    _check_partitioned_transfer(mf, pf, mc, pc, cbf, cbc, n_ranks)


# ----------------------------------------------------------------------------------------------- local smoothing
LS_SETUP_CASES = [("quadrant", 3, 2), ("quadrant", 4, 1), ("quadrant", 4, 3), ("annulus", 6, 2), ("circle", 4, 2), ("hypercube", 2, 3)]


def _covered(points, cell_set, level):
    """is each point strictly inside a cell of the level-`level` cell set?"""
    h = 2.0 / 2 ** level
    ijk = np.floor((points + 1.0) / h).astype(np.int64)
    return np.array([tuple(c) in cell_set for c in ijk])


@pytest.mark.parametrize("geometry,L,p", LS_SETUP_CASES)
def test_level_meshes_and_refinement_edges(geometry, L, p):
    """level l of the LS hierarchy = all ancestors at level l; edge DoFs = level DoFs on a face between a level cell and
    the part of the domain not covered by level cells (MGTools::extract_inner_interface_dofs), checked geometrically
    with probe points around every DoF; edge cells = cells touching such a DoF (include/operator.h:71-99)"""
    mesh = mg.Mesh(geometry, L)
    cells = mesh.cells().astype(np.int64)
    for l in range(mesh.n_global_levels):
        lm = mesh.level_mesh(l)
        lc = lm.cells().astype(np.int64)
        fine = cells[cells[:, 0] >= l]
        anc = {tuple(int(v) >> int(c[0] - l) for v in c[1:]) for c in fine}
        assert {tuple(c[1:]) for c in lc} == anc and np.all(lc[:, 0] == l) and len(lc) == len(anc)
        d = mg.DoFMap(lm, p)
        e = mg.LevelEdges(lm, d)
        maps = ref.LevelMaps.from_dofmap(d)
        xyz = maps.dof_coordinates()
        assert not np.isnan(xyz).any()
        eps = 1e-3 * 2.0 / 2 ** l
        signs = np.array([[sx, sy, sz] for sz in (-1, 1) for sy in (-1, 1) for sx in (-1, 1)])
        cov = np.stack([_covered(xyz + eps * s, anc, l) for s in signs], axis=1)               # (n_dofs, 8)
        inside = np.stack([np.all(np.abs(xyz + eps * s) < 1.0, axis=1) for s in signs], axis=1)
        on_edge = np.zeros(len(xyz), dtype=bool)
        for a in range(8):
            for dd in range(3):
                b = a ^ (1 << dd)
                on_edge |= inside[:, a] & inside[:, b] & (cov[:, a] != cov[:, b])
        np.testing.assert_array_equal(np.nonzero(on_edge)[0], e.edge_dofs)
        assert e.any_global == bool(on_edge.any())
        touching = np.nonzero(on_edge[maps.own_idx].any(axis=1))[0]
        np.testing.assert_array_equal(touching, e.edge_cells)
        assert len(maps.constrained) == np.count_nonzero(np.any(np.abs(np.abs(xyz) - 1.0) < 1e-12, axis=1))  # Dirichlet only


@pytest.mark.parametrize("geometry,L,p", LS_SETUP_CASES)
def test_copy_indices_cover_the_active_dofs_once(geometry, L, p):
    """copy_indices: pairs connect DoFs at the same point; every active DoF that is not a hanging node is in
    exactly one level's list, hanging nodes in none; level DoFs on the refinement edge are never copied"""
    mesh = mg.Mesh(geometry, L)
    act = mg.DoFMap(mesh, p)
    amaps = ref.LevelMaps.from_dofmap(act)
    axyz = amaps.dof_coordinates()
    count = np.zeros(amaps.n_dofs, dtype=int)
    for l in range(mesh.n_global_levels):
        lm = mesh.level_mesh(l)
        d = mg.DoFMap(lm, p)
        ci = mg.CopyIndices(act, lm, d)
        lxyz = ref.LevelMaps.from_dofmap(d).dof_coordinates()
        np.testing.assert_allclose(axyz[ci.active_idx], lxyz[ci.level_idx], rtol=0, atol=1e-13)
        assert len(np.unique(ci.level_idx)) == len(ci.level_idx)
        assert not np.intersect1d(ci.level_idx, mg.LevelEdges(lm, d).edge_dofs).size
        count[ci.active_idx] += 1
    dirichlet = np.any(np.abs(np.abs(axyz) - 1.0) < 1e-12, axis=1)
    hanging = np.zeros(amaps.n_dofs, dtype=bool)
    hanging[amaps.constrained] = True
    hanging &= ~dirichlet
    np.testing.assert_array_equal(count[~dirichlet], np.where(hanging, 0, 1)[~dirichlet])
    assert count[dirichlet].max() <= 1  # boundary DoFs that are hanging nodes as well are in no list (their value is 0)


@pytest.mark.parametrize("geometry,L,p", [("quadrant", 3, 2), ("quadrant", 4, 1), ("annulus", 6, 1), ("hypercube", 2, 2)])
def test_ls_transfer_map(geometry, L, p):
    mesh = mg.Mesh(geometry, L)
    for l in range(1, mesh.n_global_levels):
        mf, mc = mesh.level_mesh(l), mesh.level_mesh(l - 1)
        df, dc = mg.DoFMap(mf, p), mg.DoFMap(mc, p)
        tm = mg.TransferMap(mf, df, mc, dc, local_smoothing=True)
        nch = tm.n_children().astype(int)
        assert set(nch.tolist()) <= {0, 8} and nch.sum() == mf.n_cells
        cc, fc, first = mc.cells().astype(np.int64), mf.cells().astype(np.int64), tm.fine_first()
        for c in np.nonzero(nch == 8)[0]:
            np.testing.assert_array_equal(fc[first[c]][1:] >> 1, cc[c][1:])
        fine, coarse = ref.LevelMaps.from_dofmap(df), ref.LevelMaps.from_dofmap(dc)
        # prolongation reproduces a coarse FE function that vanishes on the boundary on the refined part of the level
        P = ref.MGTwoLevelTransfer(fine, coarse, first, nch)
        f = lambda x: np.prod(1.0 - x ** 2, axis=1) * (1.0 + 0.3 * x[:, 0] - 0.2 * x[:, 1] * x[:, 2]) if p >= 2 else np.prod(1.0 - np.abs(x), axis=1)
        if p >= 4 or p == 1:
            uc = f(coarse.dof_coordinates())
            uf = np.zeros(fine.n_dofs)
            P.prolongate_and_add(uf, uc)
            if p >= 4:
                np.testing.assert_allclose(uf, f(fine.dof_coordinates()), atol=1e-12)


@pytest.mark.parametrize("first_child", [False, True])
@pytest.mark.parametrize("n_ranks", [2, 3])
@pytest.mark.parametrize("geometry,L,p", [("quadrant", 4, 2), ("annulus", 6, 1)])
def test_partitioned_copy_indices_are_consistent(geometry, L, p, n_ranks, first_child):
    """several ranks: the pairs held by the owners of the active DoFs, mapped to global indices, are exactly the serial pairs;
    level cells split evenly, or owned by the owner of their first child (deal.II's level ownership)"""
    mesh = mg.Mesh(geometry, L)
    act = mg.DoFMap(mesh, p)
    act_begin = mg.partition_sequence("CellWeightPolicy-2.0" if first_child else "DefaultPolicy", [mesh], n_ranks)[0]
    for l in range(mesh.n_global_levels):
        lm = mesh.level_mesh(l)
        lvl_begin = mg.first_child_boundaries(lm, mesh, act_begin) if first_child else None
        serial = mg.CopyIndices(act, lm, mg.DoFMap(lm, p))
        expect = set(zip(serial.active_idx.tolist(), serial.level_idx.tolist()))
        got = set()
        for r in range(n_ranks):
            ar, lr = mg.DoFMap(mesh, p, n_ranks, r, cell_begin=act_begin), mg.DoFMap(lm, p, n_ranks, r, cell_begin=lvl_begin)
            ci = mg.CopyIndices(ar, lm, lr, active_mesh=mesh)
            assert ci.distributed and ci.n_owned == lr.n_owned and ci.owned_begin == lr.owned_begin
            assert np.all(ci.active_idx < ar.n_owned)
            lg = np.where(ci.level_idx < ci.n_owned, ci.level_idx.astype(np.int64) + ci.owned_begin,
                          ci.ghost_global[np.maximum(ci.level_idx.astype(np.int64) - ci.n_owned, 0)] if ci.n_ghost else 0)
            pairs = set(zip((ci.active_idx.astype(np.int64) + ar.owned_begin).tolist(), lg.tolist()))
            assert not (pairs & got)
            got |= pairs
        assert got == expect


# ----------------------------------------------------------------------------------------------- partition policies
def _first_leaf_owner(parents, children, child_begin):
    """independent restatement: owner of the leaf of `children` whose start corner is the parent's start corner"""
    pc, cc = parents.cells().astype(np.int64), children.cells().astype(np.int64)
    lmax = int(cc[:, 0].max())
    start = {tuple((cc[i, 1:] << (lmax - cc[i, 0])).tolist()): i for i in range(len(cc))}
    leaf = np.array([start[tuple((c[1:] << (lmax - c[0])).tolist())] for c in pc])
    return np.searchsorted(np.asarray(child_begin)[1:], leaf, side="right")


@pytest.mark.parametrize("n_ranks", [3, 8])
@pytest.mark.parametrize("geometry,L", [("quadrant", 4), ("annulus", 6), ("hypercube", 3)])
def test_partition_policies(geometry, L, n_ranks):
    """PartitionerName policies (multigrid_throughput.cc:2066-2174) on a global-coarsening sequence"""
    seq = mesh_sequence(geometry, L)
    nc = np.array([m.n_cells for m in seq])
    owner = lambda cb, n: np.searchsorted(cb[1:], np.arange(n), side="right")

    d = mg.partition_sequence("DefaultPolicy", seq, n_ranks)
    for k, m in enumerate(seq):
        used = max(1, min(n_ranks, nc[k]))                                # fewer cells than ranks: the first ranks only
        np.testing.assert_array_equal(d[k], (nc[k] * np.minimum(np.arange(n_ranks + 1), used)) // used)
    np.testing.assert_array_equal(mg.partition_sequence("", seq, n_ranks), d)

    g = mg.partition_sequence("MinimalGranularityPolicy-40", seq, n_ranks)
    for k in range(len(seq)):
        used = max(1, min(n_ranks, nc[k] // 40))
        sizes = np.diff(g[k])
        assert np.count_nonzero(sizes) == used and np.all(sizes[used:] == 0) and sizes[:used].max() - sizes[:used].min() <= 1

    hang = seq[-1].hanging_cells()
    np.testing.assert_array_equal(hang, (mg.DoFMap(seq[-1], 1).mask() >> 3) != 0)
    w = mg.partition_sequence("CellWeightPolicy-2.5", seq, n_ranks)
    for k, m in enumerate(seq):
        weights = np.where(m.hanging_cells(), 2.5, 1.0)
        per_rank = np.bincount(owner(w[k], nc[k]), weights=weights, minlength=n_ranks)
        assert np.all(np.diff(w[k]) >= 0) and w[k][0] == 0 and w[k][-1] == nc[k]
        assert per_rank.max() - weights.sum() / n_ranks <= 2.5 + 1e-9     # no rank exceeds its share by more than one cell
    if hang.any() and n_ranks == 3:                                      # (8 ranks = the 8 octants: symmetric meshes split evenly anyway)
        assert not np.array_equal(w[-1], d[-1])
    np.testing.assert_array_equal(mg.partition_sequence("CellWeightPolicy-1.0", seq, n_ranks), d)

    f = mg.partition_sequence("FirstChildPolicy", seq, n_ranks)
    np.testing.assert_array_equal(f[-1], d[-1])
    for k in range(len(seq) - 1):
        np.testing.assert_array_equal(owner(f[k], nc[k]), _first_leaf_owner(seq[k], seq[k + 1], f[k + 1]))
        np.testing.assert_array_equal(f[k], mg.first_child_boundaries(seq[k], seq[k + 1], f[k + 1]))
    np.testing.assert_array_equal(mg.partition_sequence("FirstChildPolicy-2.5", seq, n_ranks)[-1], w[-1])

    b = mg.partition_sequence("BalancedGranularityPartitionPolicy", seq, n_ranks)
    procs = n_ranks
    np.testing.assert_array_equal(b[-1], d[-1])
    for k in range(len(seq) - 2, -1, -1):
        grain = min(200, 8 * nc[k] // procs + 1)                         # multigrid_throughput.cc:397-398
        procs = max(1, min(n_ranks, nc[k] // grain))
        assert np.count_nonzero(np.diff(b[k])) == procs

    with pytest.raises(RuntimeError):
        mg.partition_sequence("NoSuchPolicy", seq, n_ranks)


def test_level_cells_live_with_their_first_child():
    """local smoothing: level owner of a cell = owner of its first child, recursively"""
    mesh = mg.Mesh("quadrant", 4)
    act = mg.partition_sequence("DefaultPolicy", [mesh], 3)[0]
    for l in range(mesh.n_global_levels):
        lm = mesh.level_mesh(l)
        cb = mg.first_child_boundaries(lm, mesh, act)
        np.testing.assert_array_equal(np.searchsorted(cb[1:], np.arange(lm.n_cells), side="right"), _first_leaf_owner(lm, mesh, act))


@pytest.mark.parametrize("first_child", [False, True])
@pytest.mark.parametrize("n_ranks", [2, 3])
@pytest.mark.parametrize("geometry,L,p", [("quadrant", 4, 2), ("annulus", 6, 1)])
def test_partitioned_ls_transfer_maps_are_consistent(geometry, L, p, n_ranks, first_child):
    """MGTransferMatrixFree level transfers on several ranks, level cells split evenly or living with their first child"""
    mesh = mg.Mesh(geometry, L)
    act_begin = mg.partition_sequence("DefaultPolicy", [mesh], n_ranks)[0]
    lms = [mesh.level_mesh(l) for l in range(mesh.n_global_levels)]
    cbs = [mg.first_child_boundaries(lm, mesh, act_begin) if first_child else None for lm in lms]
    for l in range(1, len(lms)):
        _check_partitioned_transfer(lms[l], p, lms[l - 1], p, cbs[l], cbs[l - 1], n_ranks, ls=True)


def _boxes(mesh):
    c = mesh.cells().astype(np.int64)
    lmax = int(c[:, 0].max())
    lo = c[:, 1:] << (lmax - c[:, 0])[:, None]
    return lo, lo + (1 << (lmax - c[:, 0]))[:, None], lmax


def _brute_statistics(meshes, cbs, n_ranks):
    """independent restatement of include/mg_tools.h:9-36, 117-191, 214-233, 267-470 with O(n^2) neighbour tests"""
    owners = [np.searchsorted(np.asarray(cb)[1:], np.arange(m.n_cells), side="right") for m, cb in zip(meshes, cbs)]
    counts = np.array([np.bincount(o, minlength=n_ranks) for o in owners])
    out = {"workload_path_max": counts.max(axis=1).sum(), "workload_eff": counts.sum() / n_ranks / counts.max(axis=1).sum()}
    v_local = v_remote = 0
    for k in range(len(meshes) - 1):
        cc, fc = meshes[k].cells().astype(np.int64), meshes[k + 1].cells().astype(np.int64)
        parent = {tuple(c.tolist()): i for i, c in enumerate(cc)}
        for j, c in enumerate(fc):
            i = parent.get((c[0] - 1, c[1] >> 1, c[2] >> 1, c[3] >> 1)) if c[0] > 0 else None
            if i is not None:
                if owners[k][i] == owners[k + 1][j]:
                    v_local += 1
                else:
                    v_remote += 1
    out["vertical_eff"] = v_local / (v_local + v_remote) if v_local + v_remote else 1.0
    h_local = h_remote = 0
    for m, o in zip(meshes, owners):
        lo, hi, _ = _boxes(m)
        touch = np.all((lo[:, None, :] <= hi[None, :, :]) & (lo[None, :, :] <= hi[:, None, :]), axis=2)
        h_local += m.n_cells
        for r in range(n_ranks):
            mine = o == r
            h_remote += int(np.count_nonzero(touch[mine].any(axis=0) & ~mine)) if mine.any() else 0
    out["horizontal_eff"] = (h_local + h_remote // 2) / (h_local + h_remote)
    out["h_remote"], out["v_remote"] = h_remote, v_remote
    return out


@pytest.mark.parametrize("policy", ["DefaultPolicy", "FirstChildPolicy", "CellWeightPolicy-2.0"])
@pytest.mark.parametrize("n_ranks", [1, 3, 8])
@pytest.mark.parametrize("geometry,L", [("quadrant", 3), ("circle", 4)])
def test_hierarchy_statistics_global_coarsening(geometry, L, n_ranks, policy):
    """MGTools::print_multigrid_statistics on a coarsening sequence against a brute-force restatement"""
    seq = mesh_sequence(geometry, L)
    cbs = mg.partition_sequence(policy, seq, n_ranks)
    st, ref_ = mg.hierarchy_statistics(seq, cbs, n_ranks), _brute_statistics(seq, cbs, n_ranks)
    for k in ("workload_eff", "workload_path_max", "vertical_eff", "horizontal_eff"):
        assert st[k] == pytest.approx(ref_[k], rel=1e-14), k
    assert sum(l["horizontal_remote"] for l in st["levels"]) == ref_["h_remote"]
    assert sum(l["vertical_remote"] for l in st["levels"]) == ref_["v_remote"]
    assert [l["n_cells_sum"] for l in st["levels"]] == [m.n_cells for m in seq]
    if n_ranks == 1:
        assert st["workload_eff"] == st["vertical_eff"] == st["horizontal_eff"] == 1.0
    if policy == "FirstChildPolicy" and n_ranks > 1:
        # a coarse cell lives with its first child: at least one child of every refined cell is local
        assert st["vertical_eff"] >= 1.0 / 8.0


@pytest.mark.parametrize("n_ranks", [2, 5])
def test_hierarchy_statistics_local_smoothing(n_ranks):
    """the same for the level meshes of local smoothing with deal.II's level ownership"""
    mesh = mg.Mesh("quadrant", 4)
    act = mg.partition_sequence("DefaultPolicy", [mesh], n_ranks)[0]
    lms = [mesh.level_mesh(l) for l in range(mesh.n_global_levels)]
    cbs = [mg.first_child_boundaries(lm, mesh, act) for lm in lms]
    st, ref_ = mg.hierarchy_statistics(lms, cbs, n_ranks), _brute_statistics(lms, cbs, n_ranks)
    for k in ("workload_eff", "workload_path_max", "vertical_eff", "horizontal_eff"):
        assert st[k] == pytest.approx(ref_[k], rel=1e-14), k
    assert st["workload_eff"] < 1.0   # local smoothing: the refined region's owners carry the fine levels alone


def test_dofmap_does_not_depend_on_the_number_of_setup_threads():
    """DoFMap::build analyses the cells on worker threads (MGB_SETUP_THREADS; csrc/host/dofmap.cc) and sorts what they append,
    TransferMap::build counts the valences with atomic increments: the maps are bitwise the same for 1, 3 and 7 threads, on meshes above the threading threshold, with hanging nodes and
    on several ranks (thread count is fixed per process, hence subprocesses)"""
    import hashlib
    import os
    import subprocess
    import sys
    script = r'''
import hashlib, sys
import numpy as np
import dealii_multigrid_b200 as mg
h = hashlib.sha256()
for geometry, L, p, n_ranks, rank in (("annulus", 6, 2, 1, 0), ("quadrant", 5, 3, 3, 1), ("hypercube", 4, 4, 2, 1), ("quadrant", 5, 1, 4, 0)):
    d = mg.DoFMap(mg.Mesh(geometry, L), p, n_ranks, rank)
    assert d.n_cells * n_ranks >= 4096 or geometry == "hypercube"
    part = d.partitioner()
    for a in (d.eval_table(), d.own_table(), d.mask(), d.cell_level(), d.cell_xyz(), d.constrained(), d.ghost_global()):
        h.update(np.ascontiguousarray(a).tobytes())
    for k in sorted(part):
        h.update(np.ascontiguousarray(part[k]).tobytes())
mf = mg.Mesh("hypercube", 5)
mc = mf.coarsen_global()
df, dc = mg.DoFMap(mf, 2), mg.DoFMap(mc, 2)
assert dc.n_cells >= 4096
tm = mg.TransferMap(mf, df, mc, dc)
for a in (tm.fine_first(), tm.n_children(), tm.weight_code(), tm.fine_own()):
    h.update(np.ascontiguousarray(a).tobytes())
print(h.hexdigest())
'''
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    digests = set()
    for threads in ("1", "3", "7"):
        env = dict(os.environ, MGB_SETUP_THREADS=threads, PYTHONPATH=root)
        out = subprocess.run([sys.executable, "-c", script], env=env, capture_output=True, text=True, timeout=600)
        assert out.returncode == 0, out.stderr[-2000:]
        digests.add(out.stdout.strip().splitlines()[-1])
    assert len(digests) == 1
