"""The hierarchy wired by the HOST PROGRAM from standalone handles, as mg_solve does with deal.II objects
(multigrid_throughput.cc:1082-1133): level operators, Chebyshev smoothers, two-level transfers, coarse CG, Multigrid with
set_edge_in_matrix, PreconditionMG (copy_to_mg / copy_from_mg: plain top-level copy for global coarsening :1531-1533,
copy_indices for local smoothing :1800-1817), SolverCG with ReductionControl -- every piece through the C ABI
(mgb_multigrid_*, mgb_level_copy_*, mgb_transfer_prolongate, mgb_cg_solve), against the oracle."""
import numpy as np
import pytest

from oracle import reference as ref
from util import OracleHierarchy, OracleHierarchyLS, rel_l2, seeded

pytestmark = pytest.mark.gpu


def _vec(op, values=None):
    v = op.initialize_dof_vector()
    if values is not None:
        v.from_numpy(values)
    return v


@pytest.mark.parametrize("geometry,L,p,mg_type,coarse", [("quadrant", 4, 2, "HMG-global", "cg"), ("annulus", 5, 4, "HPMG", "cg_with_chebyshev"),
                                                          ("hypercube", 3, 3, "PMG", "cg")])
def test_global_coarsening_composed_from_handles(gpu, geometry, L, p, mg_type, coarse):
    mg = gpu
    H = OracleHierarchy(geometry, L, p, mg_type, 3, coarse)
    ops = [mg.Operator(d) for d in H.dofmaps]
    smoothers = [None] + [mg.PreconditionChebyshev(op, 3, 20.0, 20) for op in ops[1:]]
    transfers = [None] + [mg.MGTwoLevelTransfer(H.dofmaps[l], H.dofmaps[l - 1], H.tmaps[l], ops[l], ops[l - 1]) for l in range(1, len(ops))]
    cp = mg.PreconditionChebyshev(ops[0], 3, 20.0, 20) if coarse == "cg_with_chebyshev" else None
    M = mg.Multigrid(ops, smoothers, transfers, coarse_precond=cp)
    n = H.fine_maps.n_dofs
    r = seeded(n)
    r[H.fine_maps.constrained] = 0.0
    src, dst = _vec(ops[-1], r), _vec(ops[-1])
    M.vmult(dst, src)
    assert rel_l2(dst.to_numpy(), H.mg.vmult(r)) < 1e-11
    # MGTwoLevelTransfer::prolongate (no accumulation) next to prolongate_and_add
    uc = seeded(H.maps[-2].n_dofs, 3)
    vc, vf = _vec(ops[-2], uc), _vec(ops[-1], seeded(n, 4))
    transfers[-1].prolongate(vf, vc)
    e = np.zeros(n)
    H.transfers[-1].prolongate_and_add(e, uc)
    assert rel_l2(vf.to_numpy(), e) < 1e-12
    # SolverCG + ReductionControl, preconditioned and not
    b = ref.assemble_rhs_constant(H.fine_maps)
    x, vb = _vec(ops[-1]), _vec(ops[-1], b)
    its, res = mg.cg_solve(ops[-1], x, vb, M, reltol=1e-4)
    xo, its_o, ctl = H.solve(b)
    assert its == its_o and rel_l2(x.to_numpy(), xo) < 1e-9 and abs(res - ctl.last_value) < 1e-8 * ctl.last_value
    x2 = np.zeros(n)
    ctl2 = ref.ReductionControl(10000, 1e-20, 1e-2)
    its_plain_o = ref.solver_cg(H.fine_op.vmult, x2, b, None, ctl2)
    its_plain, _ = mg.cg_solve(ops[-1], x, vb, None, reltol=1e-2)
    assert its_plain == its_plain_o
    with pytest.raises(mg.MgbError):
        mg.cg_solve(ops[-1], x, vb, None, maxiter=2, reltol=1e-10)    # SolverControl::NoConvergence
    # the seven stage timers (multigrid_throughput.cc:1163-1234)
    M.enable_timers(True)
    M.vmult(dst, src)
    t = M.collect_times()
    assert t.shape == (len(ops), 7) and np.all(t >= 0) and t[1:, 0].min() > 0 and t[0, 3] > 0   # pre-smoothing on the levels, coarse solve on 0
    M.enable_timers(False)


@pytest.mark.parametrize("geometry,L,p", [("quadrant", 4, 2), ("annulus", 5, 1), ("hypercube", 3, 2)])
def test_local_smoothing_composed_from_handles(gpu, geometry, L, p):
    mg = gpu
    H = OracleHierarchyLS(geometry, L, p, "HMG-local", 3, "cg")
    nl = len(H.level_meshes)
    ops = [mg.Operator(d, edges=e) for d, e in zip(H.level_dofmaps, H.level_edges)]
    smoothers = [None] + [mg.PreconditionChebyshev(op, 3, 20.0, 20) for op in ops[1:]]
    transfers = [None] + [mg.MGTwoLevelTransfer(H.level_dofmaps[l], H.level_dofmaps[l - 1], H.ls_tmaps[l], ops[l], ops[l - 1]) for l in range(1, nl)]
    copies = [mg.LevelCopy(c, op) for c, op in zip(H.copy, ops)]
    active_op = mg.Operator(H.active_dofmaps[-1])
    n = H.fine_maps.n_dofs
    x = seeded(n)
    # copy_to_mg / copy_from_mg level by level (MGLevelGlobalTransfer)
    va = _vec(active_op, x)
    back = _vec(active_op)
    for l, (c, op) in enumerate(zip(copies, ops)):
        vl = _vec(op, seeded(H.level_maps[l].n_dofs, 9))        # non-zero: copy_to_mg must clear what it does not set
        c.copy_to_mg(vl, va)
        e = np.zeros(H.level_maps[l].n_dofs)
        e[H.copy[l].level_idx] = x[H.copy[l].active_idx]
        np.testing.assert_array_equal(vl.to_numpy(), e)
        c.copy_from_mg(back, vl)
    covered = np.zeros(n, dtype=bool)
    for c in H.copy:
        covered[c.active_idx] = True
    np.testing.assert_array_equal(back.to_numpy()[covered], x[covered])
    assert np.all(back.to_numpy()[~covered] == 0.0)
    has_hanging = len(set(H.mesh.cells()[:, 0].tolist())) > 1
    M = mg.Multigrid(ops, smoothers, transfers, edge_in=has_hanging, level_copies=copies)
    r = x.copy()
    r[H.fine_maps.constrained] = 0.0
    src, dst = _vec(active_op, r), _vec(active_op)
    M.vmult(dst, src)
    assert rel_l2(dst.to_numpy(), H.mg.vmult(r)) < 1e-11
    b = ref.assemble_rhs_constant(H.fine_maps)
    vx, vb = _vec(active_op), _vec(active_op, b)
    its, _ = mg.cg_solve(active_op, vx, vb, M)
    xo, its_o, _ = H.solve(b)
    assert its == its_o and rel_l2(vx.to_numpy(), xo) < 1e-9


def test_transfer_and_level_copy_created_from_plain_arrays(gpu):
    """mgb_transfer_create_from_arrays / mgb_level_copy_create_from_arrays: what include/mgb200_dealii.h calls with the arrays it
    extracts from deal.II (MGTwoLevelTransfer::reinit, MGLevelGlobalTransfer::copy_indices) must behave like the objects built from
    the generator's handles"""
    mg = gpu
    H = OracleHierarchyLS("quadrant", 4, 2, "HMG-local", 3, "cg")
    l = len(H.level_meshes) - 1
    opf, opc = mg.Operator(H.level_dofmaps[l], edges=H.level_edges[l]), mg.Operator(H.level_dofmaps[l - 1], edges=H.level_edges[l - 1])
    T_handles = mg.MGTwoLevelTransfer(H.level_dofmaps[l], H.level_dofmaps[l - 1], H.ls_tmaps[l], opf, opc)
    T_arrays = mg.MGTwoLevelTransfer(None, None, H.ls_tmaps[l], opf, opc)          # the view's plain arrays only
    uc, uf = seeded(H.level_maps[l - 1].n_dofs, 1), seeded(H.level_maps[l].n_dofs, 2)
    for T in (T_handles, T_arrays):
        vc, vf = _vec(opc, uc), _vec(opf, uf)
        T.prolongate_and_add(vf, vc)
        e = uf.copy()
        H.ls_transfers[l].prolongate_and_add(e, uc)
        assert rel_l2(vf.to_numpy(), e) < 1e-12
        vc, vf = _vec(opc, uc), _vec(opf, uf)
        T.restrict_and_add(vc, vf)
        e = uc.copy()
        H.ls_transfers[l].restrict_and_add(e, uf)
        assert rel_l2(vc.to_numpy(), e) < 1e-12
    active_op = mg.Operator(H.active_dofmaps[-1])
    x = seeded(H.fine_maps.n_dofs, 3)
    va = _vec(active_op, x)
    c = mg.LevelCopy(None, opf, active_idx=H.copy[l].active_idx, level_idx=H.copy[l].level_idx)
    vl = _vec(opf, seeded(H.level_maps[l].n_dofs, 4))
    c.copy_to_mg(vl, va)
    e = np.zeros(H.level_maps[l].n_dofs)
    e[H.copy[l].level_idx] = x[H.copy[l].active_idx]
    np.testing.assert_array_equal(vl.to_numpy(), e)
    back = _vec(active_op)
    c.copy_from_mg(back, vl)
    np.testing.assert_array_equal(back.to_numpy()[H.copy[l].active_idx], x[H.copy[l].active_idx])


def test_array_based_creation_rejects_indices_outside_the_vectors(gpu):
    """fine_own, level_idx and active_idx become addresses in k_transfer / k_idx_pair_copy: checked on the host against the vectors
    they address (csrc/host/capi.cc check_entity_table / check_index_list)"""
    import ctypes as C
    mg = gpu
    H = OracleHierarchyLS("quadrant", 4, 2, "HMG-local", 3, "cg")
    l = len(H.level_meshes) - 1
    opf, opc = mg.Operator(H.level_dofmaps[l], edges=H.level_edges[l]), mg.Operator(H.level_dofmaps[l - 1], edges=H.level_edges[l - 1])
    tm = H.ls_tmaps[l]
    view = type(tm.view)()
    C.pointer(view)[0] = tm.view
    bad = tm.fine_own().copy()
    flat = bad.reshape(-1)
    flat[np.flatnonzero((flat & 0x80000000) == 0)[-1]] = H.level_maps[l].n_dofs - 1   # a vertex would fit; edges and faces do not
    flat[np.flatnonzero((flat & 0x80000000) == 0)[0]] = H.level_maps[l].n_dofs
    view.fine_own = bad.ctypes.data_as(C.POINTER(C.c_uint32))

    class BadMap:
        pass
    BadMap.view = view
    with pytest.raises(mg.MgbError, match="fine_own: cell"):
        mg.MGTwoLevelTransfer(None, None, BadMap, opf, opc)
    idx = H.copy[l].level_idx.copy()
    idx[0] = H.level_maps[l].n_dofs
    with pytest.raises(mg.MgbError, match="level_idx: entry 0"):
        mg.LevelCopy(None, opf, active_idx=H.copy[l].active_idx, level_idx=idx)
    c = mg.LevelCopy(None, opf, active_idx=H.copy[l].active_idx, level_idx=H.copy[l].level_idx)
    short = mg.Vector(int(H.copy[l].active_idx.max()))                                # one entry too short
    vl = _vec(opf)
    with pytest.raises(mg.MgbError, match="copy indices address entry"):
        c.copy_to_mg(vl, short)
    with pytest.raises(mg.MgbError, match="copy indices address entry"):
        c.copy_from_mg(short, vl)
