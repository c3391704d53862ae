"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.
Bar (north_star): index maps bit-exact (shared inputs), vmult and one V-cycle within 1e-12 relative l2 (FP64),
CG iteration counts identical."""
import numpy as np
import pytest

from oracle import reference as ref
from util import OracleHierarchy, PARITY_RTOL, mesh_sequence, params, rel_l2, seeded

pytestmark = pytest.mark.gpu

OP_CASES = [("hypercube", 2, 1), ("hypercube", 3, 2), ("hypercube", 2, 3), ("hypercube", 3, 4), ("hypercube", 1, 5), ("hypercube", 1, 6),
            ("quadrant", 4, 1), ("quadrant", 3, 2), ("quadrant", 4, 4), ("annulus", 5, 3), ("annulus", 6, 4), ("circle", 4, 4)]


@pytest.mark.parametrize("geometry,L,p", OP_CASES)
def test_vmult_and_diagonal(gpu, geometry, L, p):
    mg = gpu
    d = mg.DoFMap(mg.Mesh(geometry, L), p)
    maps = ref.LevelMaps.from_dofmap(d)
    oracle = ref.Operator(maps)
    op = mg.Operator(d)
    assert op.m() == maps.n_dofs
    src, dst = op.initialize_dof_vector(), op.initialize_dof_vector()
    x = seeded(maps.n_dofs)
    src.from_numpy(x)
    op.vmult(dst, src)
    y = dst.to_numpy()
    assert rel_l2(y, oracle.vmult(x)) < PARITY_RTOL
    np.testing.assert_array_equal(y[maps.constrained], x[maps.constrained])       # identity rows exactly
    np.testing.assert_array_equal(src.to_numpy(), x)                               # src restored
    op.Tvmult(dst, src)
    assert rel_l2(dst.to_numpy(), oracle.vmult(x)) < PARITY_RTOL
    op.vmult_interface_down(dst, src)
    assert rel_l2(dst.to_numpy(), oracle.vmult_interface_down(x)) < PARITY_RTOL
    op.vmult_interface_up(dst, src)                                                # no edge DoFs on active meshes
    assert np.all(dst.to_numpy() == 0.0)
    diag = op.initialize_dof_vector()
    op.compute_inverse_diagonal(diag)
    assert rel_l2(diag.to_numpy(), oracle.compute_inverse_diagonal()) < PARITY_RTOL
    # symmetry of the device operator itself: x.(A z) == z.(A x)
    z = seeded(maps.n_dofs, 3)
    vz, Az, Ax = op.initialize_dof_vector(), op.initialize_dof_vector(), op.initialize_dof_vector()
    vz.from_numpy(z)
    op.vmult(Az, vz)
    op.vmult(Ax, src)
    assert abs(src.dot(Az) - vz.dot(Ax)) < 1e-10 * abs(src.dot(Az))


def test_operator_from_raw_arrays_and_full_index_map(gpu):
    """the boundary takes plain index maps: entity table or the full lexicographic MatrixFree-style map"""
    mg = gpu
    d = mg.DoFMap(mg.Mesh("quadrant", 3), 3)
    maps = ref.LevelMaps.from_dofmap(d)
    x = seeded(maps.n_dofs)
    expect = ref.Operator(maps).vmult(x)
    for kw in (dict(entity_table=d.eval_table()), dict(dof_indices=d.expand("eval"))):
        op = mg.Operator(degree=3, cell_level=d.cell_level(), n_owned=d.n_owned, mask=d.mask(), constrained=d.constrained(), **kw)
        src, dst = op.initialize_dof_vector(), op.initialize_dof_vector()
        src.from_numpy(x)
        op.vmult(dst, src)
        assert rel_l2(dst.to_numpy(), expect) < PARITY_RTOL
    bad = d.expand("eval").copy()
    bad[0, 5], bad[0, 6] = bad[0, 6], bad[0, 5]
    with pytest.raises(mg.MgbError, match="not contiguous"):
        mg.Operator(degree=3, cell_level=d.cell_level(), n_owned=d.n_owned, dof_indices=bad)


def test_vector_blas1(gpu):
    mg = gpu
    n = 100003
    a, b = seeded(n, 1), seeded(n, 2)
    x, y = mg.Vector(n), mg.Vector(n)
    x.from_numpy(a)
    y.from_numpy(b)
    assert abs(x.dot(y) - a @ b) < 1e-12 * np.sqrt(n)
    assert abs(x.l2_norm() - np.linalg.norm(a)) < 1e-12 * np.linalg.norm(a)
    y.add(0.5, x)
    np.testing.assert_allclose(y.to_numpy(), b + 0.5 * a, rtol=1e-15)
    y.sadd(2.0, -1.0, x)
    np.testing.assert_allclose(y.to_numpy(), 2 * (b + 0.5 * a) - a, rtol=1e-14, atol=1e-15)
    y.equ(3.0, x)
    np.testing.assert_allclose(y.to_numpy(), 3 * a, rtol=1e-15)
    g = (3 * a) + 0.25 * a
    assert abs(y.add_and_dot(0.25, x) - g @ g) < 1e-11 * (g @ g)
    # deterministic reductions: bitwise identical on repetition
    assert x.dot(x) == x.dot(x)
    e = mg.Vector(0)
    assert e.dot(e) == 0.0                                                         # empty vector


@pytest.mark.parametrize("geometry,L,p,degree", [("hypercube", 3, 2, 3), ("quadrant", 4, 4, 3), ("annulus", 5, 1, 5), ("hypercube", 2, 4, 1)])
def test_chebyshev(gpu, geometry, L, p, degree):
    mg = gpu
    d = mg.DoFMap(mg.Mesh(geometry, L), p)
    maps = ref.LevelMaps.from_dofmap(d)
    oracle = ref.PreconditionChebyshev(ref.Operator(maps), degree, 20.0, 20)
    op = mg.Operator(d)
    cheb = mg.PreconditionChebyshev(op, degree, 20.0, 20)
    lmin, lmax = cheb.eigenvalues()
    oracle.estimate_eigenvalues()
    assert abs(lmax - oracle.lmax) < 1e-10 * oracle.lmax
    b = seeded(maps.n_dofs)
    rhs, x = op.initialize_dof_vector(), op.initialize_dof_vector()
    rhs.from_numpy(b)
    cheb.vmult(x, rhs)
    assert rel_l2(x.to_numpy(), oracle.vmult(b)) < 1e-11
    x0 = seeded(maps.n_dofs, 5)
    x.from_numpy(x0)
    cheb.step(x, rhs)
    assert rel_l2(x.to_numpy(), oracle.step(x0.copy(), b)) < 1e-11


@pytest.mark.parametrize("geometry,L,p,degree,number", [("hypercube", 3, 4, 3, "double"), ("quadrant", 4, 3, 4, "double"),
                                                       ("quadrant", 3, 4, 3, "float"), ("hypercube", 3, 1, 5, "double")])
def test_chebyshev_zeroing_inside_the_cell_loop(gpu, monkeypatch, geometry, L, p, degree, number):
    """the smoother keeps A x alternately in two temporaries and lets the cell loop that fills one clear the other
    (k_laplace_apply<P,T,true>; split plain / hanging launches on the quadrant); MGB_ZERO_IN_CELL_LOOP=0 is the memset before
    every cell loop.  Both against the oracle, several calls in a row (the temporaries carry old products), odd vector
    lengths included (the 16-byte zeroing has a scalar tail)"""
    mg = gpu
    d = mg.DoFMap(mg.Mesh(geometry, L), p)
    maps = ref.LevelMaps.from_dofmap(d)
    oracle = ref.PreconditionChebyshev(ref.Operator(maps), degree, 20.0, 20)
    oracle.estimate_eigenvalues()
    ntype, dt, tol = (mg.DOUBLE, np.float64, 1e-11) if number == "double" else (mg.FLOAT, np.float32, 1e-4)
    b, x0 = seeded(maps.n_dofs), seeded(maps.n_dofs, 5)
    want_vmult, want_step = oracle.vmult(b), oracle.step(x0.copy(), b)
    for env in ("1", "0"):
        monkeypatch.setenv("MGB_ZERO_IN_CELL_LOOP", env)
        op = mg.Operator(d, ntype)
        cheb = mg.PreconditionChebyshev(op, degree, 20.0, 20)
        rhs, x = op.initialize_dof_vector(), op.initialize_dof_vector()
        rhs.from_numpy(b.astype(dt))
        for _ in range(3):
            cheb.vmult(x, rhs)
            assert rel_l2(x.to_numpy().astype(np.float64), want_vmult) < tol
            x.from_numpy(x0.astype(dt))
            cheb.step(x, rhs)
            assert rel_l2(x.to_numpy().astype(np.float64), want_step) < tol


TRANSFER_CASES = [("hypercube", 2, 1, "h"), ("hypercube", 3, 4, "h"), ("quadrant", 4, 2, "h"), ("annulus", 5, 3, "h"), ("circle", 4, 4, "h"),
                  ("quadrant", 3, 4, "p"), ("annulus", 5, 2, "p"), ("hypercube", 2, 3, "p"), ("quadrant", 3, 6, "p")]


@pytest.mark.parametrize("geometry,L,p,kind", TRANSFER_CASES)
def test_two_level_transfer(gpu, geometry, L, p, kind):
    mg = gpu
    seq = mesh_sequence(geometry, L)
    (mf, pf), (mc, pc) = ((seq[-1], p), (seq[-2], p)) if kind == "h" else ((seq[-1], p), (seq[-1], max(p // 2, 1)))
    df, dc = mg.DoFMap(mf, pf), mg.DoFMap(mc, pc)
    tm = mg.TransferMap(mf, df, mc, dc)
    fine, coarse = ref.LevelMaps.from_dofmap(df), ref.LevelMaps.from_dofmap(dc)
    oracle = ref.MGTwoLevelTransfer(fine, coarse, tm.fine_first(), tm.n_children())
    opf, opc = mg.Operator(df), mg.Operator(dc)
    T = mg.MGTwoLevelTransfer(df, dc, tm, opf, opc)
    uc, uf = seeded(coarse.n_dofs, 1), seeded(fine.n_dofs, 2)
    vc, vf = opc.initialize_dof_vector(), opf.initialize_dof_vector()
    # prolongate_and_add adds into a non-zero destination
    vc.from_numpy(uc)
    vf.from_numpy(uf)
    T.prolongate_and_add(vf, vc)
    expect = uf.copy()
    oracle.prolongate_and_add(expect, uc)
    assert rel_l2(vf.to_numpy(), expect) < PARITY_RTOL
    vc.from_numpy(uc)
    vf.from_numpy(uf)
    T.restrict_and_add(vc, vf)
    expect = uc.copy()
    oracle.restrict_and_add(expect, uf)
    assert rel_l2(vc.to_numpy(), expect) < PARITY_RTOL


SOLVE_CASES = [("hypercube", 3, 1, "HMG-global", "cg"), ("hypercube", 2, 4, "HMG-global", "cg"), ("hypercube", 3, 2, "HMG-global", "cg_with_chebyshev"),
               ("quadrant", 3, 2, "HMG-global", "cg"), ("quadrant", 3, 4, "HPMG", "cg"), ("annulus", 5, 2, "PMG", "cg"),
               ("annulus", 5, 4, "HPMG", "cg_with_chebyshev"), ("circle", 4, 1, "HMG-global", "cg")]


@pytest.mark.parametrize("geometry,L,p,mg_type,coarse", SOLVE_CASES)
def test_vcycle_and_cg_iteration_counts(gpu, geometry, L, p, mg_type, coarse):
    mg = gpu
    prob = mg.Problem(params(geometry, L, p, mg_type, coarse))
    H = OracleHierarchy(geometry, L, p, mg_type, 3, coarse)
    assert prob.n_levels == len(H.levels)
    for l in range(prob.n_levels):
        assert prob.level_n_dofs(l) == H.maps[l].n_dofs and prob.level_degree(l) == H.levels[l][1]
    n = H.maps[-1].n_dofs
    # one V-cycle (PreconditionMG::vmult) on a seeded defect
    r = seeded(n)
    r[H.maps[-1].constrained] = 0.0
    src, dst = mg.Vector(n), mg.Vector(n)
    src.from_numpy(r)
    prob.vcycle(dst, src)
    assert rel_l2(dst.to_numpy(), H.mg.vmult(r)) < 1e-11
    # right-hand side of the problem and the full solve
    b = prob.rhs().to_numpy()
    x_ref, its_ref, ctl = H.solve(b.copy())
    res = prob.solve_once()
    assert res["n_iterations"] == its_ref
    assert rel_l2(prob.solution().to_numpy(), x_ref) < 1e-9
    assert abs(res["final_residual"] - ctl.last_value) < 1e-8 * ctl.initial_value
    assert res["n_dofs"] == n and res["n_levels"] == len(H.levels)


def test_rhs_matches_oracle_integration(gpu):
    """Operator::rhs for f = 1 (include/operator.h:362-413)"""
    mg = gpu
    prob = mg.Problem(params("quadrant", 3, 3))
    maps = ref.LevelMaps.from_dofmap(prob.level_dofmap(prob.n_levels - 1))
    m1 = maps.fe.M.sum(axis=1)
    loc = np.einsum("k,j,i->kji", m1, m1, m1).ravel()
    rhs = np.zeros(maps.n_dofs)
    for m in np.unique(maps.mask):
        c = np.nonzero(maps.mask == m)[0]
        r = np.tile(loc, (len(c), 1)) * (maps.h[c] ** 3)[:, None]
        if m >> 3:
            r = r @ maps.C(m)
        np.add.at(rhs, maps.eval_idx[c][~maps.eval_flag[c]], r[~maps.eval_flag[c]])
    assert rel_l2(prob.rhs().to_numpy(), rhs) < 1e-13


def test_float_level_number_type(gpu):
    """MGNumberType=float (the reference default): float V-cycle inside the double CG still converges"""
    mg = gpu
    res_d = mg.Problem(params("hypercube", 3, 2, number="double")).solve_once()
    res_f = mg.Problem(params("hypercube", 3, 2, number="float")).solve_once()
    assert abs(res_f["n_iterations"] - res_d["n_iterations"]) <= 1
    assert res_f["final_residual"] <= 1e-4 * res_f["initial_residual"]


def test_unsupported_inputs_raise(gpu):
    mg = gpu
    with pytest.raises(mg.MgbError, match="out of scope"):
        mg.Problem(params(coarse="amg"))
    with pytest.raises(mg.MgbError, match="unknown GeometryType"):
        mg.Problem(params(geometry="torus"))
    with pytest.raises(mg.MgbError, match="Type"):
        mg.Problem(params(mg_type="AMG"))


def test_cuda_graph_replay_matches_eager(gpu):
    """on a non-default stream the V-cycle is captured into a CUDA graph after the first solve; replays must reproduce the
    eager launches (same kernels, same order) and the single-kernel coarse CG must keep the oracle's iteration count"""
    import torch
    mg = gpu
    stream = torch.cuda.Stream()
    try:
        with torch.cuda.stream(stream):
            mg.set_stream(stream.cuda_stream)
            prob = mg.Problem(params("quadrant", 4, 3, "HMG-global", "cg"))
            H = OracleHierarchy("quadrant", 4, 3, "HMG-global", 3, "cg")
            _, its_ref, _ = H.solve(prob.rhs().to_numpy())
            r1 = prob.solve_once()                      # eager (lazy initialisation)
            x1 = prob.solution().to_numpy()
            n0 = mg.launch_count()
            r2 = prob.solve_once()                      # captures, then replays
            r3 = prob.solve_once()
            x3 = prob.solution().to_numpy()
            assert r1["n_iterations"] == r2["n_iterations"] == r3["n_iterations"] == its_ref
            assert rel_l2(x3, x1) < 1e-12
            assert mg.launch_count() - n0 > 100         # replayed launches are counted
            n = x1.size
            v = seeded(n)
            v[H.maps[-1].constrained] = 0.0
            src, dst = prob.fine_operator().initialize_dof_vector(), prob.fine_operator().initialize_dof_vector()
            src.from_numpy(v)
            prob.vcycle(dst, src)
            prob.vcycle(dst, src)
            assert rel_l2(dst.to_numpy(), H.mg.vmult(v)) < 1e-11
            mg.synchronize()
    finally:
        mg.set_stream(0)


# ----------------------------------------------------------------------------------------------- local smoothing
LS_OP_CASES = [("quadrant", 3, 2), ("quadrant", 4, 1), ("quadrant", 4, 4), ("annulus", 6, 2), ("circle", 4, 3)]


@pytest.mark.parametrize("geometry,L,p", LS_OP_CASES)
def test_level_operator_edge_handling(gpu, geometry, L, p):
    """level operators of the LS hierarchy: vmult with refinement-edge DoFs as Dirichlet, vmult_interface_down/up,
    inverse diagonal zeroed at the edge (include/operator.h:152-242)"""
    mg = gpu
    mesh = mg.Mesh(geometry, L)
    seen_edges = False
    for l in range(1, mesh.n_global_levels):
        lm = mesh.level_mesh(l)
        d = mg.DoFMap(lm, p)
        e = mg.LevelEdges(lm, d)
        maps = ref.LevelMaps.from_dofmap(d)
        oracle = ref.Operator(maps, e.edge_dofs, e.edge_cells, e.any_global)
        op = mg.Operator(d, edges=e)
        seen_edges |= e.any_global
        x = seeded(maps.n_dofs, 7 + l)
        src, dst = op.initialize_dof_vector(), op.initialize_dof_vector()
        src.from_numpy(x)
        dst.from_numpy(seeded(maps.n_dofs, 1))
        op.vmult(dst, src)
        y = dst.to_numpy()
        assert rel_l2(y, oracle.vmult(x)) < PARITY_RTOL
        np.testing.assert_array_equal(y[e.edge_dofs], x[e.edge_dofs])               # dst_edge = src_edge exactly
        np.testing.assert_array_equal(src.to_numpy(), x)                             # src restored
        op.vmult_interface_down(dst, src)
        assert rel_l2(dst.to_numpy(), oracle.vmult_interface_down(x)) < PARITY_RTOL
        op.vmult_interface_up(dst, src)
        up = oracle.vmult_interface_up(x)
        if e.any_global:
            assert rel_l2(dst.to_numpy(), up) < PARITY_RTOL
        else:
            assert np.all(dst.to_numpy() == 0.0)
        diag = op.initialize_dof_vector()
        op.compute_inverse_diagonal(diag)
        dd = diag.to_numpy()
        assert rel_l2(dd, oracle.compute_inverse_diagonal()) < PARITY_RTOL
        assert np.all(dd[e.edge_dofs] == 1.0)
    assert seen_edges


@pytest.mark.parametrize("geometry,L,p", [("quadrant", 3, 2), ("quadrant", 4, 4), ("annulus", 6, 1), ("circle", 4, 3), ("hypercube", 2, 2)])
def test_local_smoothing_transfer(gpu, geometry, L, p):
    """MGTransferMatrixFree::prolongate_and_add / restrict_and_add between consecutive LS levels (refined parents only)"""
    mg = gpu
    mesh = mg.Mesh(geometry, L)
    for l in range(1, mesh.n_global_levels):
        mf, mc = mesh.level_mesh(l), mesh.level_mesh(l - 1)
        df, dc = mg.DoFMap(mf, p), mg.DoFMap(mc, p)
        tm = mg.TransferMap(mf, df, mc, dc, local_smoothing=True)
        assert set(tm.n_children().tolist()) <= {0, 8}
        fine, coarse = ref.LevelMaps.from_dofmap(df), ref.LevelMaps.from_dofmap(dc)
        oracle = ref.MGTwoLevelTransfer(fine, coarse, tm.fine_first(), tm.n_children())
        opf, opc = mg.Operator(df), mg.Operator(dc)
        T = mg.MGTwoLevelTransfer(df, dc, tm, opf, opc)
        uc, uf = seeded(coarse.n_dofs, 1), seeded(fine.n_dofs, 2)
        vc, vf = opc.initialize_dof_vector(), opf.initialize_dof_vector()
        vc.from_numpy(uc)
        vf.from_numpy(uf)
        T.prolongate_and_add(vf, vc)
        expect = uf.copy()
        oracle.prolongate_and_add(expect, uc)
        assert rel_l2(vf.to_numpy(), expect) < PARITY_RTOL
        vc.from_numpy(uc)
        vf.from_numpy(uf)
        T.restrict_and_add(vc, vf)
        expect = uc.copy()
        oracle.restrict_and_add(expect, uf)
        assert rel_l2(vc.to_numpy(), expect) < PARITY_RTOL


LS_SOLVE_CASES = [("quadrant", 3, 2, "HMG-local", "cg"), ("quadrant", 4, 1, "HMG-local", "cg"), ("quadrant", 3, 4, "HMG-local", "cg_with_chebyshev"),
                  ("hypercube", 3, 2, "HMG-local", "cg"), ("annulus", 6, 2, "HMG-local", "cg"), ("quadrant", 3, 4, "HPMG-local", "cg"),
                  ("circle", 4, 2, "HPMG-local", "cg")]


@pytest.mark.parametrize("geometry,L,p,mg_type,coarse", LS_SOLVE_CASES)
def test_local_smoothing_vcycle_and_cg(gpu, geometry, L, p, mg_type, coarse):
    """solve_with_local_smoothing (multigrid_throughput.cc:1673-1873): one V-cycle with copy_to_mg / copy_from_mg and the
    edge_in matrix, then the full CG solve with identical iteration counts"""
    from util import OracleHierarchyLS
    mg = gpu
    prob = mg.Problem(params(geometry, L, p, mg_type, coarse))
    H = OracleHierarchyLS(geometry, L, p, mg_type, 3, coarse)
    assert prob.n_levels == len(H.maps)
    for l in range(prob.n_levels):
        assert prob.level_n_dofs(l) == H.maps[l].n_dofs
    n = H.fine_maps.n_dofs
    assert prob.fine_dofmap().n_dofs_global == n
    r = seeded(n)
    r[H.fine_maps.constrained] = 0.0
    src, dst = mg.Vector(n), mg.Vector(n)
    src.from_numpy(r)
    prob.vcycle(dst, src)
    assert rel_l2(dst.to_numpy(), H.mg.vmult(r)) < 1e-11
    b = prob.rhs().to_numpy()
    x_ref, its_ref, ctl = H.solve(b.copy())
    res = prob.solve_once()
    assert res["n_iterations"] == its_ref
    assert rel_l2(prob.solution().to_numpy(), x_ref) < 1e-9
    assert res["n_dofs"] == n
    res = prob.solve_once(True)                                                  # with stage timers (edge_pro column)
    assert res["n_iterations"] == its_ref


@pytest.mark.parametrize("mg_type", ["HMG-local", "HPMG-local"])
def test_cuda_graph_replay_local_smoothing(gpu, mg_type):
    """the LS V-cycle (copy_to_mg, edge_in, copy_from_mg; for HPMG-local nested as the coarse solver) replayed from a CUDA graph"""
    import torch
    from util import OracleHierarchyLS
    mg = gpu
    stream = torch.cuda.Stream()
    try:
        with torch.cuda.stream(stream):
            mg.set_stream(stream.cuda_stream)
            prob = mg.Problem(params("quadrant", 4, 2, mg_type, "cg"))
            H = OracleHierarchyLS("quadrant", 4, 2, mg_type, 3, "cg")
            _, its_ref, _ = H.solve(prob.rhs().to_numpy())
            r1 = prob.solve_once()
            x1 = prob.solution().to_numpy()
            r2 = prob.solve_once()
            r3 = prob.solve_once()
            assert r1["n_iterations"] == r2["n_iterations"] == r3["n_iterations"] == its_ref
            assert rel_l2(prob.solution().to_numpy(), x1) < 1e-12
            v = seeded(x1.size)
            v[H.fine_maps.constrained] = 0.0
            src, dst = prob.fine_operator().initialize_dof_vector(), prob.fine_operator().initialize_dof_vector()
            src.from_numpy(v)
            prob.vcycle(dst, src)
            prob.vcycle(dst, src)
            assert rel_l2(dst.to_numpy(), H.mg.vmult(v)) < 1e-11
            mg.synchronize()
    finally:
        mg.set_stream(0)


# ----------------------------------------------------------------------------------------------- general geometry (MappingQ1)
Q1_CASES = [("quadrant", 3, 1, "affine", "double"), ("quadrant", 3, 2, "affine", "double"), ("quadrant", 3, 4, "affine", "double"),
            ("annulus", 5, 3, "affine", "double"), ("hypercube", 3, 2, "smooth", "double"), ("hypercube", 2, 4, "smooth", "double"),
            ("hypercube", 1, 5, "smooth", "double"), ("hypercube", 1, 6, "smooth", "double"), ("hypercube", 2, 3, "cube", "double"),
            ("quadrant", 3, 2, "affine", "float")]


@pytest.mark.parametrize("geometry,L,p,deformation,number", Q1_CASES)
def test_general_geometry_operator(gpu, geometry, L, p, deformation, number):
    """non-Cartesian cells (8 vertices per cell, trilinear map): vmult, interface operator, inverse diagonal and the
    Chebyshev smoother built on them against the assembled-matrix oracle; undeformed vertices reproduce the Cartesian kernel"""
    from util import affine_map, smooth_map
    mg = gpu
    d = mg.DoFMap(mg.Mesh(geometry, L), p)
    maps = ref.LevelMaps.from_dofmap(d)
    verts = ref.cell_vertices(maps, {"affine": affine_map, "smooth": smooth_map, "cube": None}[deformation])
    oracle = ref.MappedOperator(maps, verts)
    ntype, tol = (mg.DOUBLE, PARITY_RTOL) if number == "double" else (mg.FLOAT, 2e-5)
    op = mg.Operator(d, ntype)
    op.set_cell_vertices(verts)
    with pytest.raises(RuntimeError):
        op.set_cell_vertices(verts[:-1])                                           # wrong cell count is refused
    src, dst = op.initialize_dof_vector(), op.initialize_dof_vector()
    x = seeded(maps.n_dofs)
    src.from_numpy(x)
    op.vmult(dst, src)
    y = dst.to_numpy()
    assert rel_l2(y, oracle.vmult(x)) < tol
    np.testing.assert_array_equal(y[maps.constrained], src.to_numpy()[maps.constrained])
    op.vmult_interface_down(dst, src)
    assert rel_l2(dst.to_numpy(), oracle.vmult(x)) < tol
    diag = op.initialize_dof_vector()
    op.compute_inverse_diagonal(diag)
    assert rel_l2(diag.to_numpy(), oracle.compute_inverse_diagonal()) < tol
    if deformation == "cube":                                                      # same operator as the Cartesian kernel
        plain = mg.Operator(d, ntype)
        plain.vmult(dst, src)
        assert rel_l2(dst.to_numpy(), y) < tol
    if number == "double":
        ocheb = ref.PreconditionChebyshev(oracle, 3, 20.0, 20)
        cheb = mg.PreconditionChebyshev(op, 3, 20.0, 20)
        ocheb.estimate_eigenvalues()
        assert abs(cheb.eigenvalues()[1] - ocheb.lmax) < 1e-9 * ocheb.lmax
        b = seeded(maps.n_dofs, 7)
        rhs = op.initialize_dof_vector()
        rhs.from_numpy(b)
        cheb.vmult(dst, rhs)
        assert rel_l2(dst.to_numpy(), ocheb.vmult(b)) < 1e-10


# ------------------------------------------------------------------------------------ MGNumberType = float (the reference's default)
F32 = np.float32


@pytest.mark.parametrize("geometry,L,p", [("hypercube", 3, 4), ("quadrant", 4, 2), ("annulus", 5, 3), ("quadrant", 3, 1)])
def test_float_kernels_against_the_float_oracle(gpu, geometry, L, p):
    """Cartesian vmult, inverse diagonal, Chebyshev and the two-level transfers of the float instantiation against the
    oracle with float32-rounded level vectors (oracle/reference.py round_f32): 1e-5 relative"""
    mg = gpu
    seq = mesh_sequence(geometry, L)
    df, dc = mg.DoFMap(seq[-1], p), mg.DoFMap(seq[-2], p)
    fine, coarse = ref.LevelMaps.from_dofmap(df), ref.LevelMaps.from_dofmap(dc)
    oracle = ref.Operator(fine, rnd=ref.round_f32)
    op, opc = mg.Operator(df, number_type=mg.FLOAT), mg.Operator(dc, number_type=mg.FLOAT)
    x = ref.round_f32(seeded(fine.n_dofs))
    src, dst = op.initialize_dof_vector(), op.initialize_dof_vector()
    src.from_numpy(x.astype(F32))
    op.vmult(dst, src)
    assert rel_l2(dst.to_numpy().astype(np.float64), oracle.vmult(x)) < 1e-5
    diag = op.initialize_dof_vector()
    op.compute_inverse_diagonal(diag)
    assert rel_l2(diag.to_numpy().astype(np.float64), oracle.compute_inverse_diagonal()) < 1e-5
    # Chebyshev: eigenvalue estimate and both entry points
    cheb_o = ref.PreconditionChebyshev(oracle, 3, 20.0, 20)
    cheb = mg.PreconditionChebyshev(op, 3, 20.0, 20)
    cheb_o.estimate_eigenvalues()
    assert abs(cheb.eigenvalues()[1] - cheb_o.lmax) < 1e-4 * cheb_o.lmax
    cheb.vmult(dst, src)
    assert rel_l2(dst.to_numpy().astype(np.float64), cheb_o.vmult(x)) < 1e-4
    x0 = ref.round_f32(seeded(fine.n_dofs, 5))
    dst.from_numpy(x0.astype(F32))
    cheb.step(dst, src)
    assert rel_l2(dst.to_numpy().astype(np.float64), cheb_o.step(x0.copy(), x)) < 1e-4
    # h-transfer between float vectors
    tm = mg.TransferMap(seq[-1], df, seq[-2], dc)
    T_o = ref.MGTwoLevelTransfer(fine, coarse, tm.fine_first(), tm.n_children())
    T_o.rnd = ref.round_f32
    T = mg.MGTwoLevelTransfer(df, dc, tm, op, opc)
    uc, uf = ref.round_f32(seeded(coarse.n_dofs, 1)), ref.round_f32(seeded(fine.n_dofs, 2))
    vc, vf = opc.initialize_dof_vector(), op.initialize_dof_vector()
    vc.from_numpy(uc.astype(F32)), vf.from_numpy(uf.astype(F32))
    T.prolongate_and_add(vf, vc)
    e = uf.copy()
    T_o.prolongate_and_add(e, uc)
    assert rel_l2(vf.to_numpy().astype(np.float64), e) < 1e-5
    vc.from_numpy(uc.astype(F32)), vf.from_numpy(uf.astype(F32))
    T.restrict_and_add(vc, vf)
    e = uc.copy()
    T_o.restrict_and_add(e, uf)
    assert rel_l2(vc.to_numpy().astype(np.float64), e) < 1e-5


@pytest.mark.parametrize("geometry,L,p,mg_type", [("hypercube", 3, 4, "HMG-global"), ("quadrant", 4, 2, "HPMG"), ("annulus", 5, 2, "PMG")])
def test_float_vcycle_and_solve_against_the_float_oracle(gpu, geometry, L, p, mg_type):
    """MGNumberType=float: the double -> float -> double casts around the V-cycle (multigrid_throughput.cc:528-550), the float
    V-cycle within 1e-4 of the float32-rounded oracle, and IDENTICAL iteration counts of the double-precision outer CG"""
    mg = gpu
    prob = mg.Problem(params(geometry, L, p, mg_type, "cg", number="float"))
    H = OracleHierarchy(geometry, L, p, mg_type, 3, "cg", number="float")
    n = H.fine_maps.n_dofs
    r = seeded(n)
    r[H.fine_maps.constrained] = 0.0
    op = prob.fine_operator()                      # double: the outer operator
    src, dst = op.initialize_dof_vector(), op.initialize_dof_vector()
    src.from_numpy(r)
    prob.vcycle(dst, src)                          # double in, double out, float inside
    y = dst.to_numpy()
    assert y.dtype == np.float64
    assert rel_l2(y, H.mg.vmult(r)) < 1e-4
    np.testing.assert_array_equal(y, ref.round_f32(y))   # the result went through float storage
    res = prob.solve_once()
    x, its, _ = H.solve(prob.rhs().to_numpy())
    assert res["n_iterations"] == its
    assert rel_l2(prob.solution().to_numpy(), x) < 1e-3


@pytest.mark.parametrize("geometry,L,p,mg_type", [("hypercube", 3, 2, "HMG-global"), ("quadrant", 4, 2, "HMG-global"), ("quadrant", 3, 2, "HMG-local")])
def test_iteration_count_at_a_borderline_tolerance(gpu, geometry, L, p, mg_type):
    """north_star: CG iteration counts identical.  The relative tolerance is put a hair above and a hair below the residual
    reduction the oracle reaches after k iterations (ReductionControl, multigrid_throughput.cc:1625-1635): the count must flip
    from k to k+1 in the GPU solve exactly as in the oracle, although the atomic scatter orders its additions differently."""
    from util import make_oracle
    mg = gpu
    H = make_oracle(geometry, L, p, mg_type, 3, "cg")
    b = ref.assemble_rhs_constant(H.fine_maps)
    _, k, ctl = H.solve(b, reltol=1e-4)
    rho = ctl.last_value / ctl.initial_value
    for tol, want in ((rho * (1 + 1e-7), k), (rho * (1 - 1e-7), k + 1)):
        _, its_oracle, _ = H.solve(b, reltol=tol)
        assert its_oracle == want
        res = mg.Problem(params(geometry, L, p, mg_type, "cg", reltol=tol)).solve_once()
        assert res["n_iterations"] == want
