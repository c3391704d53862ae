"""The chunked cell loop (csrc/device/chunk.cuh, opt-in with MGB_CHUNKED=1 when an operator is created) against the oracle:
vmult, the fused epilogues (residual, the two Chebyshev updates: north_star "the smoother's vector updates fused with vmult",
reference smoother multigrid_throughput.cc:866-883 on the operator of include/operator.h:152-183), the smoother built on them,
the whole V-cycle-preconditioned CG, and bitwise reproducibility (no atomics on owned entries)."""
import os

import numpy as np
import pytest

from oracle import reference as ref
from util import PARITY_RTOL, make_oracle, params, rel_l2, seeded

pytestmark = pytest.mark.gpu

CASES = [("hypercube", 2, 1), ("hypercube", 3, 2), ("hypercube", 3, 3), ("hypercube", 4, 4), ("hypercube", 2, 5), ("hypercube", 2, 6),
         ("quadrant", 4, 1), ("quadrant", 4, 2), ("quadrant", 5, 4), ("annulus", 5, 3), ("annulus", 6, 4), ("circle", 4, 4)]


@pytest.fixture()
def chunked(gpu):
    old = os.environ.get("MGB_CHUNKED")
    os.environ["MGB_CHUNKED"] = "1"
    yield gpu
    if old is None:
        del os.environ["MGB_CHUNKED"]
    else:
        os.environ["MGB_CHUNKED"] = old


@pytest.mark.parametrize("geometry,L,p", CASES)
def test_chunked_vmult_and_fused_epilogues(chunked, geometry, L, p):
    mg = chunked
    d = mg.DoFMap(mg.Mesh(geometry, L), p)
    maps = ref.LevelMaps.from_dofmap(d)
    oracle = ref.Operator(maps)
    op = mg.Operator(d)
    assert op.fused_available()
    n = maps.n_dofs
    x, b, xold = seeded(n, 1), seeded(n, 2), seeded(n, 3)
    dinv = oracle.compute_inverse_diagonal()
    Ax = oracle.vmult(x)
    vx, vb, vo, vd, out = (op.initialize_dof_vector() for _ in range(5))
    vx.from_numpy(x), vb.from_numpy(b), vo.from_numpy(xold), vd.from_numpy(dinv)
    op.vmult(out, vx)
    y = out.to_numpy()
    assert rel_l2(y, Ax) < PARITY_RTOL
    np.testing.assert_array_equal(y[maps.constrained], x[maps.constrained])
    np.testing.assert_array_equal(vx.to_numpy(), x)
    # the same launch again: bitwise identical where no late (atomic) contribution exists, 1e-15 otherwise
    op.vmult(out, vx)
    late = d.chunk_plan_check(8)["foreign_late"]
    if late == 0:
        np.testing.assert_array_equal(out.to_numpy(), y)
    else:
        assert rel_l2(out.to_numpy(), y) < 1e-15
    op.vmult_fused(out, vx, 0)
    assert rel_l2(out.to_numpy(), Ax) < PARITY_RTOL
    op.vmult_fused(out, vx, 1, b=vb)
    assert rel_l2(out.to_numpy(), b - Ax) < PARITY_RTOL
    f1, f2 = 0.37, 0.011
    op.vmult_fused(out, vx, 2, b=vb, dinv=vd, f1=f1, f2=f2)
    assert rel_l2(out.to_numpy(), (1 + f1) * x + f2 * dinv * (b - Ax)) < PARITY_RTOL
    op.vmult_fused(vo, vx, 3, b=vb, dinv=vd, xold=vo, f1=f1, f2=f2)   # out aliases xold, as the smoother does
    assert rel_l2(vo.to_numpy(), (1 + f1) * x - f1 * xold + f2 * dinv * (b - Ax)) < PARITY_RTOL
    # the other kernels of the operator are untouched by the switch
    diag = op.initialize_dof_vector()
    op.compute_inverse_diagonal(diag)
    assert rel_l2(diag.to_numpy(), dinv) < PARITY_RTOL


def test_chunked_float_operator(chunked):
    mg = chunked
    d = mg.DoFMap(mg.Mesh("quadrant", 4), 4)
    maps = ref.LevelMaps.from_dofmap(d)
    op = mg.Operator(d, number_type=mg.FLOAT)
    x = seeded(maps.n_dofs)
    src, dst = op.initialize_dof_vector(), op.initialize_dof_vector()
    src.from_numpy(x.astype(np.float32))
    op.vmult(dst, src)
    assert rel_l2(dst.to_numpy().astype(np.float64), ref.Operator(maps).vmult(x.astype(np.float32).astype(np.float64))) < 2e-5


@pytest.mark.parametrize("geometry,L,p,degree", [("hypercube", 3, 4, 3), ("quadrant", 4, 2, 3), ("annulus", 5, 4, 5), ("quadrant", 4, 3, 2),
                                                  ("hypercube", 3, 2, 4)])
def test_chunked_chebyshev_runs_fused(chunked, geometry, L, p, degree):
    mg = chunked
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
    n0 = mg.launch_count()
    cheb.vmult(x, rhs)
    n_fused = mg.launch_count() - n0
    assert rel_l2(x.to_numpy(), oracle.vmult(b)) < 1e-11
    # zero guess: one scaling pass + (degree-1) fused cell loops, each followed by at most one identity-row kernel (a single
    # update cannot land in dst without a copy and runs unfused)
    if degree >= 3:
        assert n_fused <= 1 + 2 * (degree - 1)
    x0 = seeded(maps.n_dofs, 5)
    x.from_numpy(x0)
    cheb.step(x, rhs)
    assert rel_l2(x.to_numpy(), oracle.step(x0.copy(), b)) < 1e-11
    np.testing.assert_array_equal(rhs.to_numpy(), b)


@pytest.mark.parametrize("geometry,L,p,mg_type", [("hypercube", 3, 4, "HMG-global"), ("quadrant", 4, 4, "HPMG"), ("annulus", 5, 2, "PMG"),
                                                   ("quadrant", 3, 2, "HMG-local")])
def test_chunked_vcycle_and_solve(chunked, geometry, L, p, mg_type):
    mg = chunked
    prob = mg.Problem(params(geometry, L, p, mg_type, "cg"))
    H = make_oracle(geometry, L, p, mg_type, 3, "cg")
    n = H.fine_maps.n_dofs
    r = seeded(n)
    r[H.fine_maps.constrained] = 0.0
    op = prob.fine_operator()
    src, dst = op.initialize_dof_vector(), op.initialize_dof_vector()
    src.from_numpy(r)
    prob.vcycle(dst, src)
    assert rel_l2(dst.to_numpy(), H.mg.vmult(r)) < 1e-11
    res = prob.solve_once()
    _, its, _ = H.solve(prob.rhs().to_numpy())
    assert res["n_iterations"] == its
