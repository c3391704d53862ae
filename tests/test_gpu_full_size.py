"""The bench workload itself (BASELINE.json configs[1] at the size bench.py runs: hypercube L=7, p=4, 513^3 = 135 005 697 DoFs,
2 097 152 cells) through properties that need no oracle of that size: symmetry and linearity of the operator
(include/operator.h:152-183), adjointness of the finest two-level transfer pair (multigrid_throughput.cc:1600-1604),
symmetry and positivity of the V-cycle as a preconditioner (:1091-1133), an independent residual of the CG solve and its
mesh-independent iteration count (4, as at L=5 and L=6 where the C restatement confirms it).  32-bit index overflow, a
wrong chunk of a 1 GB vector or a lost atomic at this size would show here; the small-size tests cannot see them."""
import numpy as np
import pytest

from util import params

pytestmark = pytest.mark.gpu

L, P, N = 7, 4, 513 ** 3


@pytest.fixture(scope="module")
def problem(gpu):
    import torch
    if torch.cuda.get_device_properties(0).total_memory < 60e9:
        pytest.skip("needs ~40 GB of device memory")
    prob = gpu.Problem(params("hypercube", L, P, "HMG-global", "cg"))
    yield prob
    del prob


def _random(op, seed):
    v = op.initialize_dof_vector()
    v.from_numpy(np.random.default_rng(seed).uniform(-1.0, 1.0, N))
    return v


def test_operator_is_symmetric_and_linear_at_full_size(problem):
    op = problem.fine_operator()
    assert problem.level_n_dofs(problem.n_levels - 1) == N
    u, v, Au, Av = _random(op, 1), _random(op, 2), op.initialize_dof_vector(), op.initialize_dof_vector()
    op.vmult(Au, u)
    op.vmult(Av, v)
    a, b = v.dot(Au), u.dot(Av)
    assert abs(a - b) < 1e-12 * v.l2_norm() * Au.l2_norm()
    assert u.dot(Au) > 0.0
    # A (2 u - 3 v) = 2 A u - 3 A v
    w, Aw = op.initialize_dof_vector(), op.initialize_dof_vector()
    w.equ(2.0, u)
    w.add(-3.0, v)
    op.vmult(Aw, w)
    Aw.add(-2.0, Au)
    Aw.add(3.0, Av)
    assert Aw.l2_norm() < 1e-13 * (2.0 * Au.l2_norm() + 3.0 * Av.l2_norm())


def test_finest_transfer_pair_is_adjoint_at_full_size(problem):
    top = problem.n_levels - 1
    fine, coarse, T = problem.level_operator(top), problem.level_operator(top - 1), problem.level_transfer(top)
    nc = problem.level_n_dofs(top - 1)
    assert nc == 257 ** 3
    uf, uc = _random(fine, 3), coarse.initialize_dof_vector()
    uc.from_numpy(np.random.default_rng(4).uniform(-1.0, 1.0, nc))
    Puc, Ruf = fine.initialize_dof_vector(), coarse.initialize_dof_vector()
    T.prolongate_and_add(Puc, uc)
    T.restrict_and_add(Ruf, uf)
    a, b = uf.dot(Puc), uc.dot(Ruf)
    assert abs(a - b) < 1e-12 * Puc.l2_norm() * uf.l2_norm()


def test_vcycle_is_a_symmetric_positive_preconditioner_at_full_size(problem):
    op = problem.fine_operator()
    dm = problem.fine_dofmap()
    r, s = _random(op, 5), _random(op, 6)
    # the V-cycle acts on residuals, which vanish on constrained rows
    constrained = np.asarray(dm.constrained(), dtype=np.int64)
    assert constrained.size == 513 ** 3 - 511 ** 3
    for vec in (r, s):
        a = vec.to_numpy()
        a[constrained] = 0.0
        vec.from_numpy(a)
    Mr, Ms = op.initialize_dof_vector(), op.initialize_dof_vector()
    problem.vcycle(Mr, r)
    problem.vcycle(Ms, s)
    a, b = s.dot(Mr), r.dot(Ms)
    # symmetric up to the coarse CG, which stops at a relative residual of 1e-4 on the one-cell level
    assert abs(a - b) < 1e-6 * (Mr.l2_norm() * s.l2_norm())
    assert r.dot(Mr) > 0.0 and s.dot(Ms) > 0.0


def test_solve_residual_and_iteration_count_at_full_size(problem):
    op = problem.fine_operator()
    res = problem.solve_once()
    assert res["n_iterations"] == 4
    b, x, Ax = problem.rhs(), problem.solution(), op.initialize_dof_vector()
    op.vmult(Ax, x)
    Ax.sadd(-1.0, 1.0, b)  # b - A x
    assert Ax.l2_norm() <= 1e-4 * b.l2_norm()
    # the recurrence residual the solver reports and the true residual agree
    assert abs(Ax.l2_norm() - res["final_residual"]) < 1e-6 * b.l2_norm()
    assert abs(res["initial_residual"] - b.l2_norm()) < 1e-12 * b.l2_norm()
