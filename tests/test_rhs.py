"""Operator::rhs (include/operator.h:362-447) for both SimulationTypes of the reference (multigrid_throughput.cc:2286-2303):
the oracle restatement is pinned mathematically (the discrete solution converges to the analytic GaussianSolution at the
rate h^(p+1)), the product's host routine is compared with the oracle, and on the GPU the whole Gaussian problem (config 5
of BASELINE.json) is solved with the oracle's iteration count."""
import numpy as np
import pytest
import scipy.sparse.linalg as sla

import dealii_multigrid_b200 as mg
from oracle import reference as ref
from util import make_oracle, params, rel_l2


def _nodal_error(geometry, L, p):
    d = mg.DoFMap(mg.Mesh(geometry, L), p)
    maps = ref.LevelMaps.from_dofmap(d)
    b = ref.assemble_rhs_gaussian(maps)
    op = ref.Operator(maps)
    u, info = sla.cg(sla.LinearOperator((maps.n_dofs,) * 2, matvec=op.vmult), b, rtol=1e-12, maxiter=4000)
    assert info == 0
    free = np.setdiff1d(np.arange(maps.n_dofs), maps.constrained)
    return np.max(np.abs(u[free] - ref.gaussian_solution(maps.dof_coordinates()[free])))


def test_oracle_gaussian_problem_converges_to_the_analytic_solution():
    """-laplace u = GaussianRightHandSide, u = GaussianSolution on the boundary: max nodal error falls like h^(p+1)"""
    e2 = [_nodal_error("hypercube", L, 2) for L in (3, 4)]
    assert e2[0] / e2[1] > 8.0                      # p = 2: at least h^3
    e4 = [_nodal_error("hypercube", L, 4) for L in (2, 3)]
    assert e4[0] / e4[1] > 24.0                     # p = 4: h^5 = 32 asymptotically
    # the locally refined mesh resolves the Gaussian like the uniform one (hanging nodes included)
    assert abs(_nodal_error("quadrant", 4, 2) / e2[1] - 1.0) < 1e-3


def test_gaussian_functions_are_consistent():
    """GaussianRightHandSide = -laplace GaussianSolution (multigrid_throughput.cc:60-127), by finite differences"""
    x = np.array([[-0.45, -0.52, -0.4], [-0.6, -0.5, -0.55], [0.1, 0.2, -0.3]])
    e, lap = 1e-4, np.zeros(3)
    for d in range(3):
        dx = np.zeros(3)
        dx[d] = e
        lap += (ref.gaussian_solution(x + dx) - 2 * ref.gaussian_solution(x) + ref.gaussian_solution(x - dx)) / e ** 2
    assert np.allclose(-lap, ref.gaussian_rhs(x), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("geometry,L,p", [("quadrant", 3, 2), ("quadrant", 4, 4), ("hypercube", 3, 3), ("annulus", 5, 2), ("circle", 4, 4),
                                          ("quadrant", 4, 1)])
@pytest.mark.parametrize("simulation", ["Constant", "Gaussian"])
def test_host_rhs_matches_oracle(geometry, L, p, simulation):
    mesh = mg.Mesh(geometry, L)
    d = mg.DoFMap(mesh, p)
    maps = ref.LevelMaps.from_dofmap(d)
    expect = ref.assemble_rhs_gaussian(maps) if simulation == "Gaussian" else ref.assemble_rhs_constant(maps)
    assert rel_l2(d.assemble_rhs(mesh, simulation), expect) < 1e-13
    assert np.all(expect[maps.constrained] == 0.0)


def test_host_rhs_on_partitioned_mesh_sums_to_the_serial_one():
    mesh = mg.Mesh("quadrant", 4)
    serial = mg.DoFMap(mesh, 2)
    expect = ref.assemble_rhs_gaussian(ref.LevelMaps.from_dofmap(serial))
    total = np.zeros(serial.n_dofs_global)
    for rank in range(3):
        d = mg.DoFMap(mesh, 2, 3, rank)
        loc = d.assemble_rhs(mesh, "Gaussian")
        glob = np.concatenate([np.arange(d.owned_begin, d.owned_begin + d.n_owned), d.ghost_global()])
        np.add.at(total, glob, loc)   # compress(add)
    assert rel_l2(total, expect) < 1e-13


@pytest.mark.gpu
@pytest.mark.parametrize("geometry,L,p,mg_type", [("quadrant", 4, 2, "HMG-global"), ("quadrant", 3, 4, "HPMG"), ("hypercube", 3, 2, "HMG-local")])
def test_gpu_gaussian_problem(gpu, geometry, L, p, mg_type):
    prob = gpu.Problem(params(geometry, L, p, mg_type, "cg", simulation="Gaussian"))
    H = make_oracle(geometry, L, p, mg_type, 3, "cg")
    b = ref.assemble_rhs_gaussian(H.fine_maps)
    assert rel_l2(prob.rhs().to_numpy(), b) < 1e-13
    x, its, _ = H.solve(b)
    res = prob.solve_once()
    assert res["n_iterations"] == its
    assert rel_l2(prob.solution().to_numpy(), x) < 1e-9
