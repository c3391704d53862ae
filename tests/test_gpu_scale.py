"""Parity at sizes beyond the toy cases (index arithmetic over millions of cell records is where a 32-bit overflow
would hide): hypercube L=5 p=4 (2 146 689 DoFs, 32 768 cells) against the threaded C restatement oracle/mgcpu.c, and the
locally refined quadrant L=5 p=4 (hanging nodes) against the numpy oracle -- vmult <= 1e-12, V-cycle <= 1e-11, identical CG
iteration counts (reference: include/operator.h:152-183, multigrid_throughput.cc:1091-1133, 1247-1248).  bench.py repeats
the first comparison at L=6 (17 M DoFs) in its `parity_check` object on every driver run."""
import numpy as np
import pytest

from oracle import reference as ref
from util import OracleHierarchy, mesh_sequence, params, rel_l2, seeded

pytestmark = pytest.mark.gpu


def test_hypercube_L5_p4_against_the_c_restatement(gpu):
    from oracle.cpu_solver import CpuMultigrid
    mg = gpu
    L, p = 5, 4
    meshes = mesh_sequence("hypercube", L)
    dofmaps = [mg.DoFMap(m, p) for m in meshes]
    tmaps = [None] + [mg.TransferMap(meshes[l], dofmaps[l], meshes[l - 1], dofmaps[l - 1]) for l in range(1, len(meshes))]
    M = CpuMultigrid(dofmaps, tmaps, 3, 20.0, 20)
    n = M.maps[-1].n_dofs
    assert n == 129 ** 3
    prob = mg.Problem(params("hypercube", L, p, "HMG-global", "cg"))
    op = prob.fine_operator()
    x = seeded(n, 7)
    src, dst = op.initialize_dof_vector(), op.initialize_dof_vector()
    src.from_numpy(x)
    op.vmult(dst, src)
    assert rel_l2(dst.to_numpy(), M.vmult(len(M.maps) - 1, x)) < 1e-12
    r = x.copy()
    r[M.maps[-1].constrained] = 0.0
    src.from_numpy(r)
    prob.vcycle(dst, src)
    assert rel_l2(dst.to_numpy(), M.vcycle(r)) < 1e-11
    b = ref.assemble_rhs_constant(M.maps[-1])
    assert rel_l2(prob.rhs().to_numpy(), b) < 1e-13
    xs, its, _, _ = M.solve(b)
    res = prob.solve_once()
    assert res["n_iterations"] == its
    assert rel_l2(prob.solution().to_numpy(), xs) < 1e-9


def test_quadrant_L5_p4_with_hanging_nodes_against_the_numpy_oracle(gpu):
    mg = gpu
    H = OracleHierarchy("quadrant", 5, 4, "HMG-global", 3, "cg")
    n = H.fine_maps.n_dofs
    assert n > 250_000 and np.any(H.fine_maps.mask >> 3)
    prob = mg.Problem(params("quadrant", 5, 4, "HMG-global", "cg"))
    op = prob.fine_operator()
    x = seeded(n, 9)
    src, dst = op.initialize_dof_vector(), op.initialize_dof_vector()
    src.from_numpy(x)
    op.vmult(dst, src)
    assert rel_l2(dst.to_numpy(), H.fine_op.vmult(x)) < 1e-12
    r = x.copy()
    r[H.fine_maps.constrained] = 0.0
    src.from_numpy(r)
    prob.vcycle(dst, src)
    assert rel_l2(dst.to_numpy(), H.mg.vmult(r)) < 1e-11
    _, its, _ = H.solve(prob.rhs().to_numpy())
    assert prob.solve_once()["n_iterations"] == its
