"""mgb_problem_solve_host_batch: a sequence of host-buffer solves whose uploads / downloads run behind the neighbouring solves
must return exactly what the one-call-per-solve entry point (mgb_problem_solve_host) returns for the same right-hand sides
(the solve itself is the reference's SolverCG + PreconditionMG of multigrid_throughput.cc:1247-1248)."""
import numpy as np
import pytest

from util import params, rel_l2

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("geometry,L,p,mg_type", [("quadrant", 4, 2, "HMG-global"), ("hypercube", 3, 4, "HPMG"), ("quadrant", 3, 2, "HMG-local")])
def test_host_batch_matches_single_calls(gpu, geometry, L, p, mg_type):
    import torch
    mg = gpu
    prob = mg.Problem(params(geometry, L, p, mg_type, "cg"))
    b = prob.rhs().to_numpy()
    n = len(b)
    scales = [1.0, -2.5, 0.125, 3.0]
    rhs = [torch.from_numpy(s * b).pin_memory() for s in scales]
    single, batch = [torch.empty(n, dtype=torch.float64).pin_memory() for _ in scales], [torch.full((n,), 7.0, dtype=torch.float64).pin_memory() for _ in scales]
    its_single = [prob.solve_host(r.data_ptr(), o.data_ptr())["n_iterations"] for r, o in zip(rhs, single)]
    res = prob.solve_host_batch([r.data_ptr() for r in rhs], [o.data_ptr() for o in batch])
    assert [r["n_iterations"] for r in res] == its_single
    for s, a, c in zip(scales, single, batch):
        assert rel_l2(c.numpy(), a.numpy()) < 1e-10
        assert rel_l2(c.numpy(), s * single[0].numpy()) < 1e-9     # the solver is linear in the right-hand side
    # a second batch reuses the staging buffers; one buffer may serve several steps
    res2 = prob.solve_host_batch([rhs[1].data_ptr()] * 3, [batch[0].data_ptr()] * 3)
    assert [r["n_iterations"] for r in res2] == [its_single[1]] * 3
    assert rel_l2(batch[0].numpy(), single[1].numpy()) < 1e-10
    assert prob.solve_host_batch([], []) == []
