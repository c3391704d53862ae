"""Chunk plan of the cell loop (csrc/host/chunks.h, kernel csrc/device/chunk.cuh), host side: the plan is walked on the CPU
in the kernel's own order (chunks descending, DoF window, per-DoF contribution lists, side buffers, pull / late / ghost
contributions) with a diagonal stand-in for the cell matrices and compared with the direct scatter-add over the entity
table of the reference's cell loop (include/operator.h:450-493).  No GPU."""
import pytest

import dealii_multigrid_b200 as mg
from util import mesh_sequence

CASES = [("hypercube", 3, 1), ("hypercube", 3, 2), ("hypercube", 3, 3), ("hypercube", 3, 4), ("hypercube", 2, 5), ("hypercube", 2, 6),
         ("hypercube", 4, 4), ("quadrant", 3, 1), ("quadrant", 3, 2), ("quadrant", 4, 4), ("quadrant", 3, 3), ("annulus", 4, 2),
         ("annulus", 5, 4), ("quadrant", 3, 6), ("quadrant", 6, 4), ("annulus", 6, 4), ("circle", 4, 2)]


@pytest.mark.parametrize("geometry,nref,degree", CASES)
@pytest.mark.parametrize("value_bytes", [8, 4])
def test_plan_reproduces_scatter_add(geometry, nref, degree, value_bytes):
    mesh = mg.Mesh(geometry, nref)
    r = mg.DoFMap(mesh, degree).chunk_plan_check(value_bytes)
    assert r["max_err"] < 1e-13, r
    assert r["n_chunks"] >= 1


def test_uniform_mesh_shares_tables_and_has_no_late_contributions():
    dm = mg.DoFMap(mg.Mesh("hypercube", 5), 4)
    r = dm.chunk_plan_check(8)
    assert r["max_err"] < 1e-13
    assert r["foreign_late"] == 0 and r["foreign_ghost"] == 0
    # equal configurations share one table: far fewer tables than chunks
    assert r["n_compute_tables"] < 0.5 * r["n_chunks"], r
    # the side buffer holds about a fifth of the DoFs
    assert r["side_entries"] < 0.35 * dm.n_owned


@pytest.mark.parametrize("geometry,nref,degree,n_ranks", [("hypercube", 3, 4, 2), ("quadrant", 4, 2, 3), ("annulus", 4, 4, 4),
                                                            ("hypercube", 4, 1, 8)])
def test_plan_on_partitioned_meshes(geometry, nref, degree, n_ranks):
    mesh = mg.Mesh(geometry, nref)
    for rank in range(n_ranks):
        dm = mg.DoFMap(mesh, degree, n_ranks, rank)
        r = dm.chunk_plan_check(8)
        assert r["max_err"] < 1e-13, (rank, r)
        if dm.n_ghost and dm.n_cells:
            assert r["foreign_ghost"] > 0


def test_plan_on_local_smoothing_level_meshes():
    mesh = mg.Mesh("quadrant", 4)
    for level in range(mesh.n_global_levels):
        lm = mesh.level_mesh(level)
        r = mg.DoFMap(lm, 3).chunk_plan_check(8)
        assert r["max_err"] < 1e-13, (level, r)
