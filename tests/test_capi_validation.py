"""The array-based entry points of include/mgb200.h take index maps from another program (the deal.II adapter,
include/mgb200_dealii.h); every index becomes an address inside a kernel, so mgb_operator_create,
mgb_transfer_create_from_arrays and mgb_level_copy_create_from_arrays check them on the host before anything is uploaded.
These checks run without a GPU (the error is raised before the first CUDA call)."""
import numpy as np
import pytest

import dealii_multigrid_b200 as mg


@pytest.fixture(scope="module")
def dofs():
    return mg.DoFMap(mg.Mesh("quadrant", 3), 3)


def _create(d, **override):
    kw = dict(degree=3, cell_level=d.cell_level(), n_owned=d.n_owned, mask=d.mask(), constrained=d.constrained(),
              entity_table=d.eval_table())
    kw.update(override)
    return mg.Operator(**kw)


def test_entity_table_beyond_the_local_vector_is_rejected(dofs):
    bad = dofs.eval_table().copy()
    flat = bad.reshape(-1)
    k = int(np.argmax(flat & 0x7FFFFFFF))          # the entity with the largest base: push it past the end
    flat[k] = (flat[k] & 0x80000000) | (dofs.n_owned - 1)
    if flat[k] & 0x80000000:                       # a Dirichlet entity is never dereferenced: take an unflagged one
        k = int(np.flatnonzero((flat & 0x80000000) == 0)[-1])
        flat[k] = dofs.n_owned - 1
    with pytest.raises(mg.MgbError, match="exceed the local vector"):
        _create(dofs, entity_table=bad)
    with pytest.raises(mg.MgbError, match="exceed the local vector"):
        _create(dofs, n_owned=dofs.n_owned // 2, constrained=None)


def test_index_lists_are_checked_against_their_ranges(dofs):
    c = dofs.constrained().copy()
    c[-1] = dofs.n_owned
    with pytest.raises(mg.MgbError, match=r"constrained: entry \d+ = \d+ is outside"):
        _create(dofs, constrained=c)
    with pytest.raises(mg.MgbError, match="edge_cells"):
        _create(dofs, edge_cells=np.array([dofs.n_cells], dtype=np.uint32))
    with pytest.raises(mg.MgbError, match="edge_constrained"):
        _create(dofs, edge_constrained=np.array([2 ** 31], dtype=np.uint32))
    lv = dofs.cell_level().copy()
    lv[0] = 31
    with pytest.raises(mg.MgbError, match="cell_level"):
        _create(dofs, cell_level=lv)


def test_partitioner_arrays_are_checked(dofs):
    part = dict(n_ghost=4, ghost_global=np.arange(4) + dofs.n_owned, ghost_ranks=[1], ghost_offsets=[0, 3],
                import_ranks=[1], import_offsets=[0, 2], import_indices=[0, 1])
    with pytest.raises(mg.MgbError, match="ghost_offsets: offsets must run from 0 to 4"):
        _create(dofs, partitioner=part, n_dofs_global=dofs.n_owned + 4)
    part.update(ghost_offsets=[0, 4], import_indices=[0, dofs.n_owned])
    with pytest.raises(mg.MgbError, match="import_indices"):
        _create(dofs, partitioner=part, n_dofs_global=dofs.n_owned + 4)
