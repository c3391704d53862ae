#!/usr/bin/env python
"""make_golden.py -- (re)generate tests/golden/*.npz from the CPU oracle.

The reference ships no golden vectors for this path, and deal.II cannot run here, so these fixtures
freeze the ORACLE's answers on seeded inputs: they guard the oracle against silent drift (CPU suite) and give the GPU
suite a file-based known answer that does not depend on executing oracle code on the GPU box.
Index maps come from the product's host generator (they are inputs of both sides); run from the repo root:
    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

from util import make_oracle, seeded  # noqa: E402

CASES = [("hypercube", 2, 4, "HMG-global", "cg"), ("quadrant", 3, 2, "HMG-global", "cg"), ("annulus", 4, 2, "HPMG", "cg_with_chebyshev"),
         ("quadrant", 3, 4, "PMG", "cg"), ("quadrant", 3, 2, "HMG-local", "cg"), ("circle", 4, 2, "HPMG-local", "cg")]


def case_name(c):
    return "%s_L%d_p%d_%s_%s" % c


def generate(c):
    from oracle import reference as ref
    geometry, L, p, mg_type, coarse = c
    H = make_oracle(geometry, L, p, mg_type, 3, coarse)
    n = H.fine_maps.n_dofs
    x = seeded(n)
    r = x.copy()
    r[H.fine_maps.constrained] = 0.0
    b = ref.assemble_rhs_constant(H.fine_maps)
    sol, its, ctl = H.solve(b.copy())
    for s in H.smoothers:
        s.estimate_eigenvalues()
    return dict(n_dofs=n, level_dofs=np.array([m.n_dofs for m in H.maps]), vmult=H.fine_op.vmult(x),
                inverse_diagonal=H.fine_op.compute_inverse_diagonal(), vcycle=H.mg.vmult(r), rhs=b, solution=sol,
                n_iterations=its, final_residual=ctl.last_value, lambda_max=np.array([s.lmax for s in H.smoothers]))


# general (non-Cartesian) geometry: operator-level fixtures in tests/golden/mapped/ (vmult and inverse diagonal of the
# mapped-cell oracle on seeded input; the deformation is applied to the cell vertices, util.affine_map / smooth_map)
MAPPED_CASES = [("quadrant", 3, 4, "affine"), ("quadrant", 3, 1, "affine"), ("hypercube", 2, 4, "smooth"), ("hypercube", 3, 3, "smooth")]


def mapped_case_name(c):
    return "%s_L%d_p%d_%s" % c


def generate_mapped(c):
    import dealii_multigrid_b200 as mg
    from oracle import reference as ref
    from util import affine_map, smooth_map
    geometry, L, p, deformation = c
    maps = ref.LevelMaps.from_dofmap(mg.DoFMap(mg.Mesh(geometry, L), p))
    verts = ref.cell_vertices(maps, {"affine": affine_map, "smooth": smooth_map}[deformation])
    op = ref.MappedOperator(maps, verts)
    return dict(n_dofs=maps.n_dofs, vertices=verts, vmult=op.vmult(seeded(maps.n_dofs)), inverse_diagonal=op.compute_inverse_diagonal())


if __name__ == "__main__":
    os.makedirs(os.path.join(HERE, "mapped"), exist_ok=True)
    for c in MAPPED_CASES:
        if os.path.exists(os.path.join(HERE, "mapped", mapped_case_name(c) + ".npz")) and "--all" not in sys.argv:
            continue
        np.savez_compressed(os.path.join(HERE, "mapped", mapped_case_name(c) + ".npz"), **generate_mapped(c))
    for c in CASES:
        if os.path.exists(os.path.join(HERE, case_name(c) + ".npz")) and "--all" not in sys.argv:
            continue  # existing fixtures stay frozen; pass --all to regenerate everything
        np.savez_compressed(os.path.join(HERE, case_name(c) + ".npz"), **generate(c))
        print("wrote", case_name(c))
