"""helpers shared by the tests: build oracle objects from the product's index maps (inputs of both sides)"""
import numpy as np

import dealii_multigrid_b200 as mg
from oracle import reference as ref

PARITY_RTOL = 1e-12  # north_star: vmult and V-cycle within 1e-12 relative in l2 (FP64)


def rel_l2(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1e-300)


def seeded(n, seed=42):
    return np.random.default_rng(seed).uniform(-1.0, 1.0, n)


def mesh_sequence(geometry, n_ref_global, n_ref_local=0):
    """coarse -> fine list of meshes (create_geometric_coarsening_sequence)"""
    seq = [mg.Mesh(geometry, n_ref_global, n_ref_local)]
    while seq[-1].n_global_levels > 1:
        seq.append(seq[-1].coarsen_global())
    return seq[::-1]


def gc_hierarchy(geometry, n_ref_global, degree, mg_type="HMG-global"):
    """levels (mesh, degree) of solve_with_global_coarsening (multigrid_throughput.cc:1506-1576)"""
    meshes = mesh_sequence(geometry, n_ref_global)
    degrees = ref.polynomial_coarsening_sequence_bisect(degree)
    if mg_type == "PMG":
        return [(meshes[-1], d) for d in degrees]
    if mg_type == "HMG-global":
        return [(m, degree) for m in meshes]
    if mg_type == "HPMG":
        return [(m, degrees[0]) for m in meshes] + [(meshes[-1], d) for d in degrees[1:]]
    raise ValueError(mg_type)


class OracleHierarchy:
    """oracle multigrid built from product-generated maps for a GC hierarchy"""

    def __init__(self, geometry, n_ref_global, degree, mg_type="HMG-global", smoother_degree=3, coarse="cg"):
        self.levels = gc_hierarchy(geometry, n_ref_global, degree, mg_type)
        self.dofmaps = [mg.DoFMap(m, d) for m, d in self.levels]
        self.maps = [ref.LevelMaps.from_dofmap(d) for d in self.dofmaps]
        self.ops = [ref.Operator(m) for m in self.maps]
        self.tmaps = [None] + [mg.TransferMap(self.levels[l][0], self.dofmaps[l], self.levels[l - 1][0], self.dofmaps[l - 1])
                               for l in range(1, len(self.levels))]
        self.transfers = [None] + [ref.MGTwoLevelTransfer(self.maps[l], self.maps[l - 1], self.tmaps[l].fine_first(),
                                                          self.tmaps[l].n_children())
                                   for l in range(1, len(self.levels))]
        self.smoothers = [ref.PreconditionChebyshev(op, smoother_degree, 20.0, 20) for op in self.ops]
        cheb = ref.PreconditionChebyshev(self.ops[0], smoother_degree, 20.0, 20) if coarse == "cg_with_chebyshev" else None
        self.coarse = ref.CoarseGridCG(self.ops[0], cheb)
        self.mg = ref.Multigrid(self.ops, self.smoothers, self.transfers, self.coarse)

    def solve(self, rhs, reltol=1e-4):
        x = np.zeros_like(rhs)
        ctl = ref.ReductionControl(10000, 1e-20, reltol)
        its = ref.solver_cg(self.ops[-1].vmult, x, rhs, self.mg.vmult, ctl)
        return x, its, ctl


def params(geometry="hypercube", n_ref=3, degree=2, mg_type="HMG-global", coarse="cg", number="double", smoother_degree=3,
           simulation="Constant", reltol=1e-4):
    return {"Type": mg_type, "GeometryType": geometry, "NRefGlobal": str(n_ref), "NRefLocal": "0", "Degree": str(degree),
            "Paraview": False, "Verbosity": False, "PartitionerName": "DefaultPolicy", "MinLevel": "0", "MinNCells": "0",
            "CoarseGridSolverType": coarse, "SmootherDegree": smoother_degree, "CoarseSolverNCycles": 1,
            "RelativeTolerance": reltol, "MGNumberType": number, "SimulationType": simulation}
