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

    def __init__(self, geometry, n_ref_global, degree, mg_type="HMG-global", smoother_degree=3, coarse="cg", number="double"):
        """number="float": MGNumberType=float -- every level vector is rounded to float32 when it is written; the outer
        operator of the CG (self.fine_op) stays in double (multigrid_throughput.cc:2432-2433, 528-550)"""
        rnd = ref.round_f32 if number == "float" else None
        self.levels = gc_hierarchy(geometry, n_ref_global, degree, mg_type)
        self.dofmaps = [mg.DoFMap(m, d) for m, d in self.levels]
        self.maps = [ref.LevelMaps.from_dofmap(d) for d in self.dofmaps]
        self.ops = [ref.Operator(m, rnd=rnd) for m in self.maps]
        self.tmaps = [None] + [mg.TransferMap(self.levels[l][0], self.dofmaps[l], self.levels[l - 1][0], self.dofmaps[l - 1])
                               for l in range(1, len(self.levels))]
        self.transfers = [None] + [ref.MGTwoLevelTransfer(self.maps[l], self.maps[l - 1], self.tmaps[l].fine_first(),
                                                          self.tmaps[l].n_children())
                                   for l in range(1, len(self.levels))]
        self.smoothers = [ref.PreconditionChebyshev(op, smoother_degree, 20.0, 20) for op in self.ops]
        cheb = ref.PreconditionChebyshev(self.ops[0], smoother_degree, 20.0, 20) if coarse == "cg_with_chebyshev" else None
        self.coarse = ref.CoarseGridCG(self.ops[0], cheb)
        self.mg = ref.Multigrid(self.ops, self.smoothers, self.transfers, self.coarse)
        if rnd is not None:
            for t in self.transfers[1:]:
                t.rnd = rnd

        self.fine_maps = self.maps[-1]
        self.fine_op = ref.Operator(self.maps[-1]) if rnd is not None else self.ops[-1]   # the outer CG runs in double

    def solve(self, rhs, reltol=1e-4):
        x = np.zeros_like(rhs)
        ctl = ref.ReductionControl(10000, 1e-20, reltol)
        its = ref.solver_cg(self.fine_op.vmult, x, rhs, self.mg.vmult, ctl)
        return x, its, ctl


def make_oracle(geometry, n_ref_global, degree, mg_type, smoother_degree=3, coarse="cg"):
    """oracle hierarchy for any multigrid Type of the reference's JSON input"""
    cls = OracleHierarchyLS if mg_type.endswith("-local") else OracleHierarchy
    return cls(geometry, n_ref_global, degree, mg_type, smoother_degree, coarse)


class OracleHierarchyLS:
    """oracle multigrid for solve_with_local_smoothing (multigrid_throughput.cc:1673-1873): HMG-local, or HPMG-local with
    the LS V-cycle nested as the coarse solver of the p-multigrid on the active mesh"""

    def __init__(self, geometry, n_ref_global, degree, mg_type="HMG-local", smoother_degree=3, coarse="cg"):
        self.mesh = mg.Mesh(geometry, n_ref_global)
        nl = self.mesh.n_global_levels
        degrees = ref.polynomial_coarsening_sequence_bisect(degree) if mg_type == "HPMG-local" else [degree]
        has_hanging = len(set(self.mesh.cells()[:, 0].tolist())) > 1
        self.level_meshes = [self.mesh.level_mesh(l) for l in range(nl)]
        self.level_dofmaps = [mg.DoFMap(m, degrees[0]) for m in self.level_meshes]
        self.level_edges = [mg.LevelEdges(m, d) for m, d in zip(self.level_meshes, self.level_dofmaps)]
        self.level_maps = [ref.LevelMaps.from_dofmap(d) for d in self.level_dofmaps]
        self.level_ops = [ref.Operator(m, e.edge_dofs, e.edge_cells, e.any_global) for m, e in zip(self.level_maps, self.level_edges)]
        self.active_dofmaps = [mg.DoFMap(self.mesh, d) for d in degrees]
        self.active_maps = [ref.LevelMaps.from_dofmap(d) for d in self.active_dofmaps]
        self.copy = [mg.CopyIndices(self.active_dofmaps[0], m, d) for m, d in zip(self.level_meshes, self.level_dofmaps)]
        self.ls_tmaps = [None] + [mg.TransferMap(self.level_meshes[l], self.level_dofmaps[l], self.level_meshes[l - 1],
                                                 self.level_dofmaps[l - 1], local_smoothing=True) for l in range(1, nl)]
        self.ls_transfers = [None] + [ref.MGTwoLevelTransfer(self.level_maps[l], self.level_maps[l - 1], self.ls_tmaps[l].fine_first(),
                                                             self.ls_tmaps[l].n_children()) for l in range(1, nl)]
        cheb = lambda op: ref.PreconditionChebyshev(op, smoother_degree, 20.0, 20)
        self.level_smoothers = [cheb(op) for op in self.level_ops]
        self.coarse = ref.CoarseGridCG(self.level_ops[0], cheb(self.level_ops[0]) if coarse == "cg_with_chebyshev" else None)
        self.mg_ls = ref.Multigrid(self.level_ops, self.level_smoothers, self.ls_transfers, self.coarse, edge_in=has_hanging,
                                   copy_indices=[(c.active_idx.astype(np.int64), c.level_idx.astype(np.int64)) for c in self.copy])
        if mg_type == "HPMG-local":
            self.active_ops = [ref.Operator(m) for m in self.active_maps]
            self.p_tmaps = [None] + [mg.TransferMap(self.mesh, self.active_dofmaps[i], self.mesh, self.active_dofmaps[i - 1])
                                     for i in range(1, len(degrees))]
            self.p_transfers = [None] + [ref.MGTwoLevelTransfer(self.active_maps[i], self.active_maps[i - 1], self.p_tmaps[i].fine_first(),
                                                                self.p_tmaps[i].n_children()) for i in range(1, len(degrees))]
            self.active_smoothers = [cheb(op) for op in self.active_ops]
            self.mg = ref.Multigrid(self.active_ops, self.active_smoothers, self.p_transfers, self.mg_ls.vmult)
            self.fine_op = self.active_ops[-1]
        else:
            self.mg = self.mg_ls
            self.fine_op = ref.Operator(self.active_maps[-1])
        self.maps = self.level_maps + (self.active_maps if mg_type == "HPMG-local" else [])
        self.ops = self.level_ops + (self.active_ops if mg_type == "HPMG-local" else [])
        self.smoothers = self.level_smoothers + (self.active_smoothers if mg_type == "HPMG-local" else [])
        self.fine_maps = self.active_maps[-1]

    def solve(self, rhs, reltol=1e-4):
        x = np.zeros_like(rhs)
        ctl = ref.ReductionControl(10000, 1e-20, reltol)
        its = ref.solver_cg(self.fine_op.vmult, x, rhs, self.mg.vmult, ctl)
        return x, its, ctl


def params(geometry="hypercube", n_ref=3, degree=2, mg_type="HMG-global", coarse="cg", number="double", smoother_degree=3,
           simulation="Constant", reltol=1e-4):
    return {"Type": mg_type, "GeometryType": geometry, "NRefGlobal": str(n_ref), "NRefLocal": "0", "Degree": str(degree),
            "Paraview": False, "Verbosity": False, "PartitionerName": "DefaultPolicy", "MinLevel": "0", "MinNCells": "0",
            "CoarseGridSolverType": coarse, "SmootherDegree": smoother_degree, "CoarseSolverNCycles": 1,
            "RelativeTolerance": reltol, "MGNumberType": number, "SimulationType": simulation}


# ------------------------------------------------------------------------------------ general (non-Cartesian) geometry
AFFINE_A = np.array([[1.10, 0.25, -0.10], [0.05, 0.90, 0.30], [-0.20, 0.15, 1.20]])


def affine_map(xyz):
    """shear + stretch + shift: keeps hanging nodes conforming (midpoints map to midpoints)"""
    return xyz @ AFFINE_A.T + np.array([0.3, -0.2, 0.1])


def smooth_map(xyz):
    """smooth non-affine deformation of [-1,1]^3 (vertices only; trilinear inside the cells) for meshes without hanging nodes"""
    x, y, z = xyz[:, 0], xyz[:, 1], xyz[:, 2]
    return np.stack([x + 0.08 * np.sin(np.pi * y) * np.cos(0.5 * np.pi * z), y + 0.06 * np.sin(np.pi * z) * x,
                     z + 0.07 * np.cos(np.pi * x) * np.sin(0.5 * np.pi * y)], axis=1)
