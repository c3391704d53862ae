"""Python mirror of the reference's operator interface over the C ABI (include/mgb200.h).

Class and method names follow the deal.II concepts the reference composes:
  Operator.vmult / Tvmult / vmult_interface_down / vmult_interface_up / compute_inverse_diagonal
                                                    -> include/operator.h:152-242
  PreconditionChebyshev.vmult / step                -> multigrid_throughput.cc:847-883
  MGTwoLevelTransfer.prolongate_and_add / restrict_and_add -> multigrid_throughput.cc:1600-1604
  Problem (JSON) .run / .solve_once / .vcycle       -> run() + solve_with_global_coarsening + mg_solve
All compute happens in libmgb200.so on the GPU; errors surface as MgbError (the reference throws C++
exceptions caught in main(), multigrid_throughput.cc:2444-2468).
"""
import ctypes as C
import json

import numpy as np

from . import _lib

DOUBLE, FLOAT = 0, 1
FLAG_BIT = 0x80000000
INDEX_MASK = 0x7FFFFFFF
_NP = {DOUBLE: np.float64, FLOAT: np.float32}


class MgbError(RuntimeError):
    pass


def _check(rc):
    if rc != 0:
        raise MgbError(_lib.load().mgb_last_error().decode())


def _arr(ptr, n, dtype):
    """copy n entries from a borrowed C pointer into a numpy array"""
    if n == 0:
        return np.zeros(0, dtype=dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True)


def init(device=0):
    _check(_lib.load().mgb_init(device))


def set_stream(cuda_stream_ptr):
    _check(_lib.load().mgb_set_stream(C.c_void_p(cuda_stream_ptr)))


def synchronize():
    _check(_lib.load().mgb_synchronize())


def launch_count():
    return int(_lib.load().mgb_launch_count())


def profile_start(n_cells):
    _check(_lib.load().mgb_profile_start(int(n_cells)))


def profile_stop():
    """-> (total milliseconds, number of launches) of the profiled cell-loop kernel"""
    ms, n = C.c_double(0), C.c_int64(0)
    _check(_lib.load().mgb_profile_stop(C.byref(ms), C.byref(n)))
    return ms.value, n.value


def comm_unique_id(libnccl_path=None):
    buf = C.create_string_buffer(128)
    _check(_lib.load().mgb_comm_unique_id(buf, libnccl_path.encode() if libnccl_path else None))
    return buf.raw


def comm_init(unique_id, rank, size, libnccl_path=None):
    buf = C.create_string_buffer(bytes(unique_id), 128)
    _check(_lib.load().mgb_comm_init(buf, rank, size, libnccl_path.encode() if libnccl_path else None))


def comm_enable_peer_exchange():
    """collective: back the partitioners created so far with peer-mapped staging memory (NVLink ghost exchange)"""
    on = C.c_int(0)
    _check(_lib.load().mgb_comm_enable_peer_exchange(C.byref(on)))
    return bool(on.value)


def comm_finalize():
    _check(_lib.load().mgb_comm_finalize())


def fe_1d(degree):
    """1-D FE_Q(degree)/QGauss(degree+1) data of the product (for comparison with the oracle's own)."""
    n = degree + 1
    out = {k: np.zeros(s) for k, s in (("nodes", n), ("M", n * n), ("K", n * n), ("S0", n * n), ("S1", n * n),
                                       ("Ph", (2 * degree + 1) * n))}
    ptr = lambda a: a.ctypes.data_as(_lib.c_f64p)
    _check(_lib.load().mgb_fe_1d(degree, ptr(out["nodes"]), ptr(out["M"]), ptr(out["K"]), ptr(out["S0"]), ptr(out["S1"]),
                                 ptr(out["Ph"])))
    for k in ("M", "K", "S0", "S1"):
        out[k] = out[k].reshape(n, n)
    out["Ph"] = out["Ph"].reshape(2 * degree + 1, n)
    return out


def fe_1d_fdm(degree):
    """fast-diagonalisation data of the cell kernels: (Tm, lambda) with K = Tm^T diag(lambda) Tm, M = Tm^T Tm"""
    n = degree + 1
    Tm, lam = np.zeros(n * n), np.zeros(n)
    _check(_lib.load().mgb_fe_1d_fdm(degree, Tm.ctypes.data_as(_lib.c_f64p), lam.ctypes.data_as(_lib.c_f64p)))
    return Tm.reshape(n, n), lam


def fe_p_embedding(degree_fine, degree_coarse):
    P = np.zeros((degree_fine + 1) * (degree_coarse + 1))
    _check(_lib.load().mgb_fe_p_embedding(degree_fine, degree_coarse, P.ctypes.data_as(_lib.c_f64p)))
    return P.reshape(degree_fine + 1, degree_coarse + 1)


class Mesh:
    """Octree stand-in for parallel::distributed::Triangulation (multigrid_throughput.cc:2041-2062)."""

    def __init__(self, geometry_type=None, n_ref_global=0, n_ref_local=0, _handle=None):
        self._lib = _lib.load()
        if _handle is not None:
            self._h = _handle
        else:
            h = C.c_void_p()
            _check(self._lib.mgb_mesh_create(geometry_type.encode(), n_ref_global, n_ref_local, C.byref(h)))
            self._h = h

    def coarsen_global(self):
        h = C.c_void_p()
        _check(self._lib.mgb_mesh_coarsen(self._h, C.byref(h)))
        return Mesh(_handle=h)

    def level_mesh(self, level):
        """local-smoothing level: all cells of that refinement level, active or not (multigrid_throughput.cc:2044-2046)"""
        h = C.c_void_p()
        _check(self._lib.mgb_mesh_level(self._h, level, C.byref(h)))
        return Mesh(_handle=h)

    @property
    def n_cells(self):
        return int(self._lib.mgb_mesh_n_cells(self._h))

    @property
    def n_global_levels(self):
        return int(self._lib.mgb_mesh_n_levels(self._h))

    def cells(self):
        """(n_cells, 4) array of (level, x, y, z) in Morton order"""
        out = np.zeros((self.n_cells, 4), dtype=np.uint32)
        _check(self._lib.mgb_mesh_get_cells(self._h, out.ctypes.data_as(_lib.c_u32p)))
        return out

    def hanging_cells(self):
        """per leaf: does the cell carry a hanging-node constraint (Helper::is_constrained, multigrid_throughput.cc:231-268)"""
        out = np.zeros(self.n_cells, dtype=np.uint8)
        _check(self._lib.mgb_mesh_hanging_cells(self._h, out.ctypes.data_as(C.POINTER(C.c_uint8))))
        return out.astype(bool)

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.mgb_mesh_destroy(self._h)
            self._h = None


def partition_sequence(policy_name, meshes, n_ranks, min_cells_per_rank=1):
    """Morton ranges of every rank on every mesh of a coarse -> fine sequence under one of the reference's
    `PartitionerName` policies (multigrid_throughput.cc:2066-2174); returns an (n_meshes, n_ranks+1) int64 array"""
    lib = _lib.load()
    hs = (C.c_void_p * len(meshes))(*[m._h for m in meshes])
    out = np.zeros((len(meshes), n_ranks + 1), dtype=np.int64)
    _check(lib.mgb_partition_sequence(policy_name.encode(), hs, len(meshes), n_ranks, min_cells_per_rank,
                                      out.ctypes.data_as(C.POINTER(C.c_int64))))
    return out


_LEVEL_COLUMNS = ("n_cells_min", "n_cells_max", "n_cells_sum", "vertical_local", "vertical_remote", "horizontal_local",
                  "horizontal_remote")


def _statistics_dict(st, per_level):
    out = {k: getattr(st, k) for k in ("workload_eff", "workload_path_max", "vertical_eff", "horizontal_eff", "mem_total")}
    out["levels"] = [dict(zip(_LEVEL_COLUMNS, map(int, row))) for row in per_level]
    return out


def hierarchy_statistics(meshes, cell_begin, n_ranks):
    """MGTools::print_multigrid_statistics (include/mg_tools.h:267-513) of a partitioned hierarchy: meshes coarse -> fine
    (a coarsening sequence, or the level meshes of local smoothing), cell_begin from partition_sequence"""
    lib = _lib.load()
    hs = (C.c_void_p * len(meshes))(*[m._h for m in meshes])
    cb = np.ascontiguousarray(cell_begin, dtype=np.int64).reshape(len(meshes), n_ranks + 1)
    st, per = _lib.Statistics(), np.zeros((len(meshes), len(_LEVEL_COLUMNS)), dtype=np.int64)
    _check(lib.mgb_hierarchy_statistics(hs, len(meshes), n_ranks, cb.ctypes.data_as(C.POINTER(C.c_int64)), C.byref(st),
                                        per.ctypes.data_as(C.POINTER(C.c_int64))))
    return _statistics_dict(st, per)


def first_child_boundaries(parents, children, child_begin):
    """ownership with the first child (FirstChildPolicy, level cells of local smoothing)"""
    lib = _lib.load()
    cb = np.ascontiguousarray(child_begin, dtype=np.int64)
    out = np.zeros_like(cb)
    _check(lib.mgb_partition_first_child(parents._h, children._h, len(cb) - 1, cb.ctypes.data_as(C.POINTER(C.c_int64)),
                                         out.ctypes.data_as(C.POINTER(C.c_int64))))
    return out


class DoFMap:
    """FE_Q(degree) DoFs + constraints + partitioner (DoFHandler/AffineConstraints/MatrixFree stand-in)."""

    def __init__(self, mesh=None, degree=1, n_ranks=1, rank=0, cell_begin=None, _handle=None, _owner=None):
        self._lib = _lib.load()
        self._owner = _owner  # keeps a Problem alive for borrowed handles
        if _handle is not None:
            self._h = _handle
            self._borrowed = True
        elif cell_begin is not None:  # explicit partition (partition_sequence / first_child_boundaries)
            cb = np.ascontiguousarray(cell_begin, dtype=np.int64)
            if cb.shape != (n_ranks + 1,):
                raise ValueError("cell_begin must hold n_ranks+1 offsets")
            h = C.c_void_p()
            _check(self._lib.mgb_dofmap_create_partitioned(mesh._h, degree, n_ranks, rank, cb.ctypes.data_as(C.POINTER(C.c_int64)),
                                                           C.byref(h)))
            self._h = h
            self._borrowed = False
        else:
            h = C.c_void_p()
            _check(self._lib.mgb_dofmap_create(mesh._h, degree, n_ranks, rank, C.byref(h)))
            self._h = h
            self._borrowed = False
        v = _lib.DofmapView()
        _check(self._lib.mgb_dofmap_view_get(self._h, C.byref(v)))
        self.view = v
        for name in ("degree", "n_ranks", "rank", "mesh_max_level", "n_cells_global", "cell_begin", "n_cells",
                     "n_cells_hanging", "n_dofs_global", "owned_begin", "n_owned", "n_ghost", "n_constrained"):
            setattr(self, name, int(getattr(v, name)))

    # numpy copies of the maps (inputs shared by the product and the oracle)
    def eval_table(self):
        return _arr(self.view.eval_table, self.n_cells * 27, np.uint32).reshape(self.n_cells, 27)

    def own_table(self):
        return _arr(self.view.own_table, self.n_cells * 27, np.uint32).reshape(self.n_cells, 27)

    def mask(self):
        return _arr(self.view.mask, self.n_cells, np.uint16)

    def cell_level(self):
        return _arr(self.view.cell_level, self.n_cells, np.uint8)

    def cell_xyz(self):
        return _arr(self.view.cell_xyz, self.n_cells * 3, np.uint32).reshape(self.n_cells, 3)

    def constrained(self):
        return _arr(self.view.constrained, self.n_constrained, np.uint32)

    def ghost_global(self):
        return _arr(self.view.ghost_global, self.n_ghost, np.int64)

    def partitioner(self):
        v = self.view
        return dict(
            n_global=self.n_dofs_global, owned_begin=self.owned_begin, n_owned=self.n_owned, n_ghost=self.n_ghost,
            ghost_global=self.ghost_global(),
            ghost_ranks=_arr(v.ghost_ranks, v.n_ghost_ranks, np.int32),
            ghost_offsets=_arr(v.ghost_offsets, v.n_ghost_ranks + 1, np.uint32),
            import_ranks=_arr(v.import_ranks, v.n_import_ranks, np.int32),
            import_offsets=_arr(v.import_offsets, v.n_import_ranks + 1, np.uint32),
            import_indices=_arr(v.import_indices, v.n_import_indices, np.uint32))

    def expand(self, which="eval"):
        """lexicographic (p+1)^3 index map per cell, flag in bit 31 (MatrixFree's dof_indices)"""
        n3 = (self.degree + 1) ** 3
        out = np.zeros((self.n_cells, n3), dtype=np.uint32)
        _check(self._lib.mgb_dofmap_expand(self._h, 0 if which == "eval" else 1, out.ctypes.data_as(_lib.c_u32p)))
        return out

    def assemble_rhs(self, mesh, simulation_type="Constant"):
        """Operator::rhs on the host (n_owned + n_ghost entries, ghost contributions not compressed)"""
        out = np.zeros(self.n_owned + self.n_ghost)
        _check(self._lib.mgb_assemble_rhs(mesh._h, self._h, simulation_type.encode(), out.ctypes.data_as(C.POINTER(C.c_double))))
        return out

    def chunk_plan_check(self, value_bytes=8):
        """host-only check of the cell loop's chunk plan (host/chunks.h): dict with the deviation of the CPU walk through
        the plan from the direct scatter-add and the plan's statistics"""
        st = (C.c_double * 11)()
        _check(self._lib.mgb_dofmap_chunk_plan_check(self._h, value_bytes, st))
        keys = ("max_err", "n_chunks", "n_compute_tables", "n_pull_tables", "table_words", "foreign_low", "foreign_late",
                "foreign_ghost", "side_entries", "max_slots", "build_seconds")
        return dict(zip(keys, [float(x) for x in st]))

    def __del__(self):
        if getattr(self, "_h", None) and not self._borrowed:
            self._lib.mgb_dofmap_destroy(self._h)
        self._h = None


class LevelEdges:
    """refinement-edge DoFs and cells of one local-smoothing level (include/operator.h:49-99)"""

    def __init__(self, level_mesh, level_dofs):
        self._lib = _lib.load()
        h = C.c_void_p()
        _check(self._lib.mgb_level_edges_create(level_mesh._h, level_dofs._h, C.byref(h)))
        self._h = h
        v = _lib.LevelEdgesView()
        _check(self._lib.mgb_level_edges_view_get(self._h, C.byref(v)))
        self.edge_dofs = _arr(v.edge_dofs, v.n_edge_dofs, np.uint32)
        self.edge_cells = _arr(v.edge_cells, v.n_edge_cells, np.uint32)
        self.any_global = bool(v.any_global)

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.mgb_level_edges_destroy(self._h)
            self._h = None


class CopyIndices:
    """copy_indices[level] of MGTransferMatrixFree: (active-vector index, level-vector index) pairs"""

    def __init__(self, active_dofs, level_mesh, level_dofs, active_mesh=None):
        lib = _lib.load()
        h = C.c_void_p()
        _check(lib.mgb_copy_indices_create(active_dofs._h, active_mesh._h if active_mesh is not None else None, level_mesh._h,
                                           level_dofs._h, C.byref(h)))
        v = _lib.CopyIndicesView()
        _check(lib.mgb_copy_indices_view_get(h, C.byref(v)))
        self.active_idx = _arr(v.active_idx, v.n, np.uint32)
        self.level_idx = _arr(v.level_idx, v.n, np.uint32)
        self.distributed = bool(v.distributed)
        self.owned_begin, self.n_owned, self.n_ghost = int(v.owned_begin), int(v.n_owned), int(v.n_ghost)
        self.ghost_global = _arr(v.ghost_global, v.n_ghost, np.int64)
        self._lib, self._h = lib, h

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.mgb_copy_indices_destroy(self._h)
        self._h = None


class TransferMap:
    def __init__(self, mesh_fine, dofs_fine, mesh_coarse, dofs_coarse, local_smoothing=False):
        self._lib = _lib.load()
        h = C.c_void_p()
        create = self._lib.mgb_transfer_map_create_ls if local_smoothing else self._lib.mgb_transfer_map_create
        _check(create(mesh_fine._h, dofs_fine._h, mesh_coarse._h, dofs_coarse._h, C.byref(h)))
        self._h = h
        v = _lib.TransferMapView()
        _check(self._lib.mgb_transfer_map_view_get(self._h, C.byref(v)))
        self.view = v
        self.degree_fine, self.degree_coarse = int(v.degree_fine), int(v.degree_coarse)
        self.n_coarse_cells, self.n_fine_cells = int(v.n_coarse_cells), int(v.n_fine_cells)
        self.n_fine_slots, self.distributed = int(v.n_fine_slots), bool(v.distributed)

    def fine_own(self):
        """fine-side own table indexed by fine_first (+child); several ranks: numbering of partitioner()"""
        return _arr(self.view.fine_own, self.n_fine_slots * 27, np.uint32).reshape(self.n_fine_slots, 27)

    def partitioner(self):
        """layout of the transfer's internal ghosted fine vector (several ranks only)"""
        v = self.view
        return dict(
            owned_begin=int(v.owned_begin), n_owned=int(v.n_owned), n_ghost=int(v.n_ghost),
            ghost_global=_arr(v.ghost_global, v.n_ghost, np.int64),
            ghost_ranks=_arr(v.ghost_ranks, v.n_ghost_ranks, np.int32),
            ghost_offsets=_arr(v.ghost_offsets, v.n_ghost_ranks + 1, np.uint32) if v.distributed else np.zeros(1, np.uint32),
            import_ranks=_arr(v.import_ranks, v.n_import_ranks, np.int32),
            import_offsets=_arr(v.import_offsets, v.n_import_ranks + 1, np.uint32) if v.distributed else np.zeros(1, np.uint32),
            import_indices=_arr(v.import_indices, v.n_import_indices, np.uint32))

    def fine_first(self):
        return _arr(self.view.fine_first, self.n_coarse_cells, np.uint32)

    def n_children(self):
        return _arr(self.view.n_children, self.n_coarse_cells, np.uint8)

    def weight_code(self):
        return _arr(self.view.weight_code, self.n_fine_slots * 27, np.uint8).reshape(self.n_fine_slots, 27)

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.mgb_transfer_map_destroy(self._h)
            self._h = None


class Vector:
    """LinearAlgebra::distributed::Vector<Number> in device memory, [owned | ghost]."""

    def __init__(self, n_owned=0, n_ghost=0, number_type=DOUBLE, _handle=None, _borrowed=False, _owner=None):
        self._lib = _lib.load()
        self.number_type = number_type
        self._owner = _owner
        self._borrowed = _borrowed
        if _handle is not None:
            self._h = _handle
        else:
            h = C.c_void_p()
            _check(self._lib.mgb_vector_create(n_owned, n_ghost, number_type, C.byref(h)))
            self._h = h
        self.n_owned = int(self._lib.mgb_vector_n_owned(self._h))

    @classmethod
    def wrap(cls, device_ptr, n_owned, n_ghost=0, number_type=DOUBLE, keepalive=None):
        lib = _lib.load()
        h = C.c_void_p()
        _check(lib.mgb_vector_wrap(C.c_void_p(device_ptr), n_owned, n_ghost, number_type, C.byref(h)))
        return cls(number_type=number_type, _handle=h, _owner=keepalive)

    @property
    def data_ptr(self):
        return int(self._lib.mgb_vector_data(self._h))

    def from_numpy(self, a):
        a = np.ascontiguousarray(a, dtype=_NP[self.number_type])
        _check(self._lib.mgb_vector_copy_from_host(self._h, a.ctypes.data_as(C.c_void_p), a.size))
        return self

    def to_numpy(self, n=None):
        n = self.n_owned if n is None else n
        out = np.zeros(n, dtype=_NP[self.number_type])
        _check(self._lib.mgb_vector_copy_to_host(self._h, out.ctypes.data_as(C.c_void_p), n))
        return out

    def set(self, a):
        _check(self._lib.mgb_vector_set(self._h, a))

    def add(self, a, x):
        _check(self._lib.mgb_vector_add(self._h, a, x._h))

    def sadd(self, s, a, x):
        _check(self._lib.mgb_vector_sadd(self._h, s, a, x._h))

    def equ(self, a, x):
        _check(self._lib.mgb_vector_equ(self._h, a, x._h))

    def dot(self, y):
        r = C.c_double()
        _check(self._lib.mgb_vector_dot(self._h, y._h, C.byref(r)))
        return r.value

    def l2_norm(self):
        r = C.c_double()
        _check(self._lib.mgb_vector_l2_norm(self._h, C.byref(r)))
        return r.value

    def add_and_dot(self, a, h):
        r = C.c_double()
        _check(self._lib.mgb_vector_add_and_dot(self._h, a, h._h, C.byref(r)))
        return r.value

    def update_ghost_values(self):
        _check(self._lib.mgb_vector_update_ghost_values(self._h))

    def compress_add(self):
        _check(self._lib.mgb_vector_compress_add(self._h))

    def zero_out_ghost_values(self):
        _check(self._lib.mgb_vector_zero_out_ghost_values(self._h))

    def __del__(self):
        if getattr(self, "_h", None) and not self._borrowed:
            self._lib.mgb_vector_destroy(self._h)
        self._h = None


def _ptr(a, ctype):
    return a.ctypes.data_as(C.POINTER(ctype)) if a is not None and a.size else None


class Operator:
    """Operator<3,1,Number> of include/operator.h (matrix-free Laplace with identity rows on constrained DoFs)."""

    def __init__(self, dofs=None, number_type=DOUBLE, _handle=None, _owner=None, edges=None, **arrays):
        self._lib = _lib.load()
        self.number_type = number_type
        self._owner = _owner
        self._borrowed = _handle is not None
        if _handle is not None:
            self._h = _handle
            return
        h = C.c_void_p()
        if dofs is not None and edges is not None:
            # level operator of the local-smoothing hierarchy (Operator::reinit with mg_level)
            _check(self._lib.mgb_operator_create_level(dofs._h, edges._h, number_type, C.byref(h)))
        elif dofs is not None and not arrays:
            _check(self._lib.mgb_operator_create_from_dofmap(dofs._h, number_type, C.byref(h)))
        else:
            _check(self._lib.mgb_operator_create(C.byref(self._make_desc(number_type, **arrays)), C.byref(h)))
        self._h = h

    def _make_desc(self, number_type, degree, cell_level, n_owned, entity_table=None, dof_indices=None, mask=None,
                   constrained=None, edge_constrained=None, has_edge_constrained_global=False, edge_cells=None,
                   n_dofs_global=None, partitioner=None, root_length=2.0):
        d = _lib.OperatorDesc()
        keep = []

        def arr(a, dtype):
            if a is None:
                return None
            a = np.ascontiguousarray(a, dtype=dtype)
            keep.append(a)
            return a

        et, di = arr(entity_table, np.uint32), arr(dof_indices, np.uint32)
        lv, mk = arr(cell_level, np.uint8), arr(mask, np.uint16)
        cs, ec, ecl = arr(constrained, np.uint32), arr(edge_constrained, np.uint32), arr(edge_cells, np.uint32)
        d.degree, d.number_type, d.n_cells = degree, number_type, lv.size
        d.entity_table, d.dof_indices = _ptr(et, C.c_uint32), _ptr(di, C.c_uint32)
        d.mask, d.cell_level, d.root_length = _ptr(mk, C.c_uint16), _ptr(lv, C.c_uint8), root_length
        d.constrained, d.n_constrained = _ptr(cs, C.c_uint32), 0 if cs is None else cs.size
        d.edge_constrained, d.n_edge_constrained = _ptr(ec, C.c_uint32), 0 if ec is None else ec.size
        d.has_edge_constrained_global = int(has_edge_constrained_global)
        d.edge_cells, d.n_edge_cells = _ptr(ecl, C.c_uint32), 0 if ecl is None else ecl.size
        p = partitioner or {}
        d.n_owned, d.n_ghost = n_owned, p.get("n_ghost", 0)
        d.n_dofs_global = n_dofs_global if n_dofs_global is not None else n_owned
        d.owned_begin = p.get("owned_begin", 0)
        gg, gr, go = arr(p.get("ghost_global"), np.int64), arr(p.get("ghost_ranks"), np.int32), arr(p.get("ghost_offsets"), np.uint32)
        ir, io, ii = arr(p.get("import_ranks"), np.int32), arr(p.get("import_offsets"), np.uint32), arr(p.get("import_indices"), np.uint32)
        d.ghost_global = _ptr(gg, C.c_int64)
        d.ghost_ranks, d.ghost_offsets, d.n_ghost_ranks = _ptr(gr, C.c_int32), _ptr(go, C.c_uint32), 0 if gr is None else gr.size
        d.import_ranks, d.import_offsets, d.n_import_ranks = _ptr(ir, C.c_int32), _ptr(io, C.c_uint32), 0 if ir is None else ir.size
        d.import_indices = _ptr(ii, C.c_uint32)
        self._keep = keep
        return d

    def m(self):
        return int(self._lib.mgb_operator_m(self._h))

    def initialize_dof_vector(self):
        h = C.c_void_p()
        _check(self._lib.mgb_operator_initialize_dof_vector(self._h, C.byref(h)))
        return Vector(number_type=self.number_type, _handle=h)

    def vmult(self, dst, src):
        _check(self._lib.mgb_operator_vmult(self._h, dst._h, src._h))

    def Tvmult(self, dst, src):
        _check(self._lib.mgb_operator_tvmult(self._h, dst._h, src._h))

    def vmult_interface_down(self, dst, src):
        _check(self._lib.mgb_operator_vmult_interface_down(self._h, dst._h, src._h))

    def vmult_interface_up(self, dst, src):
        _check(self._lib.mgb_operator_vmult_interface_up(self._h, dst._h, src._h))

    def compute_inverse_diagonal(self, diag):
        _check(self._lib.mgb_operator_compute_inverse_diagonal(self._h, diag._h))

    def fused_available(self):
        """True if the operator runs the chunked cell loop (created with MGB_CHUNKED=1) and can fuse vector updates"""
        return bool(self._lib.mgb_operator_fused_available(self._h))

    def vmult_fused(self, out, x, mode, b=None, dinv=None, xold=None, f1=0.0, f2=0.0):
        """cell loop with the vector update in the same pass: mode 0 out = A x, 1 out = b - A x, 2 / 3 the Chebyshev updates"""
        h = lambda v: v._h if v is not None else None
        _check(self._lib.mgb_operator_vmult_fused(self._h, out._h, x._h, mode, h(b), h(dinv), h(xold), f1, f2))

    def set_cell_vertices(self, xyz):
        """general (non-Cartesian) cells: (n_cells, 8, 3) vertex coordinates, vertices lexicographic with x fastest, trilinear
        geometry (MappingQ1); every later cell loop evaluates the metric at the Gauss points from them"""
        a = np.ascontiguousarray(xyz, dtype=np.float64)
        if a.ndim != 3 or a.shape[1:] != (8, 3):
            raise ValueError("xyz must have shape (n_cells, 8, 3)")
        _check(self._lib.mgb_operator_set_cell_vertices(self._h, a.ctypes.data_as(_lib.c_f64p), a.shape[0]))

    def __del__(self):
        if getattr(self, "_h", None) and not self._borrowed:
            self._lib.mgb_operator_destroy(self._h)
        self._h = None


class PreconditionChebyshev:
    """PreconditionChebyshev<Operator, Vector, DiagonalMatrix> (multigrid_throughput.cc:847-883)."""

    def __init__(self, op=None, degree=3, smoothing_range=20.0, eig_cg_n_iterations=20, _handle=None, _owner=None):
        self._lib = _lib.load()
        self._owner = _owner if _owner is not None else op
        self._borrowed = _handle is not None
        if _handle is not None:
            self._h = _handle
        else:
            h = C.c_void_p()
            _check(self._lib.mgb_chebyshev_create(op._h, degree, smoothing_range, eig_cg_n_iterations, C.byref(h)))
            self._h = h

    def vmult(self, dst, rhs):
        _check(self._lib.mgb_chebyshev_vmult(self._h, dst._h, rhs._h))

    def step(self, dst, rhs):
        _check(self._lib.mgb_chebyshev_step(self._h, dst._h, rhs._h))

    def eigenvalues(self):
        a, b = C.c_double(), C.c_double()
        _check(self._lib.mgb_chebyshev_eigenvalues(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def __del__(self):
        if getattr(self, "_h", None) and not self._borrowed:
            self._lib.mgb_chebyshev_destroy(self._h)
        self._h = None


class MGTwoLevelTransfer:
    """MGTwoLevelTransfer<3, Vector>: reinit(dof_fine, dof_coarse, constraints...) + prolongate/restrict."""

    def __init__(self, dofs_fine=None, dofs_coarse=None, tmap=None, op_fine=None, op_coarse=None, _handle=None, _owner=None):
        self._lib = _lib.load()
        self._owner = _owner if _owner is not None else (dofs_fine, dofs_coarse, tmap, op_fine, op_coarse)
        self._borrowed = _handle is not None
        if _handle is not None:
            self._h = _handle
        elif dofs_fine is None:   # plain arrays: the view of the transfer map is passed as the adapter would fill it
            h = C.c_void_p()
            _check(self._lib.mgb_transfer_create_from_arrays(C.byref(tmap.view), op_fine._h, op_coarse._h, C.byref(h)))
            self._h = h
        else:
            h = C.c_void_p()
            _check(self._lib.mgb_transfer_create(dofs_fine._h, dofs_coarse._h, tmap._h, op_fine._h, op_coarse._h, C.byref(h)))
            self._h = h

    def prolongate_and_add(self, dst_fine, src_coarse):
        _check(self._lib.mgb_transfer_prolongate_and_add(self._h, dst_fine._h, src_coarse._h))

    def restrict_and_add(self, dst_coarse, src_fine):
        _check(self._lib.mgb_transfer_restrict_and_add(self._h, dst_coarse._h, src_fine._h))

    def prolongate(self, dst_fine, src_coarse):
        """MGTwoLevelTransfer::prolongate: dst = P src (no accumulation)"""
        _check(self._lib.mgb_transfer_prolongate(self._h, dst_fine._h, src_coarse._h))

    def __del__(self):
        if getattr(self, "_h", None) and not self._borrowed:
            self._lib.mgb_transfer_destroy(self._h)
        self._h = None


NO_ALLOC_MAX_N = 16777216  # MGB_NO_ALLOC_MAX_N of include/mgb200.h


def dot_no_alloc(x_ptr, y_ptr, n, result_ptr, stream_ptr=None):
    """mgb_vector_dot_no_alloc: *result_ptr = x . y for FP64 device arrays given as raw device addresses; asynchronous on the
    stream (None: the library's stream); allocates nothing -- the reduction's scratch is a static __device__ array"""
    _check(_lib.load().mgb_vector_dot_no_alloc(C.c_void_p(x_ptr), C.c_void_p(y_ptr), n, C.c_void_p(result_ptr),
                                               C.c_void_p(stream_ptr) if stream_ptr else None))


class LevelCopy:
    """device copy of copy_indices[level]: MGLevelGlobalTransfer::copy_to_mg / copy_from_mg for one level"""

    def __init__(self, copy_indices, level_op, active_idx=None, level_idx=None):
        """copy_indices: a CopyIndices object, or None with the (active, level) local index pairs given as arrays (what a
        deal.II adapter extracts from MGLevelGlobalTransfer::copy_indices)"""
        self._lib = _lib.load()
        self._owner = (copy_indices, level_op)
        h = C.c_void_p()
        if copy_indices is not None:
            _check(self._lib.mgb_level_copy_create(copy_indices._h, level_op._h, C.byref(h)))
        else:
            a = np.ascontiguousarray(active_idx, dtype=np.uint32)
            l = np.ascontiguousarray(level_idx, dtype=np.uint32)
            if a.shape != l.shape or a.ndim != 1:
                raise ValueError("active_idx and level_idx must be 1-D arrays of equal length")
            _check(self._lib.mgb_level_copy_create_from_arrays(len(a), a.ctypes.data_as(_lib.c_u32p), l.ctypes.data_as(_lib.c_u32p),
                                                               level_op._h, C.byref(h)))
        self._h = h

    def copy_to_mg(self, dst_level, src_active):
        _check(self._lib.mgb_level_copy_to_mg(self._h, dst_level._h, src_active._h))

    def copy_from_mg(self, dst_active, src_level):
        _check(self._lib.mgb_level_copy_from_mg(self._h, dst_active._h, src_level._h))

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.mgb_level_copy_destroy(self._h)
        self._h = None


class Multigrid:
    """Multigrid + PreconditionMG over borrowed level operators, smoothers and two-level transfers (mg_solve,
    multigrid_throughput.cc:1082-1133).  operators / smoothers / transfers: lists coarse -> fine, entry 0 of smoothers and
    transfers ignored; coarse_precond: a PreconditionChebyshev on level 0 (cg_with_chebyshev) or None (cg)."""

    N_STAGES = 7

    def __init__(self, operators, smoothers, transfers, coarse_precond=None, coarse_maxiter=10000, coarse_abstol=1e-20,
                 coarse_reltol=1e-4, edge_in=False, level_copies=None):
        self._lib = _lib.load()
        self._owner = (operators, smoothers, transfers, coarse_precond, level_copies)
        n = len(operators)
        P = C.c_void_p
        arr = lambda objs: (P * n)(*[(o._h if o is not None else None) for o in objs])
        h = P()
        _check(self._lib.mgb_multigrid_create(n, arr(operators), arr(smoothers), arr(transfers),
                                              coarse_precond._h if coarse_precond is not None else None, coarse_maxiter,
                                              coarse_abstol, coarse_reltol, C.byref(h)))
        self._h, self.n_levels = h, n
        if edge_in:
            _check(self._lib.mgb_multigrid_set_edge_in(self._h, 1))
        if level_copies is not None:
            _check(self._lib.mgb_multigrid_set_level_copies(self._h, arr(level_copies)))

    def vmult(self, dst, src):
        """PreconditionMG::vmult"""
        _check(self._lib.mgb_multigrid_vcycle(self._h, dst._h, src._h))

    def enable_timers(self, on=True):
        _check(self._lib.mgb_multigrid_enable_timers(self._h, 1 if on else 0))

    def collect_times(self):
        out = np.zeros((self.n_levels, self.N_STAGES))
        _check(self._lib.mgb_multigrid_collect_times(self._h, out.ctypes.data_as(C.POINTER(C.c_double))))
        return out

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.mgb_multigrid_destroy(self._h)
        self._h = None


def cg_solve(A, x, b, preconditioner=None, maxiter=10000, abstol=1e-20, reltol=1e-4):
    """SolverCG(ReductionControl(maxiter, abstol, reltol)).solve(A, x, b, preconditioner) from x = 0; returns (n_iterations, residual)"""
    its, res = C.c_uint(0), C.c_double(0.0)
    _check(_lib.load().mgb_cg_solve(A._h, preconditioner._h if preconditioner is not None else None, x._h, b._h, maxiter, abstol, reltol,
                                    C.byref(its), C.byref(res)))
    return int(its.value), float(res.value)


class Problem:
    """One JSON input of multigrid_throughput (scripts/default.json schema): setup + V-cycle-preconditioned CG."""

    def __init__(self, params, n_ranks=1, rank=0):
        self._lib = _lib.load()
        text = params if isinstance(params, str) else json.dumps(params)
        h = C.c_void_p()
        _check(self._lib.mgb_problem_create(text.encode(), n_ranks, rank, C.byref(h)))
        self._h = h

    @property
    def n_levels(self):
        return int(self._lib.mgb_problem_n_levels(self._h))

    def level_n_dofs(self, l):
        return int(self._lib.mgb_problem_level_n_dofs(self._h, l))

    def level_n_cells(self, l):
        return int(self._lib.mgb_problem_level_n_cells(self._h, l))

    def level_degree(self, l):
        return int(self._lib.mgb_problem_level_degree(self._h, l))

    def statistics(self):
        """workload / vertical / horizontal efficiency of the partitioned hierarchy (the reference's Verbosity columns)"""
        st = _lib.Statistics()
        per = np.zeros((int(self._lib.mgb_problem_n_meshes(self._h)), len(_LEVEL_COLUMNS)), dtype=np.int64)
        _check(self._lib.mgb_problem_statistics(self._h, C.byref(st), per.ctypes.data_as(C.POINTER(C.c_int64))))
        return _statistics_dict(st, per)

    def level_dofmap(self, l):
        return DoFMap(_handle=C.c_void_p(self._lib.mgb_problem_level_dofmap(self._h, l)), _owner=self)

    def level_operator(self, l, number_type=DOUBLE):
        return Operator(number_type=number_type, _handle=C.c_void_p(self._lib.mgb_problem_level_operator(self._h, l)), _owner=self)

    def level_smoother(self, l):
        return PreconditionChebyshev(_handle=C.c_void_p(self._lib.mgb_problem_level_smoother(self._h, l)), _owner=self)

    def level_transfer(self, l):
        return MGTwoLevelTransfer(_handle=C.c_void_p(self._lib.mgb_problem_level_transfer(self._h, l)), _owner=self)

    def fine_operator(self):
        return Operator(number_type=DOUBLE, _handle=C.c_void_p(self._lib.mgb_problem_fine_operator(self._h)), _owner=self)

    def fine_dofmap(self):
        """active-mesh DoFs of the outer operator (for local smoothing this is not a multigrid level)"""
        return DoFMap(_handle=C.c_void_p(self._lib.mgb_problem_fine_dofmap(self._h)), _owner=self)

    def rhs(self):
        return Vector(_handle=C.c_void_p(self._lib.mgb_problem_rhs(self._h)), _borrowed=True, _owner=self)

    def solution(self):
        return Vector(_handle=C.c_void_p(self._lib.mgb_problem_solution(self._h)), _borrowed=True, _owner=self)

    def vcycle(self, dst, src):
        _check(self._lib.mgb_problem_vcycle(self._h, dst._h, src._h))

    def solve_once(self, with_stage_timers=False):
        r = _lib.Result()
        _check(self._lib.mgb_problem_solve_once(self._h, int(with_stage_timers), C.byref(r)))
        return r.as_dict()

    def solve_host(self, rhs_host_ptr, solution_host_ptr):
        """solve with HOST buffers (raw addresses of n_owned doubles each, ideally pinned): H2D, solve, D2H"""
        r = _lib.Result()
        _check(self._lib.mgb_problem_solve_host(self._h, C.c_void_p(rhs_host_ptr), C.c_void_p(solution_host_ptr), C.byref(r)))
        return r.as_dict()

    def solve_host_batch(self, rhs_host_ptrs, solution_host_ptrs):
        """len(rhs_host_ptrs) solves in a row with HOST buffers (raw addresses, pinned): the upload of the next right-hand
        side and the download of the previous solution run behind the current solve; returns the list of result dicts"""
        n = len(rhs_host_ptrs)
        if n != len(solution_host_ptrs):
            raise ValueError("one solution buffer per right-hand side")
        rp = (C.c_void_p * n)(*[C.c_void_p(a) for a in rhs_host_ptrs])
        sp = (C.c_void_p * n)(*[C.c_void_p(a) for a in solution_host_ptrs])
        res = (_lib.Result * n)()
        _check(self._lib.mgb_problem_solve_host_batch(self._h, n, rp, sp, res))
        return [r.as_dict() for r in res]

    def run(self, n_repetitions=5):
        r = _lib.Result()
        _check(self._lib.mgb_problem_run(self._h, n_repetitions, C.byref(r)))
        return r.as_dict()

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.mgb_problem_destroy(self._h)
            self._h = None
