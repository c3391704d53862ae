"""ctypes binding of libmgb200.so (include/mgb200.h).  The library is the product; this file only declares
signatures.  Loading fails loudly when the shared library is missing -- there is no Python/CPU fallback."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MGB200_LIB") or os.path.join(_HERE, "lib", "libmgb200.so")  # MGB200_LIB: kernel-variant experiments

c_u8p = C.POINTER(C.c_uint8)
c_u16p = C.POINTER(C.c_uint16)
c_u32p = C.POINTER(C.c_uint32)
c_i32p = C.POINTER(C.c_int32)
c_i64p = C.POINTER(C.c_int64)
c_f64p = C.POINTER(C.c_double)


class DofmapView(C.Structure):
    _fields_ = [
        ("degree", C.c_int32), ("n_ranks", C.c_int32), ("rank", C.c_int32), ("mesh_max_level", C.c_int32),
        ("n_cells_global", C.c_int64), ("cell_begin", C.c_int64), ("n_cells", C.c_int64), ("n_cells_hanging", C.c_int64),
        ("n_dofs_global", C.c_int64), ("owned_begin", C.c_int64),
        ("n_owned", C.c_uint32), ("n_ghost", C.c_uint32), ("n_constrained", C.c_uint32),
        ("n_ghost_ranks", C.c_uint32), ("n_import_ranks", C.c_uint32), ("n_import_indices", C.c_uint32),
        ("eval_table", c_u32p), ("own_table", c_u32p), ("mask", c_u16p), ("cell_level", c_u8p), ("cell_xyz", c_u32p),
        ("constrained", c_u32p), ("ghost_global", c_i64p), ("ghost_ranks", c_i32p), ("ghost_offsets", c_u32p),
        ("import_ranks", c_i32p), ("import_offsets", c_u32p), ("import_indices", c_u32p),
    ]


class TransferMapView(C.Structure):
    _fields_ = [
        ("degree_fine", C.c_int32), ("degree_coarse", C.c_int32),
        ("n_coarse_cells", C.c_int64), ("n_fine_cells", C.c_int64),
        ("fine_first", c_u32p), ("n_children", c_u8p), ("weight_code", c_u8p),
        ("fine_own", c_u32p), ("n_fine_slots", C.c_int64), ("distributed", C.c_int32),
        ("owned_begin", C.c_int64),
        ("n_owned", C.c_uint32), ("n_ghost", C.c_uint32), ("n_ghost_ranks", C.c_uint32), ("n_import_ranks", C.c_uint32),
        ("n_import_indices", C.c_uint32),
        ("ghost_global", c_i64p), ("ghost_ranks", c_i32p), ("ghost_offsets", c_u32p),
        ("import_ranks", c_i32p), ("import_offsets", c_u32p), ("import_indices", c_u32p),
    ]


class LevelEdgesView(C.Structure):
    _fields_ = [("n_edge_dofs", C.c_uint32), ("n_edge_cells", C.c_uint32), ("any_global", C.c_int32),
                ("edge_dofs", c_u32p), ("edge_cells", c_u32p)]


class CopyIndicesView(C.Structure):
    _fields_ = [("n", C.c_uint32), ("active_idx", c_u32p), ("level_idx", c_u32p), ("distributed", C.c_int32),
                ("owned_begin", C.c_int64), ("n_owned", C.c_uint32), ("n_ghost", C.c_uint32), ("ghost_global", c_i64p)]


class OperatorDesc(C.Structure):
    _fields_ = [
        ("degree", C.c_int32), ("number_type", C.c_int32),
        ("n_cells", C.c_int64),
        ("entity_table", c_u32p), ("dof_indices", c_u32p), ("mask", c_u16p), ("cell_level", c_u8p),
        ("root_length", C.c_double),
        ("constrained", c_u32p), ("n_constrained", C.c_uint32),
        ("edge_constrained", c_u32p), ("n_edge_constrained", C.c_uint32),
        ("has_edge_constrained_global", C.c_int32),
        ("edge_cells", c_u32p), ("n_edge_cells", C.c_uint32),
        ("n_dofs_global", C.c_int64),
        ("owned_begin", C.c_int64),
        ("n_owned", C.c_uint32), ("n_ghost", C.c_uint32),
        ("ghost_global", c_i64p),
        ("ghost_ranks", c_i32p), ("ghost_offsets", c_u32p), ("n_ghost_ranks", C.c_uint32),
        ("import_ranks", c_i32p), ("import_offsets", c_u32p), ("n_import_ranks", C.c_uint32),
        ("import_indices", c_u32p),
    ]


class Statistics(C.Structure):
    _fields_ = [("workload_eff", C.c_double), ("workload_path_max", C.c_double), ("vertical_eff", C.c_double),
                ("horizontal_eff", C.c_double), ("mem_total", C.c_double), ("verbose", C.c_int32)]


class Result(C.Structure):
    _fields_ = [
        ("dim", C.c_int32), ("degree", C.c_int32), ("n_ref_global", C.c_int32), ("n_ref_local", C.c_int32),
        ("n_levels", C.c_int32), ("n_iterations", C.c_int32), ("sub_comm_size", C.c_int32),
        ("n_cells", C.c_int64), ("n_cells_hn", C.c_int64), ("n_cells_n", C.c_int64), ("n_dofs", C.c_int64),
        ("time", C.c_double), ("time_cg", C.c_double), ("throughput", C.c_double),
        ("time_pre", C.c_double), ("time_residuum", C.c_double), ("time_res", C.c_double), ("time_cs", C.c_double),
        ("time_pro", C.c_double), ("time_edge_pro", C.c_double), ("time_post", C.c_double),
        ("time_to_mg", C.c_double), ("time_to_global", C.c_double),
        ("initial_residual", C.c_double), ("final_residual", C.c_double),
        ("coarse_iterations", C.c_int64),
    ]

    def as_dict(self):
        return {name: getattr(self, name) for name, _ in self._fields_}


# every symbol include/mgb200.h declares: name -> (restype, argtypes)
P = C.c_void_p
PP = C.POINTER(C.c_void_p)
SIGNATURES = {
    "mgb_last_error": (C.c_char_p, []),
    "mgb_version": (C.c_char_p, []),
    "mgb_init": (C.c_int, [C.c_int]),
    "mgb_set_stream": (C.c_int, [P]),
    "mgb_synchronize": (C.c_int, []),
    "mgb_launch_count": (C.c_int64, []),
    "mgb_profile_start": (C.c_int, [C.c_int64]),
    "mgb_profile_stop": (C.c_int, [c_f64p, c_i64p]),
    "mgb_comm_unique_id": (C.c_int, [P, C.c_char_p]),
    "mgb_comm_init": (C.c_int, [P, C.c_int, C.c_int, C.c_char_p]),
    "mgb_comm_enable_peer_exchange": (C.c_int, [C.POINTER(C.c_int)]),
    "mgb_comm_finalize": (C.c_int, []),
    "mgb_fe_1d": (C.c_int, [C.c_int, c_f64p, c_f64p, c_f64p, c_f64p, c_f64p, c_f64p]),
    "mgb_fe_1d_fdm": (C.c_int, [C.c_int, c_f64p, c_f64p]),
    "mgb_fe_p_embedding": (C.c_int, [C.c_int, C.c_int, c_f64p]),
    "mgb_mesh_create": (C.c_int, [C.c_char_p, C.c_uint, C.c_uint, PP]),
    "mgb_mesh_coarsen": (C.c_int, [P, PP]),
    "mgb_mesh_level": (C.c_int, [P, C.c_int, PP]),
    "mgb_mesh_n_cells": (C.c_int64, [P]),
    "mgb_mesh_n_levels": (C.c_int, [P]),
    "mgb_mesh_get_cells": (C.c_int, [P, c_u32p]),
    "mgb_mesh_destroy": (None, [P]),
    "mgb_dofmap_create": (C.c_int, [P, C.c_int, C.c_int, C.c_int, PP]),
    "mgb_dofmap_create_partitioned": (C.c_int, [P, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int64), PP]),
    "mgb_partition_sequence": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int64, C.POINTER(C.c_int64)]),
    "mgb_partition_first_child": (C.c_int, [P, P, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "mgb_mesh_hanging_cells": (C.c_int, [P, C.POINTER(C.c_uint8)]),
    "mgb_hierarchy_statistics": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.POINTER(C.c_int64), C.POINTER(Statistics),
                                           C.POINTER(C.c_int64)]),
    "mgb_problem_n_meshes": (C.c_int, [P]),
    "mgb_problem_statistics": (C.c_int, [P, C.POINTER(Statistics), C.POINTER(C.c_int64)]),
    "mgb_dofmap_view_get": (C.c_int, [P, C.POINTER(DofmapView)]),
    "mgb_dofmap_expand": (C.c_int, [P, C.c_int, c_u32p]),
    "mgb_assemble_rhs": (C.c_int, [P, P, C.c_char_p, C.POINTER(C.c_double)]),
    "mgb_dofmap_chunk_plan_check": (C.c_int, [P, C.c_int, C.POINTER(C.c_double)]),
    "mgb_dofmap_destroy": (None, [P]),
    "mgb_transfer_map_create": (C.c_int, [P, P, P, P, PP]),
    "mgb_transfer_map_create_ls": (C.c_int, [P, P, P, P, PP]),
    "mgb_level_edges_create": (C.c_int, [P, P, PP]),
    "mgb_level_edges_view_get": (C.c_int, [P, C.POINTER(LevelEdgesView)]),
    "mgb_level_edges_destroy": (None, [P]),
    "mgb_copy_indices_create": (C.c_int, [P, P, P, P, PP]),
    "mgb_copy_indices_view_get": (C.c_int, [P, C.POINTER(CopyIndicesView)]),
    "mgb_copy_indices_destroy": (None, [P]),
    "mgb_transfer_map_view_get": (C.c_int, [P, C.POINTER(TransferMapView)]),
    "mgb_transfer_map_destroy": (None, [P]),
    "mgb_vector_create": (C.c_int, [C.c_uint32, C.c_uint32, C.c_int, PP]),
    "mgb_vector_wrap": (C.c_int, [P, C.c_uint32, C.c_uint32, C.c_int, PP]),
    "mgb_vector_destroy": (None, [P]),
    "mgb_vector_n_owned": (C.c_uint32, [P]),
    "mgb_vector_data": (P, [P]),
    "mgb_vector_copy_from_host": (C.c_int, [P, P, C.c_uint32]),
    "mgb_vector_copy_to_host": (C.c_int, [P, P, C.c_uint32]),
    "mgb_vector_set": (C.c_int, [P, C.c_double]),
    "mgb_vector_add": (C.c_int, [P, C.c_double, P]),
    "mgb_vector_sadd": (C.c_int, [P, C.c_double, C.c_double, P]),
    "mgb_vector_equ": (C.c_int, [P, C.c_double, P]),
    "mgb_vector_dot": (C.c_int, [P, P, c_f64p]),
    "mgb_vector_l2_norm": (C.c_int, [P, c_f64p]),
    "mgb_vector_add_and_dot": (C.c_int, [P, C.c_double, P, c_f64p]),
    "mgb_vector_update_ghost_values": (C.c_int, [P]),
    "mgb_vector_compress_add": (C.c_int, [P]),
    "mgb_vector_zero_out_ghost_values": (C.c_int, [P]),
    "mgb_operator_create": (C.c_int, [C.POINTER(OperatorDesc), PP]),
    "mgb_operator_create_from_dofmap": (C.c_int, [P, C.c_int, PP]),
    "mgb_operator_create_level": (C.c_int, [P, P, C.c_int, PP]),
    "mgb_operator_destroy": (None, [P]),
    "mgb_operator_m": (C.c_int64, [P]),
    "mgb_operator_initialize_dof_vector": (C.c_int, [P, PP]),
    "mgb_operator_vmult": (C.c_int, [P, P, P]),
    "mgb_operator_tvmult": (C.c_int, [P, P, P]),
    "mgb_transfer_prolongate": (C.c_int, [P, P, P]),
    "mgb_transfer_create_from_arrays": (C.c_int, [P, P, P, PP]),
    "mgb_level_copy_create": (C.c_int, [P, P, PP]),
    "mgb_level_copy_create_from_arrays": (C.c_int, [C.c_uint32, c_u32p, c_u32p, P, PP]),
    "mgb_level_copy_destroy": (None, [P]),
    "mgb_level_copy_to_mg": (C.c_int, [P, P, P]),
    "mgb_level_copy_from_mg": (C.c_int, [P, P, P]),
    "mgb_multigrid_create": (C.c_int, [C.c_int, PP, PP, PP, P, C.c_uint, C.c_double, C.c_double, PP]),
    "mgb_multigrid_destroy": (None, [P]),
    "mgb_multigrid_set_edge_in": (C.c_int, [P, C.c_int]),
    "mgb_multigrid_set_level_copies": (C.c_int, [P, PP]),
    "mgb_multigrid_vcycle": (C.c_int, [P, P, P]),
    "mgb_multigrid_enable_timers": (C.c_int, [P, C.c_int]),
    "mgb_multigrid_collect_times": (C.c_int, [P, C.POINTER(C.c_double)]),
    "mgb_cg_solve": (C.c_int, [P, P, P, P, C.c_uint, C.c_double, C.c_double, C.POINTER(C.c_uint), C.POINTER(C.c_double)]),
    "mgb_vector_dot_no_alloc": (C.c_int, [P, P, C.c_uint64, P, P]),
    "mgb_operator_fused_available": (C.c_int, [P]),
    "mgb_operator_vmult_fused": (C.c_int, [P, P, P, C.c_int, P, P, P, C.c_double, C.c_double]),
    "mgb_operator_vmult_interface_down": (C.c_int, [P, P, P]),
    "mgb_operator_vmult_interface_up": (C.c_int, [P, P, P]),
    "mgb_operator_compute_inverse_diagonal": (C.c_int, [P, P]),
    "mgb_operator_set_cell_vertices": (C.c_int, [P, c_f64p, C.c_int64]),
    "mgb_chebyshev_create": (C.c_int, [P, C.c_int, C.c_double, C.c_uint, PP]),
    "mgb_chebyshev_destroy": (None, [P]),
    "mgb_chebyshev_vmult": (C.c_int, [P, P, P]),
    "mgb_chebyshev_step": (C.c_int, [P, P, P]),
    "mgb_chebyshev_eigenvalues": (C.c_int, [P, c_f64p, c_f64p]),
    "mgb_transfer_create": (C.c_int, [P, P, P, P, P, PP]),
    "mgb_transfer_destroy": (None, [P]),
    "mgb_transfer_prolongate_and_add": (C.c_int, [P, P, P]),
    "mgb_transfer_restrict_and_add": (C.c_int, [P, P, P]),
    "mgb_problem_create": (C.c_int, [C.c_char_p, C.c_int, C.c_int, PP]),
    "mgb_problem_destroy": (None, [P]),
    "mgb_problem_n_levels": (C.c_int, [P]),
    "mgb_problem_level_n_dofs": (C.c_int64, [P, C.c_int]),
    "mgb_problem_level_n_cells": (C.c_int64, [P, C.c_int]),
    "mgb_problem_level_degree": (C.c_int, [P, C.c_int]),
    "mgb_problem_level_dofmap": (P, [P, C.c_int]),
    "mgb_problem_level_operator": (P, [P, C.c_int]),
    "mgb_problem_level_smoother": (P, [P, C.c_int]),
    "mgb_problem_level_transfer": (P, [P, C.c_int]),
    "mgb_problem_fine_operator": (P, [P]),
    "mgb_problem_fine_dofmap": (P, [P]),
    "mgb_problem_rhs": (P, [P]),
    "mgb_problem_solution": (P, [P]),
    "mgb_problem_vcycle": (C.c_int, [P, P, P]),
    "mgb_problem_solve_once": (C.c_int, [P, C.c_int, C.POINTER(Result)]),
    "mgb_problem_solve_host": (C.c_int, [P, C.c_void_p, C.c_void_p, C.POINTER(Result)]),
    "mgb_problem_solve_host_batch": (C.c_int, [P, C.c_int, PP, PP, P]),
    "mgb_problem_run": (C.c_int, [P, C.c_int, C.POINTER(Result)]),
}

_lib = None


def load():
    """dlopen libmgb200.so and attach the signatures; raises if the library is missing (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(make -C dealii_multigrid_b200/csrc). There is no CPU/Python fallback.")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
