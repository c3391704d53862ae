"""dealii_multigrid_b200 -- B200-native hot path of peterrum/dealii-multigrid's multigrid_throughput.

The product is the CUDA/C++ shared library csrc -> lib/libmgb200.so (C ABI: include/mgb200.h).  This
package is its thin ctypes mirror of the reference's operator interface.  Importing it loads the shared
library and raises ImportError if it has not been built: there is no CPU or pure-Python compute path.
"""
from . import _lib

_lib.load()

from .api import (  # noqa: E402
    DOUBLE, FLOAT, FLAG_BIT, INDEX_MASK, MgbError, Mesh, DoFMap, TransferMap, LevelEdges, CopyIndices, Vector, Operator,
    PreconditionChebyshev, MGTwoLevelTransfer, LevelCopy, Multigrid, cg_solve, dot_no_alloc, NO_ALLOC_MAX_N, Problem, init, set_stream, synchronize, launch_count, profile_start, profile_stop,
    comm_unique_id, comm_init, comm_finalize, comm_enable_peer_exchange, partition_sequence, first_child_boundaries, hierarchy_statistics, fe_1d, fe_1d_fdm, fe_p_embedding)

__all__ = ["DOUBLE", "FLOAT", "FLAG_BIT", "INDEX_MASK", "MgbError", "Mesh", "DoFMap", "TransferMap", "LevelEdges", "CopyIndices", "Vector", "Operator",
           "PreconditionChebyshev", "MGTwoLevelTransfer", "LevelCopy", "Multigrid", "cg_solve", "dot_no_alloc", "NO_ALLOC_MAX_N", "Problem", "init", "set_stream", "synchronize", "launch_count", "profile_start", "profile_stop",
           "comm_unique_id", "comm_init", "comm_finalize", "comm_enable_peer_exchange", "partition_sequence", "first_child_boundaries", "hierarchy_statistics", "fe_1d", "fe_1d_fdm", "fe_p_embedding"]
