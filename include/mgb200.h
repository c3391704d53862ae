/* mgb200 -- C ABI of the B200-native matrix-free multigrid hot path.
 *
 * Drop-in boundary for the deal.II-style concepts peterrum/dealii-multigrid composes (the reference has no
 * FFI; these entry points are what a thin header adapter around its C++ classes binds -- INTEGRATION.md):
 *
 *   mgb_operator_*      <->  Operator<dim,1,Number>               include/operator.h:11-557
 *                            vmult :152-183, Tvmult :185-189, vmult_interface_down :191-201,
 *                            vmult_interface_up :203-226, compute_inverse_diagonal :228-242,
 *                            initialize_dof_vector :140-144, m :123-131
 *   mgb_chebyshev_*     <->  PreconditionChebyshev<Operator,Vector,DiagonalMatrix> as configured in
 *                            multigrid_throughput.cc:847-883 (vmult = pre-smoothing, step = post-smoothing)
 *   mgb_transfer_*      <->  MGTwoLevelTransfer::reinit / prolongate_and_add / restrict_and_add
 *                            multigrid_throughput.cc:1600-1604, 1809-1812
 *   mgb_vector_*        <->  LinearAlgebra::distributed::Vector<Number> (BLAS-1, update_ghost_values,
 *                            compress(add)) used implicitly by all of the above
 *   mgb_problem_*       <->  run<3,1,double,LevelNumber>() + solve_with_global_coarsening /
 *                            solve_with_local_smoothing + mg_solve   multigrid_throughput.cc:817-1404,
 *                            1443-1873, 2019-2396 (JSON keys of RunParameters::parse :1989-2014)
 *   mgb_mesh_*, mgb_dofmap_*, mgb_transfer_map_*
 *                       <->  stand-in for the deal.II setup objects (Triangulation, DoFHandler,
 *                            AffineConstraints, MatrixFree::reinit) that produce the index maps consumed by
 *                            the device objects; with deal.II present an adapter passes deal.II's own maps.
 *
 * Conventions: every function returns 0 on success, non-zero on error (message: mgb_last_error()).
 * Host arrays passed to *_create are copied; handles are destroyed with the matching *_destroy.
 * Vectors live in device memory ([owned | ghost] layout).  One CUDA stream per process (mgb_set_stream);
 * handles are not thread-safe (the reference runs 1 thread per rank, multigrid_throughput.cc:2403).
 * There is NO CPU fallback: every compute entry point fails if no CUDA device is available.
 */
#ifndef MGB200_H
#define MGB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MGB_DOUBLE 0
#define MGB_FLOAT 1
#define MGB_FLAG_BIT 0x80000000u /* entity-table flag: Dirichlet (eval table) / constrained (own table) */

typedef struct mgb_mesh         mgb_mesh;
typedef struct mgb_dofmap       mgb_dofmap;
typedef struct mgb_transfer_map mgb_transfer_map;
typedef struct mgb_level_edges  mgb_level_edges;
typedef struct mgb_copy_indices mgb_copy_indices;
typedef struct mgb_vector       mgb_vector;
typedef struct mgb_operator     mgb_operator;
typedef struct mgb_chebyshev    mgb_chebyshev;
typedef struct mgb_transfer     mgb_transfer;
typedef struct mgb_problem      mgb_problem;

const char *mgb_last_error(void);
const char *mgb_version(void);

/* ---- device / stream / communicator ------------------------------------------------------------ */
int mgb_init(int device);                 /* cudaSetDevice + scratch allocation; fails without a GPU */
int mgb_set_stream(void *cuda_stream);    /* all subsequent kernels are enqueued on this cudaStream_t */
int mgb_synchronize(void);
int64_t mgb_launch_count(void);           /* kernels enqueued by the library so far */
/* live timing of the dominant kernel: CUDA-event pairs around every cell-loop launch over exactly n_cells cells
 * between start and stop (bench.py's roofline.achieved); stop synchronises and returns the summed duration */
int mgb_profile_start(int64_t n_cells);
int mgb_profile_stop(double *total_ms, int64_t *n_launches);
/* NCCL communicator for ghost exchange and reductions; unique_id = 128-byte ncclUniqueId */
int mgb_comm_unique_id(void *out128, const char *libnccl_path);
int mgb_comm_init(const void *unique_id128, int rank, int size, const char *libnccl_path);
/* collective: back the partitioners of every operator / transfer created so far with peer-mapped staging memory
 * (cudaIpc); their update_ghost_values / compress(add) then run as NVLink peer stores instead of NCCL send/recv.
 * mgb_problem_create calls it itself.  Opt-in: environment variable MGB_P2P=1; *enabled = 0 if the ranks stay on the
 * NCCL send/recv path (variable unset, no peer access). */
int mgb_comm_enable_peer_exchange(int *enabled);
/* collective; releases the captured V-cycle graphs (they hold the communicator) and the peer-mapped staging memory first */
int mgb_comm_finalize(void);

/* ---- 1-D finite element data (host; FE_Q(degree) + QGauss(degree+1) on [0,1]) ------------------- */
/* nodes[n], M[n*n], K[n*n], S0[n*n], S1[n*n], Ph[(2p+1)*n]; any pointer may be NULL */
int mgb_fe_1d(int degree, double *nodes, double *M, double *K, double *S0, double *S1, double *Ph);
/* fast diagonalisation of the 1-D pencil used by the cell kernels: K = Tm^T diag(lambda) Tm, M = Tm^T Tm;
 * Tm[n*n] (rows = modes, symmetric ones first), lambda[n] */
int mgb_fe_1d_fdm(int degree, double *Tm, double *lambda);
/* Pp[(pf+1)*(pc+1)] */
int mgb_fe_p_embedding(int degree_fine, int degree_coarse, double *Pp);

/* ---- stand-in setup: octree mesh (include/grid_generator.h, multigrid_throughput.cc:2048-2062) --- */
int     mgb_mesh_create(const char *geometry_type, unsigned n_ref_global, unsigned n_ref_local, mgb_mesh **out);
int     mgb_mesh_coarsen(const mgb_mesh *fine, mgb_mesh **out); /* one global-coarsening step (:2219-2224) */
/* local-smoothing level of the refinement tree: ALL cells of that level, active or not
 * (Triangulation::construct_multigrid_hierarchy, multigrid_throughput.cc:2044-2046) */
int     mgb_mesh_level(const mgb_mesh *active, int level, mgb_mesh **out);
int64_t mgb_mesh_n_cells(const mgb_mesh *m);
int     mgb_mesh_n_levels(const mgb_mesh *m);
int     mgb_mesh_get_cells(const mgb_mesh *m, uint32_t *level_x_y_z /* n_cells*4 */);
void    mgb_mesh_destroy(mgb_mesh *m);

/* ---- stand-in setup: FE_Q DoFs, constraints, partitioner ----------------------------------------- */
typedef struct
{
  int32_t  degree, n_ranks, rank, mesh_max_level;
  int64_t  n_cells_global, cell_begin, n_cells, n_cells_hanging;
  int64_t  n_dofs_global, owned_begin;
  uint32_t n_owned, n_ghost, n_constrained;
  uint32_t n_ghost_ranks, n_import_ranks, n_import_indices;
  /* borrowed pointers, valid until mgb_dofmap_destroy */
  const uint32_t *eval_table;   /* n_cells*27, local index | MGB_FLAG_BIT(Dirichlet)   -- MatrixFree dof_indices */
  const uint32_t *own_table;    /* n_cells*27, local index | MGB_FLAG_BIT(constrained) -- DoFHandler cell dofs  */
  const uint16_t *mask;         /* n_cells hanging-node masks                                                   */
  const uint8_t  *cell_level;   /* n_cells                                                                      */
  const uint32_t *cell_xyz;     /* n_cells*3                                                                    */
  const uint32_t *constrained;  /* n_constrained owned local indices (Dirichlet + hanging), ascending           */
  const int64_t  *ghost_global; /* n_ghost                                                                      */
  const int32_t  *ghost_ranks;  const uint32_t *ghost_offsets;  /* n_ghost_ranks (+1)                           */
  const int32_t  *import_ranks; const uint32_t *import_offsets; /* n_import_ranks (+1)                          */
  const uint32_t *import_indices;                               /* n_import_indices                             */
} mgb_dofmap_view;

int  mgb_dofmap_create(const mgb_mesh *m, int degree, int n_ranks, int rank, mgb_dofmap **out);
/* the same with an explicit partition: rank r owns the Morton range [cell_begin[r], cell_begin[r+1]) */
int  mgb_dofmap_create_partitioned(const mgb_mesh *m, int degree, int n_ranks, int rank, const int64_t *cell_begin /* n_ranks+1 */,
                                   mgb_dofmap **out);
/* Partition policies selected by `PartitionerName` (multigrid_throughput.cc:2066-2174; RepartitioningPolicyTools and the
 * reference's BalancedGranularityPartitionPolicy, :372-419): "DefaultPolicy", "MinimalGranularityPolicy-N",
 * "BalancedGranularityPartitionPolicy", "CellWeightPolicy-W" (hanging_nodes_weighting, :276-292), "FirstChildPolicy[-W]".
 * meshes: coarse -> fine coarsening sequence; cell_begin: n_meshes * (n_ranks+1) Morton offsets.  Unknown names fail
 * like the reference's ExcNotImplemented. */
int  mgb_partition_sequence(const char *policy_name, const mgb_mesh *const *meshes, int n_meshes, int n_ranks,
                            int64_t min_cells_per_rank, int64_t *cell_begin);
/* ownership "with the first child" (FirstChildPolicy; level cells of local smoothing): cell_begin of `parents` from the
 * partition of `children` */
int  mgb_partition_first_child(const mgb_mesh *parents, const mgb_mesh *children, int n_ranks, const int64_t *child_begin,
                               int64_t *cell_begin /* n_ranks+1 */);
/* per leaf: 1 if the cell has a hanging-node constraint on a face or edge (Helper::is_constrained, :231-268) */
int  mgb_mesh_hanging_cells(const mgb_mesh *m, uint8_t *flags /* n_cells */);
int  mgb_dofmap_view_get(const mgb_dofmap *d, mgb_dofmap_view *view);
/* expand an entity table to the lexicographic (p+1)^3 index map per cell (flag kept in bit 31) */
int  mgb_dofmap_expand(const mgb_dofmap *d, int which /*0 eval, 1 own*/, uint32_t *out /* n_cells*(p+1)^3 */);
void mgb_dofmap_destroy(mgb_dofmap *d);
/* Operator::rhs on the host (include/operator.h:362-447): simulation_type "Constant" (f = 1, u_D = 0) or "Gaussian"
 * (GaussianRightHandSide / GaussianSolution of multigrid_throughput.cc:60-127 with the inhomogeneous Dirichlet lift
 * b -= A_hn x_D); rhs_local holds n_owned + n_ghost entries (ghost contributions are NOT compressed).  Host only. */
int  mgb_assemble_rhs(const mgb_mesh *m, const mgb_dofmap *d, const char *simulation_type, double *rhs_local);
/* Host-only diagnostic of the chunk plan the operators build over a DoF map (cell loop of device/chunk.cuh): builds the
 * plan for values of value_bytes (8 or 4) and walks through it on the CPU in the kernel's order with a diagonal stand-in
 * for the cell matrices.  stats[0] = largest relative deviation from the direct scatter-add (>= 1e300: structural defect),
 * [1] chunks, [2] distinct compute tables, [3] distinct pull tables, [4] table words, [5..7] foreign DoF slots owned by a
 * lower chunk / by a higher chunk (late) / by another rank, [8] side-buffer entries, [9] largest DoF window, [10] seconds spent building the plan */
int  mgb_dofmap_chunk_plan_check(const mgb_dofmap *d, int value_bytes, double *stats /* 11 */);

typedef struct
{
  int32_t         degree_fine, degree_coarse;
  int64_t         n_coarse_cells, n_fine_cells;
  const uint32_t *fine_first;  /* n_coarse_cells */
  const uint8_t  *n_children;  /* n_coarse_cells: 1 or 8 (local smoothing: 8 or 0) */
  const uint8_t  *weight_code; /* n_fine_slots*27: 0 -> weight 0, v -> 1/v */
  /* fine-side table indexed by fine_first (+child): single rank = the fine dofmap's own table; several ranks = the
   * children of the locally owned coarse cells in the numbering of the transfer's internal ghosted fine vector */
  const uint32_t *fine_own;    /* n_fine_slots*27 */
  int64_t         n_fine_slots;
  int32_t         distributed;
  /* partitioner of that internal vector (distributed only): same owned range as the fine level */
  int64_t         owned_begin;
  uint32_t        n_owned, n_ghost, n_ghost_ranks, n_import_ranks, n_import_indices;
  const int64_t  *ghost_global;
  const int32_t  *ghost_ranks;  const uint32_t *ghost_offsets;
  const int32_t  *import_ranks; const uint32_t *import_offsets;
  const uint32_t *import_indices;
} mgb_transfer_map_view;
int  mgb_transfer_map_create(const mgb_mesh *mesh_fine, const mgb_dofmap *fine, const mgb_mesh *mesh_coarse, const mgb_dofmap *coarse,
                             mgb_transfer_map **out);
/* MGTransferMatrixFree::build (multigrid_throughput.cc:1800-1802): the meshes are consecutive local-smoothing levels;
 * a coarse cell has 8 children on the fine level or none (n_children 0) */
int  mgb_transfer_map_create_ls(const mgb_mesh *level_fine, const mgb_dofmap *fine, const mgb_mesh *level_coarse,
                                const mgb_dofmap *coarse, mgb_transfer_map **out);
int  mgb_transfer_map_view_get(const mgb_transfer_map *t, mgb_transfer_map_view *view);
void mgb_transfer_map_destroy(mgb_transfer_map *t);

/* ---- stand-in setup: local smoothing (MGConstrainedDoFs, copy_indices) ------------------------------ */
typedef struct
{
  uint32_t        n_edge_dofs, n_edge_cells;
  int32_t         any_global;  /* include/operator.h:67-69 */
  const uint32_t *edge_dofs;   /* owned local indices on the refinement edge of the level (include/operator.h:49-66) */
  const uint32_t *edge_cells;  /* local cells touching an edge DoF (include/operator.h:71-99) */
} mgb_level_edges_view;
int  mgb_level_edges_create(const mgb_mesh *level_mesh, const mgb_dofmap *level_dofs, mgb_level_edges **out);
int  mgb_level_edges_view_get(const mgb_level_edges *e, mgb_level_edges_view *view);
void mgb_level_edges_destroy(mgb_level_edges *e);
typedef struct
{
  uint32_t        n;
  const uint32_t *active_idx, *level_idx; /* MGLevelGlobalTransfer::copy_indices[level] */
  /* several ranks: the pairs are held by the owner of the ACTIVE DoF; level_idx refers to an auxiliary ghosted level vector
   * [level's owned range | ghost_global] (copy_indices_global_mine / copy_indices_level_mine of deal.II) */
  int32_t         distributed;
  int64_t         owned_begin;
  uint32_t        n_owned, n_ghost;
  const int64_t  *ghost_global;
} mgb_copy_indices_view;
/* active_mesh: the mesh of active_dofs; needed on several ranks only (may be NULL on one rank) */
int  mgb_copy_indices_create(const mgb_dofmap *active_dofs, const mgb_mesh *active_mesh, const mgb_mesh *level_mesh,
                             const mgb_dofmap *level_dofs, mgb_copy_indices **out);
int  mgb_copy_indices_view_get(const mgb_copy_indices *c, mgb_copy_indices_view *view);
void mgb_copy_indices_destroy(mgb_copy_indices *c);

/* ---- vectors ----------------------------------------------------------------------------------- */
int  mgb_vector_create(uint32_t n_owned, uint32_t n_ghost, int number_type, mgb_vector **out);
int  mgb_vector_wrap(void *device_ptr, uint32_t n_owned, uint32_t n_ghost, int number_type, mgb_vector **out);
void mgb_vector_destroy(mgb_vector *v);
uint32_t mgb_vector_n_owned(const mgb_vector *v);
void *mgb_vector_data(mgb_vector *v); /* device pointer */
int  mgb_vector_copy_from_host(mgb_vector *v, const void *host, uint32_t n); /* first n local entries */
int  mgb_vector_copy_to_host(const mgb_vector *v, void *host, uint32_t n);
int  mgb_vector_set(mgb_vector *v, double a);
int  mgb_vector_add(mgb_vector *y, double a, const mgb_vector *x);           /* y += a x      */
int  mgb_vector_sadd(mgb_vector *y, double s, double a, const mgb_vector *x); /* y = s y + a x */
int  mgb_vector_equ(mgb_vector *y, double a, const mgb_vector *x);           /* y = a x       */
int  mgb_vector_dot(const mgb_vector *x, const mgb_vector *y, double *result);
int  mgb_vector_l2_norm(const mgb_vector *x, double *result);
int  mgb_vector_add_and_dot(mgb_vector *g, double a, const mgb_vector *h, double *result); /* g += a h; g.g */
int  mgb_vector_update_ghost_values(const mgb_vector *v);
int  mgb_vector_compress_add(mgb_vector *v);
int  mgb_vector_zero_out_ghost_values(mgb_vector *v);
/* Dot product that allocates nothing (no counterpart in the reference: added for
 * callers that cannot call cudaMalloc -- CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * mgb_vector_dot up to the summation order (1e-13 relative), for FP64 device arrays of the CALLER: *result_dev = x . y.
 * The scratch of the reduction (one partial sum per 4096-element block + a ticket) lives in a static __device__ array of
 * MGB_NO_ALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit; its device address is looked up once per
 * device with cudaGetSymbolAddress.  n > MGB_NO_ALLOC_MAX_N is rejected (MGB error, mgb_last_error says why).
 * cuda_stream: the cudaStream_t to enqueue on, NULL = the stream of mgb_set_stream.  The call is asynchronous (no host
 * synchronisation, no cudaMalloc / cudaFree / allocating library call).  All calls share the scratch array, which keeps the
 * last call's contents (the ticket is cleared on the stream before each launch): calls that overlap on the GPU -- other
 * streams, other host threads on the same device -- race on it and must be serialised by the caller. */
#define MGB_NO_ALLOC_MAX_N         16777216ull /* 4096 partial sums x 4096 elements per block */
#define MGB_NO_ALLOC_SCRATCH_BYTES 33024       /* (4096 + 32) doubles */
int  mgb_vector_dot_no_alloc(const double *x_dev, const double *y_dev, uint64_t n, double *result_dev, void *cuda_stream);

/* ---- operator (include/operator.h) ---------------------------------------------------------------- */
typedef struct
{
  int32_t         degree, number_type;
  int64_t         n_cells;
  const uint32_t *entity_table; /* n_cells*27 compressed dof_indices (see mgb_dofmap_view.eval_table) ...      */
  const uint32_t *dof_indices;  /* ... or the full lexicographic map n_cells*(p+1)^3 (compressed on the host;
                                   fails if a cell's entities are not contiguous). Exactly one is non-NULL.   */
  const uint16_t *mask;         /* hanging-node masks or NULL                                                  */
  const uint8_t  *cell_level;   /* cell size h = root_length / 2^level (all reference meshes are Cartesian)    */
  double          root_length;
  const uint32_t *constrained;  uint32_t n_constrained;          /* operator.h:44-47  */
  const uint32_t *edge_constrained; uint32_t n_edge_constrained; /* operator.h:49-66  */
  int32_t         has_edge_constrained_global;                   /* operator.h:67-69  */
  const uint32_t *edge_cells;   uint32_t n_edge_cells;           /* operator.h:71-99  */
  int64_t         n_dofs_global;
  /* vector partitioner (MatrixFree::get_vector_partitioner) */
  int64_t         owned_begin;
  uint32_t        n_owned, n_ghost;
  const int64_t  *ghost_global;
  const int32_t  *ghost_ranks;  const uint32_t *ghost_offsets;  uint32_t n_ghost_ranks;
  const int32_t  *import_ranks; const uint32_t *import_offsets; uint32_t n_import_ranks;
  const uint32_t *import_indices;
} mgb_operator_desc;

/* Every index of the descriptor is checked on the host before anything is uploaded (entity bases + entity sizes against
 * n_owned + n_ghost, the index lists against n_owned / n_cells, the offset arrays for 0 ... total, ascending): a map that
 * would address memory outside the vectors fails here with a message naming the entry, not inside a kernel.  The same holds
 * for mgb_transfer_create_from_arrays (fine_own) and mgb_level_copy_create_from_arrays (level_idx at creation, active_idx
 * against the active-mesh vector of every copy_to_mg / copy_from_mg call). */
int  mgb_operator_create(const mgb_operator_desc *desc, mgb_operator **out);
int  mgb_operator_create_from_dofmap(const mgb_dofmap *d, int number_type, mgb_operator **out);
/* level operator of the local-smoothing hierarchy: Operator::reinit(..., mg_level) with its refinement-edge data */
int  mgb_operator_create_level(const mgb_dofmap *level_dofs, const mgb_level_edges *edges, int number_type, mgb_operator **out);
void mgb_operator_destroy(mgb_operator *op);
int64_t mgb_operator_m(const mgb_operator *op);
int  mgb_operator_initialize_dof_vector(const mgb_operator *op, mgb_vector **out);
int  mgb_operator_vmult(const mgb_operator *op, mgb_vector *dst, const mgb_vector *src);
int  mgb_operator_tvmult(const mgb_operator *op, mgb_vector *dst, const mgb_vector *src);
int  mgb_operator_vmult_interface_down(const mgb_operator *op, mgb_vector *dst, const mgb_vector *src);
int  mgb_operator_vmult_interface_up(const mgb_operator *op, mgb_vector *dst, const mgb_vector *src);
int  mgb_operator_compute_inverse_diagonal(const mgb_operator *op, mgb_vector *diag);
/* Cell loop with the vector update in its own pass (what deal.II does with the pre/post hooks of MatrixFree::cell_loop;
 * north_star: "the smoother's vector updates fused with vmult").  mode 0: out = A x;  1: out = b - A x;
 * 2: out = (1+f1) x + f2 dinv (b - A x);  3: out = (1+f1) x - f1 xold + f2 dinv (b - A x)   (identity rows of the
 * constrained DoFs included; b / dinv / xold may be NULL where the mode does not read them; out may alias xold, never x).
 * Needs the chunked cell loop (environment MGB_CHUNKED=1 when the operator is created): mgb_operator_fused_available
 * returns 1 if this operator can run it for vectors of its own layout, else the call fails with MGB_ERROR. */
int  mgb_operator_fused_available(const mgb_operator *op);
int  mgb_operator_vmult_fused(const mgb_operator *op, mgb_vector *out, const mgb_vector *x, int mode, const mgb_vector *b,
                              const mgb_vector *dinv, const mgb_vector *xold, double f1, double f2);
/* General (non-Cartesian) cells -- what MatrixFree does for a mapping that is not Cartesian (include/operator.h:37-42 asks
 * for update_gradients; the reference's own meshes are all axis-aligned cubes): 8 vertices per cell, lexicographic with x
 * fastest, trilinear geometry (MappingQ1).  All subsequent cell loops (vmult, interface operators, diagonal) evaluate
 * the metric terms at the Gauss points from these vertices.  Call before compute_inverse_diagonal / smoother setup. */
int  mgb_operator_set_cell_vertices(mgb_operator *op, const double *xyz /* n_cells*8*3 */, int64_t n_cells /* checked */);

/* ---- Chebyshev smoother ---------------------------------------------------------------------------- */
int  mgb_chebyshev_create(const mgb_operator *op, int degree, double smoothing_range, unsigned eig_cg_n_iterations,
                          mgb_chebyshev **out);
void mgb_chebyshev_destroy(mgb_chebyshev *c);
int  mgb_chebyshev_vmult(mgb_chebyshev *c, mgb_vector *dst, const mgb_vector *rhs); /* zero initial guess     */
int  mgb_chebyshev_step(mgb_chebyshev *c, mgb_vector *dst, const mgb_vector *rhs);  /* non-zero initial guess */
int  mgb_chebyshev_eigenvalues(mgb_chebyshev *c, double *lambda_min_estimate, double *lambda_max);

/* ---- two-level transfer ------------------------------------------------------------------------------ */
int  mgb_transfer_create(const mgb_dofmap *fine, const mgb_dofmap *coarse, const mgb_transfer_map *map, const mgb_operator *op_fine,
                         const mgb_operator *op_coarse, mgb_transfer **out);
/* the same from plain arrays (what a deal.II adapter extracts in MGTwoLevelTransfer::reinit): the fields of
 * mgb_transfer_map_view; n_fine_cells is ignored; the partitioner fields are read only if distributed != 0 */
int  mgb_transfer_create_from_arrays(const mgb_transfer_map_view *arrays, const mgb_operator *op_fine, const mgb_operator *op_coarse,
                                     mgb_transfer **out);
void mgb_transfer_destroy(mgb_transfer *t);
int  mgb_transfer_prolongate_and_add(const mgb_transfer *t, mgb_vector *dst_fine, const mgb_vector *src_coarse);
int  mgb_transfer_restrict_and_add(const mgb_transfer *t, mgb_vector *dst_coarse, const mgb_vector *src_fine);

/* MGTwoLevelTransfer::prolongate / MGTransferMatrixFree::prolongate: dst_fine = P src_coarse (no accumulation) */
int  mgb_transfer_prolongate(const mgb_transfer *t, mgb_vector *dst_fine, const mgb_vector *src_coarse);

/* ---- local smoothing: MGLevelGlobalTransfer::copy_to_mg / copy_from_mg through copy_indices ------------ *
 * (MGTransferMatrixFree as built at multigrid_throughput.cc:1800-1802; the double <-> MGNumberType casts of :528-550 happen
 * here).  One object per level: the device copy of mgb_copy_indices for that level; level_op gives the level vectors'
 * layout and number type.  On several ranks the owner of the active DoF holds the pair and the exchange goes through an
 * internal ghosted level vector. */
typedef struct mgb_level_copy mgb_level_copy;
int  mgb_level_copy_create(const mgb_copy_indices *ci, const mgb_operator *level_op, mgb_level_copy **out);
/* the same from plain (active-vector local index, level-vector local index) pairs held by this rank, both DoFs local */
int  mgb_level_copy_create_from_arrays(uint32_t n, const uint32_t *active_idx, const uint32_t *level_idx, const mgb_operator *level_op,
                                       mgb_level_copy **out);
void mgb_level_copy_destroy(mgb_level_copy *c);
/* dst_level = 0 except the entries of the active DoFs that live on this level (src_active: double, active-mesh layout) */
int  mgb_level_copy_to_mg(const mgb_level_copy *c, mgb_vector *dst_level, const mgb_vector *src_active);
/* the active DoFs living on this level receive the level's values; all other entries of dst_active are left alone */
int  mgb_level_copy_from_mg(const mgb_level_copy *c, mgb_vector *dst_active, const mgb_vector *src_level);

/* ---- Multigrid + PreconditionMG + SolverCG over borrowed handles (mg_solve, multigrid_throughput.cc:1082-1133) ---- *
 * level_operators[0..n_levels-1] coarse -> fine (mg::Matrix of MGInterfaceOperator: vmult_interface_down / _up),
 * smoothers[l] and transfers[l] (between l-1 and l) for l >= 1 (entries 0 ignored), the coarse level solved by
 * SolverCG(ReductionControl(coarse_maxiter, coarse_abstol, coarse_reltol)) with coarse_preconditioner (a Chebyshev
 * smoother on level 0: CoarseGridSolverType cg_with_chebyshev) or PreconditionIdentity (NULL: cg), :887-944.
 * All handles are borrowed and must outlive the multigrid object. */
typedef struct mgb_multigrid mgb_multigrid;
#define MGB_N_STAGES 7 /* pre-smoothing, residual, restriction, coarse solve, prolongation, edge prolongation, post-smoothing */
int  mgb_multigrid_create(int n_levels, const mgb_operator *const *level_operators, mgb_chebyshev *const *smoothers,
                          const mgb_transfer *const *transfers, mgb_chebyshev *coarse_preconditioner, unsigned coarse_maxiter,
                          double coarse_abstol, double coarse_reltol, mgb_multigrid **out);
void mgb_multigrid_destroy(mgb_multigrid *m);
/* Multigrid::set_edge_in_matrix (:1101-1105): after the prolongation defect -= vmult_interface_up(solution) */
int  mgb_multigrid_set_edge_in(mgb_multigrid *m, int on);
/* local smoothing: one mgb_level_copy per level (copy_to_mg / copy_from_mg through copy_indices); NULL = global coarsening,
 * where the top level IS the outer vector (MGTransferGlobalCoarsening) */
int  mgb_multigrid_set_level_copies(mgb_multigrid *m, const mgb_level_copy *const *copies);
/* PreconditionMG::vmult: dst = M^-1 src, double vectors of the outer operator's layout */
int  mgb_multigrid_vcycle(mgb_multigrid *m, mgb_vector *dst, const mgb_vector *src);
/* the seven stage timers the reference connects to the Multigrid signals (:1163-1234): CUDA-event pairs around every
 * stage of every level while enabled; collect returns the seconds accumulated since the last call, [level][stage] */
int  mgb_multigrid_enable_timers(mgb_multigrid *m, int on);
int  mgb_multigrid_collect_times(mgb_multigrid *m, double *seconds /* n_levels * MGB_N_STAGES */);
/* SolverCG<Vector>(ReductionControl(maxiter, abstol, reltol)).solve(A, x, b, preconditioner) from x = 0 (:1247-1248, 1625-1635);
 * preconditioner NULL = PreconditionIdentity; fails (SolverControl::NoConvergence) when maxiter is reached */
int  mgb_cg_solve(const mgb_operator *A, mgb_multigrid *preconditioner, mgb_vector *x, const mgb_vector *b, unsigned maxiter,
                  double abstol, double reltol, unsigned *n_iterations, double *final_residual);

/* ---- JSON-driven problem: run() + solve_with_* + mg_solve -------------------------------------------- */
typedef struct
{
  int32_t  dim, degree, n_ref_global, n_ref_local, n_levels, n_iterations, sub_comm_size;
  int64_t  n_cells, n_cells_hn, n_cells_n, n_dofs;
  double   time, time_cg, throughput;
  double   time_pre, time_residuum, time_res, time_cs, time_pro, time_edge_pro, time_post, time_to_mg, time_to_global;
  double   initial_residual, final_residual;
  int64_t  coarse_iterations; /* total coarse-grid CG iterations of the reported solve */
} mgb_result;

/* MGTools::print_multigrid_statistics (include/mg_tools.h:267-513), the columns the reference appends with Verbosity=true
 * (multigrid_throughput.cc:1657-1665, 1863-1872).  mem_total counts the stand-in octree, not deal.II's triangulations. */
typedef struct
{
  double  workload_eff, workload_path_max, vertical_eff, horizontal_eff, mem_total;
  int32_t verbose; /* mgb_problem_statistics: the Verbosity key of the input (the reference adds the columns only then) */
} mgb_statistics;
/* per level (coarse -> fine), 7 values: n_cells_min, n_cells_max, n_cells_sum, vertical local / remote children,
 * horizontal owned / ghost cells (sums over ranks) */
#define MGB_STATISTICS_LEVEL_VALUES 7
/* meshes: a coarsening sequence (global coarsening) or the level meshes of one tree (local smoothing), coarse -> fine;
 * cell_begin: n_meshes*(n_ranks+1) Morton offsets (mgb_partition_sequence / mgb_partition_first_child) */
int  mgb_hierarchy_statistics(const mgb_mesh *const *meshes, int n_meshes, int n_ranks, const int64_t *cell_begin, mgb_statistics *out,
                              int64_t *per_level /* n_meshes*MGB_STATISTICS_LEVEL_VALUES or NULL */);

int  mgb_problem_create(const char *json_text, int n_ranks, int rank, mgb_problem **out);
void mgb_problem_destroy(mgb_problem *p);
int  mgb_problem_n_levels(const mgb_problem *p);
int64_t mgb_problem_level_n_dofs(const mgb_problem *p, int level);
int64_t mgb_problem_level_n_cells(const mgb_problem *p, int level);
int  mgb_problem_level_degree(const mgb_problem *p, int level);
/* borrowed handles of the objects the solver uses */
const mgb_dofmap   *mgb_problem_level_dofmap(const mgb_problem *p, int level);
const mgb_operator *mgb_problem_level_operator(const mgb_problem *p, int level);
mgb_chebyshev      *mgb_problem_level_smoother(mgb_problem *p, int level);
const mgb_transfer *mgb_problem_level_transfer(const mgb_problem *p, int level); /* between level-1 and level */
const mgb_operator *mgb_problem_fine_operator(const mgb_problem *p);           /* outer FP64 operator        */
const mgb_dofmap   *mgb_problem_fine_dofmap(const mgb_problem *p);             /* active-mesh DoFs of the outer operator */
mgb_vector         *mgb_problem_rhs(mgb_problem *p);
mgb_vector         *mgb_problem_solution(mgb_problem *p);
/* PreconditionMG::vmult: one V-cycle, dst = M^-1 src (double vectors of the fine operator's layout) */
int  mgb_problem_vcycle(mgb_problem *p, mgb_vector *dst, const mgb_vector *src);
/* SolverCG with the V-cycle preconditioner: one solve from a zero initial guess (no timing protocol) */
int  mgb_problem_solve_once(mgb_problem *p, int with_stage_timers, mgb_result *result);
/* the same solve with HOST buffers: rhs_host (n_owned doubles) is copied to the device, solved from a zero
 * initial guess, and the solution copied back into solution_host (n_owned doubles) */
int  mgb_problem_solve_host(mgb_problem *p, const double *rhs_host, double *solution_host, mgb_result *result);
/* n solves in a row with host buffers (n_owned doubles each; use pinned memory): the upload of right-hand side i+1 and the
 * download of solution i-1 run on copy streams of their own behind solve i, so that only the first upload and the last
 * download are exposed.  results: n entries or NULL.  Returns when every solution_hosts[i] is filled. */
int  mgb_problem_solve_host_batch(mgb_problem *p, int n, const double *const *rhs_hosts, double *const *solution_hosts, mgb_result *results);
/* the reference's protocol: 1 warm-up + n_repetitions timed solves, best one reported (:1140-1283) */
int  mgb_problem_run(mgb_problem *p, int n_repetitions, mgb_result *result);
/* statistics of the problem's partitioned hierarchy; per_level sized mgb_problem_n_meshes()*MGB_STATISTICS_LEVEL_VALUES or NULL */
int  mgb_problem_n_meshes(const mgb_problem *p);
int  mgb_problem_statistics(const mgb_problem *p, mgb_statistics *out, int64_t *per_level);

#ifdef __cplusplus
}
#endif
#endif
