// Launcher declarations for the sm_100a kernels (definitions in *.cu).  All launchers enqueue on the
// stream passed in and never synchronise, except the reductions that return a scalar to the host.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace mgb
{
  constexpr int MAX_DEGREE = 6;

  void check_cuda(cudaError_t e, const char *what);
#define MGB_CUDA(x) ::mgb::check_cuda((x), #x)

  // upload the 1-D matrices of FE_Q(degree) to constant memory (idempotent)
  void upload_fe_constants(int degree);

  // ---------------------------------------------------------------- BLAS-1 (vector.cu)
  template <typename T> void vec_set(T *x, T a, size_t n, cudaStream_t s);
  template <typename T> void vec_scale(T *x, T a, size_t n, cudaStream_t s);
  template <typename T> void vec_add(T *y, T a, const T *x, size_t n, cudaStream_t s);            // y += a x
  template <typename T> void vec_sadd(T *y, T s_, T a, const T *x, size_t n, cudaStream_t s);     // y = s y + a x
  template <typename T> void vec_equ(T *y, T a, const T *x, size_t n, cudaStream_t s);            // y = a x
  template <typename T> void vec_add_scalar(T *y, T a, size_t n, cudaStream_t s);                 // y += a
  template <typename T, typename U> void vec_convert(T *y, const U *x, size_t n, cudaStream_t s); // y = (T)x
  // deterministic reductions; result is written to *result_dev (device) by the last block
  template <typename T> void vec_dot(const T *x, const T *y, size_t n, double *result_dev, double *scratch,
                                     unsigned *ticket, cudaStream_t s);
  template <typename T> void vec_sum(const T *x, size_t n, double *result_dev, double *scratch, unsigned *ticket,
                                     cudaStream_t s);
  // g += a h; result = g.g   (SolverCG's add_and_dot with w == g)
  template <typename T> void vec_add_and_dot(T *g, T a, const T *h, size_t n, double *result_dev, double *scratch,
                                             unsigned *ticket, cudaStream_t s);
  // x += a d; g += a h; result = g.g   (the two vector updates of one CG iteration in one pass)
  template <typename T>
  void vec_cg_update(T *x, T *g, T a, const T *d, const T *h, size_t n, double *result_dev, double *scratch, unsigned *ticket,
                     cudaStream_t s);
  int reduction_scratch_size(); // number of doubles `scratch` must hold

  // gather / scatter by index lists (identity rows, edge DoFs, ghost pack/unpack, copy_indices)
  template <typename T> void idx_copy(T *dst, const T *src, const uint32_t *idx, uint32_t n, cudaStream_t s); // dst[idx]=src[idx]
  template <typename T> void idx_residual(T *dst, const T *a, const T *b, const uint32_t *idx, uint32_t n, cudaStream_t s); // dst[idx]=a[idx]-b[idx]
  template <typename T> void idx_gather(T *buf, const T *src, const uint32_t *idx, uint32_t n, cudaStream_t s); // buf[i]=src[idx[i]]
  template <typename T> void idx_scatter(T *dst, const T *buf, const uint32_t *idx, uint32_t n, cudaStream_t s); // dst[idx[i]]=buf[i]
  template <typename T> void idx_scatter_add(T *dst, const T *buf, const uint32_t *idx, uint32_t n, cudaStream_t s);
  template <typename T> void idx_set(T *dst, T a, const uint32_t *idx, uint32_t n, cudaStream_t s);            // dst[idx[i]]=a
  template <typename T, typename U>
  void idx_pair_copy(T *dst, const uint32_t *dst_idx, const U *src, const uint32_t *src_idx, uint32_t n, cudaStream_t s);

  // Chebyshev vector update (PreconditionChebyshev's vector_updates):
  //   it==0: out = f2 dinv rhs
  //   it==1: out = (1+f1) x + f2 dinv (rhs - t)
  //   it>=2: out = (1+f1) x - f1 x_old + f2 dinv (rhs - t)
  // `out` may alias x or x_old (pointwise update)
  template <typename T>
  void cheb_update(int it, T f1, T f2, const T *rhs, const T *dinv, const T *t, const T *x, const T *x_old, T *out, size_t n,
                   cudaStream_t s);
  // d = (|d| > 1e-10) ? 1/d : 1   (include/operator.h:240-241)
  template <typename T> void invert_diagonal(T *d, size_t n, cudaStream_t s);

  // ---------------------------------------------------------------- ghost exchange over NVLink peer memory (halo.cu)
  constexpr int HALO_MAX_PEERS = 16;
  struct HaloSide // the peers this rank sends to (or receives from) in one direction of the exchange
  {
    int                 n = 0;
    uint32_t            off[HALO_MAX_PEERS + 1] = {}; // element ranges of the peers in the pack list / ghost tail / staging slot
    char               *remote_stage[HALO_MAX_PEERS] = {};      // send: slot 0 of my range in the peer's staging block (mapped)
    size_t              remote_slot_bytes[HALO_MAX_PEERS] = {}; // send: distance to slot 1 there
    unsigned long long *remote_flag[HALO_MAX_PEERS] = {}; // send: the peer's "ready" flag for me; recv: the peer's "consumed" flag for me
    unsigned long long *local_flag[HALO_MAX_PEERS]  = {}; // send: my "consumed" flag from the peer; recv: my "ready" flag from the peer
  };
  struct HaloDir // one direction: owner -> ghoster (update_ghost_values) or ghoster -> owner (compress)
  {
    HaloSide            send, recv;
    char               *stage      = nullptr; // my staging block: 2 slots of 8-byte elements
    size_t              slot_bytes = 0;
    unsigned long long *epoch      = nullptr; // device counter of completed exchanges
    unsigned           *ticket     = nullptr; // [2]: last-block detection of the send and the recv kernel
  };
  template <typename T>
  void halo_exchange(const HaloDir &hd, bool ghost_direction, T *vec, uint32_t n_owned, const uint32_t *import_indices, cudaStream_t s,
                     bool fused = false); // fused: one kernel (needs the GPU to itself: the grid must be co-resident)

  // ---------------------------------------------------------------- cell kernels (operator.cu)
  struct CellData
  {
    int             degree     = 1;
    long long       n_cells    = 0;
    const uint32_t *etab       = nullptr; // n_cells*27 entity table (local indices | FLAG_BIT)
    const uint32_t *ctab       = nullptr; // n_cells*32 packed cell records: 27 entity entries, word 27 = mask | level << 16
    const uint16_t *mask       = nullptr; // n_cells hanging-node masks (may be null)
    const uint8_t  *level      = nullptr; // n_cells refinement levels; h = root_length / 2^level
    double          root_length = 2.0;
    int             any_hanging = 0;
    // general geometry (MappingQ1): n_cells*8*3 vertex coordinates, vertices lexicographic (x fastest); null = the cells
    // are axis-aligned cubes of side root_length / 2^level
    const double   *vertices    = nullptr;
    double          scale       = 1.0; // general geometry: dst += scale * A src (the Cartesian kernel folds it into root_length)
    // Cartesian kernel only: zero_n elements at zero_ptr (16-byte aligned, value type of the operator, not src or dst) are
    // set to zero by the cell loop on the side -- the destination of the next cell loop (k_laplace_apply<P, T, true>)
    void           *zero_ptr    = nullptr;
    long long       zero_n      = 0;
  };
  // dst += A_cells src   (scatter-add with FP atomics); cell_list (optional) restricts to a subset of cells
  template <typename T>
  void cell_laplace_apply(const CellData &cd, const T *src, T *dst, const uint32_t *cell_list, long long n_list,
                          cudaStream_t s);
  // Chunked cell loop (device/chunk.cuh, plan: host/chunks.h): out = E(A src) on the owned range with plain stores, E the
  // epilogue below; contributions to ghost entries are added atomically to the ghost tail of out (must be zero on entry).
  struct ChunkDevice
  {
    uint32_t        n_chunks  = 0;
    const uint32_t *desc      = nullptr, *pool = nullptr, *deps = nullptr, *ghost_idx = nullptr;
    void           *side      = nullptr; // side buffer (value type of the operator)
    uint32_t       *flags     = nullptr; // 2 * n_chunks
    uint32_t       *state     = nullptr; // 4 words: ticket, epoch, finished blocks
  };
  // mode 0: out = A x;  1: out = b - A x;  2: out = (1+f1) x + f2 dinv (b - A x);  3: out = (1+f1) x - f1 xold + f2 dinv (b - A x)
  // (the Chebyshev updates of cheb_update with t = A x); out may alias xold, never x
  template <typename T> struct FusedEpilogue
  {
    int      mode = 0;
    const T *b = nullptr, *dinv = nullptr, *xold = nullptr;
    T        f1 = T(0), f2 = T(0);
  };
  // true in the experimental build whose chunk kernel adds its sums atomically (-DMGB_CHUNK_ABLATE=3): dst must be zeroed, no
  // fused epilogues
  bool chunk_atomic_output();
  template <typename T>
  void cell_laplace_chunked(const CellData &cd, const ChunkDevice &ch, const T *x, T *out, const FusedEpilogue<T> &ep, cudaStream_t s);
  // out[idx[i]] += coef(idx[i]) * x[idx[i]]: the identity rows of the constrained DoFs after a fused cell loop (coef as the
  // epilogue applies it to A x: 1, -1, -f2 dinv)
  template <typename T>
  void idx_fused_identity(T *out, const T *x, const FusedEpilogue<T> &ep, const uint32_t *idx, uint32_t n, cudaStream_t s);

  // diag += diag(C^T A_e C) cell-wise (MatrixFreeTools::compute_diagonal)
  template <typename T> void cell_laplace_diagonal(const CellData &cd, T *diag, cudaStream_t s);

  // ---------------------------------------------------------------- coarse-grid solver (coarse.cu)
  // whole preconditioned CG (identity or Chebyshev(cheb_degree) preconditioner; cheb_degree = 0: none) on the dense,
  // symmetric n x n level matrix A in ONE thread block; adds the iteration count to *n_iterations_dev
  template <typename T>
  void coarse_cg_solve(const T *A, int n, const T *dinv, int cheb_degree, double theta, double delta, unsigned maxiter, double abstol,
                       double reltol, const T *b, T *x, int *n_iterations_dev, cudaStream_t s);

  // ---------------------------------------------------------------- transfers (transfer.cu)
  struct TransferData
  {
    int             degree_fine = 1, degree_coarse = 1;
    long long       n_coarse_cells = 0;
    const uint32_t *coarse_ctab    = nullptr; // packed 32-word evaluation records of the coarse cells (CellData::ctab)
    int             coarse_any_hanging = 0;
    bool            h_mode         = true;    // geometric (same degree, cell -> itself or 8 children) or polynomial (same cell)
    const uint32_t *fine_own       = nullptr; // own table of the fine cells
    const uint32_t *fine_first     = nullptr;
    const uint8_t  *n_children     = nullptr;
    const uint8_t  *weight_code    = nullptr; // per fine cell-entity
    double          P[(2 * MAX_DEGREE + 1) * (MAX_DEGREE + 1)] = {}; // host: 1-D embedding, NF x (pc+1) row-major
  };
  template <typename T> void transfer_prolongate_add(const TransferData &td, T *dst_fine, const T *src_coarse, cudaStream_t s);
  template <typename T> void transfer_restrict_add(const TransferData &td, T *dst_coarse, const T *src_fine, cudaStream_t s);
} // namespace mgb
