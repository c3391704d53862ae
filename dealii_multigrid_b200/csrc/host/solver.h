// Host-side mirror of the deal.II concepts the reference composes: distributed vector,
// matrix-free Operator, PreconditionChebyshev, MGTwoLevelTransfer / MGTransferGlobalCoarsening,
// MGTransferMatrixFree, Multigrid, PreconditionMG, SolverCG.  Control flow runs on the host and enqueues
// the sm_100a kernels of ../device on one CUDA stream; there is no CPU compute path.
#pragma once
#include "../device/device.h"
#include "dofmap.h"
#include "fe1d.h"

#include <array>
#include <cmath>
#include <deque>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace mgb
{
  // ------------------------------------------------------------------------------------------ context
  class Comm; // halo exchange + reductions between ranks (comm.h); nullptr = single rank
  struct Context
  {
    int          device = -1;
    cudaStream_t stream = nullptr;
    double      *red_result  = nullptr; // device scalar(s)
    double      *red_scratch = nullptr;
    unsigned    *red_ticket  = nullptr;
    double      *host_result = nullptr; // pinned
    Comm        *comm        = nullptr;
    // several ranks: second stream + events that let the ghost exchange of a cell loop run beside the interior cells
    // (fork/join with events, so the pattern is also captured into the V-cycle's CUDA graph)
    cudaStream_t comm_stream = nullptr;
    cudaEvent_t  ev_fork[2] = {nullptr, nullptr}, ev_join[2] = {nullptr, nullptr};
    bool         overlap_ready();
    long long    launches    = 0; // kernels enqueued by this library (bench.py's gpu_launches)
    void         ensure();
    // live timing of the dominant kernel (bench.py's roofline): CUDA-event pairs around every cell-loop
    // launch over exactly prof_cells cells while prof_enabled
    bool                                             prof_enabled = false;
    long long                                        prof_cells   = 0;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;
    size_t                                           prof_used = 0;
    void prof_begin(long long n_cells);
    void prof_end(long long n_cells);
  };
  Context &ctx();

  template <typename T> struct DeviceBuffer
  {
    T     *ptr = nullptr;
    size_t n   = 0;
    DeviceBuffer() = default;
    explicit DeviceBuffer(size_t n_) { alloc(n_); }
    DeviceBuffer(const std::vector<T> &h) { upload(h.data(), h.size()); }
    DeviceBuffer(const DeviceBuffer &) = delete;
    DeviceBuffer &operator=(const DeviceBuffer &) = delete;
    DeviceBuffer(DeviceBuffer &&o) noexcept { *this = std::move(o); }
    DeviceBuffer &operator=(DeviceBuffer &&o) noexcept
    {
      release();
      ptr   = o.ptr;
      n     = o.n;
      o.ptr = nullptr;
      o.n   = 0;
      return *this;
    }
    ~DeviceBuffer() { release(); }
    void release()
    {
      if (ptr)
        cudaFree(ptr);
      ptr = nullptr;
      n   = 0;
    }
    void alloc(size_t n_)
    {
      release();
      n = n_;
      if (n)
        MGB_CUDA(cudaMalloc((void **)&ptr, n * sizeof(T)));
    }
    void upload(const T *h, size_t n_)
    {
      alloc(n_);
      if (n)
        {
          // on the library's stream and complete on return: the kernels run on a (possibly non-blocking) user stream that a
          // copy on the legacy default stream would not be ordered against
          MGB_CUDA(cudaMemcpyAsync(ptr, h, n * sizeof(T), cudaMemcpyHostToDevice, ctx().stream));
          MGB_CUDA(cudaStreamSynchronize(ctx().stream));
        }
    }
  };

  // ------------------------------------------------------------------------------------------ partitioner
  // Device copy of Utilities::MPI::Partitioner + the exchange buffers.
  // peer-memory backing of one partitioner (host/p2p.cc, device/halo.cu): when present, ghost exchanges are two small
  // kernels storing straight into the neighbours' staging blocks over NVLink instead of NCCL send/recv groups
  struct P2PBacking
  {
    HaloDir               ghost, compress;
    std::shared_ptr<void> segment; // keeps the mapped segments alive
  };
  struct DevicePartitioner
  {
    Partitioner            host;
    DeviceBuffer<uint32_t> import_indices;
    // type-erased exchange buffers (sized for double) of the NCCL path
    DeviceBuffer<double>        send_buf, recv_buf;
    std::unique_ptr<P2PBacking> p2p;
    explicit DevicePartitioner(const Partitioner &p);
    ~DevicePartitioner();
    DevicePartitioner(const DevicePartitioner &) = delete;
    DevicePartitioner &operator=(const DevicePartitioner &) = delete;
    bool trivial() const { return host.n_ghost == 0 && host.import_indices.empty(); }
  };
  // collective over all ranks: back every partitioner created since the last call with peer-mapped staging memory;
  // returns false (NCCL path stays in use) on one rank, without peer access, or with MGB_P2P=0
  bool p2p_back_partitioners();
  void p2p_release_all(); // collective, before the communicator is destroyed
  void p2p_register(DevicePartitioner *p);
  void p2p_unregister(DevicePartitioner *p);

  // ------------------------------------------------------------------------------------------ vector
  template <typename T> struct DVec
  {
    T                                 *data    = nullptr;
    uint32_t                           n_owned = 0, n_ghost = 0;
    bool                               owns = false;
    std::shared_ptr<DevicePartitioner> part;

    DVec() = default;
    DVec(const DVec &) = delete;
    DVec &operator=(const DVec &) = delete;
    ~DVec() { release(); }
    void release()
    {
      if (owns && data)
        cudaFree(data);
      data = nullptr;
      owns = false;
    }
    size_t size_local() const { return size_t(n_owned) + n_ghost; }
    void   reinit(uint32_t owned, uint32_t ghost, std::shared_ptr<DevicePartitioner> p = nullptr)
    {
      if (!(owns && owned == n_owned && ghost == n_ghost))
        {
          release();
          n_owned = owned;
          n_ghost = ghost;
          if (size_local())
            MGB_CUDA(cudaMalloc((void **)&data, size_local() * sizeof(T)));
          owns = true;
        }
      part = p;
      zero();
    }
    void reinit_like(const DVec &o) { reinit(o.n_owned, o.n_ghost, o.part); }
    void wrap(T *ptr, uint32_t owned, uint32_t ghost, std::shared_ptr<DevicePartitioner> p = nullptr)
    {
      release();
      data    = ptr;
      n_owned = owned;
      n_ghost = ghost;
      part    = p;
    }
    void swap_data(DVec &o)
    {
      if (owns && o.owns)
        std::swap(data, o.data);
      else
        {
          // one side wraps foreign memory: exchange contents instead of pointers
          DeviceBuffer<T> tmp(n_owned);
          MGB_CUDA(cudaMemcpyAsync(tmp.ptr, data, n_owned * sizeof(T), cudaMemcpyDeviceToDevice, ctx().stream));
          MGB_CUDA(cudaMemcpyAsync(data, o.data, n_owned * sizeof(T), cudaMemcpyDeviceToDevice, ctx().stream));
          MGB_CUDA(cudaMemcpyAsync(o.data, tmp.ptr, n_owned * sizeof(T), cudaMemcpyDeviceToDevice, ctx().stream));
          MGB_CUDA(cudaStreamSynchronize(ctx().stream));
        }
    }
    // BLAS-1 on owned entries
    void zero()
    {
      if (size_local())
        {
          vec_set<T>(data, T(0), size_local(), ctx().stream);
          ++ctx().launches;
        }
    }
    void set(T a) { vec_set<T>(data, a, n_owned, ctx().stream), ++ctx().launches; }
    void add(T a, const DVec &x) { vec_add<T>(data, a, x.data, n_owned, ctx().stream), ++ctx().launches; }
    void sadd(T s, T a, const DVec &x) { vec_sadd<T>(data, s, a, x.data, n_owned, ctx().stream), ++ctx().launches; }
    void equ(T a, const DVec &x) { vec_equ<T>(data, a, x.data, n_owned, ctx().stream), ++ctx().launches; }
    void add_scalar(T a) { vec_add_scalar<T>(data, a, n_owned, ctx().stream), ++ctx().launches; }
    void copy_from(const DVec &x)
    {
      if (n_owned)
        MGB_CUDA(cudaMemcpyAsync(data, x.data, size_t(n_owned) * sizeof(T), cudaMemcpyDeviceToDevice, ctx().stream));
    }
    double dot(const DVec &y) const;
    double norm_sqr() const { return dot(*this); }
    double l2_norm() const { return std::sqrt(norm_sqr()); }
    double add_and_dot(T a, const DVec &h); // this += a h ; returns this.this
    // x += a d; this += a h; returns this.this  (this = CG residual g)
    double cg_update(DVec &x, T a, const DVec &d, const DVec &h);
    // same layout as o, zero-filled, without reallocating when the storage already fits (also for wrapped memory)
    void reinit_keep(const DVec &o)
    {
      if (data && n_owned == o.n_owned && n_ghost == o.n_ghost)
        {
          part = o.part;
          zero();
        }
      else
        reinit_like(o);
    }
    void reinit_keep_layout(const std::shared_ptr<DevicePartitioner> &p) // zero-filled vector of the partitioner's layout
    {
      if (data && owns && n_owned == p->host.n_owned && n_ghost == p->host.n_ghost)
        {
          part = p;
          zero();
        }
      else
        reinit(p->host.n_owned, p->host.n_ghost, p);
    }
    void ensure_like(const DVec &o) // as reinit_keep, but existing contents are kept (ghost entries are zero by invariant)
    {
      if (data && n_owned == o.n_owned && n_ghost == o.n_ghost)
        part = o.part;
      else
        reinit_like(o);
    }
    double mean_value(long long n_global) const;
    // ghost exchange (on the library's stream, or on `on` when given)
    void update_ghost_values(cudaStream_t on = nullptr) const;
    void compress_add(cudaStream_t on = nullptr);
    void zero_out_ghost_values() const
    {
      if (n_ghost)
        vec_set<T>(data + n_owned, T(0), n_ghost, ctx().stream), ++ctx().launches;
    }
  };

  // ------------------------------------------------------------------------------------------ solver control
  struct SolverControl
  {
    enum Kind
    {
      Reduction,       // ReductionControl(maxiter, abstol, reltol)
      IterationNumber  // IterationNumberControl(maxiter, tol): success at it == maxiter or res <= tol
    } kind = Reduction;
    unsigned maxiter = 10000;
    double   abstol = 1e-20, reltol = 1e-4;
    // state
    double   initial_value = 0, reduced_tol = 0, last_value = 0;
    unsigned last_step = 0;
    enum State
    {
      iterate,
      success,
      failure
    };
    State check(unsigned step, double value)
    {
      last_step  = step;
      last_value = value;
      if (kind == Reduction)
        {
          if (step == 0)
            {
              initial_value = value;
              reduced_tol   = value * reltol;
            }
          if (value <= reduced_tol)
            return success;
        }
      else if (step >= maxiter)
        return success; // IterationNumberControl: reaching maxiter is success
      if (value <= abstol)
        return success;
      if (step >= maxiter || std::isnan(value))
        return failure;
      return iterate;
    }
  };

  // SolverCG::solve.  A(dst,src), P(dst,src) are callables; identity_precond skips P.
  // Returns the number of iterations; optionally records the Lanczos coefficients for eigenvalue estimates.
  template <typename T> struct CGWork
  {
    DVec<T> g, d, h;
  };
  template <typename T, class Mat, class Prec>
  unsigned cg_solve(Mat &&A, Prec &&P, const bool identity_precond, DVec<T> &x, const DVec<T> &b, SolverControl &ctl, CGWork<T> &w,
                    const bool x_is_zero, std::vector<double> *diagonal = nullptr, std::vector<double> *offdiagonal = nullptr)
  {
    w.g.reinit_like(x);
    w.d.reinit_like(x);
    w.h.reinit_like(x);
    DVec<T> &g = w.g, &d = w.d, &h = w.h;
    if (!x_is_zero)
      {
        A(g, x);
        g.add(T(-1), b);
      }
    else
      g.equ(T(-1), b);
    double   res = g.l2_norm();
    unsigned it  = 0;
    if (ctl.check(0, res) != SolverControl::iterate)
      return 0;
    double gh = 0, eigen_beta_alpha = 0, prev_alpha = 0;
    while (true)
      {
        ++it;
        if (it > 1)
          {
            const double gh_old = gh;
            if (!identity_precond)
              {
                P(h, g);
                gh = g.dot(h);
              }
            else
              gh = res * res;
            const double beta = gh / gh_old;
            d.sadd(T(beta), T(-1), identity_precond ? g : h);
            // Lanczos entry of iteration it-1 (deal.II pushes it after that iteration's failed convergence
            // check): T_kk = 1/alpha_k + beta_k/alpha_{k-1},  T_{k,k+1} = sqrt(beta_{k+1})/alpha_k
            if (diagonal)
              {
                diagonal->push_back(1.0 / prev_alpha + eigen_beta_alpha);
                offdiagonal->push_back(std::sqrt(beta) / prev_alpha);
                eigen_beta_alpha = beta / prev_alpha;
              }
          }
        else
          {
            if (!identity_precond)
              {
                P(h, g);
                d.equ(T(-1), h);
                gh = g.dot(h);
              }
            else
              {
                d.equ(T(-1), g);
                gh = res * res;
              }
          }
        A(h, d);
        double alpha = d.dot(h);
        alpha        = gh / alpha;
        prev_alpha   = alpha;
        res = std::sqrt(std::fabs(g.cg_update(x, T(alpha), d, h)));
        if (ctl.check(it, res) != SolverControl::iterate)
          break;
      }
    return it;
  }

  // largest eigenvalue of the symmetric tridiagonal Lanczos matrix (LAPACK in deal.II; bisection here)
  double tridiagonal_max_eigenvalue(const std::vector<double> &diag, const std::vector<double> &offdiag);
  double tridiagonal_min_eigenvalue(const std::vector<double> &diag, const std::vector<double> &offdiag);

  // ------------------------------------------------------------------------------------------ operator
  // Operator<dim,1,Number> of include/operator.h
  struct OperatorSetup
  {
    int                   degree = 1;
    long long             n_cells = 0;
    const uint32_t       *entity_table = nullptr; // n_cells*27
    const uint16_t       *mask         = nullptr; // n_cells (may be null)
    const uint8_t        *cell_level   = nullptr; // n_cells
    double                root_length  = 2.0;
    const uint32_t       *constrained  = nullptr;
    uint32_t              n_constrained = 0;
    const uint32_t       *edge_constrained = nullptr; // LS levels (include/operator.h:49-66)
    uint32_t              n_edge_constrained = 0;
    bool                  has_edge_constrained_global = false;
    const uint32_t       *edge_cells = nullptr; // cells touching an edge DoF (include/operator.h:71-99)
    uint32_t              n_edge_cells = 0;
    long long             n_dofs_global = 0;
    const Partitioner    *partitioner = nullptr;
    long long             first_cell_global = 0; // position of cell 0 in the global Morton order (aligns the chunk plan)
  };

  template <typename T> class OperatorT
  {
  public:
    explicit OperatorT(const OperatorSetup &s);
    long long m() const { return n_dofs_global_; }
    void      initialize_dof_vector(DVec<T> &v) const { v.reinit(part_->host.n_owned, part_->host.n_ghost, part_); }
    void      vmult(DVec<T> &dst, const DVec<T> &src) const;                // include/operator.h:152-183
    void      Tvmult(DVec<T> &dst, const DVec<T> &src) const { vmult(dst, src); }
    // vmult for callers that own a pair of temporaries (the Chebyshev loop): dst_is_zero = the caller knows that dst holds
    // zeros (owned entries and ghost tail), so the memset is skipped; also_zero (may be null; neither dst nor src, same
    // layout) is cleared by the cell loop itself where the Cartesian atomic kernel runs (k_laplace_apply<P, T, true>).
    // Returns whether also_zero now holds zeros.
    bool      vmult_zeroing(DVec<T> &dst, const DVec<T> &src, bool dst_is_zero, DVec<T> *also_zero) const;
    void      vmult_interface_down(DVec<T> &dst, const DVec<T> &src) const; // :191-201
    // dst = b - A src with A = vmult_interface_down: Multigrid::level_v_step's residual (matrix->vmult followed by
    // t.sadd(-1, 1, defect)) in one cell loop
    void      residual_interface_down(DVec<T> &dst, const DVec<T> &b, const DVec<T> &src) const;
    void      vmult_interface_up(DVec<T> &dst, const DVec<T> &src) const;   // :203-226
    void      compute_inverse_diagonal(DVec<T> &diag) const;                // :228-242
    // general (non-Cartesian) geometry: 8 vertices per cell (lexicographic, x fastest), trilinear map (MappingQ1); the
    // cell loops then run the quadrature-point kernel (device/operator_q1.cuh) instead of the Cartesian one
    void      set_cell_vertices(const double *xyz /* n_cells*8*3 */);
    // Fused cell loop (device/chunk.cuh): out = E(A x), E = FusedEpilogue (residual or Chebyshev update), identity rows of
    // the constrained DoFs included.  Available for Cartesian cells without refinement-edge DoFs on vectors without ghost
    // exchange; otherwise the caller runs vmult followed by the separate vector update.  out must not alias x.
    bool      fused_available(const DVec<T> &out, const DVec<T> &x) const;
    void      vmult_fused(DVec<T> &out, const DVec<T> &x, const FusedEpilogue<T> &ep) const;
    bool      chunked() const { return chunked_; }
    bool      has_edge_dofs() const { return edge_.n != 0; }
    const std::shared_ptr<DevicePartitioner> &partitioner() const { return part_; }
    int       degree() const { return cd_.degree; }
    const CellData &cell_data() const { return cd_; }
    uint32_t  n_constrained() const { return (uint32_t)constrained_.n; }
    const uint32_t *constrained_dev() const { return constrained_.ptr; }

  private:
    // also_zero: a vector (not dst, not src) that the cell loop may clear on the side; *did_zero tells whether it did
    void cell_loop(DVec<T> &dst, const DVec<T> &src, bool zero_dst, bool edge_cells_only, double alpha = 1.0, DVec<T> *also_zero = nullptr,
                   bool *did_zero = nullptr) const;
    CellData                           cd_;
    DeviceBuffer<uint32_t>             etab_, ctab_;
    DeviceBuffer<uint16_t>             mask_;
    DeviceBuffer<uint8_t>              level_;
    DeviceBuffer<double>               vertices_;
    DeviceBuffer<uint32_t>             constrained_, edge_, edge_cells_;
    // several ranks: cells that read ghost entries (boundary) and the others (interior), for overlapping the exchange
    DeviceBuffer<uint32_t>             boundary_cells_, interior_cells_;
    DeviceBuffer<uint32_t>             plain_cells_, hanging_cells_; // MGB_SPLIT_HANGING=1: cells without / with hanging-node constraints
    mutable DeviceBuffer<T>            edge_src_vals_;
    mutable DVec<T>                    src_copy_;
    // chunk plan of the cell loop (host/chunks.h) on the device
    bool                               chunked_ = false;
    ChunkDevice                        chd_;
    DeviceBuffer<uint32_t>             ch_desc_, ch_pool_, ch_deps_, ch_ghost_, ch_flags_, ch_state_;
    DeviceBuffer<T>                    ch_side_;
    bool chunk_path_ok(const T *out, const T *x) const;
    bool                               has_edge_global_ = false;
    long long                          n_dofs_global_   = 0;
    std::shared_ptr<DevicePartitioner> part_;
  };

  // ------------------------------------------------------------------------------------------ Chebyshev
  // PreconditionChebyshev<Operator, Vector, DiagonalMatrix>
  template <typename T> class ChebyshevT
  {
  public:
    ChebyshevT(const OperatorT<T> &op, int degree, double smoothing_range, unsigned eig_cg_n_iterations);
    void vmult(DVec<T> &dst, const DVec<T> &rhs);
    void step(DVec<T> &dst, const DVec<T> &rhs);
    void iterate(DVec<T> &dst, const DVec<T> &rhs, bool zero_guess);
    void estimate_eigenvalues();
    double lambda_max() const { return lambda_max_; }
    double theta() const { return theta_; }
    double delta() const { return delta_; }
    int    degree() const { return degree_; }
    double lambda_min_estimate() const { return lambda_min_est_; }
    bool   eigenvalues_done() const { return eig_done_; }
    const DVec<T> &inverse_diagonal() const { return dinv_; }

  private:
    const OperatorT<T> &op_;
    int                 degree_;
    double              smoothing_range_;
    unsigned            eig_its_;
    bool                eig_done_ = false;
    bool                fuse_     = true; // vector updates inside the cell loop (MGB_FUSE_CHEB=0: separate pass)
    double              lambda_max_ = 0, lambda_min_est_ = 0, theta_ = 1, delta_ = 1;
    DVec<T>             dinv_, solution_old_, temp1_, temp2_;
    // the unfused loop alternates its A x between temp1_ and temp2_; the cell loop that fills one clears the other
    // (opt-in MGB_ZERO_IN_CELL_LOOP=1; the default is a memset before every cell loop, which is faster)
    bool                zero_in_loop_ = false;
    CGWork<T>           cgw_;
  };

  // ------------------------------------------------------------------------------------------ transfers
  template <typename T> class TwoLevelTransferT
  {
  public:
    // geometric or polynomial two-level transfer between two DoF maps
    TwoLevelTransferT(const DoFMap &fine, const DoFMap &coarse, const TransferMap &tm, const OperatorT<T> &op_fine,
                      const OperatorT<T> &op_coarse);
    void prolongate_and_add(DVec<T> &dst_fine, const DVec<T> &src_coarse) const;
    void restrict_and_add(DVec<T> &dst_coarse, const DVec<T> &src_fine) const;

  private:
    TransferData           td_;
    DeviceBuffer<uint32_t> fine_own_, fine_first_;
    std::shared_ptr<DevicePartitioner> fine_part_; // several ranks: layout of the internal ghosted fine vector
    mutable DVec<T>                    vec_fine_;
    DeviceBuffer<uint8_t>  n_children_, weight_code_;
    const OperatorT<T>    &op_fine_, &op_coarse_;
  };

  template <typename T> class MGTransferBase
  {
  public:
    virtual ~MGTransferBase() = default;
    virtual void prolongate_and_add(unsigned to_level, DVec<T> &dst, const DVec<T> &src) const = 0;
    virtual void restrict_and_add(unsigned from_level, DVec<T> &dst, const DVec<T> &src) const = 0;
    // double-precision outer vector <-> level vectors
    // skip_top: the top-level defect already aliases src (PreconditionMGT), only the lower levels are (re)initialised
    virtual void copy_to_mg(std::vector<DVec<T>> &defect, unsigned min_level, unsigned max_level, const DVec<double> &src,
                            bool skip_top = false) const = 0;
    // layout of the level vectors on `level` (initialize_dof_vector of that level's operator)
    virtual void level_layout(unsigned level, uint32_t &n_owned, uint32_t &n_ghost, std::shared_ptr<DevicePartitioner> &part) const = 0;
    virtual void copy_from_mg(DVec<double> &dst, const std::vector<DVec<T>> &solution, unsigned min_level, unsigned max_level) const = 0;
    // true if copy_to_mg / copy_from_mg only touch the top level and are plain copies there (global coarsening): the
    // V-cycle may then work directly on the caller's vectors
    virtual bool top_level_is_plain_copy() const { return true; }
  };

  // MGTransferGlobalCoarsening: vector of two-level transfers; transfers[l] connects l-1 and l
  template <typename T> class GCTransferT : public MGTransferBase<T>
  {
  public:
    unsigned                                        min_level = 0;
    std::vector<std::unique_ptr<TwoLevelTransferT<T>>> transfers; // index l - min_level (entry 0 unused)
    std::vector<const OperatorT<T> *>               operators;   // index l - min_level
    void prolongate_and_add(unsigned to_level, DVec<T> &dst, const DVec<T> &src) const override
    {
      transfers[to_level - min_level]->prolongate_and_add(dst, src);
    }
    void restrict_and_add(unsigned from_level, DVec<T> &dst, const DVec<T> &src) const override
    {
      transfers[from_level - min_level]->restrict_and_add(dst, src);
    }
    void copy_to_mg(std::vector<DVec<T>> &defect, unsigned minl, unsigned maxl, const DVec<double> &src, bool skip_top = false) const override;
    void level_layout(unsigned level, uint32_t &n_owned, uint32_t &n_ghost, std::shared_ptr<DevicePartitioner> &part) const override
    {
      part    = operators[level - min_level]->partitioner();
      n_owned = part->host.n_owned;
      n_ghost = part->host.n_ghost;
    }
    void copy_from_mg(DVec<double> &dst, const std::vector<DVec<T>> &solution, unsigned minl, unsigned maxl) const override;
  };

  // MGTransferMatrixFree (local smoothing; multigrid_throughput.cc:1800-1802): level-to-level transfers
  // over refined parents only + copy_indices between the active-mesh vector and the level vectors
  template <typename T> class LSTransferT : public MGTransferBase<T>
  {
  public:
    std::vector<std::unique_ptr<TwoLevelTransferT<T>>> transfers; // index l (entry 0 unused)
    std::vector<const OperatorT<T> *>                  operators; // index l
    struct Copy
    {
      DeviceBuffer<uint32_t> active_idx, level_idx;
      // several ranks: layout of the auxiliary ghosted level vector the pairs refer to (CopyIndices::part)
      std::shared_ptr<DevicePartitioner> part;
      mutable DVec<T>                    tmp;
    };
    std::deque<Copy> copy; // index l (deque: Copy holds a non-movable vector)
    void              set_copy_indices(unsigned level, const CopyIndices &ci)
    {
      if (copy.size() <= level)
        copy.resize(level + 1);
      copy[level].active_idx.upload(ci.active_idx.data(), ci.active_idx.size());
      copy[level].level_idx.upload(ci.level_idx.data(), ci.level_idx.size());
      if (ci.distributed)
        copy[level].part = std::make_shared<DevicePartitioner>(ci.part);
    }
    void prolongate_and_add(unsigned to_level, DVec<T> &dst, const DVec<T> &src) const override
    {
      transfers[to_level]->prolongate_and_add(dst, src);
    }
    void restrict_and_add(unsigned from_level, DVec<T> &dst, const DVec<T> &src) const override
    {
      transfers[from_level]->restrict_and_add(dst, src);
    }
    void copy_to_mg(std::vector<DVec<T>> &defect, unsigned minl, unsigned maxl, const DVec<double> &src, bool skip_top = false) const override;
    void level_layout(unsigned level, uint32_t &n_owned, uint32_t &n_ghost, std::shared_ptr<DevicePartitioner> &part) const override
    {
      part    = operators[level]->partitioner();
      n_owned = part->host.n_owned;
      n_ghost = part->host.n_ghost;
    }
    void copy_from_mg(DVec<double> &dst, const std::vector<DVec<T>> &solution, unsigned minl, unsigned maxl) const override;
    bool top_level_is_plain_copy() const override { return false; }
  };

  // ------------------------------------------------------------------------------------------ multigrid
  template <typename T> class MGCoarseBase
  {
  public:
    virtual ~MGCoarseBase() = default;
    virtual void operator()(unsigned level, DVec<T> &dst, const DVec<T> &src) = 0;
  };

  // MGCoarseGridIterativeSolver<SolverCG, Operator, Identity|Chebyshev> (multigrid_throughput.cc:911-944)
  template <typename T> class CoarseCGT : public MGCoarseBase<T>
  {
  public:
    CoarseCGT(const OperatorT<T> &op, ChebyshevT<T> *precond, unsigned maxiter, double abstol, double reltol)
      : op_(op)
      , precond_(precond)
    {
      ctl_.kind    = SolverControl::Reduction;
      ctl_.maxiter = maxiter;
      ctl_.abstol  = abstol;
      ctl_.reltol  = reltol;
    }
    void operator()(unsigned, DVec<T> &dst, const DVec<T> &src) override;
    // iterations of the current solve accumulated on the device by the single-kernel path (read once per outer solve)
    long long collect_device_iterations();
    unsigned  last_iterations  = 0;
    long long total_iterations = 0;
    static constexpr int max_dense_size = 512; // coarse problems up to this many DoFs are solved by one thread block

  private:
    const OperatorT<T> &op_;
    ChebyshevT<T>      *precond_;
    SolverControl       ctl_;
    CGWork<T>           work_;
    // dense path
    bool                dense_checked_ = false, dense_ = false;

  public:
    bool                single_kernel() const { return dense_checked_ && dense_; }

  private:
    DeviceBuffer<T>     dense_A_;
    DeviceBuffer<int>   dense_its_;
    void                setup_dense();
  };

  enum MGStage
  {
    STAGE_PRE = 0,
    STAGE_RESIDUAL,
    STAGE_RESTRICTION,
    STAGE_COARSE,
    STAGE_PROLONGATION,
    STAGE_EDGE_PROLONGATION,
    STAGE_POST,
    N_STAGES
  };
  // stage timer: CUDA-event pairs recorded on the stream, resolved after the solve
  struct StageTimers
  {
    bool enabled = false;
    struct Rec
    {
      cudaEvent_t a, b;
      int         level, stage;
    };
    std::vector<Rec>         recs;
    std::vector<size_t>      open; // indices of the records whose end event is still to come
    std::vector<cudaEvent_t> pool;
    size_t                   pool_next = 0;
    cudaEvent_t              get();
    void                     begin(int level, int stage);
    void                     end();
    // accumulate into times[level][stage] (seconds) and reset
    void collect(std::vector<std::array<double, N_STAGES + 2>> &times, int min_level);
    ~StageTimers();
  };

  template <typename T> class MultigridT
  {
  public:
    unsigned                            min_level = 0, max_level = 0;
    std::vector<const OperatorT<T> *>   matrices;   // index l - min_level  (mg::Matrix of MGInterfaceOperator -> vmult_interface_down)
    std::vector<ChebyshevT<T> *>        smoothers;  // index l - min_level
    const MGTransferBase<T>            *transfer = nullptr;
    MGCoarseBase<T>                    *coarse   = nullptr;
    bool                                edge_in  = false; // set_edge_in_matrix (multigrid_throughput.cc:1101-1105)
    std::vector<DVec<T>>                defect, solution, t; // index l - min_level
    StageTimers                        *timers = nullptr;

    void resize_levels()
    {
      const size_t n = max_level - min_level + 1;
      if (defect.size() != n)
        {
          defect   = std::vector<DVec<T>>(n);
          solution = std::vector<DVec<T>>(n);
          t        = std::vector<DVec<T>>(n);
        }
    }
    void cycle(); // Multigrid::cycle (V-cycle)
    void level_v_step(unsigned level);
  };

  // PreconditionMG: copy_to_mg, cycle, copy_from_mg
  template <typename T> class PreconditionMGT
  {
  public:
    MultigridT<T>           *mg       = nullptr;
    const MGTransferBase<T> *transfer = nullptr;
    StageTimers             *timers   = nullptr;
    void                     vmult(DVec<double> &dst, const DVec<double> &src);
    // CUDA-graph replay of the V-cycle: the level structure is static, so once every lazy initialisation has happened
    // (eigenvalue estimates, dense coarse matrix, vector allocations: the owner sets allow_graph after a first solve)
    // the whole launch sequence between two fixed vectors is captured once and replayed.  Not used while stage timers
    // or the per-kernel profiling events are active, on the legacy default stream, or with several ranks.
    bool allow_graph = false;
    PreconditionMGT();
    ~PreconditionMGT();
    PreconditionMGT(const PreconditionMGT &) = delete;
    PreconditionMGT &operator=(const PreconditionMGT &) = delete;
    void release_graphs(); // destroy the captured graphs (they are re-captured on demand)

  private:
    void vmult_eager(DVec<double> &dst, const DVec<double> &src);
    struct GraphEntry
    {
      const void     *src, *dst;
      cudaGraphExec_t exec;
      long long       launches;
    };
    std::vector<GraphEntry> graphs_;
  };

  // graphs that captured NCCL calls keep the communicator referenced: release them all before it is destroyed
  void release_all_graphs();

  // MGCoarseGridApplyPreconditioner (multigrid_throughput.cc:422-564): nested V-cycle as coarse solver
  template <typename T> class CoarseApplyMGT : public MGCoarseBase<T>
  {
  public:
    PreconditionMGT<T> *pmg = nullptr;
    DVec<double>        src_d, dst_d;
    void                operator()(unsigned level, DVec<T> &dst, const DVec<T> &src) override;
  };
} // namespace mgb
