#include "solver.h"

#include <cstdio>
#include <cstdlib>

#include "chunks.h"
#include "comm.h"

#include <algorithm>
#include <cmath>
#include <cstring>

namespace mgb
{
  // ------------------------------------------------------------------------------------------ context
  Context &ctx()
  {
    static Context c;
    return c;
  }

  void Context::ensure()
  {
    if (red_result)
      return;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0)
      throw std::runtime_error("mgb200: no CUDA device available (this library has no CPU path)");
    if (device < 0)
      {
        device = 0;
        MGB_CUDA(cudaSetDevice(device));
      }
    MGB_CUDA(cudaMalloc((void **)&red_result, 8 * sizeof(double)));
    MGB_CUDA(cudaMalloc((void **)&red_scratch, reduction_scratch_size() * sizeof(double)));
    MGB_CUDA(cudaMalloc((void **)&red_ticket, sizeof(unsigned)));
    MGB_CUDA(cudaMemset(red_ticket, 0, sizeof(unsigned)));
    MGB_CUDA(cudaMallocHost((void **)&host_result, 8 * sizeof(double)));
  }

  bool Context::overlap_ready()
  {
    static const bool enabled = [] {
      // opt-in: measured on 2 and 8 B200 (hypercube L=7, p=4) the split cell loop hides nothing that matters -- the
      // exchange latency that limits strong scaling sits on the coarse levels, whose interior kernels are shorter than
      // one exchange (profiles/r1h_multigpu_notes.txt)
      const char *e = std::getenv("MGB_OVERLAP");
      return e && e[0] == '1';
    }();
    if (!enabled || !comm || comm->size() < 2 || stream == nullptr)
      return false;
    if (!comm_stream)
      {
        MGB_CUDA(cudaStreamCreateWithFlags(&comm_stream, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i)
          {
            MGB_CUDA(cudaEventCreateWithFlags(&ev_fork[i], cudaEventDisableTiming));
            MGB_CUDA(cudaEventCreateWithFlags(&ev_join[i], cudaEventDisableTiming));
          }
      }
    return true;
  }

  void Context::prof_begin(long long n_cells)
  {
    if (!prof_enabled || n_cells != prof_cells)
      return;
    if (prof_used == prof_events.size())
      {
        cudaEvent_t a, b;
        MGB_CUDA(cudaEventCreate(&a));
        MGB_CUDA(cudaEventCreate(&b));
        prof_events.emplace_back(a, b);
      }
    MGB_CUDA(cudaEventRecord(prof_events[prof_used].first, stream));
  }
  void Context::prof_end(long long n_cells)
  {
    if (!prof_enabled || n_cells != prof_cells)
      return;
    MGB_CUDA(cudaEventRecord(prof_events[prof_used].second, stream));
    ++prof_used;
  }

  DevicePartitioner::DevicePartitioner(const Partitioner &p)
    : host(p)
  {
    ctx().ensure();
    if (!p.import_indices.empty())
      {
        import_indices.upload(p.import_indices.data(), p.import_indices.size());
        send_buf.alloc(std::max<size_t>(p.import_indices.size(), p.n_ghost));
        recv_buf.alloc(std::max<size_t>(p.import_indices.size(), p.n_ghost));
      }
    else if (p.n_ghost)
      {
        send_buf.alloc(p.n_ghost);
        recv_buf.alloc(p.n_ghost);
      }
    if (ctx().comm && ctx().comm->size() > 1)
      p2p_register(this);
  }
  DevicePartitioner::~DevicePartitioner() { p2p_unregister(this); }

  // ------------------------------------------------------------------------------------------ vector
  namespace
  {
    bool p2p_fused()
    {
      static const bool on = [] {
        const char *e = std::getenv("MGB_P2P_FUSED");
        return !(e && e[0] == '0');
      }();
      return on;
    }
    // Where a reduction kernel writes its scalar.  One rank: straight into the pinned host word (cudaMallocHost memory is
    // device-accessible under unified addressing), so reading it back needs no copy -- a cudaMemcpyAsync of 8 bytes would
    // queue on the device-to-host copy engine behind whatever bulk download is in flight (mgb_problem_solve_host_batch).
    // Several ranks: device memory, because ncclAllReduce works on it.
    double *reduction_target()
    {
      Context &c = ctx();
      return (c.comm && c.comm->size() > 1) ? c.red_result : c.host_result;
    }
    double finish_reduction()
    {
      Context &c = ctx();
      if (c.comm && c.comm->size() > 1)
        {
          c.comm->allreduce_sum(c.red_result, 1, c.stream);
          MGB_CUDA(cudaMemcpyAsync(c.host_result, c.red_result, sizeof(double), cudaMemcpyDeviceToHost, c.stream));
        }
      MGB_CUDA(cudaStreamSynchronize(c.stream));
      return c.host_result[0];
    }
  } // namespace

  template <typename T> double DVec<T>::dot(const DVec &y) const
  {
    Context &c = ctx();
    vec_dot<T>(data, y.data, n_owned, reduction_target(), c.red_scratch, c.red_ticket, c.stream);
    ++c.launches;
    return finish_reduction();
  }
  template <typename T> double DVec<T>::add_and_dot(T a, const DVec &h)
  {
    Context &c = ctx();
    vec_add_and_dot<T>(data, a, h.data, n_owned, reduction_target(), c.red_scratch, c.red_ticket, c.stream);
    ++c.launches;
    return finish_reduction();
  }
  template <typename T> double DVec<T>::cg_update(DVec &x, T a, const DVec &d, const DVec &h)
  {
    Context &c = ctx();
    vec_cg_update<T>(x.data, data, a, d.data, h.data, n_owned, reduction_target(), c.red_scratch, c.red_ticket, c.stream);
    ++c.launches;
    return finish_reduction();
  }
  template <typename T> double DVec<T>::mean_value(long long n_global) const
  {
    Context &c = ctx();
    vec_sum<T>(data, n_owned, reduction_target(), c.red_scratch, c.red_ticket, c.stream);
    ++c.launches;
    return finish_reduction() / double(n_global);
  }

  template <typename T> void DVec<T>::update_ghost_values(cudaStream_t on) const
  {
    if (!part || part->trivial())
      return;
    Context           &c      = ctx();
    const cudaStream_t stream = on ? on : c.stream;
    if (!c.comm)
      throw std::runtime_error("update_ghost_values: vector has ghosts but no communicator is set");
    if (part->p2p)
      {
        // one fused kernel on the library's own stream; the send/recv pair when the exchange runs beside cell kernels on
        // the second stream (MGB_OVERLAP) or with MGB_P2P_FUSED=0
        const bool fused = on == nullptr && p2p_fused();
        halo_exchange<T>(part->p2p->ghost, true, data, n_owned, part->import_indices.ptr, stream, fused);
        c.launches += fused ? 1 : 2;
        return;
      }
    const Partitioner &p   = part->host;
    T                 *snd = reinterpret_cast<T *>(part->send_buf.ptr);
    if (!p.import_indices.empty())
      {
        idx_gather<T>(snd, data, part->import_indices.ptr, (uint32_t)p.import_indices.size(), stream);
        ++c.launches;
      }
    std::vector<Msg> sends, recvs;
    for (size_t i = 0; i < p.import_ranks.size(); ++i)
      sends.push_back(Msg{p.import_ranks[i], snd + p.import_offsets[i], size_t(p.import_offsets[i + 1] - p.import_offsets[i]) * sizeof(T)});
    for (size_t i = 0; i < p.ghost_ranks.size(); ++i)
      recvs.push_back(
        Msg{p.ghost_ranks[i], data + n_owned + p.ghost_offsets[i], size_t(p.ghost_offsets[i + 1] - p.ghost_offsets[i]) * sizeof(T)});
    c.comm->exchange(sends, recvs, stream);
  }

  template <typename T> void DVec<T>::compress_add(cudaStream_t on)
  {
    if (!part || part->trivial())
      return;
    Context           &c      = ctx();
    const cudaStream_t stream = on ? on : c.stream;
    if (!c.comm)
      throw std::runtime_error("compress: vector has ghosts but no communicator is set");
    if (part->p2p)
      {
        // the send kernel zeroes the ghost tail after reading it (compress ends with zero_out_ghost_values)
        const bool fused = on == nullptr && p2p_fused();
        halo_exchange<T>(part->p2p->compress, false, data, n_owned, part->import_indices.ptr, stream, fused);
        c.launches += fused ? 1 : 2;
        return;
      }
    const Partitioner &p   = part->host;
    T                 *rcv = reinterpret_cast<T *>(part->recv_buf.ptr);
    std::vector<Msg>   sends, recvs;
    for (size_t i = 0; i < p.ghost_ranks.size(); ++i)
      sends.push_back(
        Msg{p.ghost_ranks[i], data + n_owned + p.ghost_offsets[i], size_t(p.ghost_offsets[i + 1] - p.ghost_offsets[i]) * sizeof(T)});
    for (size_t i = 0; i < p.import_ranks.size(); ++i)
      recvs.push_back(Msg{p.import_ranks[i], rcv + p.import_offsets[i], size_t(p.import_offsets[i + 1] - p.import_offsets[i]) * sizeof(T)});
    c.comm->exchange(sends, recvs, stream);
    if (!p.import_indices.empty())
      {
        idx_scatter_add<T>(data, rcv, part->import_indices.ptr, (uint32_t)p.import_indices.size(), stream);
        ++c.launches;
      }
    if (n_ghost)
      vec_set<T>(data + n_owned, T(0), n_ghost, stream), ++c.launches;
  }

  // ------------------------------------------------------------------------------------------ tridiagonal eigenvalues
  namespace
  {
    // number of eigenvalues of the symmetric tridiagonal matrix smaller than x (Sturm count)
    int sturm_count(const std::vector<double> &d, const std::vector<double> &e, double x)
    {
      int    count = 0;
      double q     = 1.0;
      for (size_t i = 0; i < d.size(); ++i)
        {
          const double e2 = (i == 0) ? 0.0 : e[i - 1] * e[i - 1];
          q               = d[i] - x - (i == 0 ? 0.0 : e2 / q);
          if (q == 0.0)
            q = 1e-300;
          if (q < 0)
            ++count;
        }
      return count;
    }
    double kth_eigenvalue(const std::vector<double> &d, const std::vector<double> &e, int k)
    {
      double lo = 1e300, hi = -1e300;
      for (size_t i = 0; i < d.size(); ++i)
        {
          const double r = (i > 0 ? std::fabs(e[i - 1]) : 0.0) + (i + 1 < d.size() ? std::fabs(e[i]) : 0.0);
          lo             = std::min(lo, d[i] - r);
          hi             = std::max(hi, d[i] + r);
        }
      for (int it = 0; it < 200; ++it)
        {
          const double mid = 0.5 * (lo + hi);
          if (mid <= lo || mid >= hi)
            break;
          if (sturm_count(d, e, mid) > k)
            hi = mid;
          else
            lo = mid;
        }
      return 0.5 * (lo + hi);
    }
  } // namespace

  double tridiagonal_max_eigenvalue(const std::vector<double> &diag, const std::vector<double> &offdiag)
  {
    return kth_eigenvalue(diag, offdiag, int(diag.size()) - 1);
  }
  double tridiagonal_min_eigenvalue(const std::vector<double> &diag, const std::vector<double> &offdiag)
  {
    return kth_eigenvalue(diag, offdiag, 0);
  }

  // ------------------------------------------------------------------------------------------ operator
  template <typename T> OperatorT<T>::OperatorT(const OperatorSetup &s)
  {
    ctx().ensure();
    upload_fe_constants(s.degree);
    etab_.upload(s.entity_table, size_t(s.n_cells) * 27);
    level_.upload(s.cell_level, size_t(s.n_cells));
    int any_hanging = 0;
    if (s.mask)
      {
        for (long long c = 0; c < s.n_cells; ++c)
          any_hanging |= (s.mask[c] >> 3) != 0;
        mask_.upload(s.mask, size_t(s.n_cells));
      }
    {
      // packed 32-word cell records of the vmult kernel (operator.cuh)
      std::vector<uint32_t> rec(size_t(s.n_cells) * 32, 0u);
      for (long long c = 0; c < s.n_cells; ++c)
        {
          std::memcpy(&rec[size_t(c) * 32], s.entity_table + size_t(c) * 27, 27 * sizeof(uint32_t));
          rec[size_t(c) * 32 + 27] = uint32_t(s.mask ? s.mask[c] : 0) | (uint32_t(s.cell_level[c]) << 16);
        }
      ctab_.upload(rec.data(), rec.size());
    }
    // Cells with hanging-node constraints in a launch of their own: a warp of the cell kernel takes the shared-memory detour
    // as soon as ONE of its cells has a constrained face or edge, so mixing them in slows the unconstrained majority down
    // (quadrant L=7 p=4: 66.7 -> 78.8 GDoF/s, profiles/r2a_split_hanging.txt).  On by default for levels large enough that a
    // second launch does not matter; MGB_SPLIT_HANGING=0 / 1 forces it off / on.
    const char *split_env = std::getenv("MGB_SPLIT_HANGING");
    const bool  split     = split_env ? split_env[0] == '1' : s.n_cells >= 8192;
    if (any_hanging && split)
      {
        std::vector<uint32_t> plain, hanging;
        for (long long c = 0; c < s.n_cells; ++c)
          ((s.mask[c] >> 3) ? hanging : plain).push_back(uint32_t(c));
        plain_cells_.upload(plain.data(), plain.size());
        hanging_cells_.upload(hanging.data(), hanging.size());
      }
    {
      // chunk plan of the cell loop (device/chunk.cuh): opt-in with MGB_CHUNKED=1.  The chunked loop needs no atomics on
      // owned entries (bitwise reproducible results), no zeroing of dst, and carries the smoother's vector update in its
      // epilogue, but measured on B200 (profiles/r2f_*) it is 1.6-1.9x slower than the atomic kernel (latency of the
      // per-chunk phases), so the atomic kernel stays the default
      const char *e = std::getenv("MGB_CHUNKED");
      if (e && e[0] == '1' && s.n_cells > 0)
        try
          {
            const ChunkPlan plan = build_chunk_plan(s.degree, s.n_cells, s.entity_table, s.mask, s.cell_level, s.partitioner ? s.partitioner->n_owned : 0u,
                                                    s.partitioner ? s.partitioner->n_ghost : 0u, (int)sizeof(T), s.first_cell_global);
            bool deps_fit = true;
            for (uint32_t k = 0; k < plan.n_chunks && deps_fit; ++k)
              {
                const uint32_t *cb = &plan.pool[plan.desc[size_t(k) * CHUNK_DESC_WORDS + CD_BLOB]];
                const uint32_t *pb = &plan.pool[plan.desc[size_t(k) * CHUNK_DESC_WORDS + CD_PULL_BLOB]];
                deps_fit = cb[CB_N_DEP_LOW] + cb[CB_N_DEP_LATE] <= 96u && pb[PB_N_DEPS] <= 96u;
              }
            if (const char *v = std::getenv("MGB_CHUNK_VERBOSE"); v && v[0] == '1')
              std::fprintf(stderr, "mgb200: chunk plan p=%d cells=%lld chunks=%u tables=%u+%u (%zu words) foreign low/late/ghost=%llu/%llu/%llu deps_fit=%d\n",
                           s.degree, s.n_cells, plan.n_chunks, plan.n_compute_blobs, plan.n_pull_blobs, plan.pool.size(),
                           (unsigned long long)plan.n_foreign_low, (unsigned long long)plan.n_foreign_late,
                           (unsigned long long)plan.n_foreign_ghost, int(deps_fit));
            if (deps_fit && plan.n_chunks > 0)
              {
                ch_desc_.upload(plan.desc.data(), plan.desc.size());
                ch_pool_.upload(plan.pool.data(), plan.pool.size());
                ch_deps_.upload(plan.deps.data(), plan.deps.size());
                ch_ghost_.upload(plan.ghost_idx.data(), plan.ghost_idx.size());
                ch_side_.alloc(size_t(plan.side_size) + 2);
                std::vector<uint32_t> zeros(size_t(plan.n_chunks) * 2 + 4, 0u);
                ch_flags_.upload(zeros.data(), size_t(plan.n_chunks) * 2);
                ch_state_.upload(zeros.data(), 4);
                chd_.n_chunks  = plan.n_chunks;
                chd_.desc      = ch_desc_.ptr;
                chd_.pool      = ch_pool_.ptr;
                chd_.deps      = ch_deps_.ptr;
                chd_.ghost_idx = ch_ghost_.ptr;
                chd_.side      = ch_side_.ptr;
                chd_.flags     = ch_flags_.ptr;
                chd_.state     = ch_state_.ptr;
                chunked_       = true;
              }
          }
        catch (const std::runtime_error &err)
          {
            chunked_ = false; // meshes the planner cannot chunk stay on the atomic kernel
            if (const char *v = std::getenv("MGB_CHUNK_VERBOSE"); v && v[0] == '1')
              std::fprintf(stderr, "mgb200: no chunk plan (%s)\n", err.what());
          }
    }
    cd_.ctab        = ctab_.ptr;
    cd_.degree      = s.degree;
    cd_.n_cells     = s.n_cells;
    cd_.etab        = etab_.ptr;
    cd_.mask        = mask_.ptr;
    cd_.level       = level_.ptr;
    cd_.root_length = s.root_length;
    cd_.any_hanging = any_hanging;
    constrained_.upload(s.constrained, s.n_constrained);
    edge_.upload(s.edge_constrained, s.n_edge_constrained);
    edge_cells_.upload(s.edge_cells, s.n_edge_cells);
    edge_src_vals_.alloc(s.n_edge_constrained);
    has_edge_global_ = s.has_edge_constrained_global;
    n_dofs_global_   = s.n_dofs_global;
    if (!s.partitioner)
      throw std::runtime_error("Operator: partitioner required");
    part_ = std::make_shared<DevicePartitioner>(*s.partitioner);
    if (s.partitioner->n_ghost > 0)
      {
        // cells reading ghost entries must wait for the exchange; all the others can run beside it
        std::vector<uint32_t> boundary, interior;
        for (long long c = 0; c < s.n_cells; ++c)
          {
            bool touches = false;
            for (int e = 0; e < 27 && !touches; ++e)
              touches = (s.entity_table[size_t(c) * 27 + e] & INDEX_MASK) >= s.partitioner->n_owned;
            (touches ? boundary : interior).push_back(uint32_t(c));
          }
        boundary_cells_.upload(boundary.data(), boundary.size());
        interior_cells_.upload(interior.data(), interior.size());
      }
  }

  template <typename T>
  void OperatorT<T>::cell_loop(DVec<T> &dst, const DVec<T> &src, bool zero_dst, bool edge_only, double alpha, DVec<T> *also_zero,
                               bool *did_zero) const
  {
    Context &c  = ctx();
    CellData cd = cd_;
    cd.root_length *= alpha; // the cell matrices are linear in h = root_length / 2^level
    cd.scale = alpha;
    if (did_zero)
      *did_zero = false;
    const bool overlap = !edge_only && !c.prof_enabled && boundary_cells_.n > 0 && interior_cells_.n > 0 && c.overlap_ready();
    if (overlap)
      {
        // ghost exchange of src on the second stream beside the first half of the interior cells; boundary cells once the
        // ghosts have arrived; compress of dst beside the second half of the interior cells (the unpack adds atomically)
        const long long n_int = (long long)interior_cells_.n, half = n_int / 2;
        MGB_CUDA(cudaEventRecord(c.ev_fork[0], c.stream));
        MGB_CUDA(cudaStreamWaitEvent(c.comm_stream, c.ev_fork[0], 0));
        src.update_ghost_values(c.comm_stream);
        MGB_CUDA(cudaEventRecord(c.ev_join[0], c.comm_stream));
        if (zero_dst)
          dst.zero();
        if (half > 0)
          cell_laplace_apply<T>(cd, src.data, dst.data, interior_cells_.ptr, half, c.stream), ++c.launches;
        MGB_CUDA(cudaStreamWaitEvent(c.stream, c.ev_join[0], 0));
        cell_laplace_apply<T>(cd, src.data, dst.data, boundary_cells_.ptr, (long long)boundary_cells_.n, c.stream), ++c.launches;
        MGB_CUDA(cudaEventRecord(c.ev_fork[1], c.stream));
        MGB_CUDA(cudaStreamWaitEvent(c.comm_stream, c.ev_fork[1], 0));
        dst.compress_add(c.comm_stream);
        MGB_CUDA(cudaEventRecord(c.ev_join[1], c.comm_stream));
        cell_laplace_apply<T>(cd, src.data, dst.data, interior_cells_.ptr + half, n_int - half, c.stream), ++c.launches;
        MGB_CUDA(cudaStreamWaitEvent(c.stream, c.ev_join[1], 0));
        src.zero_out_ghost_values();
        return;
      }
    if (zero_dst && !edge_only && alpha == 1.0 && chunk_path_ok(dst.data, src.data))
      {
        // chunked cell loop: every owned entry of dst is written once (no zeroing, no atomics); only the contributions to
        // ghost entries are added atomically to the (zeroed) ghost tail
        src.update_ghost_values();
        if (chunk_atomic_output())
          dst.zero();
        else if (dst.n_ghost)
          dst.zero_out_ghost_values();
        c.prof_begin(cd_.n_cells);
        cell_laplace_chunked<T>(cd, chd_, src.data, dst.data, FusedEpilogue<T>(), c.stream), ++c.launches;
        c.prof_end(cd_.n_cells);
        dst.compress_add();
        src.zero_out_ghost_values();
        return;
      }
    src.update_ghost_values();
    if (zero_dst)
      dst.zero();
    if (edge_only)
      {
        if (edge_cells_.n)
          cell_laplace_apply<T>(cd, src.data, dst.data, edge_cells_.ptr, (long long)edge_cells_.n, c.stream), ++c.launches;
      }
    else
      {
        CellData cz = cd; // the first (or only) launch also clears also_zero
        if (also_zero && !cd.vertices && also_zero->data && also_zero->data != dst.data && also_zero->data != src.data)
          {
            cz.zero_ptr = also_zero->data;
            cz.zero_n   = (long long)also_zero->n_owned + also_zero->n_ghost;
            if (did_zero)
              *did_zero = true;
          }
        c.prof_begin(cd_.n_cells);
        if (plain_cells_.n > 0 && hanging_cells_.n > 0)
          {
            CellData cp    = cz;
            cp.any_hanging = 0;
            cell_laplace_apply<T>(cp, src.data, dst.data, plain_cells_.ptr, (long long)plain_cells_.n, c.stream), ++c.launches;
            cell_laplace_apply<T>(cd, src.data, dst.data, hanging_cells_.ptr, (long long)hanging_cells_.n, c.stream), ++c.launches;
          }
        else
          cell_laplace_apply<T>(cz, src.data, dst.data, nullptr, 0, c.stream), ++c.launches;
        c.prof_end(cd_.n_cells);
      }
    dst.compress_add();
    src.zero_out_ghost_values();
  }

  template <typename T> bool OperatorT<T>::chunk_path_ok(const T *out, const T *x) const
  {
    return chunked_ && !cd_.vertices && out != x && (reinterpret_cast<uintptr_t>(x) & 15u) == 0;
  }

  template <typename T> bool OperatorT<T>::fused_available(const DVec<T> &out, const DVec<T> &x) const
  {
    return chunk_path_ok(out.data, x.data) && edge_.n == 0 && part_->trivial() && !chunk_atomic_output();
  }

  template <typename T> void OperatorT<T>::vmult_fused(DVec<T> &out, const DVec<T> &x, const FusedEpilogue<T> &ep) const
  {
    if (!fused_available(out, x))
      throw std::runtime_error("Operator::vmult_fused: not available for this operator");
    Context &c = ctx();
    c.prof_begin(cd_.n_cells);
    cell_laplace_chunked<T>(cd_, chd_, x.data, out.data, ep, c.stream), ++c.launches;
    c.prof_end(cd_.n_cells);
    if (constrained_.n)
      idx_fused_identity<T>(out.data, x.data, ep, constrained_.ptr, (uint32_t)constrained_.n, c.stream), ++c.launches;
  }

  template <typename T> void OperatorT<T>::set_cell_vertices(const double *xyz)
  {
    if (!xyz)
      throw std::runtime_error("Operator::set_cell_vertices: null pointer");
    for (size_t i = 0; i < size_t(cd_.n_cells) * 24; ++i)
      if (!std::isfinite(xyz[i]))
        throw std::runtime_error("Operator::set_cell_vertices: non-finite coordinate");
    vertices_.upload(xyz, size_t(cd_.n_cells) * 24);
    cd_.vertices = vertices_.ptr;
  }

  template <typename T> void OperatorT<T>::vmult(DVec<T> &dst, const DVec<T> &src) const
  {
    Context       &c  = ctx();
    const uint32_t ne = (uint32_t)edge_.n;
    if (ne)
      {
        // save values for edge constrained dofs and set them to 0 in src vector (operator.h:156-164)
        idx_gather<T>(edge_src_vals_.ptr, src.data, edge_.ptr, ne, c.stream);
        idx_set<T>(const_cast<T *>(src.data), T(0), edge_.ptr, ne, c.stream);
        c.launches += 2;
      }
    cell_loop(dst, src, true, false);
    if (constrained_.n)
      idx_copy<T>(dst.data, src.data, constrained_.ptr, (uint32_t)constrained_.n, c.stream), ++c.launches;
    if (ne)
      {
        idx_scatter<T>(const_cast<T *>(src.data), edge_src_vals_.ptr, edge_.ptr, ne, c.stream);
        idx_scatter<T>(dst.data, edge_src_vals_.ptr, edge_.ptr, ne, c.stream);
        c.launches += 2;
      }
  }

  template <typename T> bool OperatorT<T>::vmult_zeroing(DVec<T> &dst, const DVec<T> &src, const bool dst_is_zero, DVec<T> *also_zero) const
  {
    Context       &c  = ctx();
    const uint32_t ne = (uint32_t)edge_.n;
    if (ne) // as vmult
      {
        idx_gather<T>(edge_src_vals_.ptr, src.data, edge_.ptr, ne, c.stream);
        idx_set<T>(const_cast<T *>(src.data), T(0), edge_.ptr, ne, c.stream);
        c.launches += 2;
      }
    bool did_zero = false;
    cell_loop(dst, src, !dst_is_zero, false, 1.0, also_zero, &did_zero);
    if (constrained_.n)
      idx_copy<T>(dst.data, src.data, constrained_.ptr, (uint32_t)constrained_.n, c.stream), ++c.launches;
    if (ne)
      {
        idx_scatter<T>(const_cast<T *>(src.data), edge_src_vals_.ptr, edge_.ptr, ne, c.stream);
        idx_scatter<T>(dst.data, edge_src_vals_.ptr, edge_.ptr, ne, c.stream);
        c.launches += 2;
      }
    return did_zero;
  }

  template <typename T> void OperatorT<T>::vmult_interface_down(DVec<T> &dst, const DVec<T> &src) const
  {
    Context &c = ctx();
    cell_loop(dst, src, true, false);
    if (constrained_.n)
      idx_copy<T>(dst.data, src.data, constrained_.ptr, (uint32_t)constrained_.n, c.stream), ++c.launches;
  }

  template <typename T> void OperatorT<T>::residual_interface_down(DVec<T> &dst, const DVec<T> &b, const DVec<T> &src) const
  {
    Context &c = ctx();
    if (fused_available(dst, src) && dst.data != b.data)
      {
        FusedEpilogue<T> ep;
        ep.mode = 1;
        ep.b    = b.data;
        vmult_fused(dst, src, ep); // dst = b - A src in the cell loop's own pass, identity rows included
        return;
      }
    dst.copy_from(b);
    cell_loop(dst, src, false, false, -1.0);
    if (constrained_.n)
      idx_residual<T>(dst.data, b.data, src.data, constrained_.ptr, (uint32_t)constrained_.n, c.stream), ++c.launches;
  }

  template <typename T> void OperatorT<T>::vmult_interface_up(DVec<T> &dst, const DVec<T> &src) const
  {
    Context &c = ctx();
    dst.zero();
    if (!has_edge_global_)
      return;
    if (src_copy_.n_owned != src.n_owned || src_copy_.n_ghost != src.n_ghost)
      src_copy_.reinit_like(src);
    else
      src_copy_.zero();
    src_copy_.part = src.part;
    if (edge_.n)
      idx_copy<T>(src_copy_.data, src.data, edge_.ptr, (uint32_t)edge_.n, c.stream), ++c.launches;
    cell_loop(dst, src_copy_, false, true);
  }

  template <typename T> void OperatorT<T>::compute_inverse_diagonal(DVec<T> &diag) const
  {
    Context &c = ctx();
    initialize_dof_vector(diag);
    cell_laplace_diagonal<T>(cd_, diag.data, c.stream);
    ++c.launches;
    diag.compress_add();
    if (edge_.n)
      idx_set<T>(diag.data, T(0), edge_.ptr, (uint32_t)edge_.n, c.stream), ++c.launches;
    invert_diagonal<T>(diag.data, diag.n_owned, c.stream);
    ++c.launches;
  }

  // ------------------------------------------------------------------------------------------ Chebyshev
  template <typename T>
  ChebyshevT<T>::ChebyshevT(const OperatorT<T> &op, int degree, double smoothing_range, unsigned eig_its)
    : op_(op)
    , degree_(degree)
    , smoothing_range_(smoothing_range)
    , eig_its_(eig_its)
  {
    const char *e = std::getenv("MGB_FUSE_CHEB");
    fuse_         = !(e && e[0] == '0');
    e             = std::getenv("MGB_ZERO_IN_CELL_LOOP"); // opt-in: measured slower than the memset (profiles/r2v_zero_in_cell_loop.txt)
    zero_in_loop_ = e && e[0] == '1';
    op_.compute_inverse_diagonal(dinv_);
  }

  template <typename T> void ChebyshevT<T>::estimate_eigenvalues()
  {
    if (eig_done_)
      return;
    Context &c = ctx();
    op_.initialize_dof_vector(solution_old_);
    op_.initialize_dof_vector(temp1_);
    if (zero_in_loop_ && degree_ > 1)
      op_.initialize_dof_vector(temp2_); // here and not in iterate(): no allocation while a stream is being captured
    op_.initialize_dof_vector(temp2_);
    // initial guess (i + first_local_range) % 11 minus its mean (PreconditionChebyshev::set_initial_guess)
    {
      const Partitioner &p = op_.partitioner()->host;
      std::vector<T>     h(p.n_owned);
      for (uint32_t i = 0; i < p.n_owned; ++i)
        h[i] = T((i + p.owned_begin) % 11);
      if (p.n_owned)
        MGB_CUDA(cudaMemcpyAsync(temp1_.data, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, c.stream));
      MGB_CUDA(cudaStreamSynchronize(c.stream));
      const double mean = temp1_.mean_value(op_.m());
      temp1_.add_scalar(T(-mean));
    }
    SolverControl ctl;
    ctl.kind    = SolverControl::IterationNumber;
    ctl.maxiter = eig_its_;
    ctl.abstol  = 1e-10;
    std::vector<double> diag, off;
    const auto          A = [this](DVec<T> &d, const DVec<T> &s) { op_.vmult(d, s); };
    const auto          P = [this, &c](DVec<T> &d, const DVec<T> &s) {
      cheb_update<T>(0, T(0), T(1), s.data, dinv_.data, nullptr, nullptr, nullptr, d.data, d.n_owned, c.stream);
      ++c.launches;
    };
    cg_solve<T>(A, P, false, solution_old_, temp1_, ctl, cgw_, /*x_is_zero=*/true, &diag, &off);
    if (diag.empty())
      lambda_min_est_ = lambda_max_ = 1.0;
    else
      {
        off.resize(diag.size() - 1);
        lambda_min_est_ = tridiagonal_min_eigenvalue(diag, off);
        // include a safety factor since the CG method will in general not be converged
        lambda_max_ = 1.2 * tridiagonal_max_eigenvalue(diag, off);
      }
    const double alpha = smoothing_range_ > 1.0 ? lambda_max_ / smoothing_range_ : std::min(0.9 * lambda_max_, lambda_min_est_);
    delta_             = (lambda_max_ - alpha) * 0.5;
    theta_             = (lambda_max_ + alpha) * 0.5;
    solution_old_.zero();
    temp1_.zero();
    eig_done_ = true;
  }

  // PreconditionChebyshev::vmult (zero initial guess) and ::step, SURVEY.md A.4.  deal.II swaps the iterate with
  // solution_old after every update; here the updates alternate between dst and solution_old_ and the LAST one always
  // writes into dst (pointwise update: it may overwrite either input), so the caller's vector is never re-pointed and
  // can alias memory owned by someone else (PreconditionMG works directly on the CG vectors).
  template <typename T> void ChebyshevT<T>::iterate(DVec<T> &dst, const DVec<T> &rhs, const bool zero_guess)
  {
    estimate_eigenvalues();
    Context     &c = ctx();
    const size_t n = dst.n_owned;
    if (solution_old_.n_owned != dst.n_owned || solution_old_.n_ghost != dst.n_ghost)
      solution_old_.reinit_like(dst);
    solution_old_.part = dst.part;
    if (zero_guess)
      {
        // iteration_index 0: dst = (1/theta) D^-1 rhs
        cheb_update<T>(0, T(0), T(1.0 / theta_), rhs.data, dinv_.data, nullptr, nullptr, nullptr, dst.data, n, c.stream);
        ++c.launches;
      }
    const bool degenerate = degree_ < 2 || std::fabs(delta_) < 1e-40;
    const int  n_updates  = zero_guess ? (degenerate ? 0 : degree_ - 1) : (degenerate ? 1 : degree_);
    DVec<T>   *x = &dst, *xo = &solution_old_;
    bool       have_old = false;
    double     rhok = delta_ / theta_, sigma = theta_ / delta_;
    if (n_updates >= 2 && fuse_ && (temp1_.n_owned != dst.n_owned || temp1_.n_ghost != dst.n_ghost))
      temp1_.reinit_like(dst);
    if (n_updates >= 2 && fuse_ && op_.fused_available(solution_old_, dst) && op_.fused_available(dst, solution_old_) &&
        op_.fused_available(dst, temp1_))
      {
        // The vector update runs in the cell loop's own pass (north_star: "the smoother's vector updates fused with
        // vmult"; deal.II: cell_loop post hook).  An update may overwrite x_old but not x, and the last one has to land
        // in dst, so the iterates rotate through dst, solution_old_ and (odd counts only) temp1_:
        //   even: dst -> old -> dst -> ...        odd: dst -> old -> temp -> dst -> old -> dst ...
        temp1_.part = dst.part;
        for (int u = 0; u < n_updates; ++u)
          {
            double f1, f2;
            if (!zero_guess && u == 0)
              {
                f1 = 0.0;
                f2 = 1.0 / theta_;
              }
            else
              {
                const double rhokp = 1.0 / (2.0 * sigma - rhok);
                f1   = rhokp * rhok;
                f2   = 2.0 * rhokp / delta_;
                rhok = rhokp;
              }
            DVec<T> *out;
            if (n_updates % 2 == 1 && u <= 1)
              out = u == 0 ? &solution_old_ : &temp1_;
            else
              out = ((n_updates - 1 - u) % 2 == 0) ? &dst : &solution_old_;
            FusedEpilogue<T> ep;
            ep.mode = have_old ? 3 : 2;
            ep.b    = rhs.data;
            ep.dinv = dinv_.data;
            ep.xold = have_old ? xo->data : nullptr;
            ep.f1   = T(f1);
            ep.f2   = T(f2);
            op_.vmult_fused(*out, *x, ep);
            xo       = x;
            x        = out;
            have_old = true;
          }
        return;
      }
    // A x lands alternately in temp1_ / temp2_: the cell loop that fills one clears the other for the next product, so only
    // the first product of a call needs a memset.  Nothing is remembered between calls (a caller may replay this sequence
    // from a graph of its own, which the host code would not see).
    DVec<T> *tt[2] = {&temp1_, &temp2_};
    if (zero_in_loop_ && n_updates > 1)
      {
        if (temp2_.n_owned != dst.n_owned || temp2_.n_ghost != dst.n_ghost)
          temp2_.reinit_like(dst);
        temp2_.part = temp1_.part;
      }
    bool next_is_clean = false;
    for (int u = 0; u < n_updates; ++u)
      {
        DVec<T> *t = &temp1_;
        if (zero_in_loop_ && n_updates > 1)
          {
            t             = tt[u & 1];
            next_is_clean = op_.vmult_zeroing(*t, *x, next_is_clean, u + 1 < n_updates ? tt[(u + 1) & 1] : nullptr);
          }
        else
          op_.vmult(temp1_, *x);
        double f1, f2;
        if (!zero_guess && u == 0)
          {
            f1 = 0.0;
            f2 = 1.0 / theta_;
          }
        else
          {
            const double rhokp = 1.0 / (2.0 * sigma - rhok);
            f1   = rhokp * rhok;
            f2   = 2.0 * rhokp / delta_;
            rhok = rhokp;
          }
        const bool last = u == n_updates - 1;
        DVec<T>   *out  = last ? &dst : xo;
        cheb_update<T>(have_old ? 2 : 1, T(f1), T(f2), rhs.data, dinv_.data, t->data, x->data, xo->data, out->data, n, c.stream);
        ++c.launches;
        if (!last)
          {
            std::swap(x, xo);
            have_old = true;
          }
      }
  }

  template <typename T> void ChebyshevT<T>::vmult(DVec<T> &dst, const DVec<T> &rhs) { iterate(dst, rhs, true); }
  template <typename T> void ChebyshevT<T>::step(DVec<T> &dst, const DVec<T> &rhs) { iterate(dst, rhs, false); }

  // ------------------------------------------------------------------------------------------ transfers
  template <typename T>
  TwoLevelTransferT<T>::TwoLevelTransferT(const DoFMap &fine, const DoFMap &coarse, const TransferMap &tm, const OperatorT<T> &op_fine,
                                          const OperatorT<T> &op_coarse)
    : op_fine_(op_fine)
    , op_coarse_(op_coarse)
  {
    ctx().ensure();
    upload_fe_constants(coarse.degree);
    fine_own_.upload(tm.fine_own.data(), tm.fine_own.size());
    if (tm.distributed)
      fine_part_ = std::make_shared<DevicePartitioner>(tm.part);
    fine_first_.upload(tm.fine_first.data(), tm.fine_first.size());
    n_children_.upload(tm.n_children.data(), tm.n_children.size());
    weight_code_.upload(tm.weight_code.data(), tm.weight_code.size());
    td_.degree_fine        = fine.degree;
    td_.degree_coarse      = coarse.degree;
    td_.h_mode             = fine.degree == coarse.degree;
    td_.n_coarse_cells     = tm.n_coarse_cells;
    td_.coarse_ctab        = op_coarse.cell_data().ctab;
    td_.coarse_any_hanging = op_coarse.cell_data().any_hanging;
    td_.fine_own           = fine_own_.ptr;
    td_.fine_first         = fine_first_.ptr;
    td_.n_children         = n_children_.ptr;
    td_.weight_code        = weight_code_.ptr;
    {
      const std::vector<double> P = td_.h_mode ? FE1D(coarse.degree).Ph : p_embedding(fine.degree, coarse.degree);
      std::copy(P.begin(), P.end(), td_.P);
    }
    if (fine.degree != coarse.degree)
      for (uint8_t nchild : tm.n_children)
        if (nchild != 1)
          throw std::runtime_error("TwoLevelTransfer: simultaneous h- and p-coarsening is not supported (deal.II neither)");
  }

  template <typename T> void TwoLevelTransferT<T>::prolongate_and_add(DVec<T> &dst_fine, const DVec<T> &src_coarse) const
  {
    Context &c = ctx();
    src_coarse.update_ghost_values();
    if (fine_part_)
      {
        // children of local coarse cells may live on other ranks: add into the transfer's own ghosted fine vector,
        // send the ghost contributions to their owners, then add the owned part to dst
        vec_fine_.reinit_keep_layout(fine_part_);
        transfer_prolongate_add<T>(td_, vec_fine_.data, src_coarse.data, c.stream);
        ++c.launches;
        vec_fine_.compress_add();
        dst_fine.add(T(1), vec_fine_);
      }
    else
      {
        transfer_prolongate_add<T>(td_, dst_fine.data, src_coarse.data, c.stream);
        ++c.launches;
        dst_fine.compress_add();
      }
    src_coarse.zero_out_ghost_values();
  }

  template <typename T> void TwoLevelTransferT<T>::restrict_and_add(DVec<T> &dst_coarse, const DVec<T> &src_fine) const
  {
    Context &c = ctx();
    if (fine_part_)
      {
        vec_fine_.reinit_keep_layout(fine_part_);
        vec_fine_.copy_from(src_fine);
        vec_fine_.update_ghost_values();
        transfer_restrict_add<T>(td_, dst_coarse.data, vec_fine_.data, c.stream);
        ++c.launches;
      }
    else
      {
        src_fine.update_ghost_values();
        transfer_restrict_add<T>(td_, dst_coarse.data, src_fine.data, c.stream);
        ++c.launches;
        src_fine.zero_out_ghost_values();
      }
    dst_coarse.compress_add();
  }

  template <typename T>
  void GCTransferT<T>::copy_to_mg(std::vector<DVec<T>> &defect, unsigned minl, unsigned maxl, const DVec<double> &src, bool skip_top) const
  {
    Context &c = ctx();
    for (unsigned l = minl; l <= maxl; ++l)
      if (!(skip_top && l == maxl))
        operators[l - min_level]->initialize_dof_vector(defect[l - minl]);
    if (skip_top)
      return;
    DVec<T> &top = defect[maxl - minl];
    vec_convert<T, double>(top.data, src.data, top.n_owned, c.stream);
    ++c.launches;
  }

  template <typename T>
  void GCTransferT<T>::copy_from_mg(DVec<double> &dst, const std::vector<DVec<T>> &solution, unsigned minl, unsigned maxl) const
  {
    Context       &c   = ctx();
    const DVec<T> &top = solution[maxl - minl];
    vec_convert<double, T>(dst.data, top.data, top.n_owned, c.stream);
    ++c.launches;
  }

  template <typename T>
  void LSTransferT<T>::copy_to_mg(std::vector<DVec<T>> &defect, unsigned minl, unsigned maxl, const DVec<double> &src, bool skip_top) const
  {
    if (skip_top)
      throw std::logic_error("LSTransfer::copy_to_mg: the top level is not a plain copy");
    Context &c = ctx();
    for (unsigned l = minl; l <= maxl; ++l)
      {
        DVec<T> &d = defect[l - minl];
        d.reinit_keep_layout(operators[l]->partitioner()); // zero: restriction adds into it
        const uint32_t n = (uint32_t)copy[l].active_idx.n;
        if (copy[l].part)
          {
            // the owner of the active DoF writes into the ghosted level vector; compress delivers to the level owner
            DVec<T> &tmp = copy[l].tmp;
            tmp.reinit_keep_layout(copy[l].part);
            if (n)
              idx_pair_copy<T, double>(tmp.data, copy[l].level_idx.ptr, src.data, copy[l].active_idx.ptr, n, c.stream), ++c.launches;
            tmp.compress_add();
            d.copy_from(tmp);
          }
        else if (n)
          idx_pair_copy<T, double>(d.data, copy[l].level_idx.ptr, src.data, copy[l].active_idx.ptr, n, c.stream), ++c.launches;
      }
  }

  template <typename T>
  void LSTransferT<T>::copy_from_mg(DVec<double> &dst, const std::vector<DVec<T>> &solution, unsigned minl, unsigned maxl) const
  {
    Context &c = ctx();
    dst.zero(); // constrained (hanging) active DoFs are in no level's copy list
    for (unsigned l = minl; l <= maxl; ++l)
      {
        const uint32_t n = (uint32_t)copy[l].active_idx.n;
        if (copy[l].part)
          {
            DVec<T> &tmp = copy[l].tmp;
            tmp.reinit_keep_layout(copy[l].part);
            tmp.copy_from(solution[l - minl]);
            tmp.update_ghost_values();
            if (n)
              idx_pair_copy<double, T>(dst.data, copy[l].active_idx.ptr, tmp.data, copy[l].level_idx.ptr, n, c.stream), ++c.launches;
          }
        else if (n)
          idx_pair_copy<double, T>(dst.data, copy[l].active_idx.ptr, solution[l - minl].data, copy[l].level_idx.ptr, n, c.stream),
            ++c.launches;
      }
  }

  // ------------------------------------------------------------------------------------------ coarse solver
  template <typename T> void CoarseCGT<T>::setup_dense()
  {
    dense_checked_ = true;
    Context           &c = ctx();
    const Partitioner &p = op_.partitioner()->host;
    // the level lives entirely on this rank (or this rank owns nothing of it) and is small
    const bool local = p.n_ghost == 0 && p.import_indices.empty() && (p.n_owned == 0 || (long long)p.n_owned == op_.m());
    bool       use_dense = local && op_.m() <= max_dense_size;
    if (c.comm && c.comm->size() > 1)
      {
        // The decision must be the same on every rank: a rank that owns nothing of the level would otherwise go "dense" and
        // skip the generic CG whose dot products are collective, while the ranks that share the level wait in them.
        int              mine = use_dense ? 1 : 0;
        std::vector<int> all(size_t(c.comm->size()), 0);
        c.comm->allgather_host(&mine, all.data(), sizeof(int), c.stream);
        for (const int v : all)
          use_dense = use_dense && v != 0;
      }
    if (!use_dense)
      return;
    dense_ = true;
    dense_its_.alloc(1);
    MGB_CUDA(cudaMemsetAsync(dense_its_.ptr, 0, sizeof(int), c.stream));
    const int n = (int)p.n_owned;
    if (n == 0)
      return;
    // column j of the level matrix = A e_j (the operator with its identity rows is symmetric)
    dense_A_.alloc(size_t(n) * n);
    DVec<T> e, col;
    op_.initialize_dof_vector(e);
    op_.initialize_dof_vector(col);
    for (int j = 0; j < n; ++j)
      {
        e.zero();
        const T one = T(1);
        MGB_CUDA(cudaMemcpyAsync(e.data + j, &one, sizeof(T), cudaMemcpyHostToDevice, c.stream));
        op_.vmult(col, e);
        MGB_CUDA(cudaMemcpyAsync(dense_A_.ptr + size_t(j) * n, col.data, size_t(n) * sizeof(T), cudaMemcpyDeviceToDevice, c.stream));
        MGB_CUDA(cudaStreamSynchronize(c.stream));
      }
  }

  template <typename T> void CoarseCGT<T>::operator()(unsigned, DVec<T> &dst, const DVec<T> &src)
  {
    if (precond_)
      precond_->estimate_eigenvalues(); // also on ranks that own nothing of the level: collective
    if (!dense_checked_)
      setup_dense();
    if (dense_)
      {
        Context &c = ctx();
        if (dst.n_owned)
          {
            coarse_cg_solve<T>(dense_A_.ptr, (int)dst.n_owned, precond_ ? precond_->inverse_diagonal().data : nullptr,
                               precond_ ? precond_->degree() : 0, precond_ ? precond_->theta() : 1.0, precond_ ? precond_->delta() : 1.0,
                               ctl_.maxiter, ctl_.abstol, ctl_.reltol, src.data, dst.data, dense_its_.ptr, c.stream);
            ++c.launches;
          }
        return;
      }
    const auto A = [this](DVec<T> &d, const DVec<T> &s) { op_.vmult(d, s); };
    const auto P = [this](DVec<T> &d, const DVec<T> &s) { precond_->vmult(d, s); };
    last_iterations = cg_solve<T>(A, P, precond_ == nullptr, dst, src, ctl_, work_, /*x_is_zero=*/true);
    total_iterations += last_iterations;
  }

  template <typename T> long long CoarseCGT<T>::collect_device_iterations()
  {
    if (dense_ && dense_its_.n)
      {
        int its = 0;
        MGB_CUDA(cudaMemcpyAsync(&its, dense_its_.ptr, sizeof(int), cudaMemcpyDeviceToHost, ctx().stream));
        MGB_CUDA(cudaMemsetAsync(dense_its_.ptr, 0, sizeof(int), ctx().stream));
        MGB_CUDA(cudaStreamSynchronize(ctx().stream));
        total_iterations += its;
      }
    return total_iterations;
  }

  // ------------------------------------------------------------------------------------------ timers
  cudaEvent_t StageTimers::get()
  {
    if (pool_next == pool.size())
      {
        cudaEvent_t e;
        MGB_CUDA(cudaEventCreate(&e));
        pool.push_back(e);
      }
    return pool[pool_next++];
  }
  void StageTimers::begin(int level, int stage)
  {
    if (!enabled)
      return;
    Rec r{get(), get(), level, stage};
    MGB_CUDA(cudaEventRecord(r.a, ctx().stream));
    open.push_back(recs.size()); // stages nest when a V-cycle is the coarse solver of another one (HPMG-local)
    recs.push_back(r);
  }
  void StageTimers::end()
  {
    if (!enabled || open.empty())
      return;
    MGB_CUDA(cudaEventRecord(recs[open.back()].b, ctx().stream));
    open.pop_back();
  }
  void StageTimers::collect(std::vector<std::array<double, N_STAGES + 2>> &times, int min_level)
  {
    MGB_CUDA(cudaStreamSynchronize(ctx().stream));
    for (const Rec &r : recs)
      {
        float ms = 0;
        MGB_CUDA(cudaEventElapsedTime(&ms, r.a, r.b));
        times[r.level - min_level][r.stage] += ms * 1e-3;
      }
    recs.clear();
    open.clear();
    pool_next = 0;
  }
  StageTimers::~StageTimers()
  {
    for (cudaEvent_t e : pool)
      cudaEventDestroy(e);
  }

  // ------------------------------------------------------------------------------------------ multigrid
  template <typename T> void MultigridT<T>::cycle()
  {
    // The defect vectors have been initialised by copy_to_mg.  deal.II re-initialises solution and t on all levels to
    // zero here; only the coarse level's solution (zero initial guess of the coarse CG) needs it: on the other levels
    // the pre-smoother overwrites solution and the residual overwrites t, ghost entries are zero by invariant.
    for (unsigned l = min_level; l <= max_level; ++l)
      {
        if (l == min_level)
          solution[l - min_level].reinit_keep(defect[l - min_level]);
        else
          solution[l - min_level].ensure_like(defect[l - min_level]);
        t[l - min_level].ensure_like(defect[l - min_level]);
      }
    level_v_step(max_level);
  }

  template <typename T> void MultigridT<T>::level_v_step(unsigned level)
  {
    const unsigned i = level - min_level;
    const auto     tb = [&](int stage) {
      if (timers)
        timers->begin(int(level), stage);
    };
    const auto te = [&]() {
      if (timers)
        timers->end();
    };
    if (level == min_level)
      {
        tb(STAGE_COARSE);
        (*coarse)(level, solution[i], defect[i]);
        te();
        return;
      }
    tb(STAGE_PRE);
    smoothers[i]->vmult(solution[i], defect[i]); // pre_smooth->apply (zero initial guess)
    te();
    tb(STAGE_RESIDUAL);
    matrices[i]->residual_interface_down(t[i], defect[i], solution[i]); // t = defect - A solution
    te();
    tb(STAGE_RESTRICTION);
    transfer->restrict_and_add(level, defect[i - 1], t[i]);
    te();
    level_v_step(level - 1);
    tb(STAGE_PROLONGATION);
    transfer->prolongate_and_add(level, solution[i], solution[i - 1]);
    te();
    if (edge_in)
      {
        tb(STAGE_EDGE_PROLONGATION);
        matrices[i]->vmult_interface_up(t[i], solution[i]);
        defect[i].add(T(-1), t[i]);
        te();
      }
    tb(STAGE_POST);
    smoothers[i]->step(solution[i], defect[i]); // post_smooth->smooth
    te();
  }

  namespace
  {
    std::vector<std::pair<void *, std::function<void()>>> &graph_owners()
    {
      static std::vector<std::pair<void *, std::function<void()>>> v;
      return v;
    }
  } // namespace
  void release_all_graphs()
  {
    for (auto &o : graph_owners())
      o.second();
  }
  template <typename T> PreconditionMGT<T>::PreconditionMGT()
  {
    graph_owners().emplace_back(this, [this] { release_graphs(); });
  }
  template <typename T> void PreconditionMGT<T>::release_graphs()
  {
    for (GraphEntry &g : graphs_)
      cudaGraphExecDestroy(g.exec);
    graphs_.clear();
  }
  template <typename T> PreconditionMGT<T>::~PreconditionMGT()
  {
    release_graphs();
    auto &v = graph_owners();
    for (size_t i = 0; i < v.size(); ++i)
      if (v[i].first == this)
        {
          v.erase(v.begin() + i);
          break;
        }
  }

  // NCCL point-to-point calls can be stream-captured (every rank captures and replays the same sequence); the
  // environment variable MGB_GRAPH_MULTIRANK=0 falls back to eager launches on several ranks
  static bool graph_multirank()
  {
    static const bool on = [] {
      const char *e = std::getenv("MGB_GRAPH_MULTIRANK");
      return !(e && e[0] == '0');
    }();
    return on;
  }

  template <typename T> void PreconditionMGT<T>::vmult(DVec<double> &dst, const DVec<double> &src)
  {
    Context   &c = ctx();
    const bool plain = transfer->top_level_is_plain_copy();
    bool graph_ok = allow_graph && c.stream != nullptr && !(timers && timers->enabled) && !c.prof_enabled &&
                    !(c.comm && c.comm->size() > 1 && !graph_multirank()) && src.data != dst.data;
    if (graph_ok && plain)
      {
        // global coarsening: only the aliasing path (no allocation, no copies) is captured
        uint32_t                           no, ng;
        std::shared_ptr<DevicePartitioner> part;
        transfer->level_layout(mg->max_level, no, ng, part);
        graph_ok = std::is_same<T, double>::value && !mg->edge_in && src.n_owned == no && src.n_ghost == ng && dst.n_owned == no &&
                   dst.n_ghost == ng;
      }
    // local smoothing: copy_to_mg / copy_from_mg are index kernels between fixed buffers, captured with the cycle
    if (!graph_ok)
      {
        vmult_eager(dst, src);
        return;
      }
    for (const GraphEntry &g : graphs_)
      if (g.src == src.data && g.dst == dst.data)
        {
          MGB_CUDA(cudaGraphLaunch(g.exec, c.stream));
          c.launches += g.launches;
          return;
        }
    if (graphs_.size() >= 8)
      {
        vmult_eager(dst, src);
        return;
      }
    const long long l0 = c.launches;
    cudaGraph_t     graph = nullptr;
    MGB_CUDA(cudaStreamBeginCapture(c.stream, cudaStreamCaptureModeThreadLocal));
    try
      {
        vmult_eager(dst, src);
      }
    catch (...)
      {
        cudaStreamEndCapture(c.stream, &graph);
        if (graph)
          cudaGraphDestroy(graph);
        throw;
      }
    MGB_CUDA(cudaStreamEndCapture(c.stream, &graph));
    GraphEntry g{src.data, dst.data, nullptr, c.launches - l0};
    MGB_CUDA(cudaGraphInstantiate(&g.exec, graph, 0));
    MGB_CUDA(cudaGraphDestroy(graph));
    graphs_.push_back(g);
    MGB_CUDA(cudaGraphLaunch(g.exec, c.stream));
  }

  template <typename T> void PreconditionMGT<T>::vmult_eager(DVec<double> &dst, const DVec<double> &src)
  {
    mg->resize_levels();
    const unsigned top = mg->max_level - mg->min_level;
    // FP64 levels without edge matrices never write into defect[max]: let the V-cycle work directly on the caller's
    // vectors (defect[max] = src, solution[max] = dst) instead of copying in and out (copy_to_mg / copy_from_mg)
    bool alias = false;
    if constexpr (std::is_same<T, double>::value)
      {
        uint32_t                           no, ng;
        std::shared_ptr<DevicePartitioner> part;
        transfer->level_layout(mg->max_level, no, ng, part);
        alias = transfer->top_level_is_plain_copy() && !mg->edge_in && src.n_owned == no && src.n_ghost == ng && dst.n_owned == no && dst.n_ghost == ng && src.data != dst.data;
        if (alias)
          {
            mg->defect[top].wrap(const_cast<double *>(src.data), no, ng, part);
            mg->solution[top].wrap(dst.data, no, ng, part);
          }
        else if (!mg->defect[top].owns)
          {
            mg->defect[top].wrap(nullptr, 0, 0);
            mg->solution[top].wrap(nullptr, 0, 0);
          }
      }
    if (timers)
      timers->begin(int(mg->min_level), N_STAGES);
    transfer->copy_to_mg(mg->defect, mg->min_level, mg->max_level, src, alias);
    if (timers)
      timers->end();
    mg->cycle();
    if (timers)
      timers->begin(int(mg->min_level), N_STAGES + 1);
    if (!alias)
      transfer->copy_from_mg(dst, mg->solution, mg->min_level, mg->max_level);
    if (timers)
      timers->end();
  }

  template <typename T> void CoarseApplyMGT<T>::operator()(unsigned, DVec<T> &dst, const DVec<T> &src)
  {
    // MGCoarseGridApplyPreconditioner (multigrid_throughput.cc:528-550): cast to double, apply, cast back
    Context &c = ctx();
    if (src_d.n_owned != src.n_owned)
      {
        src_d.reinit(src.n_owned, src.n_ghost, src.part);
        dst_d.reinit(src.n_owned, src.n_ghost, src.part);
      }
    vec_convert<double, T>(src_d.data, src.data, src.n_owned, c.stream);
    ++c.launches;
    pmg->vmult(dst_d, src_d);
    vec_convert<T, double>(dst.data, dst_d.data, dst.n_owned, c.stream);
    ++c.launches;
  }

  template struct DVec<double>;
  template struct DVec<float>;
  template class OperatorT<double>;
  template class OperatorT<float>;
  template class ChebyshevT<double>;
  template class ChebyshevT<float>;
  template class TwoLevelTransferT<double>;
  template class TwoLevelTransferT<float>;
  template class GCTransferT<double>;
  template class GCTransferT<float>;
  template class LSTransferT<double>;
  template class LSTransferT<float>;
  template class MultigridT<double>;
  template class MultigridT<float>;
  template class PreconditionMGT<double>;
  template class PreconditionMGT<float>;
  template class CoarseCGT<double>;
  template class CoarseCGT<float>;
  template class CoarseApplyMGT<double>;
  template class CoarseApplyMGT<float>;
} // namespace mgb
