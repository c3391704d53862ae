// extern "C" boundary, part 2: the objects deal.II composes ABOVE operator / smoother / two-level transfer --
//   MGLevelGlobalTransfer::copy_to_mg / copy_from_mg through copy_indices     (mgb_level_copy_*)
//   MGTwoLevelTransfer / MGTransferMatrixFree::prolongate                     (mgb_transfer_prolongate)
//   Multigrid + PreconditionMG with set_edge_in_matrix, coarse CG            (mgb_multigrid_*)
//   SolverCG + ReductionControl                                              (mgb_cg_solve)
// built from BORROWED handles, so that a host program can wire its own hierarchy exactly as
// multigrid_throughput.cc:1082-1133 (mg_solve) and :1800-1817 (local smoothing) do.
#include "capi_handles.h"

#include <vector>

using namespace mgb;

namespace
{
  // copy_indices of ONE level on the device (LSTransferT::Copy as a standalone object)
  template <typename T> struct LevelCopyT
  {
    DeviceBuffer<uint32_t>             active_idx, level_idx;
    std::shared_ptr<DevicePartitioner> part; // several ranks: the auxiliary ghosted level vector (CopyIndices::part)
    mutable DVec<T>                    tmp;
    const OperatorT<T>                *op = nullptr;

    LevelCopyT(const CopyIndices &ci, const OperatorT<T> &level_op)
      : op(&level_op)
    {
      active_idx.upload(ci.active_idx.data(), ci.active_idx.size());
      level_idx.upload(ci.level_idx.data(), ci.level_idx.size());
      if (ci.distributed)
        part = std::make_shared<DevicePartitioner>(ci.part);
    }
    // dst_level = 0 except the entries of the active DoFs living on this level (MGLevelGlobalTransfer::copy_to_mg)
    void to_mg(DVec<T> &dst, const DVec<double> &src) const
    {
      Context &c = ctx();
      dst.reinit_keep_layout(op->partitioner());
      const uint32_t n = (uint32_t)active_idx.n;
      if (part)
        {
          tmp.reinit_keep_layout(part);
          if (n)
            idx_pair_copy<T, double>(tmp.data, level_idx.ptr, src.data, active_idx.ptr, n, c.stream), ++c.launches;
          tmp.compress_add();
          dst.copy_from(tmp);
        }
      else if (n)
        idx_pair_copy<T, double>(dst.data, level_idx.ptr, src.data, active_idx.ptr, n, c.stream), ++c.launches;
    }
    // the active DoFs living on this level receive the level's values; the others are left alone
    void from_mg(DVec<double> &dst, const DVec<T> &src) const
    {
      Context       &c = ctx();
      const uint32_t n = (uint32_t)active_idx.n;
      if (part)
        {
          tmp.reinit_keep_layout(part);
          tmp.copy_from(src);
          tmp.update_ghost_values();
          if (n)
            idx_pair_copy<double, T>(dst.data, active_idx.ptr, tmp.data, level_idx.ptr, n, c.stream), ++c.launches;
        }
      else if (n)
        idx_pair_copy<double, T>(dst.data, active_idx.ptr, src.data, level_idx.ptr, n, c.stream), ++c.launches;
    }
  };

  // MGTransferGlobalCoarsening / MGTransferMatrixFree over borrowed two-level transfers
  template <typename T> class BorrowedTransferT : public MGTransferBase<T>
  {
  public:
    std::vector<const TwoLevelTransferT<T> *> transfers; // [l] connects l-1 and l ([0] unused)
    std::vector<const OperatorT<T> *>         operators; // [l]
    std::vector<const LevelCopyT<T> *>        copies;    // empty: global coarsening (the top level IS the outer vector)
    void prolongate_and_add(unsigned to, DVec<T> &dst, const DVec<T> &src) const override { transfers[to]->prolongate_and_add(dst, src); }
    void restrict_and_add(unsigned from, DVec<T> &dst, const DVec<T> &src) const override { transfers[from]->restrict_and_add(dst, src); }
    void copy_to_mg(std::vector<DVec<T>> &defect, unsigned minl, unsigned maxl, const DVec<double> &src, bool skip_top) const override
    {
      Context &c = ctx();
      if (copies.empty())
        {
          for (unsigned l = minl; l <= maxl; ++l)
            if (!(skip_top && l == maxl))
              operators[l]->initialize_dof_vector(defect[l - minl]);
          if (skip_top)
            return;
          DVec<T> &top = defect[maxl - minl];
          vec_convert<T, double>(top.data, src.data, top.n_owned, c.stream);
          ++c.launches;
          return;
        }
      if (skip_top)
        throw std::logic_error("copy_to_mg: with copy_indices the top level is not a plain copy");
      for (unsigned l = minl; l <= maxl; ++l)
        copies[l]->to_mg(defect[l - minl], src);
    }
    void level_layout(unsigned level, uint32_t &n_owned, uint32_t &n_ghost, std::shared_ptr<DevicePartitioner> &part) const override
    {
      part    = operators[level]->partitioner();
      n_owned = part->host.n_owned;
      n_ghost = part->host.n_ghost;
    }
    void copy_from_mg(DVec<double> &dst, const std::vector<DVec<T>> &solution, unsigned minl, unsigned maxl) const override
    {
      Context &c = ctx();
      if (copies.empty())
        {
          const DVec<T> &top = solution[maxl - minl];
          vec_convert<double, T>(dst.data, top.data, top.n_owned, c.stream);
          ++c.launches;
          return;
        }
      dst.zero(); // constrained (hanging) active DoFs are in no level's list
      for (unsigned l = minl; l <= maxl; ++l)
        copies[l]->from_mg(dst, solution[l - minl]);
    }
    bool top_level_is_plain_copy() const override { return copies.empty(); }
  };

  template <typename T> struct MultigridHandleT
  {
    BorrowedTransferT<T>          transfer;
    std::unique_ptr<CoarseCGT<T>> coarse;
    MultigridT<T>                 mg;
    PreconditionMGT<T>            pmg;
    StageTimers                   timers;
    std::vector<std::array<double, N_STAGES + 2>> times;
  };
} // namespace

namespace mgb
{
  void check_index_list(const char *what, const uint32_t *idx, uint64_t n, uint64_t bound); // capi.cc
}
struct mgb_level_copy
{
  int                                 type = MGB_DOUBLE;
  uint64_t                            active_size_needed = 0, level_size_needed = 0; // 1 + largest index of either side
  std::unique_ptr<LevelCopyT<double>> d;
  std::unique_ptr<LevelCopyT<float>>  f;
};
struct mgb_multigrid
{
  int                                       type = MGB_DOUBLE;
  int                                       n_levels = 0;
  std::unique_ptr<MultigridHandleT<double>> d;
  std::unique_ptr<MultigridHandleT<float>>  f;
  CGWork<double>                            cg_work;
};

namespace
{
  template <typename T>
  std::unique_ptr<MultigridHandleT<T>> make_multigrid(int n_levels, const mgb_operator *const *ops, mgb_chebyshev *const *smoothers,
                                                      const mgb_transfer *const *transfers, mgb_chebyshev *coarse_precond,
                                                      unsigned maxiter, double abstol, double reltol,
                                                      OperatorT<T> *mgb_operator::*op_member, ChebyshevT<T> *mgb_chebyshev::*ch_member,
                                                      TwoLevelTransferT<T> *mgb_transfer::*tr_member)
  {
    auto h = std::make_unique<MultigridHandleT<T>>();
    for (int l = 0; l < n_levels; ++l)
      {
        if (!ops[l] || !(ops[l]->*op_member))
          throw std::runtime_error("mgb_multigrid_create: level operator missing or of another number type");
        h->transfer.operators.push_back(ops[l]->*op_member);
        h->mg.matrices.push_back(ops[l]->*op_member);
        if (l > 0)
          {
            if (!smoothers[l] || !(smoothers[l]->*ch_member) || !transfers[l] || !(transfers[l]->*tr_member))
              throw std::runtime_error("mgb_multigrid_create: smoother / transfer of a level missing or of another number type");
            h->mg.smoothers.push_back(smoothers[l]->*ch_member);
            h->transfer.transfers.push_back(transfers[l]->*tr_member);
          }
        else
          {
            h->mg.smoothers.push_back(nullptr);
            h->transfer.transfers.push_back(nullptr);
          }
      }
    h->coarse = std::make_unique<CoarseCGT<T>>(*h->mg.matrices[0], coarse_precond ? coarse_precond->*ch_member : nullptr, maxiter, abstol, reltol);
    h->mg.min_level   = 0;
    h->mg.max_level   = unsigned(n_levels - 1);
    h->mg.transfer    = &h->transfer;
    h->mg.coarse      = h->coarse.get();
    h->mg.timers      = &h->timers;
    h->pmg.mg         = &h->mg;
    h->pmg.transfer   = &h->transfer;
    h->pmg.timers     = &h->timers;
    return h;
  }
} // namespace

extern "C" {

int mgb_level_copy_create(const mgb_copy_indices *ci, const mgb_operator *level_op, mgb_level_copy **out)
{
  return guarded([&] {
    auto h  = std::make_unique<mgb_level_copy>();
    h->type = level_op->type;
    if (level_op->type == MGB_DOUBLE)
      h->d = std::make_unique<LevelCopyT<double>>(ci->c, *level_op->pd);
    else
      h->f = std::make_unique<LevelCopyT<float>>(ci->c, *level_op->pf);
    *out = h.release();
  });
}
int mgb_level_copy_create_from_arrays(uint32_t n, const uint32_t *active_idx, const uint32_t *level_idx, const mgb_operator *level_op,
                                      mgb_level_copy **out)
{
  return guarded([&] {
    if (n && (!active_idx || !level_idx))
      throw std::runtime_error("mgb_level_copy_create_from_arrays: index arrays required");
    const Partitioner &lp = (level_op->type == MGB_DOUBLE ? level_op->pd->partitioner() : level_op->pf->partitioner())->host;
    check_index_list("mgb_level_copy_create_from_arrays: level_idx", level_idx, n, uint64_t(lp.n_owned) + lp.n_ghost);
    CopyIndices ci;
    ci.active_idx.assign(active_idx, active_idx + n);
    ci.level_idx.assign(level_idx, level_idx + n);
    auto h  = std::make_unique<mgb_level_copy>();
    h->type = level_op->type;
    for (uint32_t i = 0; i < n; ++i)
      {
        h->active_size_needed = std::max<uint64_t>(h->active_size_needed, uint64_t(active_idx[i]) + 1);
        h->level_size_needed  = std::max<uint64_t>(h->level_size_needed, uint64_t(level_idx[i]) + 1);
      }
    if (level_op->type == MGB_DOUBLE)
      h->d = std::make_unique<LevelCopyT<double>>(ci, *level_op->pd);
    else
      h->f = std::make_unique<LevelCopyT<float>>(ci, *level_op->pf);
    *out = h.release();
  });
}
void mgb_level_copy_destroy(mgb_level_copy *c) { delete c; }
int  mgb_level_copy_to_mg(const mgb_level_copy *c, mgb_vector *dst_level, const mgb_vector *src_active)
{
  return guarded([&] {
    require_same_type(c->type, dst_level->type);
    require_same_type(MGB_DOUBLE, src_active->type);
    if (c->active_size_needed > uint64_t(src_active->pd->n_owned) + src_active->pd->n_ghost)
      throw std::runtime_error("mgb_level_copy_to_mg: the copy indices address entry " + std::to_string(c->active_size_needed - 1) +
                               " of an active-mesh vector with " + std::to_string(uint64_t(src_active->pd->n_owned) + src_active->pd->n_ghost) +
                               " local entries");
    if (c->type == MGB_DOUBLE)
      c->d->to_mg(*dst_level->pd, *src_active->pd);
    else
      c->f->to_mg(*dst_level->pf, *src_active->pd);
  });
}
int mgb_level_copy_from_mg(const mgb_level_copy *c, mgb_vector *dst_active, const mgb_vector *src_level)
{
  return guarded([&] {
    require_same_type(c->type, src_level->type);
    require_same_type(MGB_DOUBLE, dst_active->type);
    if (c->active_size_needed > uint64_t(dst_active->pd->n_owned) + dst_active->pd->n_ghost)
      throw std::runtime_error("mgb_level_copy_from_mg: the copy indices address entry " + std::to_string(c->active_size_needed - 1) +
                               " of an active-mesh vector with " + std::to_string(uint64_t(dst_active->pd->n_owned) + dst_active->pd->n_ghost) +
                               " local entries");
    if (c->type == MGB_DOUBLE)
      c->d->from_mg(*dst_active->pd, *src_level->pd);
    else
      c->f->from_mg(*dst_active->pd, *src_level->pf);
  });
}

int mgb_transfer_prolongate(const mgb_transfer *t, mgb_vector *dst_fine, const mgb_vector *src_coarse)
{
  return guarded([&] {
    require_same_type(t->type, dst_fine->type);
    require_same_type(t->type, src_coarse->type);
    if (t->type == MGB_DOUBLE)
      {
        dst_fine->pd->zero();
        t->pd->prolongate_and_add(*dst_fine->pd, *src_coarse->pd);
      }
    else
      {
        dst_fine->pf->zero();
        t->pf->prolongate_and_add(*dst_fine->pf, *src_coarse->pf);
      }
  });
}

int mgb_multigrid_create(int n_levels, const mgb_operator *const *level_operators, mgb_chebyshev *const *smoothers,
                         const mgb_transfer *const *transfers, mgb_chebyshev *coarse_preconditioner, unsigned coarse_maxiter,
                         double coarse_abstol, double coarse_reltol, mgb_multigrid **out)
{
  return guarded([&] {
    if (n_levels < 1 || !level_operators || !level_operators[0])
      throw std::runtime_error("mgb_multigrid_create: at least one level required");
    if (n_levels > 1 && (!smoothers || !transfers))
      throw std::runtime_error("mgb_multigrid_create: smoothers and transfers required above the coarse level");
    auto h      = std::make_unique<mgb_multigrid>();
    h->type     = level_operators[0]->type;
    h->n_levels = n_levels;
    if (h->type == MGB_DOUBLE)
      h->d = make_multigrid<double>(n_levels, level_operators, smoothers, transfers, coarse_preconditioner, coarse_maxiter, coarse_abstol,
                                    coarse_reltol, &mgb_operator::pd, &mgb_chebyshev::pd, &mgb_transfer::pd);
    else
      h->f = make_multigrid<float>(n_levels, level_operators, smoothers, transfers, coarse_preconditioner, coarse_maxiter, coarse_abstol,
                                   coarse_reltol, &mgb_operator::pf, &mgb_chebyshev::pf, &mgb_transfer::pf);
    *out = h.release();
  });
}
void mgb_multigrid_destroy(mgb_multigrid *m) { delete m; }
int  mgb_multigrid_set_edge_in(mgb_multigrid *m, int on)
{
  return guarded([&] {
    if (m->d)
      m->d->mg.edge_in = on != 0;
    else
      m->f->mg.edge_in = on != 0;
  });
}
int mgb_multigrid_set_level_copies(mgb_multigrid *m, const mgb_level_copy *const *copies)
{
  return guarded([&] {
    if (m->d)
      m->d->transfer.copies.clear();
    else
      m->f->transfer.copies.clear();
    if (!copies)
      return;
    for (int l = 0; l < m->n_levels; ++l)
      {
        if (!copies[l] || copies[l]->type != m->type)
          throw std::runtime_error("mgb_multigrid_set_level_copies: one copy object of the levels' number type per level required");
        if (m->d)
          m->d->transfer.copies.push_back(copies[l]->d.get());
        else
          m->f->transfer.copies.push_back(copies[l]->f.get());
      }
  });
}
int mgb_multigrid_vcycle(mgb_multigrid *m, mgb_vector *dst, const mgb_vector *src)
{
  return guarded([&] {
    require_same_type(MGB_DOUBLE, dst->type);
    require_same_type(MGB_DOUBLE, src->type);
    if (m->d)
      m->d->pmg.vmult(*dst->pd, *src->pd);
    else
      m->f->pmg.vmult(*dst->pd, *src->pd);
  });
}
int mgb_multigrid_enable_timers(mgb_multigrid *m, int on)
{
  return guarded([&] { (m->d ? m->d->timers : m->f->timers).enabled = on != 0; });
}
int mgb_multigrid_collect_times(mgb_multigrid *m, double *seconds)
{
  return guarded([&] {
    StageTimers &t     = m->d ? m->d->timers : m->f->timers;
    auto        &times = m->d ? m->d->times : m->f->times;
    MGB_CUDA(cudaStreamSynchronize(ctx().stream));
    times.assign(size_t(m->n_levels), std::array<double, N_STAGES + 2>{});
    t.collect(times, 0);
    for (int l = 0; l < m->n_levels; ++l)
      for (int s = 0; s < MGB_N_STAGES; ++s)
        seconds[l * MGB_N_STAGES + s] = times[size_t(l)][size_t(s)];
  });
}

int mgb_cg_solve(const mgb_operator *A, mgb_multigrid *preconditioner, mgb_vector *x, const mgb_vector *b, unsigned maxiter,
                 double abstol, double reltol, unsigned *n_iterations, double *final_residual)
{
  return guarded([&] {
    require_same_type(MGB_DOUBLE, A->type);
    require_same_type(MGB_DOUBLE, x->type);
    require_same_type(MGB_DOUBLE, b->type);
    SolverControl ctl;
    ctl.kind    = SolverControl::Reduction;
    ctl.maxiter = maxiter;
    ctl.abstol  = abstol;
    ctl.reltol  = reltol;
    static CGWork<double> plain_work;
    CGWork<double>       &w    = preconditioner ? preconditioner->cg_work : plain_work;
    const auto            Aop  = [&](DVec<double> &d, const DVec<double> &s) { A->pd->vmult(d, s); };
    const auto            Prec = [&](DVec<double> &d, const DVec<double> &s) {
      if (preconditioner->d)
        preconditioner->d->pmg.vmult(d, s);
      else
        preconditioner->f->pmg.vmult(d, s);
    };
    x->pd->zero(); // the reference always starts from a zero vector (multigrid_throughput.cc:1247-1248)
    const unsigned its = cg_solve<double>(Aop, Prec, preconditioner == nullptr, *x->pd, *b->pd, ctl, w, /*x_is_zero=*/true);
    if (ctl.last_step >= maxiter && ctl.last_value > std::max(abstol, ctl.reduced_tol))
      throw std::runtime_error("mgb_cg_solve: no convergence within maxiter iterations (SolverControl::NoConvergence)");
    if (n_iterations)
      *n_iterations = its;
    if (final_residual)
      *final_residual = ctl.last_value;
  });
}

} // extern "C"
