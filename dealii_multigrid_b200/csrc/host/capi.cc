// extern "C" boundary (include/mgb200.h): plain pointers and sizes in, opaque handles out.
#include "capi_handles.h"

#include "chunks.h"
#include "comm.h"
#include "partition.h"
#include "problem.h"

#include <chrono>
#include <cstring>
#include <map>

using namespace mgb;

namespace mgb
{
  std::string &capi_last_error()
  {
    thread_local std::string e;
    return e;
  }
} // namespace mgb

struct mgb_problem
{
  std::unique_ptr<ProblemBase>                         p;
  std::map<int, std::unique_ptr<mgb_dofmap>>           dofmaps;
  std::map<int, std::unique_ptr<mgb_operator>>         operators;
  std::map<int, std::unique_ptr<mgb_chebyshev>>        smoothers;
  std::map<int, std::unique_ptr<mgb_transfer>>         transfers;
  std::unique_ptr<mgb_operator>                        fine_op;
  std::unique_ptr<mgb_dofmap>                          fine_dofmap;
  std::unique_ptr<mgb_vector>                          rhs, solution;
  // mgb_problem_solve_host_batch: staging buffers, copy streams and events (created on first use)
  double      *stage_rhs = nullptr, *stage_sol = nullptr;
  cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;
  cudaEvent_t  ev_h2d = nullptr, ev_rhs_free = nullptr, ev_sol_ready = nullptr, ev_d2h_done = nullptr;
  ~mgb_problem()
  {
    if (stage_rhs)
      cudaFree(stage_rhs);
    if (stage_sol)
      cudaFree(stage_sol);
    for (cudaStream_t st : {h2d_stream, d2h_stream})
      if (st)
        cudaStreamDestroy(st);
    for (cudaEvent_t ev : {ev_h2d, ev_rhs_free, ev_sol_ready, ev_d2h_done})
      if (ev)
        cudaEventDestroy(ev);
  }
};

namespace
{
  void fill_result(const Result &r, mgb_result *o)
  {
    std::memset(o, 0, sizeof(*o));
    o->dim               = r.dim;
    o->degree            = r.degree;
    o->n_ref_global      = r.n_ref_global;
    o->n_ref_local       = r.n_ref_local;
    o->n_levels          = r.n_levels;
    o->n_iterations      = r.n_iterations;
    o->sub_comm_size     = r.sub_comm_size;
    o->n_cells           = r.n_cells;
    o->n_cells_hn        = r.n_cells_hn;
    o->n_cells_n         = r.n_cells_n;
    o->n_dofs            = r.n_dofs;
    o->time              = r.time;
    o->time_cg           = r.time_cg;
    o->throughput        = r.throughput;
    o->time_pre          = r.stage[STAGE_PRE];
    o->time_residuum     = r.stage[STAGE_RESIDUAL];
    o->time_res          = r.stage[STAGE_RESTRICTION];
    o->time_cs           = r.stage[STAGE_COARSE];
    o->time_pro          = r.stage[STAGE_PROLONGATION];
    o->time_edge_pro     = r.stage[STAGE_EDGE_PROLONGATION];
    o->time_post         = r.stage[STAGE_POST];
    o->time_to_mg        = r.stage[N_STAGES];
    o->time_to_global    = r.stage[N_STAGES + 1];
    o->initial_residual  = r.initial_residual;
    o->final_residual    = r.final_residual;
    o->coarse_iterations = r.coarse_iterations;
  }
} // namespace

namespace
{
  template <typename T>
  void call_fused(const OperatorT<T> &op, DVec<T> &out, const DVec<T> &x, int mode, const DVec<T> *b, const DVec<T> *dinv, const DVec<T> *xold,
                  double f1, double f2)
  {
    if (mode < 0 || mode > 3)
      throw std::runtime_error("mgb_operator_vmult_fused: mode must be 0..3");
    if ((mode >= 1 && !b) || (mode >= 2 && !dinv) || (mode == 3 && !xold))
      throw std::runtime_error("mgb_operator_vmult_fused: a vector this mode reads is NULL");
    if (!op.fused_available(out, x))
      throw std::runtime_error("mgb_operator_vmult_fused: not available (create the operator with MGB_CHUNKED=1; Cartesian cells, no "
                               "refinement-edge DoFs, one rank, out != x)");
    FusedEpilogue<T> ep;
    ep.mode = mode;
    ep.b    = b ? b->data : nullptr;
    ep.dinv = dinv ? dinv->data : nullptr;
    ep.xold = xold ? xold->data : nullptr;
    ep.f1   = T(f1);
    ep.f2   = T(f2);
    op.vmult_fused(out, x, ep);
  }
} // namespace

namespace mgb
{
  // The index arrays of the array-based entry points come from another program (the deal.II adapter); every one of them ends
  // up as an address inside a kernel, so they are checked here, on the host, before anything is uploaded.
  void check_index_list(const char *what, const uint32_t *idx, uint64_t n, uint64_t bound)
  {
    if (n && !idx)
      throw std::runtime_error(std::string(what) + ": null pointer for a non-empty index list");
    for (uint64_t i = 0; i < n; ++i)
      if (idx[i] >= bound)
        throw std::runtime_error(std::string(what) + ": entry " + std::to_string(i) + " = " + std::to_string(idx[i]) +
                                 " is outside [0, " + std::to_string(bound) + ")");
  }
  // 27 entity bases per cell: an unflagged entity of s DoFs occupies [base, base + s) of the local vector
  void check_entity_table(const char *what, const uint32_t *table, int64_t n_cells, int degree, uint64_t n_local)
  {
    uint32_t esize[27];
    for (int e = 0; e < 27; ++e)
      {
        esize[e] = 1;
        for (int k = 0; k < entity_dim(e); ++k)
          esize[e] *= uint32_t(degree - 1);
      }
    for (int64_t c = 0; c < n_cells; ++c)
      for (int e = 0; e < 27; ++e)
        {
          const uint32_t t = table[size_t(c) * 27 + e];
          if (esize[e] == 0 || (t & FLAG_BIT))
            continue;
          if (uint64_t(t) + esize[e] > n_local)
            throw std::runtime_error(std::string(what) + ": cell " + std::to_string(c) + ", entity " + std::to_string(e) + ": DoFs [" +
                                     std::to_string(t) + ", " + std::to_string(uint64_t(t) + esize[e]) + ") exceed the local vector of " +
                                     std::to_string(n_local) + " entries (owned + ghost)");
        }
  }
  void check_offsets(const char *what, const uint32_t *off, uint32_t n_ranks, uint64_t total)
  {
    if (n_ranks == 0)
      return;
    if (!off)
      throw std::runtime_error(std::string(what) + ": offsets required");
    if (off[0] != 0 || off[n_ranks] != total)
      throw std::runtime_error(std::string(what) + ": offsets must run from 0 to " + std::to_string(total));
    for (uint32_t r = 0; r < n_ranks; ++r)
      if (off[r] > off[r + 1])
        throw std::runtime_error(std::string(what) + ": offsets must ascend");
  }
} // namespace mgb

extern "C" {

const char *mgb_last_error(void) { return capi_last_error().c_str(); }
const char *mgb_version(void) { return "mgb200 0.1 (sm_100a)"; }

int mgb_init(int device)
{
  return guarded([&] {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0)
      throw std::runtime_error("mgb200: no CUDA device available (this library has no CPU path)");
    MGB_CUDA(cudaSetDevice(device));
    ctx().device = device;
    ctx().ensure();
  });
}
int mgb_set_stream(void *s)
{
  return guarded([&] { ctx().stream = (cudaStream_t)s; });
}
int mgb_synchronize(void)
{
  return guarded([&] { MGB_CUDA(cudaStreamSynchronize(ctx().stream)); });
}
int64_t mgb_launch_count(void) { return ctx().launches; }

int mgb_profile_start(int64_t n_cells)
{
  return guarded([&] {
    Context &c     = ctx();
    c.prof_enabled = true;
    c.prof_cells   = n_cells;
    c.prof_used    = 0;
  });
}
int mgb_profile_stop(double *total_ms, int64_t *n_launches)
{
  return guarded([&] {
    Context &c = ctx();
    MGB_CUDA(cudaStreamSynchronize(c.stream));
    double sum = 0;
    for (size_t i = 0; i < c.prof_used; ++i)
      {
        float ms = 0;
        MGB_CUDA(cudaEventElapsedTime(&ms, c.prof_events[i].first, c.prof_events[i].second));
        sum += ms;
      }
    if (total_ms)
      *total_ms = sum;
    if (n_launches)
      *n_launches = (int64_t)c.prof_used;
    c.prof_enabled = false;
    c.prof_used    = 0;
  });
}

int mgb_comm_unique_id(void *out128, const char *libnccl_path)
{
  return guarded([&] { nccl_get_unique_id(out128, libnccl_path); });
}
int mgb_comm_init(const void *id, int rank, int size, const char *libnccl_path)
{
  return guarded([&] {
    ctx().ensure();
    if (ctx().comm)
      throw std::runtime_error("communicator already initialised");
    ctx().comm = create_nccl_comm(id, rank, size, libnccl_path);
  });
}
int mgb_comm_enable_peer_exchange(int *enabled)
{
  return guarded([&] {
    const bool on = p2p_back_partitioners();
    if (enabled)
      *enabled = on ? 1 : 0;
  });
}
int mgb_comm_finalize(void)
{
  return guarded([&] {
    MGB_CUDA(cudaDeviceSynchronize());
    release_all_graphs(); // graphs that captured NCCL calls hold the communicator (ncclCommDestroy would wait for them)
    p2p_release_all();    // collective: peers may still acknowledge into our staging segments
    delete ctx().comm;
    ctx().comm = nullptr;
  });
}

int mgb_fe_1d(int degree, double *nodes, double *M, double *K, double *S0, double *S1, double *Ph)
{
  return guarded([&] {
    FE1D         fe(degree);
    const size_t n = fe.n;
    if (nodes)
      std::memcpy(nodes, fe.nodes.data(), n * sizeof(double));
    if (M)
      std::memcpy(M, fe.M.data(), n * n * sizeof(double));
    if (K)
      std::memcpy(K, fe.K.data(), n * n * sizeof(double));
    if (S0)
      std::memcpy(S0, fe.S[0].data(), n * n * sizeof(double));
    if (S1)
      std::memcpy(S1, fe.S[1].data(), n * n * sizeof(double));
    if (Ph)
      std::memcpy(Ph, fe.Ph.data(), fe.Ph.size() * sizeof(double));
  });
}
int mgb_fe_1d_fdm(int degree, double *Tm, double *lambda)
{
  return guarded([&] {
    FE1D fe(degree);
    if (Tm)
      std::memcpy(Tm, fe.Tm.data(), fe.Tm.size() * sizeof(double));
    if (lambda)
      std::memcpy(lambda, fe.lambda.data(), fe.lambda.size() * sizeof(double));
  });
}
int mgb_fe_p_embedding(int pf, int pc, double *Pp)
{
  return guarded([&] {
    const auto P = p_embedding(pf, pc);
    std::memcpy(Pp, P.data(), P.size() * sizeof(double));
  });
}

// ---- mesh
int mgb_mesh_create(const char *geometry, unsigned n_ref_global, unsigned n_ref_local, mgb_mesh **out)
{
  return guarded([&] {
    auto h = std::make_unique<mgb_mesh>();
    h->m   = Mesh::create(geometry, n_ref_global, n_ref_local);
    h->m.rebuild_index();
    *out = h.release();
  });
}
int mgb_mesh_coarsen(const mgb_mesh *fine, mgb_mesh **out)
{
  return guarded([&] {
    auto h = std::make_unique<mgb_mesh>();
    h->m   = fine->m.coarsen_global();
    h->m.rebuild_index();
    *out = h.release();
  });
}
int mgb_mesh_level(const mgb_mesh *active, int level, mgb_mesh **out)
{
  return guarded([&] {
    if (level < 0 || level >= active->m.n_global_levels())
      throw std::runtime_error("mgb_mesh_level: no such level");
    auto h = std::make_unique<mgb_mesh>();
    h->m   = active->m.level_mesh(level);
    *out   = h.release();
  });
}
int64_t mgb_mesh_n_cells(const mgb_mesh *m) { return (int64_t)m->m.cells.size(); }
int     mgb_mesh_n_levels(const mgb_mesh *m) { return m->m.n_global_levels(); }
int     mgb_mesh_get_cells(const mgb_mesh *m, uint32_t *o)
{
  return guarded([&] {
    for (size_t i = 0; i < m->m.cells.size(); ++i)
      {
        o[4 * i + 0] = m->m.cells[i].level;
        o[4 * i + 1] = m->m.cells[i].x;
        o[4 * i + 2] = m->m.cells[i].y;
        o[4 * i + 3] = m->m.cells[i].z;
      }
  });
}
void mgb_mesh_destroy(mgb_mesh *m) { delete m; }

// ---- dofmap
int mgb_dofmap_create(const mgb_mesh *m, int degree, int n_ranks, int rank, mgb_dofmap **out)
{
  return guarded([&] {
    if (degree > MAX_DEGREE)
      throw std::runtime_error("degree > MAX_DEGREE");
    auto h   = std::make_unique<mgb_dofmap>();
    h->owned = std::make_unique<DoFMap>(DoFMap::build(m->m, degree, n_ranks, rank));
    h->d     = h->owned.get();
    *out     = h.release();
  });
}
int mgb_dofmap_create_partitioned(const mgb_mesh *m, int degree, int n_ranks, int rank, const int64_t *cell_begin, mgb_dofmap **out)
{
  return guarded([&] {
    if (degree > MAX_DEGREE)
      throw std::runtime_error("degree > MAX_DEGREE");
    if (!cell_begin || n_ranks < 1)
      throw std::runtime_error("mgb_dofmap_create_partitioned: cell_begin[n_ranks+1] required");
    auto h   = std::make_unique<mgb_dofmap>();
    h->owned = std::make_unique<DoFMap>(DoFMap::build(m->m, degree, n_ranks, rank, std::vector<int64_t>(cell_begin, cell_begin + n_ranks + 1)));
    h->d     = h->owned.get();
    *out     = h.release();
  });
}
int mgb_partition_sequence(const char *policy_name, const mgb_mesh *const *meshes, int n_meshes, int n_ranks, int64_t min_cells_per_rank,
                           int64_t *cell_begin)
{
  return guarded([&] {
    // the policy only reads the meshes; copy them into the contiguous sequence its interface takes
    std::vector<Mesh> seq;
    seq.reserve(n_meshes);
    for (int i = 0; i < n_meshes; ++i)
      seq.push_back(meshes[i]->m);
    for (auto &m : seq)
      m.rebuild_index();
    const auto b = PartitionPolicy::parse(policy_name ? policy_name : "").sequence(seq, n_ranks, std::max<int64_t>(1, min_cells_per_rank));
    for (int i = 0; i < n_meshes; ++i)
      std::copy(b[i].begin(), b[i].end(), cell_begin + size_t(i) * (n_ranks + 1));
  });
}
int mgb_partition_first_child(const mgb_mesh *parents, const mgb_mesh *children, int n_ranks, const int64_t *child_begin, int64_t *cell_begin)
{
  return guarded([&] {
    const auto b = first_child_boundaries(parents->m, children->m, std::vector<int64_t>(child_begin, child_begin + n_ranks + 1));
    std::copy(b.begin(), b.end(), cell_begin);
  });
}
int mgb_mesh_hanging_cells(const mgb_mesh *m, uint8_t *flags)
{
  return guarded([&] {
    const auto f = cells_with_hanging_nodes(m->m);
    std::copy(f.begin(), f.end(), flags);
  });
}
namespace
{
  void export_statistics(const HierarchyStatistics &st, mgb_statistics *out, int64_t *per_level)
  {
    out->workload_eff      = st.workload_eff;
    out->workload_path_max = st.workload_path_max;
    out->vertical_eff      = st.vertical_eff;
    out->horizontal_eff    = st.horizontal_eff;
    out->mem_total         = st.mem_total;
    out->verbose           = 0;
    if (per_level)
      for (size_t l = 0; l < st.levels.size(); ++l)
        {
          const auto &L = st.levels[l];
          int64_t    *o = per_level + l * MGB_STATISTICS_LEVEL_VALUES;
          o[0] = L.n_cells_min, o[1] = L.n_cells_max, o[2] = L.n_cells_sum, o[3] = L.vertical_local, o[4] = L.vertical_remote;
          o[5] = L.horizontal_local, o[6] = L.horizontal_remote;
        }
  }
} // namespace
int mgb_hierarchy_statistics(const mgb_mesh *const *meshes, int n_meshes, int n_ranks, const int64_t *cell_begin, mgb_statistics *out,
                             int64_t *per_level)
{
  return guarded([&] {
    std::vector<const Mesh *>         ms;
    std::vector<std::vector<int64_t>> cb;
    for (int i = 0; i < n_meshes; ++i)
      {
        ms.push_back(&meshes[i]->m);
        cb.emplace_back(cell_begin + size_t(i) * (n_ranks + 1), cell_begin + size_t(i + 1) * (n_ranks + 1));
        if (cb.back().front() != 0 || cb.back().back() != (int64_t)meshes[i]->m.cells.size())
          throw std::runtime_error("mgb_hierarchy_statistics: cell_begin does not span the mesh");
      }
    export_statistics(hierarchy_statistics(ms, cb, n_ranks), out, per_level);
  });
}
int mgb_dofmap_view_get(const mgb_dofmap *h, mgb_dofmap_view *v)
{
  return guarded([&] {
    const DoFMap &d     = *h->d;
    v->degree           = d.degree;
    v->n_ranks          = d.n_ranks;
    v->rank             = d.rank;
    v->mesh_max_level   = d.mesh_max_level;
    v->n_cells_global   = d.n_cells_global;
    v->cell_begin       = d.cell_begin;
    v->n_cells          = d.n_cells;
    v->n_cells_hanging  = d.n_cells_hanging;
    v->n_dofs_global    = d.part.n_global;
    v->owned_begin      = d.part.owned_begin;
    v->n_owned          = d.part.n_owned;
    v->n_ghost          = d.part.n_ghost;
    v->n_constrained    = (uint32_t)d.constrained.size();
    v->n_ghost_ranks    = (uint32_t)d.part.ghost_ranks.size();
    v->n_import_ranks   = (uint32_t)d.part.import_ranks.size();
    v->n_import_indices = (uint32_t)d.part.import_indices.size();
    v->eval_table       = d.eval_table.data();
    v->own_table        = d.own_table.data();
    v->mask             = d.mask.data();
    v->cell_level       = d.cell_level.data();
    v->cell_xyz         = d.cell_xyz.data();
    v->constrained      = d.constrained.data();
    v->ghost_global     = d.part.ghost_global.data();
    v->ghost_ranks      = d.part.ghost_ranks.data();
    v->ghost_offsets    = d.part.ghost_offsets.data();
    v->import_ranks     = d.part.import_ranks.data();
    v->import_offsets   = d.part.import_offsets.data();
    v->import_indices   = d.part.import_indices.data();
  });
}
int mgb_dofmap_expand(const mgb_dofmap *h, int which, uint32_t *out)
{
  return guarded([&] {
    std::vector<uint32_t> tmp;
    h->d->expand(which == 0 ? h->d->eval_table : h->d->own_table, tmp);
    std::memcpy(out, tmp.data(), tmp.size() * sizeof(uint32_t));
  });
}
int mgb_assemble_rhs(const mgb_mesh *m, const mgb_dofmap *h, const char *simulation_type, double *rhs_local)
{
  return guarded([&] {
    std::vector<double> b;
    assemble_rhs(m->m, *h->d, simulation_type ? simulation_type : "Constant", b);
    std::memcpy(rhs_local, b.data(), b.size() * sizeof(double));
  });
}
int mgb_dofmap_chunk_plan_check(const mgb_dofmap *h, int value_bytes, double *stats)
{
  return guarded([&] {
    const DoFMap   &d    = *h->d;
    const auto      t0   = std::chrono::steady_clock::now();
    const ChunkPlan plan = build_chunk_plan(d.degree, d.n_cells, d.eval_table.data(), d.mask.data(), d.cell_level.data(), d.part.n_owned,
                                            d.part.n_ghost, value_bytes, d.cell_begin);
    stats[10] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    stats[0] = check_chunk_plan(plan, d.n_cells, d.eval_table.data(), d.part.n_owned, d.part.n_ghost);
    stats[1] = plan.n_chunks;
    stats[2] = plan.n_compute_blobs;
    stats[3] = plan.n_pull_blobs;
    stats[4] = double(plan.pool.size());
    stats[5] = double(plan.n_foreign_low);
    stats[6] = double(plan.n_foreign_late);
    stats[7] = double(plan.n_foreign_ghost);
    stats[8] = double(plan.side_size);
    stats[9] = plan.max_slots;
  });
}
void mgb_dofmap_destroy(mgb_dofmap *d) { delete d; }

int mgb_transfer_map_create(const mgb_mesh *mf, const mgb_dofmap *f, const mgb_mesh *mc, const mgb_dofmap *c, mgb_transfer_map **out)
{
  return guarded([&] {
    auto h          = std::make_unique<mgb_transfer_map>();
    h->t            = TransferMap::build(mf->m, *f->d, mc->m, *c->d);
    h->n_fine_cells = f->d->n_cells;
    *out            = h.release();
  });
}
int mgb_transfer_map_create_ls(const mgb_mesh *mf, const mgb_dofmap *f, const mgb_mesh *mc, const mgb_dofmap *c, mgb_transfer_map **out)
{
  return guarded([&] {
    auto h          = std::make_unique<mgb_transfer_map>();
    h->t            = TransferMap::build(mf->m, *f->d, mc->m, *c->d, /*local_smoothing=*/true);
    h->n_fine_cells = f->d->n_cells;
    *out            = h.release();
  });
}
int mgb_level_edges_create(const mgb_mesh *level_mesh, const mgb_dofmap *level_dofs, mgb_level_edges **out)
{
  return guarded([&] {
    auto h = std::make_unique<mgb_level_edges>();
    h->e   = LevelEdges::build(level_mesh->m, *level_dofs->d);
    *out   = h.release();
  });
}
int mgb_level_edges_view_get(const mgb_level_edges *e, mgb_level_edges_view *v)
{
  return guarded([&] {
    v->n_edge_dofs  = (uint32_t)e->e.edge_dofs.size();
    v->n_edge_cells = (uint32_t)e->e.edge_cells.size();
    v->any_global   = e->e.any_global ? 1 : 0;
    v->edge_dofs    = e->e.edge_dofs.data();
    v->edge_cells   = e->e.edge_cells.data();
  });
}
void mgb_level_edges_destroy(mgb_level_edges *e) { delete e; }
int  mgb_copy_indices_create(const mgb_dofmap *active_dofs, const mgb_mesh *active_mesh, const mgb_mesh *level_mesh,
                             const mgb_dofmap *level_dofs, mgb_copy_indices **out)
{
  return guarded([&] {
    auto h = std::make_unique<mgb_copy_indices>();
    h->c   = CopyIndices::build(*active_dofs->d, level_mesh->m, *level_dofs->d, active_mesh ? &active_mesh->m : nullptr);
    *out   = h.release();
  });
}
int mgb_copy_indices_view_get(const mgb_copy_indices *c, mgb_copy_indices_view *v)
{
  return guarded([&] {
    v->n            = (uint32_t)c->c.active_idx.size();
    v->active_idx   = c->c.active_idx.data();
    v->level_idx    = c->c.level_idx.data();
    v->distributed  = c->c.distributed ? 1 : 0;
    v->owned_begin  = c->c.part.owned_begin;
    v->n_owned      = c->c.part.n_owned;
    v->n_ghost      = c->c.part.n_ghost;
    v->ghost_global = c->c.part.ghost_global.data();
  });
}
void mgb_copy_indices_destroy(mgb_copy_indices *c) { delete c; }
int  mgb_transfer_map_view_get(const mgb_transfer_map *t, mgb_transfer_map_view *v)
{
  return guarded([&] {
    v->degree_fine    = t->t.degree_fine;
    v->degree_coarse  = t->t.degree_coarse;
    v->n_coarse_cells = t->t.n_coarse_cells;
    v->n_fine_cells   = t->n_fine_cells;
    v->fine_first     = t->t.fine_first.data();
    v->n_children     = t->t.n_children.data();
    v->weight_code    = t->t.weight_code.data();
    v->fine_own       = t->t.fine_own.data();
    v->n_fine_slots   = t->t.n_fine_slots;
    v->distributed    = t->t.distributed ? 1 : 0;
    const Partitioner &p = t->t.part;
    v->n_owned          = p.n_owned;
    v->n_ghost          = p.n_ghost;
    v->owned_begin      = p.owned_begin;
    v->ghost_global     = p.ghost_global.data();
    v->n_ghost_ranks    = (uint32_t)p.ghost_ranks.size();
    v->ghost_ranks      = p.ghost_ranks.data();
    v->ghost_offsets    = p.ghost_offsets.data();
    v->n_import_ranks   = (uint32_t)p.import_ranks.size();
    v->import_ranks     = p.import_ranks.data();
    v->import_offsets   = p.import_offsets.data();
    v->n_import_indices = (uint32_t)p.import_indices.size();
    v->import_indices   = p.import_indices.data();
  });
}
void mgb_transfer_map_destroy(mgb_transfer_map *t) { delete t; }

// ---- vectors
int mgb_vector_create(uint32_t n_owned, uint32_t n_ghost, int type, mgb_vector **out)
{
  return guarded([&] {
    ctx().ensure();
    auto h  = std::make_unique<mgb_vector>();
    h->type = type;
    if (type == MGB_DOUBLE)
      {
        h->od = std::make_unique<DVec<double>>();
        h->od->reinit(n_owned, n_ghost);
        h->pd = h->od.get();
      }
    else if (type == MGB_FLOAT)
      {
        h->of = std::make_unique<DVec<float>>();
        h->of->reinit(n_owned, n_ghost);
        h->pf = h->of.get();
      }
    else
      throw std::runtime_error("unknown number type");
    *out = h.release();
  });
}
int mgb_vector_wrap(void *ptr, uint32_t n_owned, uint32_t n_ghost, int type, mgb_vector **out)
{
  return guarded([&] {
    ctx().ensure();
    auto h  = std::make_unique<mgb_vector>();
    h->type = type;
    if (type == MGB_DOUBLE)
      {
        h->od = std::make_unique<DVec<double>>();
        h->od->wrap((double *)ptr, n_owned, n_ghost);
        h->pd = h->od.get();
      }
    else if (type == MGB_FLOAT)
      {
        h->of = std::make_unique<DVec<float>>();
        h->of->wrap((float *)ptr, n_owned, n_ghost);
        h->pf = h->of.get();
      }
    else
      throw std::runtime_error("unknown number type");
    *out = h.release();
  });
}
void     mgb_vector_destroy(mgb_vector *v) { delete v; }
uint32_t mgb_vector_n_owned(const mgb_vector *v) { return v->type == MGB_DOUBLE ? v->pd->n_owned : v->pf->n_owned; }
void    *mgb_vector_data(mgb_vector *v) { return v->type == MGB_DOUBLE ? (void *)v->pd->data : (void *)v->pf->data; }

#define VEC_DISPATCH(v, EXPR_D, EXPR_F) \
  do                                    \
    {                                   \
      if ((v)->type == MGB_DOUBLE)      \
        {                               \
          auto &V = *(v)->pd;           \
          (void)V;                      \
          EXPR_D;                       \
        }                               \
      else                              \
        {                               \
          auto &V = *(v)->pf;           \
          (void)V;                      \
          EXPR_F;                       \
        }                               \
    }                                   \
  while (0)

int mgb_vector_copy_from_host(mgb_vector *v, const void *host, uint32_t n)
{
  return guarded([&] {
    VEC_DISPATCH(v,
                 { if (n > V.size_local()) throw std::runtime_error("copy_from_host: n too large");
                   MGB_CUDA(cudaMemcpyAsync(V.data, host, size_t(n) * sizeof(double), cudaMemcpyHostToDevice, ctx().stream)); },
                 { if (n > V.size_local()) throw std::runtime_error("copy_from_host: n too large");
                   MGB_CUDA(cudaMemcpyAsync(V.data, host, size_t(n) * sizeof(float), cudaMemcpyHostToDevice, ctx().stream)); });
    MGB_CUDA(cudaStreamSynchronize(ctx().stream));
  });
}
int mgb_vector_copy_to_host(const mgb_vector *v, void *host, uint32_t n)
{
  return guarded([&] {
    VEC_DISPATCH(v,
                 { if (n > V.size_local()) throw std::runtime_error("copy_to_host: n too large");
                   MGB_CUDA(cudaMemcpyAsync(host, V.data, size_t(n) * sizeof(double), cudaMemcpyDeviceToHost, ctx().stream)); },
                 { if (n > V.size_local()) throw std::runtime_error("copy_to_host: n too large");
                   MGB_CUDA(cudaMemcpyAsync(host, V.data, size_t(n) * sizeof(float), cudaMemcpyDeviceToHost, ctx().stream)); });
    MGB_CUDA(cudaStreamSynchronize(ctx().stream));
  });
}
int mgb_vector_set(mgb_vector *v, double a)
{
  return guarded([&] { VEC_DISPATCH(v, V.set(a), V.set(float(a))); });
}
int mgb_vector_add(mgb_vector *y, double a, const mgb_vector *x)
{
  return guarded([&] {
    require_same_type(y->type, x->type);
    VEC_DISPATCH(y, V.add(a, *x->pd), V.add(float(a), *x->pf));
  });
}
int mgb_vector_sadd(mgb_vector *y, double s, double a, const mgb_vector *x)
{
  return guarded([&] {
    require_same_type(y->type, x->type);
    VEC_DISPATCH(y, V.sadd(s, a, *x->pd), V.sadd(float(s), float(a), *x->pf));
  });
}
int mgb_vector_equ(mgb_vector *y, double a, const mgb_vector *x)
{
  return guarded([&] {
    require_same_type(y->type, x->type);
    VEC_DISPATCH(y, V.equ(a, *x->pd), V.equ(float(a), *x->pf));
  });
}
int mgb_vector_dot(const mgb_vector *x, const mgb_vector *y, double *result)
{
  return guarded([&] {
    require_same_type(y->type, x->type);
    VEC_DISPATCH(x, *result = V.dot(*y->pd), *result = V.dot(*y->pf));
  });
}
int mgb_vector_l2_norm(const mgb_vector *x, double *result)
{
  return guarded([&] { VEC_DISPATCH(x, *result = V.l2_norm(), *result = V.l2_norm()); });
}
int mgb_vector_add_and_dot(mgb_vector *g, double a, const mgb_vector *h, double *result)
{
  return guarded([&] {
    require_same_type(g->type, h->type);
    VEC_DISPATCH(g, *result = V.add_and_dot(a, *h->pd), *result = V.add_and_dot(float(a), *h->pf));
  });
}
int mgb_vector_update_ghost_values(const mgb_vector *v)
{
  return guarded([&] { VEC_DISPATCH(v, V.update_ghost_values(), V.update_ghost_values()); });
}
int mgb_vector_compress_add(mgb_vector *v)
{
  return guarded([&] { VEC_DISPATCH(v, V.compress_add(), V.compress_add()); });
}
int mgb_vector_zero_out_ghost_values(mgb_vector *v)
{
  return guarded([&] { VEC_DISPATCH(v, V.zero_out_ghost_values(), V.zero_out_ghost_values()); });
}

// ---- operator
namespace
{
  std::unique_ptr<mgb_operator> make_operator(const OperatorSetup &s, int type)
  {
    auto h  = std::make_unique<mgb_operator>();
    h->type = type;
    if (type == MGB_DOUBLE)
      {
        h->od = std::make_unique<OperatorT<double>>(s);
        h->pd = h->od.get();
      }
    else if (type == MGB_FLOAT)
      {
        h->of = std::make_unique<OperatorT<float>>(s);
        h->pf = h->of.get();
      }
    else
      throw std::runtime_error("unknown number type");
    return h;
  }

  // compress a full lexicographic index map into the 27-entry entity table
  std::vector<uint32_t> compress_dof_indices(const uint32_t *idx, long long n_cells, int p)
  {
    const int        n = p + 1, n3 = n * n * n;
    std::vector<int> ent(n3), off(n3);
    for (int k = 0; k < n; ++k)
      for (int j = 0; j < n; ++j)
        for (int i = 0; i < n; ++i)
          node_entity(p, i, j, k, ent[(k * n + j) * n + i], off[(k * n + j) * n + i]);
    std::vector<uint32_t> tab(size_t(n_cells) * 27, 0);
    for (long long c = 0; c < n_cells; ++c)
      {
        for (int q = 0; q < n3; ++q)
          if (off[q] == 0)
            tab[c * 27 + ent[q]] = idx[c * n3 + q];
        for (int q = 0; q < n3; ++q)
          {
            const uint32_t t = tab[c * 27 + ent[q]];
            if (idx[c * n3 + q] != (((t & INDEX_MASK) + uint32_t(off[q])) | (t & FLAG_BIT)))
              throw std::runtime_error("dof_indices of a cell are not contiguous per mesh entity; cannot build the entity table");
          }
      }
    return tab;
  }
} // namespace

int mgb_operator_create(const mgb_operator_desc *d, mgb_operator **out)
{
  return guarded([&] {
    if (!d || !out)
      throw std::runtime_error("mgb_operator_create: null argument");
    if ((d->entity_table == nullptr) == (d->dof_indices == nullptr))
      throw std::runtime_error("exactly one of entity_table / dof_indices must be given");
    if (d->degree < 1 || d->degree > MAX_DEGREE)
      throw std::runtime_error("degree out of range");
    if (d->n_cells < 0 || (d->n_cells > 0 && !d->cell_level))
      throw std::runtime_error("mgb_operator_create: n_cells >= 0 and cell_level are required");
    const uint64_t n_local = uint64_t(d->n_owned) + d->n_ghost;
    if (n_local >= uint64_t(INDEX_MASK))
      throw std::runtime_error("mgb_operator_create: more than 2^31 local DoFs");
    check_index_list("mgb_operator_create: constrained", d->constrained, d->n_constrained, d->n_owned);
    check_index_list("mgb_operator_create: edge_constrained", d->edge_constrained, d->n_edge_constrained, d->n_owned);
    check_index_list("mgb_operator_create: edge_cells", d->edge_cells, d->n_edge_cells, uint64_t(d->n_cells));
    check_offsets("mgb_operator_create: ghost_offsets", d->ghost_offsets, d->n_ghost_ranks, d->n_ghost);
    if (d->n_ghost && (!d->ghost_global || d->n_ghost_ranks == 0 || !d->ghost_ranks))
      throw std::runtime_error("mgb_operator_create: ghost entries need ghost_global, ghost_ranks and ghost_offsets");
    if (d->n_import_ranks)
      {
        if (!d->import_offsets || !d->import_ranks)
          throw std::runtime_error("mgb_operator_create: import_ranks and import_offsets required");
        check_offsets("mgb_operator_create: import_offsets", d->import_offsets, d->n_import_ranks, d->import_offsets[d->n_import_ranks]);
        check_index_list("mgb_operator_create: import_indices", d->import_indices, d->import_offsets[d->n_import_ranks], d->n_owned);
      }
    for (int64_t c = 0; c < d->n_cells; ++c)
      if (d->cell_level[c] > 30)
        throw std::runtime_error("mgb_operator_create: cell_level above 30");
    std::vector<uint32_t> compressed;
    OperatorSetup         s;
    s.degree  = d->degree;
    s.n_cells = d->n_cells;
    if (d->entity_table)
      s.entity_table = d->entity_table;
    else
      {
        compressed     = compress_dof_indices(d->dof_indices, d->n_cells, d->degree);
        s.entity_table = compressed.data();
      }
    check_entity_table("mgb_operator_create", s.entity_table, d->n_cells, d->degree, n_local);
    s.mask                        = d->mask;
    s.cell_level                  = d->cell_level;
    s.root_length                 = d->root_length;
    s.constrained                 = d->constrained;
    s.n_constrained               = d->n_constrained;
    s.edge_constrained            = d->edge_constrained;
    s.n_edge_constrained          = d->n_edge_constrained;
    s.has_edge_constrained_global = d->has_edge_constrained_global != 0;
    s.edge_cells                  = d->edge_cells;
    s.n_edge_cells                = d->n_edge_cells;
    s.n_dofs_global               = d->n_dofs_global;
    Partitioner part;
    part.n_global    = d->n_dofs_global;
    part.owned_begin = d->owned_begin;
    part.n_owned     = d->n_owned;
    part.n_ghost     = d->n_ghost;
    if (d->n_ghost)
      part.ghost_global.assign(d->ghost_global, d->ghost_global + d->n_ghost);
    part.ghost_ranks.assign(d->ghost_ranks, d->ghost_ranks + d->n_ghost_ranks);
    if (d->ghost_offsets)
      part.ghost_offsets.assign(d->ghost_offsets, d->ghost_offsets + d->n_ghost_ranks + 1);
    else
      part.ghost_offsets.assign(1, 0u);
    part.import_ranks.assign(d->import_ranks, d->import_ranks + d->n_import_ranks);
    if (d->import_offsets)
      part.import_offsets.assign(d->import_offsets, d->import_offsets + d->n_import_ranks + 1);
    else
      part.import_offsets.assign(1, 0u);
    if (d->n_import_ranks)
      part.import_indices.assign(d->import_indices, d->import_indices + part.import_offsets.back());
    s.partitioner = &part;
    *out          = make_operator(s, d->number_type).release();
  });
}
int mgb_operator_create_from_dofmap(const mgb_dofmap *dm, int type, mgb_operator **out)
{
  return guarded([&] {
    const DoFMap &d = *dm->d;
    OperatorSetup s;
    s.degree        = d.degree;
    s.n_cells       = d.n_cells;
    s.entity_table  = d.eval_table.data();
    s.mask          = d.mask.data();
    s.cell_level    = d.cell_level.data();
    s.constrained   = d.constrained.data();
    s.n_constrained = (uint32_t)d.constrained.size();
    s.n_dofs_global = d.part.n_global;
    s.partitioner   = &d.part;
    *out            = make_operator(s, type).release();
  });
}
int mgb_operator_create_level(const mgb_dofmap *dm, const mgb_level_edges *edges, int type, mgb_operator **out)
{
  return guarded([&] {
    const DoFMap &d = *dm->d;
    OperatorSetup s;
    s.degree        = d.degree;
    s.n_cells       = d.n_cells;
    s.entity_table  = d.eval_table.data();
    s.mask          = d.mask.data();
    s.cell_level    = d.cell_level.data();
    s.constrained   = d.constrained.data();
    s.n_constrained = (uint32_t)d.constrained.size();
    s.n_dofs_global = d.part.n_global;
    s.partitioner   = &d.part;
    const LevelEdges &e            = edges->e;
    s.edge_constrained             = e.edge_dofs.data();
    s.n_edge_constrained           = (uint32_t)e.edge_dofs.size();
    s.has_edge_constrained_global  = e.any_global;
    s.edge_cells                   = e.edge_cells.data();
    s.n_edge_cells                 = (uint32_t)e.edge_cells.size();
    *out                           = make_operator(s, type).release();
  });
}
void    mgb_operator_destroy(mgb_operator *op) { delete op; }
int64_t mgb_operator_m(const mgb_operator *op) { return op->type == MGB_DOUBLE ? op->pd->m() : op->pf->m(); }

int mgb_operator_initialize_dof_vector(const mgb_operator *op, mgb_vector **out)
{
  return guarded([&] {
    auto h  = std::make_unique<mgb_vector>();
    h->type = op->type;
    if (op->type == MGB_DOUBLE)
      {
        h->od = std::make_unique<DVec<double>>();
        op->pd->initialize_dof_vector(*h->od);
        h->pd = h->od.get();
      }
    else
      {
        h->of = std::make_unique<DVec<float>>();
        op->pf->initialize_dof_vector(*h->of);
        h->pf = h->of.get();
      }
    *out = h.release();
  });
}

#define OP_CALL(METHOD)                                                                \
  return guarded([&] {                                                                 \
    require_same_type(op->type, dst->type);                                            \
    require_same_type(op->type, src->type);                                            \
    if (op->type == MGB_DOUBLE)                                                        \
      op->pd->METHOD(*dst->pd, *src->pd);                                              \
    else                                                                               \
      op->pf->METHOD(*dst->pf, *src->pf);                                              \
  })

int mgb_operator_vmult(const mgb_operator *op, mgb_vector *dst, const mgb_vector *src) { OP_CALL(vmult); }
int mgb_operator_tvmult(const mgb_operator *op, mgb_vector *dst, const mgb_vector *src) { OP_CALL(Tvmult); }
int mgb_operator_vmult_interface_down(const mgb_operator *op, mgb_vector *dst, const mgb_vector *src) { OP_CALL(vmult_interface_down); }
int mgb_operator_vmult_interface_up(const mgb_operator *op, mgb_vector *dst, const mgb_vector *src) { OP_CALL(vmult_interface_up); }
int mgb_operator_fused_available(const mgb_operator *op)
{
  // vectors of the operator's own layout (16-byte aligned allocations, distinct out / x)
  return (op->type == MGB_DOUBLE ? op->pd->chunked() && op->pd->partitioner()->trivial() && !op->pd->has_edge_dofs() :
                                   op->pf->chunked() && op->pf->partitioner()->trivial() && !op->pf->has_edge_dofs()) ?
           1 :
           0;
}
int mgb_operator_vmult_fused(const mgb_operator *op, mgb_vector *out, const mgb_vector *x, int mode, const mgb_vector *b,
                             const mgb_vector *dinv, const mgb_vector *xold, double f1, double f2)
{
  return guarded([&] {
    require_same_type(op->type, out->type);
    require_same_type(op->type, x->type);
    for (const mgb_vector *v : {b, dinv, xold})
      if (v)
        require_same_type(op->type, v->type);
    if (op->type == MGB_DOUBLE)
      call_fused<double>(*op->pd, *out->pd, *x->pd, mode, b ? b->pd : nullptr, dinv ? dinv->pd : nullptr, xold ? xold->pd : nullptr, f1, f2);
    else
      call_fused<float>(*op->pf, *out->pf, *x->pf, mode, b ? b->pf : nullptr, dinv ? dinv->pf : nullptr, xold ? xold->pf : nullptr, f1, f2);
  });
}
int mgb_operator_compute_inverse_diagonal(const mgb_operator *op, mgb_vector *diag)
{
  return guarded([&] {
    require_same_type(op->type, diag->type);
    if (op->type == MGB_DOUBLE)
      op->pd->compute_inverse_diagonal(*diag->pd);
    else
      op->pf->compute_inverse_diagonal(*diag->pf);
  });
}

int mgb_operator_set_cell_vertices(mgb_operator *op, const double *xyz, int64_t n_cells)
{
  return guarded([&] {
    const long long have = op->type == MGB_DOUBLE ? op->pd->cell_data().n_cells : op->pf->cell_data().n_cells;
    if (n_cells != have)
      throw std::runtime_error("mgb_operator_set_cell_vertices: the operator has " + std::to_string(have) + " cells, got " +
                               std::to_string((long long)n_cells));
    if (op->type == MGB_DOUBLE)
      op->pd->set_cell_vertices(xyz);
    else
      op->pf->set_cell_vertices(xyz);
  });
}

// ---- Chebyshev
int mgb_chebyshev_create(const mgb_operator *op, int degree, double smoothing_range, unsigned eig_its, mgb_chebyshev **out)
{
  return guarded([&] {
    auto h  = std::make_unique<mgb_chebyshev>();
    h->type = op->type;
    if (op->type == MGB_DOUBLE)
      {
        h->od = std::make_unique<ChebyshevT<double>>(*op->pd, degree, smoothing_range, eig_its);
        h->pd = h->od.get();
      }
    else
      {
        h->of = std::make_unique<ChebyshevT<float>>(*op->pf, degree, smoothing_range, eig_its);
        h->pf = h->of.get();
      }
    *out = h.release();
  });
}
void mgb_chebyshev_destroy(mgb_chebyshev *c) { delete c; }
int  mgb_chebyshev_vmult(mgb_chebyshev *c, mgb_vector *dst, const mgb_vector *rhs)
{
  return guarded([&] {
    require_same_type(c->type, dst->type);
    require_same_type(c->type, rhs->type);
    if (c->type == MGB_DOUBLE)
      c->pd->vmult(*dst->pd, *rhs->pd);
    else
      c->pf->vmult(*dst->pf, *rhs->pf);
  });
}
int mgb_chebyshev_step(mgb_chebyshev *c, mgb_vector *dst, const mgb_vector *rhs)
{
  return guarded([&] {
    require_same_type(c->type, dst->type);
    require_same_type(c->type, rhs->type);
    if (c->type == MGB_DOUBLE)
      c->pd->step(*dst->pd, *rhs->pd);
    else
      c->pf->step(*dst->pf, *rhs->pf);
  });
}
int mgb_chebyshev_eigenvalues(mgb_chebyshev *c, double *lmin, double *lmax)
{
  return guarded([&] {
    if (c->type == MGB_DOUBLE)
      {
        c->pd->estimate_eigenvalues();
        *lmin = c->pd->lambda_min_estimate();
        *lmax = c->pd->lambda_max();
      }
    else
      {
        c->pf->estimate_eigenvalues();
        *lmin = c->pf->lambda_min_estimate();
        *lmax = c->pf->lambda_max();
      }
  });
}

// ---- transfer
int mgb_transfer_create(const mgb_dofmap *fine, const mgb_dofmap *coarse, const mgb_transfer_map *map, const mgb_operator *opf,
                        const mgb_operator *opc, mgb_transfer **out)
{
  return guarded([&] {
    require_same_type(opf->type, opc->type);
    auto h  = std::make_unique<mgb_transfer>();
    h->type = opf->type;
    if (h->type == MGB_DOUBLE)
      {
        h->od = std::make_unique<TwoLevelTransferT<double>>(*fine->d, *coarse->d, map->t, *opf->pd, *opc->pd);
        h->pd = h->od.get();
      }
    else
      {
        h->of = std::make_unique<TwoLevelTransferT<float>>(*fine->d, *coarse->d, map->t, *opf->pf, *opc->pf);
        h->pf = h->of.get();
      }
    *out = h.release();
  });
}
int mgb_transfer_create_from_arrays(const mgb_transfer_map_view *v, const mgb_operator *opf, const mgb_operator *opc, mgb_transfer **out)
{
  return guarded([&] {
    require_same_type(opf->type, opc->type);
    if (!v || !v->fine_first || !v->n_children || !v->weight_code || !v->fine_own || v->n_coarse_cells < 0 || v->n_fine_slots < 0)
      throw std::runtime_error("mgb_transfer_create_from_arrays: fine_first, n_children, weight_code and fine_own are required");
    TransferMap tm;
    tm.degree_fine    = v->degree_fine;
    tm.degree_coarse  = v->degree_coarse;
    tm.n_coarse_cells = v->n_coarse_cells;
    tm.fine_first.assign(v->fine_first, v->fine_first + v->n_coarse_cells);
    tm.n_children.assign(v->n_children, v->n_children + v->n_coarse_cells);
    tm.n_fine_slots = v->n_fine_slots;
    tm.weight_code.assign(v->weight_code, v->weight_code + size_t(v->n_fine_slots) * 27);
    tm.fine_own.assign(v->fine_own, v->fine_own + size_t(v->n_fine_slots) * 27);
    for (int64_t c = 0; c < v->n_coarse_cells; ++c)
      {
        if (int64_t(tm.fine_first[c]) + tm.n_children[c] > v->n_fine_slots)
          throw std::runtime_error("mgb_transfer_create_from_arrays: fine_first + n_children exceeds n_fine_slots");
        if (tm.n_children[c] != 0 && tm.n_children[c] != 1 && tm.n_children[c] != 8)
          throw std::runtime_error("mgb_transfer_create_from_arrays: n_children must be 0, 1 or 8");
      }
    if (v->degree_fine < 1 || v->degree_fine > MAX_DEGREE || v->degree_coarse < 1 || v->degree_coarse > v->degree_fine)
      throw std::runtime_error("mgb_transfer_create_from_arrays: degrees out of range");
    {
      // fine_own addresses the fine operator's vector, or (distributed) the transfer's own ghosted fine vector
      const Partitioner &fp      = (opf->type == MGB_DOUBLE ? opf->pd->partitioner() : opf->pf->partitioner())->host;
      const uint64_t     n_local = v->distributed ? uint64_t(v->n_owned) + v->n_ghost : uint64_t(fp.n_owned) + fp.n_ghost;
      check_entity_table("mgb_transfer_create_from_arrays: fine_own", v->fine_own, v->n_fine_slots, v->degree_fine, n_local);
      if (v->distributed)
        {
          check_offsets("mgb_transfer_create_from_arrays: ghost_offsets", v->ghost_offsets, v->n_ghost_ranks, v->n_ghost);
          check_offsets("mgb_transfer_create_from_arrays: import_offsets", v->import_offsets, v->n_import_ranks, v->n_import_indices);
          check_index_list("mgb_transfer_create_from_arrays: import_indices", v->import_indices, v->n_import_indices, v->n_owned);
        }
    }
    tm.distributed = v->distributed != 0;
    if (tm.distributed)
      {
        Partitioner &pt = tm.part;
        pt.owned_begin  = v->owned_begin;
        pt.n_owned      = v->n_owned;
        pt.n_ghost      = v->n_ghost;
        pt.ghost_global.assign(v->ghost_global, v->ghost_global + v->n_ghost);
        pt.ghost_ranks.assign(v->ghost_ranks, v->ghost_ranks + v->n_ghost_ranks);
        pt.ghost_offsets.assign(v->ghost_offsets, v->ghost_offsets + v->n_ghost_ranks + 1);
        pt.import_ranks.assign(v->import_ranks, v->import_ranks + v->n_import_ranks);
        pt.import_offsets.assign(v->import_offsets, v->import_offsets + v->n_import_ranks + 1);
        pt.import_indices.assign(v->import_indices, v->import_indices + v->n_import_indices);
      }
    DoFMap fine, coarse; // only the degrees are read
    fine.degree   = v->degree_fine;
    coarse.degree = v->degree_coarse;
    auto h        = std::make_unique<mgb_transfer>();
    h->type       = opf->type;
    if (h->type == MGB_DOUBLE)
      {
        h->od = std::make_unique<TwoLevelTransferT<double>>(fine, coarse, tm, *opf->pd, *opc->pd);
        h->pd = h->od.get();
      }
    else
      {
        h->of = std::make_unique<TwoLevelTransferT<float>>(fine, coarse, tm, *opf->pf, *opc->pf);
        h->pf = h->of.get();
      }
    *out = h.release();
  });
}
void mgb_transfer_destroy(mgb_transfer *t) { delete t; }
int  mgb_transfer_prolongate_and_add(const mgb_transfer *t, mgb_vector *dst, const mgb_vector *src)
{
  return guarded([&] {
    require_same_type(t->type, dst->type);
    require_same_type(t->type, src->type);
    if (t->type == MGB_DOUBLE)
      t->pd->prolongate_and_add(*dst->pd, *src->pd);
    else
      t->pf->prolongate_and_add(*dst->pf, *src->pf);
  });
}
int mgb_transfer_restrict_and_add(const mgb_transfer *t, mgb_vector *dst, const mgb_vector *src)
{
  return guarded([&] {
    require_same_type(t->type, dst->type);
    require_same_type(t->type, src->type);
    if (t->type == MGB_DOUBLE)
      t->pd->restrict_and_add(*dst->pd, *src->pd);
    else
      t->pf->restrict_and_add(*dst->pf, *src->pf);
  });
}

// ---- problem
int mgb_problem_create(const char *json_text, int n_ranks, int rank, mgb_problem **out)
{
  return guarded([&] {
    ctx().ensure();
    auto h = std::make_unique<mgb_problem>();
    h->p.reset(create_problem(json_text, n_ranks, rank));
    const int type = h->p->level_number_type();
    for (int l = 0; l < h->p->n_levels(); ++l)
      {
        auto dm = std::make_unique<mgb_dofmap>();
        dm->d   = &h->p->level_dofmap(l);
        h->dofmaps[l] = std::move(dm);
        auto op  = std::make_unique<mgb_operator>();
        op->type = type;
        if (type == MGB_DOUBLE)
          op->pd = static_cast<OperatorT<double> *>(h->p->level_operator(l));
        else
          op->pf = static_cast<OperatorT<float> *>(h->p->level_operator(l));
        h->operators[l] = std::move(op);
        auto sm  = std::make_unique<mgb_chebyshev>();
        sm->type = type;
        if (type == MGB_DOUBLE)
          sm->pd = static_cast<ChebyshevT<double> *>(h->p->level_smoother(l));
        else
          sm->pf = static_cast<ChebyshevT<float> *>(h->p->level_smoother(l));
        h->smoothers[l] = std::move(sm);
        if (l > 0 && h->p->level_transfer(l) != nullptr)
          {
            auto tr  = std::make_unique<mgb_transfer>();
            tr->type = type;
            if (type == MGB_DOUBLE)
              tr->pd = static_cast<TwoLevelTransferT<double> *>(h->p->level_transfer(l));
            else
              tr->pf = static_cast<TwoLevelTransferT<float> *>(h->p->level_transfer(l));
            h->transfers[l] = std::move(tr);
          }
      }
    h->fine_dofmap    = std::make_unique<mgb_dofmap>();
    h->fine_dofmap->d = &h->p->fine_dofmap();
    h->fine_op       = std::make_unique<mgb_operator>();
    h->fine_op->type = MGB_DOUBLE;
    h->fine_op->pd   = &h->p->fine_operator();
    h->rhs           = std::make_unique<mgb_vector>();
    h->rhs->type     = MGB_DOUBLE;
    h->rhs->pd       = &h->p->rhs();
    h->solution      = std::make_unique<mgb_vector>();
    h->solution->type = MGB_DOUBLE;
    h->solution->pd  = &h->p->solution();
    *out             = h.release();
  });
}
void    mgb_problem_destroy(mgb_problem *p) { delete p; }
int     mgb_problem_n_levels(const mgb_problem *p) { return p->p->n_levels(); }
int64_t mgb_problem_level_n_dofs(const mgb_problem *p, int l) { return p->p->level_dofmap(l).part.n_global; }
int64_t mgb_problem_level_n_cells(const mgb_problem *p, int l) { return p->p->level_dofmap(l).n_cells_global; }
int     mgb_problem_level_degree(const mgb_problem *p, int l) { return p->p->level_dofmap(l).degree; }
const mgb_dofmap   *mgb_problem_level_dofmap(const mgb_problem *p, int l) { return p->dofmaps.at(l).get(); }
const mgb_operator *mgb_problem_level_operator(const mgb_problem *p, int l) { return p->operators.at(l).get(); }
mgb_chebyshev      *mgb_problem_level_smoother(mgb_problem *p, int l) { return p->smoothers.at(l).get(); }
const mgb_transfer *mgb_problem_level_transfer(const mgb_problem *p, int l)
{
  auto it = p->transfers.find(l);
  return it == p->transfers.end() ? nullptr : it->second.get();
}
const mgb_operator *mgb_problem_fine_operator(const mgb_problem *p) { return p->fine_op.get(); }
const mgb_dofmap   *mgb_problem_fine_dofmap(const mgb_problem *p) { return p->fine_dofmap.get(); }
int mgb_problem_n_meshes(const mgb_problem *p) { return (int)p->p->statistics_levels(); }
int mgb_problem_statistics(const mgb_problem *p, mgb_statistics *out, int64_t *per_level)
{
  return guarded([&] {
    export_statistics(p->p->statistics(), out, per_level);
    out->verbose = p->p->prm.verbose ? 1 : 0;
  });
}
mgb_vector         *mgb_problem_rhs(mgb_problem *p) { return p->rhs.get(); }
mgb_vector         *mgb_problem_solution(mgb_problem *p) { return p->solution.get(); }

int mgb_problem_vcycle(mgb_problem *p, mgb_vector *dst, const mgb_vector *src)
{
  return guarded([&] {
    if (dst->type != MGB_DOUBLE || src->type != MGB_DOUBLE)
      throw std::runtime_error("vcycle: outer vectors are double");
    p->p->vcycle(*dst->pd, *src->pd);
  });
}
int mgb_problem_solve_once(mgb_problem *p, int with_timers, mgb_result *result)
{
  return guarded([&] { fill_result(p->p->solve_once(with_timers != 0), result); });
}
int mgb_problem_solve_host(mgb_problem *p, const double *rhs_host, double *solution_host, mgb_result *result)
{
  return guarded([&] {
    DVec<double> &b = p->p->rhs();
    DVec<double> &x = p->p->solution();
    MGB_CUDA(cudaMemcpyAsync(b.data, rhs_host, size_t(b.n_owned) * sizeof(double), cudaMemcpyHostToDevice, ctx().stream));
    fill_result(p->p->solve_once(false), result);
    MGB_CUDA(cudaMemcpyAsync(solution_host, x.data, size_t(x.n_owned) * sizeof(double), cudaMemcpyDeviceToHost, ctx().stream));
    MGB_CUDA(cudaStreamSynchronize(ctx().stream));
  });
}
int mgb_problem_solve_host_batch(mgb_problem *p, int n, const double *const *rhs_hosts, double *const *solution_hosts, mgb_result *results)
{
  return guarded([&] {
    if (n <= 0)
      return;
    if (!rhs_hosts || !solution_hosts)
      throw std::runtime_error("mgb_problem_solve_host_batch: null pointer arrays");
    DVec<double> &b     = p->p->rhs();
    DVec<double> &x     = p->p->solution();
    const size_t  bytes = size_t(b.n_owned) * sizeof(double);
    cudaStream_t  main  = ctx().stream;
    if (!p->h2d_stream)
      {
        MGB_CUDA(cudaMalloc((void **)&p->stage_rhs, std::max<size_t>(bytes, 8)));
        MGB_CUDA(cudaMalloc((void **)&p->stage_sol, std::max<size_t>(bytes, 8)));
        MGB_CUDA(cudaStreamCreateWithFlags(&p->h2d_stream, cudaStreamNonBlocking));
        MGB_CUDA(cudaStreamCreateWithFlags(&p->d2h_stream, cudaStreamNonBlocking));
        for (cudaEvent_t *ev : {&p->ev_h2d, &p->ev_rhs_free, &p->ev_sol_ready, &p->ev_d2h_done})
          MGB_CUDA(cudaEventCreateWithFlags(ev, cudaEventDisableTiming));
      }
    // upload of right-hand side i+1 and download of solution i-1 run on their own streams (both PCIe directions at once)
    // behind solve i; the staging buffers decouple them from the vectors the solver works on
    MGB_CUDA(cudaMemcpyAsync(p->stage_rhs, rhs_hosts[0], bytes, cudaMemcpyHostToDevice, p->h2d_stream));
    MGB_CUDA(cudaEventRecord(p->ev_h2d, p->h2d_stream));
    for (int i = 0; i < n; ++i)
      {
        MGB_CUDA(cudaStreamWaitEvent(main, p->ev_h2d, 0));
        MGB_CUDA(cudaMemcpyAsync(b.data, p->stage_rhs, bytes, cudaMemcpyDeviceToDevice, main));
        MGB_CUDA(cudaEventRecord(p->ev_rhs_free, main));
        if (i + 1 < n)
          {
            MGB_CUDA(cudaStreamWaitEvent(p->h2d_stream, p->ev_rhs_free, 0));
            MGB_CUDA(cudaMemcpyAsync(p->stage_rhs, rhs_hosts[i + 1], bytes, cudaMemcpyHostToDevice, p->h2d_stream));
            MGB_CUDA(cudaEventRecord(p->ev_h2d, p->h2d_stream));
          }
        const Result r = p->p->solve_once(false);
        if (results)
          fill_result(r, results + i);
        if (i > 0)
          MGB_CUDA(cudaStreamWaitEvent(main, p->ev_d2h_done, 0)); // the staging buffer of the previous solution is free again
        MGB_CUDA(cudaMemcpyAsync(p->stage_sol, x.data, bytes, cudaMemcpyDeviceToDevice, main));
        MGB_CUDA(cudaEventRecord(p->ev_sol_ready, main));
        MGB_CUDA(cudaStreamWaitEvent(p->d2h_stream, p->ev_sol_ready, 0));
        MGB_CUDA(cudaMemcpyAsync(solution_hosts[i], p->stage_sol, bytes, cudaMemcpyDeviceToHost, p->d2h_stream));
        MGB_CUDA(cudaEventRecord(p->ev_d2h_done, p->d2h_stream));
      }
    MGB_CUDA(cudaStreamSynchronize(p->d2h_stream));
    MGB_CUDA(cudaStreamSynchronize(main));
  });
}
int mgb_problem_run(mgb_problem *p, int n_repetitions, mgb_result *result)
{
  return guarded([&] { fill_result(p->p->run((unsigned)n_repetitions), result); });
}

} // extern "C"
