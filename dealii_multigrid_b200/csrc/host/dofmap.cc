#include "dofmap.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <string>
#include <thread>
#include "partition.h"

#include <algorithm>
#include <stdexcept>

namespace mgb
{
  const int ENTITY_ORDER[27] = {
    // vertices
    0, 2, 6, 8, 18, 20, 24, 26,
    // lines
    1, 3, 5, 7, 9, 11, 15, 17, 19, 21, 23, 25,
    // quads
    4, 10, 12, 14, 16, 22,
    // hex
    13};

  void node_entity(int p, int i, int j, int k, int &e, int &offset)
  {
    const int tx = i == 0 ? 0 : (i == p ? 2 : 1);
    const int ty = j == 0 ? 0 : (j == p ? 2 : 1);
    const int tz = k == 0 ? 0 : (k == p ? 2 : 1);
    e            = tx + 3 * ty + 9 * tz;
    int stride   = 1;
    offset       = 0;
    if (tx == 1)
      {
        offset += (i - 1) * stride;
        stride *= (p - 1);
      }
    if (ty == 1)
      {
        offset += (j - 1) * stride;
        stride *= (p - 1);
      }
    if (tz == 1)
      offset += (k - 1) * stride;
  }

  namespace
  {
    inline int ipow(int b, int e)
    {
      int r = 1;
      for (int i = 0; i < e; ++i)
        r *= b;
      return r;
    }

    struct EntityKeys
    {
      int Lmax;
      // key of entity e of the cell (level,x,y,z): doubled-lattice centre coordinates + extension mask
      uint64_t operator()(int level, uint32_t x, uint32_t y, uint32_t z, int e) const
      {
        const uint64_t U  = uint64_t(1) << (Lmax - level);
        const int      tx = e % 3, ty = (e / 3) % 3, tz = e / 9;
        const uint64_t cx = (2 * uint64_t(x) + tx) * U, cy = (2 * uint64_t(y) + ty) * U,
                       cz  = (2 * uint64_t(z) + tz) * U;
        const uint64_t ext = uint64_t(tx == 1) | (uint64_t(ty == 1) << 1) | (uint64_t(tz == 1) << 2);
        return (ext << 60) | (cx << 40) | (cy << 20) | cz;
      }
      bool on_boundary(uint64_t key) const
      {
        const uint64_t top = uint64_t(2) << Lmax;
        const uint64_t ext = key >> 60;
        const uint64_t c[3] = {(key >> 40) & 0xfffff, (key >> 20) & 0xfffff, key & 0xfffff};
        for (int d = 0; d < 3; ++d)
          if (!((ext >> d) & 1) && (c[d] == 0 || c[d] == top))
            return true;
        return false;
      }
    };

    // hanging-node analysis of one leaf; fills eval keys (own keys with constrained faces/edges replaced
    // by the parent's entities) and the mask
    void analyse_cell(const Mesh &mesh, const EntityKeys &K, const Cell &c, uint64_t own[27], uint64_t eval[27],
                      uint16_t &mask)
    {
      for (int e = 0; e < 27; ++e)
        own[e] = eval[e] = K(c.level, c.x, c.y, c.z, e);
      mask = 0;
      if (c.level == 0)
        return;
      const uint32_t co[3] = {c.x, c.y, c.z};
      const int      s[3]  = {int(c.x & 1), int(c.y & 1), int(c.z & 1)};
      mask                 = uint16_t(s[0] | (s[1] << 1) | (s[2] << 2));
      const int64_t N      = int64_t(1) << c.level;
      // is the same-level region at offset (o0,o1,o2) covered by a leaf one level coarser?
      const auto coarser = [&](int o0, int o1, int o2) -> bool {
        const int64_t n[3] = {int64_t(co[0]) + o0, int64_t(co[1]) + o1, int64_t(co[2]) + o2};
        for (int d = 0; d < 3; ++d)
          if (n[d] < 0 || n[d] >= N)
            return false;
        return mesh.leaf(c.level - 1, uint32_t(n[0] >> 1), uint32_t(n[1] >> 1), uint32_t(n[2] >> 1)) >= 0;
      };
      bool face[3];
      for (int d = 0; d < 3; ++d)
        {
          int o[3] = {0, 0, 0};
          o[d]     = s[d] ? 1 : -1;
          face[d]  = coarser(o[0], o[1], o[2]);
          if (face[d])
            mask |= uint16_t(MASK_FACE_X << d);
        }
      bool edge[3];
      for (int d = 0; d < 3; ++d)
        {
          const int a = (d + 1) % 3, b = (d + 2) % 3;
          edge[d] = false;
          if (!face[a] && !face[b])
            {
              int o[3] = {0, 0, 0};
              o[a]     = s[a] ? 1 : -1;
              o[b]     = s[b] ? 1 : -1;
              edge[d]  = coarser(o[0], o[1], o[2]);
              if (edge[d])
                mask |= uint16_t(MASK_EDGE_X << d);
            }
        }
      if ((mask >> 3) == 0)
        return;
      const uint32_t px = c.x >> 1, py = c.y >> 1, pz = c.z >> 1;
      for (int e = 0; e < 27; ++e)
        {
          const int t[3]    = {e % 3, (e / 3) % 3, e / 9};
          bool      replace = false;
          for (int d = 0; d < 3; ++d)
            {
              if (face[d] && t[d] == 2 * s[d])
                replace = true;
              const int a = (d + 1) % 3, b = (d + 2) % 3;
              if (edge[d] && t[a] == 2 * s[a] && t[b] == 2 * s[b])
                replace = true;
            }
          if (replace)
            eval[e] = K(c.level - 1, px, py, pz, e);
        }
    }
  } // namespace

  std::vector<char> cells_with_hanging_nodes(const Mesh &mesh)
  {
    const EntityKeys  K{mesh.max_level()};
    std::vector<char> out(mesh.cells.size());
    uint64_t          own[27], eval[27];
    for (size_t i = 0; i < mesh.cells.size(); ++i)
      {
        uint16_t mask;
        analyse_cell(mesh, K, mesh.cells[i], own, eval, mask);
        out[i] = (mask >> 3) != 0;
      }
    return out;
  }

  // worker threads of the host setup: MGB_SETUP_THREADS, else the hardware threads divided among the ranks of this node
  // (torchrun starts one process per GPU and every one of them builds its maps at the same time), at most 16; small jobs
  // stay on the calling thread
  int setup_threads(int64_t n_items)
  {
    if (n_items < 4096)
      return 1;
    static const int n = [] {
      if (const char *e = std::getenv("MGB_SETUP_THREADS"))
        return std::max(1, std::atoi(e));
      int         world = 1;
      const char *w     = std::getenv("LOCAL_WORLD_SIZE");
      if (!w)
        w = std::getenv("WORLD_SIZE");
      if (w)
        world = std::max(1, std::atoi(w));
      const int hw = std::max(1u, std::thread::hardware_concurrency());
      return std::max(1, std::min(16, hw / world));
    }();
    return n;
  }
  void run_threads(int n, const std::function<void(int)> &fn)
  {
    if (n <= 1)
      {
        fn(0);
        return;
      }
    std::vector<std::thread> pool;
    for (int t = 1; t < n; ++t)
      pool.emplace_back(fn, t);
    fn(0);
    for (std::thread &th : pool)
      th.join();
  }

  DoFMap DoFMap::build(const Mesh &mesh, int degree, int n_ranks, int rank, int64_t min_cells_per_rank)
  {
    const int64_t nc = (int64_t)mesh.cells.size();
    return build(mesh, degree, n_ranks, rank,
                 equal_boundaries(nc, n_ranks, nc / std::max<int64_t>(1, min_cells_per_rank)));
  }

  DoFMap DoFMap::build(const Mesh &mesh, int degree, int n_ranks, int rank, const std::vector<int64_t> &cell_begin)
  {
    if (degree < 1)
      throw std::runtime_error("DoFMap: degree must be >= 1");
    if (n_ranks < 1 || rank < 0 || rank >= n_ranks)
      throw std::runtime_error("DoFMap: rank outside [0, n_ranks)");
    if ((int)cell_begin.size() != n_ranks + 1 || cell_begin.front() != 0 || cell_begin.back() != (int64_t)mesh.cells.size() ||
        !std::is_sorted(cell_begin.begin(), cell_begin.end()))
      throw std::runtime_error("DoFMap: cell_begin must hold n_ranks+1 ascending offsets from 0 to n_cells");
    DoFMap dm;
    dm.degree         = degree;
    dm.n_ranks        = n_ranks;
    dm.rank           = rank;
    dm.mesh_max_level = mesh.max_level();
    const int64_t nc  = (int64_t)mesh.cells.size();
    dm.n_cells_global = nc;
    const EntityKeys K{dm.mesh_max_level};
    int              esize[27];
    for (int e = 0; e < 27; ++e)
      esize[e] = ipow(degree - 1, entity_dim(e));

    dm.cell_begin = cell_begin[rank];
    dm.n_cells    = cell_begin[rank + 1] - cell_begin[rank];

    const auto T0 = std::chrono::steady_clock::now();
    // pass 1: global numbering, Morton first touch; the bases of the local cells' own entities are kept (g_own) so that
    // pass 2 does not look them up again
    static const int PREFETCH_AHEAD = [] {
      const char *e = std::getenv("MGB_SETUP_PREFETCH");
      return e ? std::max(1, std::atoi(e)) : 1;
    }();
    FlatMap ent;
    ent.reserve(size_t(nc) * (degree > 1 ? 9 : 2) + 64);
    // direct-mapped cache of the entities seen last (256 KB, stays in L2) in front of the big table: a hit costs a few
    // cycles where the table costs a memory access
    struct Recent
    {
      uint64_t key;
      int64_t  base;
    };
    constexpr int       RECENT_BITS = 14;
    long long           n_recent_hits = 0;
    std::vector<Recent> recent(size_t(1) << RECENT_BITS, Recent{~uint64_t(0), -1});
    std::vector<int64_t> dof_begin(n_ranks + 1, 0);
    std::vector<int64_t> g_eval(size_t(dm.n_cells) * 27), g_own(size_t(dm.n_cells) * 27, -1);
    int64_t              next = 0;
    {
      int           r       = 0;
      const int64_t l_begin = cell_begin[rank], l_end = cell_begin[rank + 1];
      for (int64_t i = 0; i < nc; ++i)
        {
          while (r < n_ranks && cell_begin[r] == i)
            dof_begin[r++] = next;
          const Cell &c     = mesh.cells[i];
          int64_t    *local = (i >= l_begin && i < l_end) ? &g_own[size_t(i - l_begin) * 27] : nullptr;
          if (i + PREFETCH_AHEAD < nc) // slots of a cell further on arrive while this one is numbered
            {
              const Cell &cn = mesh.cells[i + PREFETCH_AHEAD];
              for (int e = 0; e < 27; ++e)
                if (esize[e] != 0)
                  {
                    const uint64_t key = K(cn.level, cn.x, cn.y, cn.z, e);
                    if (recent[(key * 0x9e3779b97f4a7c15ULL) >> (64 - RECENT_BITS)].key != key)
                      ent.prefetch(key);
                  }
            }
          for (int k = 0; k < 27; ++k)
            {
              const int e = ENTITY_ORDER[k];
              if (esize[e] == 0)
                continue;
              const uint64_t key = K(c.level, c.x, c.y, c.z, e);
              Recent        &rc  = recent[(key * 0x9e3779b97f4a7c15ULL) >> (64 - RECENT_BITS)];
              int64_t        base;
              if (rc.key == key)
                ++n_recent_hits, base = rc.base; // numbered by one of the last few cells: most entities of a cell in Morton order
              else
                {
                  base = ent.find_or_insert(key, next);
                  if (base == next)
                    next += esize[e]; // first touch: the entity's DoFs are numbered now
                  rc = Recent{key, base};
                }
              if (local)
                local[e] = base;
            }
        }
      while (r <= n_ranks)
        dof_begin[r++] = next;
    }
    const auto T1 = std::chrono::steady_clock::now();
    if (next >= int64_t(INDEX_MASK))
      throw std::runtime_error("DoFMap: more than 2^31 DoFs need 64-bit local indices");
    const int64_t my_begin = dof_begin[rank], my_end = dof_begin[rank + 1];
    dm.rank_dof_begin   = dof_begin;
    dm.rank_cell_begin  = cell_begin;
    dm.part.n_global    = next;
    dm.part.owned_begin = my_begin;
    dm.part.n_owned     = uint32_t(my_end - my_begin);

    // pass 2: tables with global bases; ghosts; imports; constrained entities
    struct Ent
    {
      int64_t base;
      int     size;
      bool    operator<(const Ent &o) const { return base < o.base; }
      bool    operator==(const Ent &o) const { return base == o.base; }
    };
    std::vector<uint8_t>          f_eval(size_t(dm.n_cells) * 27), f_own(size_t(dm.n_cells) * 27);
    std::vector<Ent>              ghosts, constrained_ents;
    std::vector<std::vector<Ent>> imports(n_ranks);
    dm.mask.resize(dm.n_cells);
    dm.cell_level.resize(dm.n_cells);
    dm.cell_xyz.resize(size_t(dm.n_cells) * 3);

    const int64_t pass_begin = n_ranks == 1 ? cell_begin[rank] : 0;
    const int64_t pass_end   = n_ranks == 1 ? cell_begin[rank + 1] : nc;
    // closed bounding box of the local cells on the doubled lattice of the finest level: a cell of another rank can only
    // touch an entity we own (directly or, with hanging nodes, through its parent's entities) if its PARENT's closed box
    // meets this one
    uint64_t box_lo[3] = {~uint64_t(0), ~uint64_t(0), ~uint64_t(0)}, box_hi[3] = {0, 0, 0};
    for (int64_t i = cell_begin[rank]; i < cell_begin[rank + 1]; ++i)
      {
        const Cell    &c    = mesh.cells[i];
        const uint64_t U    = uint64_t(1) << (dm.mesh_max_level - c.level);
        const uint64_t x[3] = {c.x, c.y, c.z};
        for (int d = 0; d < 3; ++d)
          {
            box_lo[d] = std::min(box_lo[d], 2 * x[d] * U);
            box_hi[d] = std::max(box_hi[d], (2 * x[d] + 2) * U);
          }
      }
    const auto may_touch_local = [&](const Cell &c) {
      const int      shift = c.level > 0 ? 1 : 0; // parent's box (the cell itself on level 0)
      const uint64_t U     = uint64_t(1) << (dm.mesh_max_level - c.level + shift);
      const uint64_t x[3]  = {c.x >> shift, c.y >> shift, c.z >> shift};
      for (int d = 0; d < 3; ++d)
        if ((2 * x[d] + 2) * U < box_lo[d] || 2 * x[d] * U > box_hi[d])
          return false;
      return true;
    };
    // The cells are independent in this pass (the entity map is only read): contiguous slices of the cell range go to
    // worker threads; everything that is appended is sorted afterwards, so the maps do not depend on the thread count.
    const int n_threads = setup_threads(pass_end - pass_begin);
    struct Partial
    {
      std::vector<Ent>              ghosts, constrained_ents;
      std::vector<std::vector<Ent>> imports;
      int64_t                       n_hanging = 0;
      std::string                   error;
    };
    std::vector<Partial> partial(n_threads);
    const auto           pass2 = [&](const int t) {
      Partial &P = partial[t];
      P.imports.resize(n_ranks);
      const int64_t span = pass_end - pass_begin, i0 = pass_begin + span * t / n_threads, i1 = pass_begin + span * (t + 1) / n_threads;
      int           q = int(std::upper_bound(cell_begin.begin(), cell_begin.end(), i0) - cell_begin.begin()) - 1;
      q               = std::min(std::max(q, 0), n_ranks - 1);
      for (int64_t i = i0; i < i1; ++i)
        {
          while (cell_begin[q + 1] <= i)
            ++q;
          const Cell &c = mesh.cells[i];
          if (q != rank && (dm.n_cells == 0 || !may_touch_local(c)))
            continue;
          uint64_t own[27], eval[27];
          uint16_t mask;
          analyse_cell(mesh, K, c, own, eval, mask);
          if (q == rank)
            {
              const int64_t l = i - cell_begin[rank];
              dm.mask[l]      = mask;
              if (mask >> 3)
                ++P.n_hanging;
              dm.cell_level[l]       = c.level;
              dm.cell_xyz[3 * l + 0] = c.x;
              dm.cell_xyz[3 * l + 1] = c.y;
              dm.cell_xyz[3 * l + 2] = c.z;
              for (int e = 0; e < 27; ++e)
                {
                  if (esize[e] == 0)
                    {
                      g_eval[l * 27 + e] = g_own[l * 27 + e] = my_begin;
                      f_eval[l * 27 + e] = f_own[l * 27 + e] = 0;
                      continue;
                    }
                  const int64_t bo = g_own[l * 27 + e]; // from pass 1
                  const int64_t be = (eval[e] == own[e]) ? bo : ent.find(eval[e]);
                  if (bo < 0 || be < 0)
                    {
                      P.error = "DoFMap: entity lookup failed (mesh not 2:1 balanced?)";
                      return;
                    }
                  const bool dir_o = K.on_boundary(own[e]), dir_e = K.on_boundary(eval[e]);
                  const bool hang  = eval[e] != own[e];
                  g_own[l * 27 + e]  = bo;
                  g_eval[l * 27 + e] = be;
                  f_own[l * 27 + e]  = (dir_o || hang) ? 1 : 0;
                  f_eval[l * 27 + e] = dir_e ? 1 : 0;
                  if (bo < my_begin || bo >= my_end)
                    P.ghosts.push_back(Ent{bo, esize[e]});
                  else if (dir_o || hang)
                    P.constrained_ents.push_back(Ent{bo, esize[e]});
                  if (hang && (be < my_begin || be >= my_end))
                    P.ghosts.push_back(Ent{be, esize[e]});
                }
            }
          else
            {
              for (int e = 0; e < 27; ++e)
                {
                  if (esize[e] == 0)
                    continue;
                  const int64_t bo = ent.find(own[e]);
                  if (bo >= my_begin && bo < my_end)
                    P.imports[q].push_back(Ent{bo, esize[e]});
                  if (eval[e] != own[e])
                    {
                      const int64_t be = ent.find(eval[e]);
                      if (be >= my_begin && be < my_end)
                        P.imports[q].push_back(Ent{be, esize[e]});
                    }
                }
            }
        }
    };
    run_threads(n_threads, [&](const int t) {
      try
        {
          pass2(t);
        }
      catch (const std::exception &e) // an exception must not leave a worker thread
        {
          partial[t].error = e.what();
        }
    });
    for (Partial &P : partial)
      {
        if (!P.error.empty())
          throw std::runtime_error(P.error);
        dm.n_cells_hanging += P.n_hanging;
        ghosts.insert(ghosts.end(), P.ghosts.begin(), P.ghosts.end());
        constrained_ents.insert(constrained_ents.end(), P.constrained_ents.begin(), P.constrained_ents.end());
        for (int r = 0; r < n_ranks; ++r)
          imports[r].insert(imports[r].end(), P.imports[r].begin(), P.imports[r].end());
      }
    // a constrained entity owned by us may only be touched by cells of other ranks' ... no: the first
    // touching cell is ours by construction of ownership, so the loop above has seen it.

    const auto T2 = std::chrono::steady_clock::now();
    const auto uniq = [](std::vector<Ent> &v) {
      std::sort(v.begin(), v.end());
      v.erase(std::unique(v.begin(), v.end()), v.end());
    };
    uniq(ghosts);
    uniq(constrained_ents);

    // ghost numbering: ascending global index => grouped by owner
    std::vector<uint32_t> ghost_off(ghosts.size() + 1, 0);
    for (size_t g = 0; g < ghosts.size(); ++g)
      ghost_off[g + 1] = ghost_off[g] + uint32_t(ghosts[g].size);
    dm.part.n_ghost = ghost_off.back();
    dm.part.ghost_global.resize(dm.part.n_ghost);
    {
      int owner = -1;
      for (size_t g = 0; g < ghosts.size(); ++g)
        {
          for (int k = 0; k < ghosts[g].size; ++k)
            dm.part.ghost_global[ghost_off[g] + k] = ghosts[g].base + k;
          const int o =
            int(std::upper_bound(dof_begin.begin(), dof_begin.end(), ghosts[g].base) - dof_begin.begin()) - 1;
          if (o != owner)
            {
              dm.part.ghost_ranks.push_back(o);
              dm.part.ghost_offsets.push_back(ghost_off[g]);
              owner = o;
            }
        }
      dm.part.ghost_offsets.push_back(dm.part.n_ghost);
    }
    dm.part.import_offsets.push_back(0);
    for (int r = 0; r < n_ranks; ++r)
      {
        if (imports[r].empty())
          continue;
        uniq(imports[r]);
        dm.part.import_ranks.push_back(r);
        for (const Ent &en : imports[r])
          for (int k = 0; k < en.size; ++k)
            dm.part.import_indices.push_back(uint32_t(en.base - my_begin + k));
        dm.part.import_offsets.push_back(uint32_t(dm.part.import_indices.size()));
      }

    const auto to_local = [&](int64_t base) -> uint32_t {
      if (base >= my_begin && base < my_end)
        return uint32_t(base - my_begin);
      const auto it = std::lower_bound(ghosts.begin(), ghosts.end(), Ent{base, 0});
      return dm.part.n_owned + ghost_off[it - ghosts.begin()];
    };
    dm.eval_table.resize(g_eval.size());
    dm.own_table.resize(g_own.size());
    {
      const int nt = setup_threads(dm.n_cells);
      run_threads(nt, [&](const int t) {
        for (size_t k = g_eval.size() * t / nt; k < g_eval.size() * (t + 1) / nt; ++k)
          {
            dm.eval_table[k] = to_local(g_eval[k]) | (f_eval[k] ? FLAG_BIT : 0u);
            dm.own_table[k]  = to_local(g_own[k]) | (f_own[k] ? FLAG_BIT : 0u);
          }
      });
    }
    for (const Ent &en : constrained_ents)
      for (int k = 0; k < en.size; ++k)
        dm.constrained.push_back(uint32_t(en.base - my_begin + k));
    const auto T3 = std::chrono::steady_clock::now();
    if (std::getenv("MGB_SETUP_TIMING"))
      std::fprintf(stderr, "DoFMap::build %lld cells: %lld entities, %lld lookups answered by the recent-entity cache; pass1 %.3f s, pass2 %.3f s, tables %.3f s\n",
                   (long long)nc, (long long)ent.size(), n_recent_hits,
                   std::chrono::duration<double>(T1 - T0).count(), std::chrono::duration<double>(T2 - T1).count(),
                   std::chrono::duration<double>(T3 - T2).count());
    if (n_ranks > 1)
      dm.global_entities = std::make_shared<FlatMap>(std::move(ent));
    return dm;
  }

  void DoFMap::expand(const std::vector<uint32_t> &table, std::vector<uint32_t> &out) const
  {
    const int p = degree, n = p + 1, n3 = n * n * n;
    out.resize(size_t(n_cells) * n3);
    std::vector<int> ent(n3), off(n3);
    for (int k = 0; k < n; ++k)
      for (int j = 0; j < n; ++j)
        for (int i = 0; i < n; ++i)
          node_entity(p, i, j, k, ent[(k * n + j) * n + i], off[(k * n + j) * n + i]);
    for (int64_t c = 0; c < n_cells; ++c)
      for (int q = 0; q < n3; ++q)
        {
          const uint32_t t         = table[c * 27 + ent[q]];
          out[size_t(c) * n3 + q] = ((t & INDEX_MASK) + uint32_t(off[q])) | (t & FLAG_BIT);
        }
  }

  TransferMap TransferMap::build(const Mesh &mesh_fine, const DoFMap &df, const Mesh &mesh_coarse, const DoFMap &dc,
                                 const bool local_smoothing)
  {
    if (df.n_ranks != dc.n_ranks || df.rank != dc.rank)
      throw std::runtime_error("TransferMap: fine and coarse DoF maps belong to different partitions");
    if (df.n_ranks > 1)
      return build_distributed(mesh_fine, df, mesh_coarse, dc, local_smoothing);
    TransferMap tm;
    tm.degree_fine    = df.degree;
    tm.degree_coarse  = dc.degree;
    tm.n_coarse_cells = dc.n_cells;
    tm.fine_first.resize(dc.n_cells);
    tm.n_children.resize(dc.n_cells);
    int64_t f = 0;
    for (int64_t c = 0; c < dc.n_cells; ++c)
      {
        const Cell &cc        = mesh_coarse.cells[c];
        const Cell *fc        = f < df.n_cells ? &mesh_fine.cells[f] : nullptr;
        const bool  same      = fc && fc->level == cc.level && fc->x == cc.x && fc->y == cc.y && fc->z == cc.z;
        const bool  first_kid = fc && fc->level == cc.level + 1 && fc->x == 2 * cc.x && fc->y == 2 * cc.y && fc->z == 2 * cc.z;
        tm.fine_first[c]      = uint32_t(std::min<int64_t>(f, std::max<int64_t>(df.n_cells - 1, 0)));
        if (same && !local_smoothing)
          {
            tm.n_children[c] = 1;
            f += 1;
          }
        else if (first_kid)
          {
            for (unsigned ch = 0; ch < 8; ++ch)
              {
                if (f + ch >= df.n_cells)
                  throw std::runtime_error("TransferMap: fine mesh is not a one-level refinement of the coarse mesh");
                const Cell &s = mesh_fine.cells[f + ch];
                if (!(s.level == cc.level + 1 && s.x == 2 * cc.x + (ch & 1) && s.y == 2 * cc.y + ((ch >> 1) & 1) &&
                      s.z == 2 * cc.z + ((ch >> 2) & 1)))
                  throw std::runtime_error("TransferMap: fine mesh is not a one-level refinement of the coarse mesh");
              }
            tm.n_children[c] = 8;
            f += 8;
          }
        else if (local_smoothing)
          tm.n_children[c] = 0; // not refined: no cells on the fine level
        else
          throw std::runtime_error(fc ? "TransferMap: fine mesh is not a one-level refinement of the coarse mesh" :
                                        "TransferMap: fine mesh exhausted");
      }
    if (f != df.n_cells)
      throw std::runtime_error("TransferMap: fine cells left over");

    // valence of every fine entity = number of coarse cells whose fine patch contains it
    int esize[27];
    for (int e = 0; e < 27; ++e)
      esize[e] = ipow(df.degree - 1, entity_dim(e));
    std::vector<uint8_t> valence(size_t(df.part.n_owned) + df.part.n_ghost + 1, 0);
    {
      // coarse cells are independent; a fine entity on the boundary of two patches is counted by both, possibly from two
      // threads: relaxed atomic increments (a sum does not depend on the order)
      const int nt = setup_threads(dc.n_cells);
      run_threads(nt, [&](const int t) {
        std::vector<uint32_t> patch;
        for (int64_t c = dc.n_cells * t / nt; c < dc.n_cells * (t + 1) / nt; ++c)
          {
            patch.clear();
            for (unsigned ch = 0; ch < tm.n_children[c]; ++ch)
              for (int e = 0; e < 27; ++e)
                if (esize[e] > 0)
                  patch.push_back(df.own_table[(size_t(tm.fine_first[c]) + ch) * 27 + e] & INDEX_MASK);
            std::sort(patch.begin(), patch.end());
            patch.erase(std::unique(patch.begin(), patch.end()), patch.end());
            for (uint32_t b : patch)
              __atomic_fetch_add(&valence[b], uint8_t(1), __ATOMIC_RELAXED);
          }
      });
      tm.weight_code.assign(size_t(df.n_cells) * 27, 0);
      run_threads(nt, [&](const int t) {
        const size_t n = tm.weight_code.size();
        for (size_t k = n * t / nt; k < n * (t + 1) / nt; ++k)
          {
            const uint32_t tab = df.own_table[k];
            tm.weight_code[k]  = (esize[k % 27] == 0 || (tab & FLAG_BIT)) ? 0 : valence[tab & INDEX_MASK];
          }
      });
    }
    tm.fine_own     = df.own_table;
    tm.n_fine_slots = df.n_cells;
    tm.distributed  = false;
    return tm;
  }

  // Several ranks: the coarse level is partitioned independently of the fine level (equal Morton chunks of each mesh, the
  // reference's DefaultPolicy), so the children of a locally owned coarse cell may belong to other ranks.  As in
  // MGTwoLevelTransfer, the transfer gets its own fine-side numbering: the fine level's owned range plus the ghost DoFs
  // of those children; every rank derives the ghost/import lists of all ranks from the replicated global numbering.
  TransferMap TransferMap::build_distributed(const Mesh &mesh_fine, const DoFMap &df, const Mesh &mesh_coarse, const DoFMap &dc,
                                             const bool local_smoothing)
  {
    if (!df.global_entities)
      throw std::runtime_error("TransferMap: the fine DoF map does not carry the global numbering");
    const int     R = df.n_ranks, me = df.rank, p = df.degree;
    const int64_t ncc = (int64_t)mesh_coarse.cells.size(), nfc = (int64_t)mesh_fine.cells.size();
    const EntityKeys K{df.mesh_max_level};
    int              esize[27];
    for (int e = 0; e < 27; ++e)
      esize[e] = ipow(p - 1, entity_dim(e));
    const int64_t my_begin = df.rank_dof_begin[me], my_end = df.rank_dof_begin[me + 1];

    TransferMap tm;
    tm.degree_fine   = df.degree;
    tm.degree_coarse = dc.degree;
    tm.distributed   = true;
    // pass 0: coarse cell -> first fine cell / number of children, globally
    std::vector<int64_t> first(ncc);
    std::vector<uint8_t> nch(ncc);
    {
      int64_t f = 0;
      for (int64_t c = 0; c < ncc; ++c)
        {
          const Cell &cc = mesh_coarse.cells[c];
          const Cell *fc = f < nfc ? &mesh_fine.cells[f] : nullptr;
          first[c]       = std::min<int64_t>(f, std::max<int64_t>(nfc - 1, 0));
          if (!local_smoothing && fc && fc->level == cc.level && fc->x == cc.x && fc->y == cc.y && fc->z == cc.z)
            nch[c] = 1;
          else if (fc && fc->level == cc.level + 1 && (fc->x >> 1) == cc.x && (fc->y >> 1) == cc.y && (fc->z >> 1) == cc.z && f + 8 <= nfc)
            nch[c] = 8;
          else if (local_smoothing)
            nch[c] = 0;
          else
            throw std::runtime_error("TransferMap: fine mesh is not a one-level refinement of the coarse mesh");
          f += nch[c];
        }
      if (f != nfc)
        throw std::runtime_error("TransferMap: fine cells left over");
    }
    struct Ent
    {
      int64_t base;
      int     size;
      bool    operator<(const Ent &o) const { return base < o.base; }
      bool    operator==(const Ent &o) const { return base == o.base; }
    };
    // global own entries (base, constrained flag) of one fine cell
    const auto fine_entries = [&](const int64_t f, int64_t base[27], bool flag[27]) {
      const Cell &c = mesh_fine.cells[f];
      uint64_t    own[27], eval[27];
      uint16_t    mask;
      analyse_cell(mesh_fine, K, c, own, eval, mask);
      for (int e = 0; e < 27; ++e)
        {
          if (esize[e] == 0)
            {
              base[e] = -1;
              flag[e] = false;
              continue;
            }
          base[e] = df.global_entities->find(own[e]);
          if (base[e] < 0)
            throw std::runtime_error("TransferMap: entity lookup failed");
          flag[e] = K.on_boundary(own[e]) || eval[e] != own[e];
        }
    };
    // pass 1: slots of the locally owned coarse cells
    const int64_t cb = dc.rank_cell_begin[me], ce = dc.rank_cell_begin[me + 1];
    tm.n_coarse_cells = ce - cb;
    tm.fine_first.resize(ce - cb);
    tm.n_children.resize(ce - cb);
    std::vector<int64_t> g_own;
    std::vector<uint8_t> f_own;
    std::vector<Ent>     ghosts;
    FlatMap              needed; // global base -> index into valence
    std::vector<uint8_t> valence;
    for (int64_t c = cb; c < ce; ++c)
      {
        tm.fine_first[c - cb] = uint32_t(g_own.size() / 27);
        tm.n_children[c - cb] = nch[c];
        for (int ch = 0; ch < nch[c]; ++ch)
          {
            int64_t base[27];
            bool    flag[27];
            fine_entries(first[c] + ch, base, flag);
            for (int e = 0; e < 27; ++e)
              {
                g_own.push_back(base[e]);
                f_own.push_back(flag[e]);
                if (base[e] < 0)
                  continue;
                if (base[e] < my_begin || base[e] >= my_end)
                  ghosts.push_back(Ent{base[e], esize[e]});
                if (needed.find(uint64_t(base[e])) < 0)
                  {
                    needed.insert(uint64_t(base[e]), int64_t(valence.size()));
                    valence.push_back(0);
                  }
              }
          }
      }
    tm.n_fine_slots = int64_t(g_own.size() / 27);
    // pass 2: all coarse cells: valences of the needed entities and what the other ranks import from us
    std::vector<std::vector<Ent>> imports(R);
    {
      // A coarse cell contributes only if its fine patch touches an entity of our own patches (inside the closed bounding
      // box of our coarse cells) or an entity we own (inside the closed bounding box of our fine cells): skip the others.
      struct Box
      {
        uint64_t lo[3] = {~uint64_t(0), ~uint64_t(0), ~uint64_t(0)}, hi[3] = {0, 0, 0};
        bool     empty = true;
        void     add(const Cell &c, int Lmax)
        {
          const uint64_t U = uint64_t(1) << (Lmax - c.level), x[3] = {c.x, c.y, c.z};
          for (int d = 0; d < 3; ++d)
            {
              lo[d] = std::min(lo[d], 2 * x[d] * U);
              hi[d] = std::max(hi[d], (2 * x[d] + 2) * U);
            }
          empty = false;
        }
        bool meets(const Cell &c, int Lmax) const
        {
          if (empty)
            return false;
          const uint64_t U = uint64_t(1) << (Lmax - c.level), x[3] = {c.x, c.y, c.z};
          for (int d = 0; d < 3; ++d)
            if ((2 * x[d] + 2) * U < lo[d] || 2 * x[d] * U > hi[d])
              return false;
          return true;
        }
      } box_coarse, box_fine;
      for (int64_t c = cb; c < ce; ++c)
        box_coarse.add(mesh_coarse.cells[c], df.mesh_max_level);
      for (int64_t f = df.rank_cell_begin[me]; f < df.rank_cell_begin[me + 1]; ++f)
        box_fine.add(mesh_fine.cells[f], df.mesh_max_level);
      int              q = 0;
      std::vector<Ent> patch;
      for (int64_t c = 0; c < ncc; ++c)
        {
          while (dc.rank_cell_begin[q + 1] <= c)
            ++q;
          if (!box_coarse.meets(mesh_coarse.cells[c], df.mesh_max_level) && !box_fine.meets(mesh_coarse.cells[c], df.mesh_max_level))
            continue;
          patch.clear();
          for (int ch = 0; ch < nch[c]; ++ch)
            {
              int64_t base[27];
              bool    flag[27];
              fine_entries(first[c] + ch, base, flag);
              for (int e = 0; e < 27; ++e)
                if (base[e] >= 0)
                  patch.push_back(Ent{base[e], esize[e]});
            }
          std::sort(patch.begin(), patch.end());
          patch.erase(std::unique(patch.begin(), patch.end()), patch.end());
          for (const Ent &en : patch)
            {
              const int64_t v = needed.find(uint64_t(en.base));
              if (v >= 0)
                ++valence[v];
              if (q != me && en.base >= my_begin && en.base < my_end)
                imports[q].push_back(en);
            }
        }
    }
    const auto uniq = [](std::vector<Ent> &v) {
      std::sort(v.begin(), v.end());
      v.erase(std::unique(v.begin(), v.end()), v.end());
    };
    uniq(ghosts);
    Partitioner &P = tm.part;
    P.n_global     = df.part.n_global;
    P.owned_begin  = my_begin;
    P.n_owned      = uint32_t(my_end - my_begin);
    std::vector<uint32_t> ghost_off(ghosts.size() + 1, 0);
    for (size_t g = 0; g < ghosts.size(); ++g)
      ghost_off[g + 1] = ghost_off[g] + uint32_t(ghosts[g].size);
    P.n_ghost = ghost_off.back();
    P.ghost_global.resize(P.n_ghost);
    {
      int owner = -1;
      for (size_t g = 0; g < ghosts.size(); ++g)
        {
          for (int k = 0; k < ghosts[g].size; ++k)
            P.ghost_global[ghost_off[g] + k] = ghosts[g].base + k;
          const int o = int(std::upper_bound(df.rank_dof_begin.begin(), df.rank_dof_begin.end(), ghosts[g].base) - df.rank_dof_begin.begin()) - 1;
          if (o != owner)
            {
              P.ghost_ranks.push_back(o);
              P.ghost_offsets.push_back(ghost_off[g]);
              owner = o;
            }
        }
      P.ghost_offsets.push_back(P.n_ghost);
    }
    P.import_offsets.push_back(0);
    for (int r = 0; r < R; ++r)
      {
        if (imports[r].empty())
          continue;
        uniq(imports[r]);
        P.import_ranks.push_back(r);
        for (const Ent &en : imports[r])
          for (int k = 0; k < en.size; ++k)
            P.import_indices.push_back(uint32_t(en.base - my_begin + k));
        P.import_offsets.push_back(uint32_t(P.import_indices.size()));
      }
    const auto to_local = [&](int64_t base) -> uint32_t {
      if (base >= my_begin && base < my_end)
        return uint32_t(base - my_begin);
      const auto it = std::lower_bound(ghosts.begin(), ghosts.end(), Ent{base, 0});
      return P.n_owned + ghost_off[it - ghosts.begin()];
    };
    tm.fine_own.resize(g_own.size());
    tm.weight_code.assign(g_own.size(), 0);
    for (size_t k = 0; k < g_own.size(); ++k)
      {
        if (g_own[k] < 0)
          {
            tm.fine_own[k] = 0;
            continue;
          }
        tm.fine_own[k] = to_local(g_own[k]) | (f_own[k] ? FLAG_BIT : 0u);
        if (!f_own[k])
          tm.weight_code[k] = valence[needed.find(uint64_t(g_own[k]))];
      }
    return tm;
  }

  // ------------------------------------------------------------------------------------------ local smoothing
  namespace
  {
    // keys of the level entities lying on a face between a level cell and the region without level cells (covered by
    // a coarser active cell): MGTools::extract_inner_interface_dofs marks all DoFs of such faces
    void refinement_edge_entities(const Mesh &lm, const EntityKeys &K, FlatMap &edge)
    {
      for (const Cell &c : lm.cells)
        {
          const int64_t  N     = int64_t(1) << c.level;
          const uint32_t co[3] = {c.x, c.y, c.z};
          for (int d = 0; d < 3; ++d)
            for (int s = 0; s < 2; ++s)
              {
                int64_t n[3] = {co[0], co[1], co[2]};
                n[d] += s ? 1 : -1;
                if (n[d] < 0 || n[d] >= N)
                  continue; // domain boundary: Dirichlet, not a refinement edge
                if (lm.leaf(c.level, uint32_t(n[0]), uint32_t(n[1]), uint32_t(n[2])) >= 0)
                  continue;
                for (int e = 0; e < 27; ++e)
                  {
                    const int t[3] = {e % 3, (e / 3) % 3, e / 9};
                    if (t[d] != 2 * s)
                      continue;
                    const uint64_t key = K(c.level, c.x, c.y, c.z, e);
                    if (edge.find(key) < 0)
                      edge.insert(key, 1);
                  }
              }
        }
    }
  } // namespace

  LevelEdges LevelEdges::build(const Mesh &lm, const DoFMap &dm)
  {
    LevelEdges       le;
    const EntityKeys K{dm.mesh_max_level};
    int              esize[27];
    for (int e = 0; e < 27; ++e)
      esize[e] = ipow(dm.degree - 1, entity_dim(e));
    FlatMap edge;
    refinement_edge_entities(lm, K, edge);
    le.any_global = edge.size() > 0;
    if (!le.any_global)
      return le;
    for (int64_t c = 0; c < dm.n_cells; ++c)
      {
        bool touches = false;
        for (int e = 0; e < 27; ++e)
          {
            if (esize[e] == 0)
              continue;
            const uint64_t key = K(dm.cell_level[c], dm.cell_xyz[3 * c], dm.cell_xyz[3 * c + 1], dm.cell_xyz[3 * c + 2], e);
            if (edge.find(key) < 0)
              continue;
            touches            = true;
            const uint32_t base = dm.own_table[c * 27 + e] & INDEX_MASK;
            if (base < dm.part.n_owned)
              for (int k = 0; k < esize[e]; ++k)
                le.edge_dofs.push_back(base + uint32_t(k));
          }
        if (touches)
          le.edge_cells.push_back(uint32_t(c));
      }
    std::sort(le.edge_dofs.begin(), le.edge_dofs.end());
    le.edge_dofs.erase(std::unique(le.edge_dofs.begin(), le.edge_dofs.end()), le.edge_dofs.end());
    return le;
  }

  CopyIndices CopyIndices::build(const DoFMap &ad, const Mesh &lm, const DoFMap &ld, const Mesh *active_mesh)
  {
    if (ad.n_ranks > 1 || ld.n_ranks > 1)
      {
        if (!active_mesh)
          throw std::runtime_error("CopyIndices: the active mesh is needed on several ranks");
        return build_distributed(ad, *active_mesh, lm, ld);
      }
    if (ad.degree != ld.degree)
      throw std::runtime_error("CopyIndices: degrees differ");
    CopyIndices      ci;
    const int        level = ld.mesh_max_level;
    const EntityKeys K{level};
    int              esize[27];
    for (int e = 0; e < 27; ++e)
      esize[e] = ipow(ad.degree - 1, entity_dim(e));
    FlatMap edge;
    refinement_edge_entities(lm, K, edge);
    struct Seg
    {
      uint32_t a, b;
      int      size;
      bool     operator<(const Seg &o) const { return a < o.a; }
      bool     operator==(const Seg &o) const { return a == o.a; }
    };
    std::vector<Seg> segs;
    for (int64_t c = 0; c < ad.n_cells; ++c)
      {
        if (ad.cell_level[c] != level)
          continue;
        const uint32_t x = ad.cell_xyz[3 * c], y = ad.cell_xyz[3 * c + 1], z = ad.cell_xyz[3 * c + 2];
        const int64_t  j = lm.leaf(level, x, y, z);
        if (j < ld.cell_begin || j >= ld.cell_begin + ld.n_cells)
          throw std::runtime_error("CopyIndices: active cell not found on its level");
        const int64_t lc = j - ld.cell_begin;
        for (int e = 0; e < 27; ++e)
          {
            if (esize[e] == 0 || edge.find(K(level, x, y, z, e)) >= 0)
              continue;
            segs.push_back(Seg{ad.own_table[c * 27 + e] & INDEX_MASK, ld.own_table[lc * 27 + e] & INDEX_MASK, esize[e]});
          }
      }
    std::sort(segs.begin(), segs.end());
    segs.erase(std::unique(segs.begin(), segs.end()), segs.end());
    for (const Seg &sg : segs)
      for (int k = 0; k < sg.size; ++k)
        {
          ci.active_idx.push_back(sg.a + uint32_t(k));
          ci.level_idx.push_back(sg.b + uint32_t(k));
        }
    return ci;
  }

  CopyIndices CopyIndices::build_distributed(const DoFMap &ad, const Mesh &am, const Mesh &lm, const DoFMap &ld)
  {
    if (!ad.global_entities || !ld.global_entities)
      throw std::runtime_error("CopyIndices: the DoF maps do not carry the global numbering");
    if (ad.degree != ld.degree || ad.n_ranks != ld.n_ranks)
      throw std::runtime_error("CopyIndices: incompatible DoF maps");
    CopyIndices ci;
    ci.distributed       = true;
    const int        R = ad.n_ranks, me = ad.rank, level = ld.mesh_max_level;
    const EntityKeys KA{ad.mesh_max_level}, KL{level};
    int              esize[27];
    for (int e = 0; e < 27; ++e)
      esize[e] = ipow(ad.degree - 1, entity_dim(e));
    FlatMap edge;
    refinement_edge_entities(lm, KL, edge);
    const int64_t a_begin = ad.rank_dof_begin[me], a_end = ad.rank_dof_begin[me + 1];
    const int64_t l_begin = ld.rank_dof_begin[me], l_end = ld.rank_dof_begin[me + 1];
    const auto    owner   = [](const std::vector<int64_t> &begin, int64_t g) {
      return int(std::upper_bound(begin.begin(), begin.end(), g) - begin.begin()) - 1;
    };
    struct Seg
    {
      int64_t a, l;
      int     size;
      bool    operator<(const Seg &o) const { return a < o.a; }
      bool    operator==(const Seg &o) const { return a == o.a; }
    };
    struct Ent
    {
      int64_t base;
      int     size;
      bool    operator<(const Ent &o) const { return base < o.base; }
      bool    operator==(const Ent &o) const { return base == o.base; }
    };
    std::vector<Seg>              mine;    // pairs whose active DoF this rank owns
    std::vector<Ent>              ghosts;  // their level DoFs owned elsewhere
    std::vector<std::vector<Ent>> imports(R); // level DoFs of mine that the pairs of rank r refer to
    FlatMap                       seen;    // active entity base -> 1 (each active entity once)
    for (const Cell &c : am.cells)
      {
        if (c.level != level)
          continue;
        for (int e = 0; e < 27; ++e)
          {
            if (esize[e] == 0 || edge.find(KL(level, c.x, c.y, c.z, e)) >= 0)
              continue;
            const int64_t ga = ad.global_entities->find(KA(c.level, c.x, c.y, c.z, e));
            const int64_t gl = ld.global_entities->find(KL(level, c.x, c.y, c.z, e));
            if (ga < 0 || gl < 0)
              throw std::runtime_error("CopyIndices: entity lookup failed");
            if (seen.find(uint64_t(ga)) >= 0)
              continue;
            seen.insert(uint64_t(ga), 1);
            const int ra = owner(ad.rank_dof_begin, ga), rl = owner(ld.rank_dof_begin, gl);
            if (ra == me)
              {
                mine.push_back(Seg{ga, gl, esize[e]});
                if (rl != me)
                  ghosts.push_back(Ent{gl, esize[e]});
              }
            else if (rl == me)
              imports[ra].push_back(Ent{gl, esize[e]});
          }
      }
    const auto uniq = [](std::vector<Ent> &v) {
      std::sort(v.begin(), v.end());
      v.erase(std::unique(v.begin(), v.end()), v.end());
    };
    uniq(ghosts);
    Partitioner &P = ci.part;
    P.n_global     = ld.part.n_global;
    P.owned_begin  = l_begin;
    P.n_owned      = uint32_t(l_end - l_begin);
    std::vector<uint32_t> ghost_off(ghosts.size() + 1, 0);
    for (size_t g = 0; g < ghosts.size(); ++g)
      ghost_off[g + 1] = ghost_off[g] + uint32_t(ghosts[g].size);
    P.n_ghost = ghost_off.back();
    P.ghost_global.resize(P.n_ghost);
    {
      int own = -1;
      for (size_t g = 0; g < ghosts.size(); ++g)
        {
          for (int k = 0; k < ghosts[g].size; ++k)
            P.ghost_global[ghost_off[g] + k] = ghosts[g].base + k;
          const int o = owner(ld.rank_dof_begin, ghosts[g].base);
          if (o != own)
            {
              P.ghost_ranks.push_back(o);
              P.ghost_offsets.push_back(ghost_off[g]);
              own = o;
            }
        }
      P.ghost_offsets.push_back(P.n_ghost);
    }
    P.import_offsets.push_back(0);
    for (int r = 0; r < R; ++r)
      {
        if (imports[r].empty())
          continue;
        uniq(imports[r]);
        P.import_ranks.push_back(r);
        for (const Ent &en : imports[r])
          for (int k = 0; k < en.size; ++k)
            P.import_indices.push_back(uint32_t(en.base - l_begin + k));
        P.import_offsets.push_back(uint32_t(P.import_indices.size()));
      }
    std::sort(mine.begin(), mine.end());
    for (const Seg &sg : mine)
      {
        uint32_t lbase;
        if (sg.l >= l_begin && sg.l < l_end)
          lbase = uint32_t(sg.l - l_begin);
        else
          {
            const auto it = std::lower_bound(ghosts.begin(), ghosts.end(), Ent{sg.l, 0});
            lbase         = P.n_owned + ghost_off[it - ghosts.begin()];
          }
        for (int k = 0; k < sg.size; ++k)
          {
            ci.active_idx.push_back(uint32_t(sg.a - a_begin + k));
            ci.level_idx.push_back(lbase + uint32_t(k));
          }
      }
    (void)a_end;
    return ci;
  }
} // namespace mgb
