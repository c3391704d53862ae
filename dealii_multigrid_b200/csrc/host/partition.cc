#include "partition.h"

#include "dofmap.h"

#include <algorithm>
#include <cmath>
#include <stdexcept>

namespace mgb
{
  std::vector<int64_t> equal_boundaries(int64_t n_cells, int n_ranks, int64_t n_active)
  {
    n_active = std::max<int64_t>(1, std::min<int64_t>(n_active, n_ranks));
    std::vector<int64_t> b(n_ranks + 1);
    for (int r = 0; r <= n_ranks; ++r)
      b[r] = (n_cells * std::min<int64_t>(r, n_active)) / n_active;
    return b;
  }

  std::vector<int64_t> weighted_boundaries(const std::vector<int64_t> &w, int n_ranks, int64_t n_active)
  {
    n_active         = std::max<int64_t>(1, std::min<int64_t>(n_active, n_ranks));
    const int64_t nc = (int64_t)w.size();
    __int128      total = 0;
    for (int64_t v : w)
      total += v;
    std::vector<int64_t> b(n_ranks + 1, nc);
    b[0] = 0;
    // rank of cell i = floor(W_before(i) * n_active / total): boundaries are where that rank increases
    __int128 before = 0;
    int64_t  r      = 0;
    for (int64_t i = 0; i < nc; ++i)
      {
        const int64_t q = total > 0 ? (int64_t)((before * n_active) / total) : 0;
        while (r < q)
          b[++r] = i;
        before += w[i];
      }
    while (r < n_ranks)
      b[++r] = nc;
    return b;
  }

  std::vector<int64_t> first_child_boundaries(const Mesh &parents, const Mesh &children, const std::vector<int64_t> &child_begin)
  {
    const int     R  = (int)child_begin.size() - 1;
    const int64_t np = (int64_t)parents.cells.size();
    std::vector<int64_t> b(R + 1, np);
    b[0]          = 0;
    int     r     = 0;
    int64_t last  = -1;
    for (int64_t i = 0; i < np; ++i)
      {
        const Cell &c = parents.cells[i];
        int         level = c.level;
        uint32_t    x = c.x, y = c.y, z = c.z;
        int64_t     leaf = -1;
        for (;;)
          {
            int rel = 0;
            leaf    = children.locate(level, x, y, z, rel);
            if (rel <= 0)
              break;
            if (++level > 30) // subdivided further: descend into the first child
              throw std::runtime_error("first_child_boundaries: the child mesh does not cover the parent mesh");
            x *= 2, y *= 2, z *= 2;
          }
        if (leaf < 0 || leaf < last)
          throw std::runtime_error("first_child_boundaries: the child mesh does not cover the parent mesh in Morton order");
        last = leaf;
        while (r < R && child_begin[r + 1] <= leaf) // first parent whose first leaf belongs to a later rank
          b[++r] = i;
      }
    while (r < R)
      b[++r] = np;
    return b;
  }

  PartitionPolicy PartitionPolicy::parse(const std::string &name)
  {
    PartitionPolicy p;
    const auto      dash   = name.find('-');
    const std::string head = name.substr(0, dash);
    const std::string arg  = dash == std::string::npos ? "" : name.substr(dash + 1);
    if (name.empty() || head == "DefaultPolicy")
      p.kind = Default;
    else if (head == "MinimalGranularityPolicy")
      {
        p.kind        = MinimalGranularity;
        p.n_min_cells = std::max<int64_t>(1, arg.empty() ? 1 : std::atoll(arg.c_str()));
      }
    else if (head == "BalancedGranularityPartitionPolicy")
      p.kind = BalancedGranularity;
    else if (head == "CellWeightPolicy")
      {
        p.kind           = CellWeight;
        p.hanging_weight = arg.empty() ? 1.0 : std::atof(arg.c_str());
      }
    else if (head == "FirstChildPolicy")
      {
        p.kind           = FirstChild;
        p.hanging_weight = arg.empty() ? 1.0 : std::atof(arg.c_str());
      }
    else
      throw std::runtime_error("PartitionerName '" + name + "' is not implemented");
    if (!(p.hanging_weight > 0.0))
      throw std::runtime_error("PartitionerName '" + name + "': the weight must be positive");
    return p;
  }

  namespace
  {
    int64_t ranks_for(int64_t n_cells, int n_ranks, int64_t min_cells)
    {
      return std::max<int64_t>(1, std::min<int64_t>(n_ranks, n_cells / std::max<int64_t>(1, min_cells)));
    }

    // reference weights: 10000 per cell, 10000*W on cells with hanging-node constraints (:284-290)
    std::vector<int64_t> hanging_weights(const Mesh &mesh, double w)
    {
      const std::vector<char> hn = cells_with_hanging_nodes(mesh);
      std::vector<int64_t>    out(hn.size());
      const int64_t           heavy = (int64_t)std::llround(10000.0 * w);
      for (size_t i = 0; i < hn.size(); ++i)
        out[i] = hn[i] ? heavy : 10000;
      return out;
    }
  } // namespace

  std::vector<int64_t> PartitionPolicy::fine(const Mesh &mesh, int n_ranks, int64_t floor_min_cells) const
  {
    const int64_t nc = (int64_t)mesh.cells.size();
    const int64_t n_active = ranks_for(nc, n_ranks, kind == MinimalGranularity ? std::max(n_min_cells, floor_min_cells) : floor_min_cells);
    if ((kind == CellWeight || kind == FirstChild) && hanging_weight != 1.0 && n_ranks > 1)
      return weighted_boundaries(hanging_weights(mesh, hanging_weight), n_ranks, n_active);
    return equal_boundaries(nc, n_ranks, n_active);
  }

  std::vector<std::vector<int64_t>> PartitionPolicy::sequence(const std::vector<Mesh> &meshes, int n_ranks, int64_t floor_min_cells) const
  {
    const size_t                      n = meshes.size();
    std::vector<std::vector<int64_t>> out(n);
    if (n == 0)
      return out;
    out[n - 1] = fine(meshes[n - 1], n_ranks, floor_min_cells);
    int64_t procs_prev = n_ranks; // BalancedGranularityPartitionPolicy: n_mpi_processes_per_level.back()
    for (size_t k = n - 1; k-- > 0;)
      {
        const Mesh   &m  = meshes[k];
        const int64_t nc = (int64_t)m.cells.size();
        switch (kind)
          {
            case FirstChild:
              out[k] = first_child_boundaries(m, meshes[k + 1], out[k + 1]);
              break;
            case BalancedGranularity:
              {
                const int64_t grain = std::min<int64_t>(200, 8 * nc / procs_prev + 1); // :397-398
                procs_prev          = ranks_for(nc, n_ranks, grain);
                out[k]              = equal_boundaries(nc, n_ranks, procs_prev);
                break;
              }
            case CellWeight: // the weight signal stays connected while the mesh is coarsened (:2147-2150)
            case MinimalGranularity:
            case Default:
              out[k] = fine(m, n_ranks, floor_min_cells);
              break;
          }
      }
    return out;
  }

  namespace
  {
    inline int owner_of(const std::vector<int64_t> &begin, int64_t cell)
    {
      return int(std::upper_bound(begin.begin() + 1, begin.end(), cell) - (begin.begin() + 1));
    }

    // leaves of `mesh` whose closure touches the closure of cell c (face, edge or corner neighbours, hanging or not)
    template <typename F>
    void for_each_touching_cell(const Mesh &mesh, const Cell &c, F &&f)
    {
      const int64_t N = int64_t(1) << c.level;
      for (int dz = -1; dz <= 1; ++dz)
        for (int dy = -1; dy <= 1; ++dy)
          for (int dx = -1; dx <= 1; ++dx)
            {
              if (dx == 0 && dy == 0 && dz == 0)
                continue;
              const int64_t n[3] = {int64_t(c.x) + dx, int64_t(c.y) + dy, int64_t(c.z) + dz};
              if (n[0] < 0 || n[1] < 0 || n[2] < 0 || n[0] >= N || n[1] >= N || n[2] >= N)
                continue;
              int           rel = 0;
              const int64_t j   = mesh.locate(c.level, uint32_t(n[0]), uint32_t(n[1]), uint32_t(n[2]), rel);
              if (j >= 0)
                {
                  f(j); // same-level or coarser neighbour
                  continue;
                }
              // finer neighbours (one level, 2:1 balance): the children of the region on the side facing c
              const int d[3] = {dx, dy, dz};
              for (int ch = 0; ch < 8; ++ch)
                {
                  bool faces = true;
                  for (int a = 0; a < 3; ++a)
                    {
                      const int bit = (ch >> a) & 1;
                      if ((d[a] == 1 && bit != 0) || (d[a] == -1 && bit != 1))
                        faces = false;
                    }
                  if (!faces)
                    continue;
                  const int64_t k = mesh.leaf(c.level + 1, uint32_t(2 * n[0] + (ch & 1)), uint32_t(2 * n[1] + ((ch >> 1) & 1)),
                                              uint32_t(2 * n[2] + ((ch >> 2) & 1)));
                  if (k >= 0)
                    f(k);
                }
            }
    }
  } // namespace

  HierarchyStatistics hierarchy_statistics(const std::vector<const Mesh *> &meshes, const std::vector<std::vector<int64_t>> &cell_begin,
                                           int n_ranks)
  {
    HierarchyStatistics st;
    const size_t        nl = meshes.size();
    st.levels.resize(nl);
    if (cell_begin.size() != nl)
      throw std::runtime_error("hierarchy_statistics: one partition per mesh required");
    // workload (mg_tools.h:9-36, 277-300)
    int64_t sum_all = 0, sum_max = 0;
    for (size_t l = 0; l < nl; ++l)
      {
        auto &L       = st.levels[l];
        L.n_cells_min = INT64_MAX;
        for (int r = 0; r < n_ranks; ++r)
          {
            const int64_t n = cell_begin[l][r + 1] - cell_begin[l][r];
            L.n_cells_min   = std::min(L.n_cells_min, n);
            L.n_cells_max   = std::max(L.n_cells_max, n);
            L.n_cells_sum += n;
          }
        sum_all += L.n_cells_sum;
        sum_max += L.n_cells_max;
        st.mem_total += double(meshes[l]->cells.size()) * (sizeof(Cell) + 16.0); // leaves + their hash index entries
      }
    st.workload_path_max = double(sum_max);
    st.workload_eff      = sum_max > 0 ? (double(sum_all) / n_ranks) / double(sum_max) : 1.0;
    // vertical communication: for every cell of level l with children on level l+1, is the child on the same rank
    int64_t v_local = 0, v_remote = 0;
    for (size_t l = 0; l + 1 < nl; ++l)
      {
        const Mesh &coarse = *meshes[l], &fine = *meshes[l + 1];
        auto       &L      = st.levels[l + 1];
        const int   fine_max = fine.max_level();
        for (int64_t i = 0; i < (int64_t)coarse.cells.size(); ++i)
          {
            const Cell &c = coarse.cells[i];
            if (c.level == fine_max)
              continue;
            const int r = owner_of(cell_begin[l], i);
            for (int ch = 0; ch < 8; ++ch)
              {
                const int64_t k = fine.leaf(c.level + 1, 2 * c.x + (ch & 1), 2 * c.y + ((ch >> 1) & 1), 2 * c.z + ((ch >> 2) & 1));
                if (k < 0)
                  continue;
                (owner_of(cell_begin[l + 1], k) == r ? L.vertical_local : L.vertical_remote) += 1;
              }
          }
        v_local += L.vertical_local;
        v_remote += L.vertical_remote;
      }
    st.vertical_eff = (v_local + v_remote) > 0 ? double(v_local) / double(v_local + v_remote) : 1.0;
    // horizontal communication: owned cells and ghost cells (cells of other ranks touching an owned cell)
    int64_t h_local = 0, h_remote = 0;
    for (size_t l = 0; l < nl; ++l)
      {
        const Mesh &m = *meshes[l];
        auto       &L = st.levels[l];
        L.horizontal_local = (int64_t)m.cells.size();
        if (n_ranks > 1)
          for (int64_t j = 0; j < (int64_t)m.cells.size(); ++j)
            {
              // cell j is a ghost of every other rank that owns one of its neighbours
              const int own = owner_of(cell_begin[l], j);
              int       seen[32];
              int       n_seen = 0;
              for_each_touching_cell(m, m.cells[j], [&](int64_t k) {
                const int q = owner_of(cell_begin[l], k);
                if (q == own)
                  return;
                for (int s = 0; s < n_seen; ++s)
                  if (seen[s] == q)
                    return;
                if (n_seen < 32)
                  seen[n_seen++] = q;
              });
              L.horizontal_remote += n_seen;
            }
        h_local += L.horizontal_local;
        h_remote += L.horizontal_remote;
      }
    st.horizontal_eff = (h_local + h_remote) > 0 ? double(h_local + h_remote / 2) / double(h_local + h_remote) : 1.0;
    return st;
  }
} // namespace mgb
