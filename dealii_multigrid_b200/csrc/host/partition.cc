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
            ++level; // subdivided further: descend into the first child
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
} // namespace mgb
