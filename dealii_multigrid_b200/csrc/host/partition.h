// Partition policies of the mesh hierarchy -- the stand-in for p4est's (weighted) Morton split of the fine mesh and for
// the RepartitioningPolicyTools classes the reference selects by `PartitionerName` (multigrid_throughput.cc:2066-2174):
//   DefaultPolicy                       every mesh of the sequence split into equal Morton chunks
//   MinimalGranularityPolicy-N          as above, on as few ranks as needed to give each at least N cells
//   BalancedGranularityPartitionPolicy  the reference's own policy (:372-419): grain 200 cells, growing by at most 8x
//                                       per coarsening step
//   CellWeightPolicy-W                  cells with hanging-node constraints weigh W (hanging_nodes_weighting, :276-292)
//   FirstChildPolicy[-W]                a coarse cell lives with its first child (no vertical communication); optional
//                                       weighted split of the fine mesh
// Every policy yields contiguous Morton ranges, so a partition is a vector `cell_begin` of n_ranks+1 offsets.
#pragma once
#include "octree.h"

#include <cstdint>
#include <string>
#include <vector>

namespace mgb
{
  // per leaf: does the cell carry a hanging-node constraint on a face or an edge (Helper::is_constrained of the
  // reference, multigrid_throughput.cc:100-262)
  std::vector<char> cells_with_hanging_nodes(const Mesh &mesh);

  // n_cells split into equal chunks over the first n_active ranks (the other ranks own nothing)
  std::vector<int64_t> equal_boundaries(int64_t n_cells, int n_ranks, int64_t n_active);
  // the same with integer cell weights: cell i goes to the rank its cumulative weight (before i) falls into
  std::vector<int64_t> weighted_boundaries(const std::vector<int64_t> &weights, int n_ranks, int64_t n_active);
  // ownership "with the first child": a cell of `parents` belongs to the rank that owns the leaf of `children` at the
  // cell's Morton start (`children` refines or equals `parents` wherever `parents` has cells).  Used by
  // FirstChildPolicy and by the level meshes of local smoothing (level owner = owner of the first child).
  std::vector<int64_t> first_child_boundaries(const Mesh &parents, const Mesh &children, const std::vector<int64_t> &child_begin);

  struct PartitionPolicy
  {
    enum Kind
    {
      Default,
      MinimalGranularity,
      BalancedGranularity,
      CellWeight,
      FirstChild
    };
    Kind    kind           = Default;
    double  hanging_weight = 1.0; // CellWeightPolicy-W, FirstChildPolicy-W
    int64_t n_min_cells    = 1;   // MinimalGranularityPolicy-N

    // "" and "DefaultPolicy" -> Default; unknown names throw (ExcNotImplemented in the reference, :2170-2173)
    static PartitionPolicy parse(const std::string &name);

    // partition of the finest mesh; floor_min_cells: meshes with fewer cells per rank use fewer ranks
    std::vector<int64_t> fine(const Mesh &mesh, int n_ranks, int64_t floor_min_cells) const;
    // cell_begin for every mesh of a coarse -> fine global-coarsening sequence
    std::vector<std::vector<int64_t>> sequence(const std::vector<Mesh> &meshes, int n_ranks, int64_t floor_min_cells) const;
  };
} // namespace mgb

namespace mgb
{
  // MGTools::print_multigrid_statistics (include/mg_tools.h:267-513): workload, vertical and horizontal communication
  // efficiency of a partitioned hierarchy.  The partition is replicated knowledge here (Morton ranges), so the numbers
  // need no communication.
  struct HierarchyStatistics
  {
    double workload_eff = 1, workload_path_max = 0, vertical_eff = 1, horizontal_eff = 1, mem_total = 0;
    struct Level
    {
      int64_t n_cells_min = 0, n_cells_max = 0, n_cells_sum = 0;
      int64_t vertical_local = 0, vertical_remote = 0;     // children of the cells of the level below, summed over ranks
      int64_t horizontal_local = 0, horizontal_remote = 0; // owned / ghost cells, summed over ranks
    };
    std::vector<Level> levels;
  };
  // meshes: coarse -> fine.  Global coarsening: the coarsening sequence (mg_tools.h:64-81, 117-191, 214-233); local
  // smoothing: the level meshes of one refinement tree (mg_tools.h:39-62, 83-114, 193-212).
  HierarchyStatistics hierarchy_statistics(const std::vector<const Mesh *> &meshes, const std::vector<std::vector<int64_t>> &cell_begin,
                                           int n_ranks);
} // namespace mgb
