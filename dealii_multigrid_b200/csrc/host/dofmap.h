// FE_Q(p) degrees of freedom on an octree mesh: the stand-in for what the reference obtains from
// DoFHandler::distribute_dofs + AffineConstraints (Dirichlet id 0 + hanging nodes) + MatrixFree::reinit
// (multigrid_throughput.cc:1578-1595, 2274, 2305-2316; include/operator.h:24-47).
//
// Numbering (documented, our own -- SURVEY.md A.2: deal.II's depends on refinement history):
//   cells are walked in Morton order; each cell numbers its not-yet-numbered mesh entities in the
//   order vertices(8), lines(12), quads(6), hex(1); the (p-1)^dim DoFs of an entity are consecutive,
//   lexicographic in the entity's own axes (x fastest).  All cubes share the root orientation, so every
//   cell sees an entity's DoFs in the same order; a cell's (p+1)^3 lexicographic DoF indices are
//   therefore described exactly by 27 entity base offsets ("entity table", ternary index
//   e = tx + 3 ty + 9 tz, t in {0: low, 1: interior, 2: high}).
//
//   `eval_table` is the table used in gather/scatter (MatrixFree's dof_indices): on a hanging face /
//   edge the slots hold the COARSE neighbour's entities (the parent's face/edge), to be interpolated
//   in place with the 1-D subface matrices selected by `mask` (deal.II "fast hanging-node algorithm",
//   SURVEY.md A.3); bit 31 marks entities on the Dirichlet boundary (read as 0, never written).
//   `own_table` holds the cell's own entities (DoFHandler::get_dof_indices); bit 31 marks entities
//   whose DoFs are constrained (Dirichlet or hanging).
//
// Parallel layout: rank r owns a Morton-contiguous chunk of cells; an entity belongs to the rank of
// the first cell (Morton order) touching it, which makes every rank's owned DoFs one contiguous global
// range; local vectors are [owned | ghosts sorted by global index] (LinearAlgebra::distributed::Vector,
// SURVEY.md A.9).
#pragma once
#include "octree.h"

#include <cstdint>
#include <memory>
#include <vector>

namespace mgb
{
  constexpr uint32_t FLAG_BIT  = 0x80000000u;
  constexpr uint32_t INDEX_MASK = 0x7fffffffu;

  // mask bits (same meaning as MatrixFree's hanging_node_constraint_masks)
  enum : uint16_t
  {
    MASK_SUBCELL_X = 1,
    MASK_SUBCELL_Y = 2,
    MASK_SUBCELL_Z = 4,
    MASK_FACE_X    = 8,
    MASK_FACE_Y    = 16,
    MASK_FACE_Z    = 32,
    MASK_EDGE_X    = 64,
    MASK_EDGE_Y    = 128,
    MASK_EDGE_Z    = 256
  };

  struct Partitioner
  {
    int64_t               n_global = 0;
    int64_t               owned_begin = 0; // global index of first owned DoF
    uint32_t              n_owned = 0, n_ghost = 0;
    std::vector<int64_t>  ghost_global;    // n_ghost, ascending
    std::vector<int>      ghost_ranks;     // owners of ghost groups, ascending
    std::vector<uint32_t> ghost_offsets;   // ghost_ranks.size()+1, offsets into the ghost tail
    std::vector<int>      import_ranks;    // ranks that need owned entries of ours
    std::vector<uint32_t> import_offsets;  // import_ranks.size()+1
    std::vector<uint32_t> import_indices;  // owned local indices to send, grouped by import_ranks
  };

  struct DoFMap
  {
    int     degree  = 1;
    int     n_ranks = 1, rank = 0;
    int64_t n_cells_global = 0;
    int64_t cell_begin     = 0; // first local cell in the global Morton order
    int64_t n_cells        = 0; // local cells
    int     mesh_max_level = 0;

    std::vector<uint32_t> eval_table; // n_cells*27
    std::vector<uint32_t> own_table;  // n_cells*27
    std::vector<uint16_t> mask;       // n_cells
    std::vector<uint8_t>  cell_level; // n_cells
    std::vector<uint32_t> cell_xyz;   // n_cells*3, integer coordinates at the cell's level
    std::vector<uint32_t> constrained; // owned local indices, ascending (Dirichlet + hanging)
    int64_t               n_cells_hanging = 0; // local cells with a face/edge constraint
    Partitioner           part;
    // n_ranks > 1 only: the replicated global numbering (entity key -> first global DoF) and the rank boundaries, kept
    // for the two-level transfers, whose coarse cells may have children owned by other ranks
    std::shared_ptr<FlatMap> global_entities;
    std::vector<int64_t>     rank_dof_begin, rank_cell_begin; // n_ranks+1 each

    // build the active-mesh DoFs of `mesh` for FE_Q(degree), partitioned over n_ranks (equal cell counts)
    // min_cells_per_rank: meshes with fewer cells per rank are split over the first n_cells/min_cells_per_rank ranks
    // only (the coarse levels of a global-coarsening hierarchy then live on few GPUs or on rank 0 alone and need no
    // halo exchange; the two-level transfers move the data between the different partitions)
    static DoFMap build(const Mesh &mesh, int degree, int n_ranks = 1, int rank = 0, int64_t min_cells_per_rank = 1);
    // the same with an explicit partition: rank r owns the Morton range [cell_begin[r], cell_begin[r+1]) (ranks may be
    // empty); see partition.h for the policies that produce such ranges
    static DoFMap build(const Mesh &mesh, int degree, int n_ranks, int rank, const std::vector<int64_t> &cell_begin);

    // expand to the lexicographic (p+1)^3 index map (what MatrixFree stores); flags kept in bit 31
    void expand(const std::vector<uint32_t> &table, std::vector<uint32_t> &out) const;
  };

  // ternary entity index helpers
  inline int entity_dim(int e) { return (e % 3 == 1) + ((e / 3) % 3 == 1) + (e / 9 == 1); }
  extern const int ENTITY_ORDER[27]; // vertices, lines, quads, hex
  // local offset of node (i,j,k) inside its entity and the entity index
  void node_entity(int p, int i, int j, int k, int &e, int &offset);

  // Two-level transfer description between two DoF maps (geometric: fine mesh = coarse mesh with some
  // cells replaced by their 8 children; polynomial: same mesh) -- MGTwoLevelTransfer::reinit,
  // multigrid_throughput.cc:1600-1604; SURVEY.md A.7.
  struct TransferMap
  {
    int                   degree_fine = 1, degree_coarse = 1;
    int64_t               n_coarse_cells = 0;
    std::vector<uint32_t> fine_first;  // per coarse cell: first fine cell (Morton-contiguous children)
    std::vector<uint8_t>  n_children;  // per coarse cell: 1 (same cell) or 8
    std::vector<uint8_t>  weight_code; // n_fine_slots*27: 0 -> weight 0 (constrained fine entity), v -> 1/v
    // fine-side DoF table the kernels index with fine_first (+child): n_fine_slots*27 own-table entries.  Single rank: a
    // copy of the fine DoFMap's own_table (slot = fine cell).  Several ranks: the children of the locally owned COARSE
    // cells (which other ranks may own), in the numbering of `part`: same owned range as the fine level, plus the ghost
    // DoFs of those children -- the internal ghosted fine vector of MGTwoLevelTransfer.
    std::vector<uint32_t> fine_own;
    int64_t               n_fine_slots = 0;
    bool                  distributed  = false;
    Partitioner           part;
    // local_smoothing: the two meshes are consecutive levels of the refinement tree (Mesh::level_mesh); a coarse cell
    // then has 8 children on the fine level or none at all (n_children 0: contributes nothing, MGTransferMatrixFree,
    // SURVEY.md A.8)
    static TransferMap build(const Mesh &mesh_fine, const DoFMap &dofs_fine, const Mesh &mesh_coarse,
                             const DoFMap &dofs_coarse, bool local_smoothing = false);

  private:
    static TransferMap build_distributed(const Mesh &mesh_fine, const DoFMap &dofs_fine, const Mesh &mesh_coarse,
                                         const DoFMap &dofs_coarse, bool local_smoothing);
  };

  // Refinement-edge data of one local-smoothing level: MGConstrainedDoFs::refinement_edge_indices /
  // MGTools::extract_inner_interface_dofs (include/operator.h:49-66, 539-556) and the cells that touch such a DoF
  // (edge_constrained_cell, include/operator.h:71-99).
  struct LevelEdges
  {
    std::vector<uint32_t> edge_dofs;  // owned local indices, ascending
    std::vector<uint32_t> edge_cells; // local cells touching an edge DoF (owned or ghost), ascending
    bool                  any_global = false; // include/operator.h:67-69
    static LevelEdges build(const Mesh &level_mesh, const DoFMap &level_dofs);
  };

  // copy_indices of MGLevelGlobalTransfer: pairs (active-vector index, level-vector index) of the DoFs of
  // the ACTIVE cells living on this level, refinement-edge DoFs of the level skipped (fill_copy_indices with
  // skip_interface_dofs = true).
  // Several ranks: the active DoF and its level DoF may belong to different ranks (copy_indices_global_mine /
  // copy_indices_level_mine in deal.II).  The owner of the ACTIVE DoF holds the pair; level_idx then refers to an
  // auxiliary ghosted level vector with the layout `part` (owned range of the level + the foreign level DoFs of this
  // rank's pairs): copy_to_mg writes into it and compresses, copy_from_mg updates its ghosts and reads.
  struct CopyIndices
  {
    std::vector<uint32_t> active_idx, level_idx;
    bool                  distributed = false;
    Partitioner           part;
    // active_mesh: the active mesh of active_dofs (several ranks only; every rank derives all ranks' lists from it)
    static CopyIndices build(const DoFMap &active_dofs, const Mesh &level_mesh, const DoFMap &level_dofs, const Mesh *active_mesh = nullptr);

  private:
    static CopyIndices build_distributed(const DoFMap &active_dofs, const Mesh &active_mesh, const Mesh &level_mesh, const DoFMap &level_dofs);
  };
} // namespace mgb
