// Chunk plan of the cell loop (device/chunk.cuh): the Morton-ordered cells of one operator are cut into chunks of
// consecutive cells that one thread block processes out of shared memory, and the DoF vector is cut into the matching
// contiguous "owned" ranges (the DoFs first touched by the chunk's cells).  A chunk
//   * loads its owned range with ONE bulk copy and the few foreign DoFs (owned by other chunks, or ghosts) one by one,
//   * applies the cell matrices out of shared memory,
//   * sums the cell contributions per DoF in a fixed order (no atomics), hands the sums of its foreign DoFs to their
//     owners through a side buffer, and writes its owned range -- complete -- with plain, coalesced stores, which is
//     what lets the vector update of the smoother run in the same pass (deal.II's cell_loop post hook; the reference's
//     smoother is built at multigrid_throughput.cc:866-883 on the operator of include/operator.h:152-183).
// Chunks are processed in DESCENDING order: with first-touch numbering every foreign DoF of a chunk belongs to a lower
// chunk, so an owner only ever waits for chunks that were started before it ("pull" dependencies).  The rare opposite
// case (a fine cell that reads the face DoFs of a coarse neighbour numbered later) is a "late" contribution: it is
// added atomically once the owner has stored.
//
// Everything a chunk needs except four offsets is expressed relative to the chunk (slots, dependency ordinals), so that
// equal configurations share one table ("blob") and the tables of a uniform mesh stay resident in L2.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace mgb
{
  // geometry of a chunk for degree p: cells per chunk and slot capacity of the shared-memory DoF window
  int chunk_warps(int degree);
  int chunk_cells(int degree);
  int chunk_slot_capacity(int degree);
  constexpr int CHUNK_ZERO_SLOTS = 32; // slots that always read 0 (Dirichlet entities)
  constexpr int CHUNK_DESC_WORDS = 8;

  // words of a chunk descriptor
  enum
  {
    CD_BLOB       = 0, // word offset of the compute blob in `pool`
    CD_PULL_BLOB  = 1, // word offset of the pull blob in `pool`
    CD_OWNED_BEGIN = 2, // first DoF of the owned range (local index)
    CD_SIDE_OFF   = 3, // first entry of the chunk's side buffer
    CD_SRC_DEPS   = 4, // offset into `deps`: owners of the foreign DoFs (pull-type first, then late-type)
    CD_PULL_DEPS  = 5, // offset into `deps`: chunks that contribute to owned DoFs of this chunk
    CD_GHOST_OFF  = 6, // offset into `ghost_idx`
    CD_N_OWNED    = 7
  };
  // words of the compute blob header
  enum
  {
    CB_N_CELLS = 0,
    CB_N_OWNED,
    CB_N_FP, // foreign slots owned by a lower chunk (side buffer)
    CB_N_FL, // foreign slots owned by a higher chunk (late, atomic)
    CB_N_FG, // foreign slots in the ghost range (atomic into the ghost tail)
    CB_N_DEP_LOW,  // src deps owning the FP slots
    CB_N_DEP_LATE, // src deps owning the FL slots
    CB_HANGING,    // any cell of the chunk has a hanging-node constraint
    CB_OFF_CELLREC,
    CB_OFF_FSRC,
    CB_N_ROUNDS,   // merge rounds: copies of one DoF in the cell storage are summed pairwise, as a binary tree
    CB_OFF_MERGE,  // n_rounds + 1 offsets (relative to this word), then the pairs: destination | source << 16 (positions of
                   // the cell storage; within one round every position occurs at most once)
    CB_OFF_SLOTPOS, // per slot the 16-bit position holding its sum after the merge rounds (0xffff: no contribution)
    CB_TOTAL_WORDS,
    CB_HEADER_WORDS = 16
  };
  // words of the pull blob header
  enum
  {
    PB_N_SLOTS = 0, // owned slots that receive contributions of other chunks
    PB_N_DEPS,
    PB_OFF_REC,  // 2 words per slot: owned slot (13 bits) | count << 13 (5 bits) | offset of the further entries << 18, first
                 // entry
    PB_OFF_LIST, // the further entries: dependency ordinal << 16 | position in that chunk's side buffer
    PB_OFF_OPOS, // per owned slot (16 bit): position of its sum in the cell storage (0x7fff: none) | 0x8000 if other
                 // chunks contribute
    PB_TOTAL_WORDS,
    PB_HEADER_WORDS = 8
  };

  struct ChunkPlan
  {
    int                   degree = 0, value_bytes = 8;
    uint32_t              n_chunks = 0;
    std::vector<uint32_t> desc;      // CHUNK_DESC_WORDS per chunk
    std::vector<uint32_t> pool;      // deduplicated blobs
    std::vector<uint32_t> deps;      // chunk ids
    std::vector<uint32_t> ghost_idx; // local indices of ghost-type foreign slots
    uint64_t              side_size = 0;
    // statistics
    uint32_t n_compute_blobs = 0, n_pull_blobs = 0;
    uint64_t n_foreign_low = 0, n_foreign_late = 0, n_foreign_ghost = 0;
    uint32_t max_slots = 0;
  };

  // etab: n_cells*27 evaluation entity table (local indices, bit 31 = Dirichlet), mask/level per cell; value_bytes 8 or 4
  // selects the shared-memory cell layout (SlabLayout) the positions refer to; first_cell_global: position of the first
  // local cell in the global Morton order (chunk boundaries are aligned globally).  Throws std::runtime_error when the mesh
  // cannot be chunked (the caller keeps the atomic cell kernel).
  ChunkPlan build_chunk_plan(int degree, long long n_cells, const uint32_t *etab, const uint16_t *mask, const uint8_t *level,
                             uint32_t n_owned, uint32_t n_ghost, int value_bytes, long long first_cell_global = 0);

  // CPU walk through the plan in the kernel's order with a diagonal stand-in for the cell operator; returns the largest
  // relative deviation from the direct scatter-add (tests/test_chunk_plan.py); >= 1e300: a structural defect
  double check_chunk_plan(const ChunkPlan &plan, long long n_cells, const uint32_t *etab, uint32_t n_owned, uint32_t n_ghost);

  // shared-memory cell layout of the chunk kernel (device/chunk.cuh ChunkLayout): plane and cell stride
  void chunk_layout(int n, int value_bytes, int &plane_stride, int &cell_stride);
} // namespace mgb
