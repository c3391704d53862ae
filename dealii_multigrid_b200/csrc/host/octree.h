// Octree mesh on the single root cube [-1,1]^3 -- the stand-in for the
// p4est-backed parallel::distributed::Triangulation the reference builds in
// multigrid_throughput.cc:2041-2062.  Every geometry of the reference is
// hyper_cube(-1,1) + isotropic refinement (include/grid_generator.h:1-141), so
// a linear octree with 2:1 balance over faces, edges and corners (what p4est
// enforces) represents all of them.
//
// Leaves are stored in Morton (z-) order, child c at offsets
// (c&1, (c>>1)&1, (c>>2)&1) -- the deal.II / p4est child order.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace mgb
{
  // Allocator of the big random-access tables of the host setup: blocks of 4 MB and more are 2 MB-aligned and advised to
  // transparent huge pages (MADV_HUGEPAGE; most hosts only give huge pages on request).  A 64 MB hash table on 4 KB pages
  // pays a TLB miss and a page walk per lookup, which software prefetch does not hide.
  void *huge_alloc(size_t bytes);
  void  huge_free(void *p);
  template <typename T> struct HugeAllocator
  {
    using value_type = T;
    HugeAllocator() = default;
    template <typename U> HugeAllocator(const HugeAllocator<U> &) {}
    T   *allocate(size_t n) { return static_cast<T *>(huge_alloc(n * sizeof(T))); }
    void deallocate(T *p, size_t) { huge_free(p); }
    template <typename U> bool operator==(const HugeAllocator<U> &) const { return true; }
    template <typename U> bool operator!=(const HugeAllocator<U> &) const { return false; }
  };

  // open-addressing hash map uint64 -> int64 (keys never removed)
  class FlatMap
  {
  public:
    void    reserve(size_t n);
    void    insert(uint64_t key, int64_t value);
    int64_t find(uint64_t key) const; // -1 if absent
    // the value stored for key; if there is none, value is stored and returned (one probe sequence instead of find + insert)
    int64_t find_or_insert(uint64_t key, int64_t value);
    size_t  size() const { return n_; }
    void    clear();
    // hint that `key` is looked up or inserted soon: the big tables of the setup are random-access bound, and the 27
    // entities of a cell can be requested together
    void prefetch(uint64_t key) const
    {
      if (!slots_.empty())
        __builtin_prefetch(&slots_[hash(key) & mask_]);
    }

  private:
    struct Slot // key and value share a cache line
    {
      uint64_t key;
      int64_t  val;
    };
    static uint64_t   hash(uint64_t k);
    void              grow();
    std::vector<Slot, HugeAllocator<Slot>> slots_;
    size_t                                 n_    = 0;
    size_t            mask_ = 0;
  };

  struct Cell
  {
    uint32_t x, y, z; // integer coordinates at the cell's own level
    uint8_t  level;
  };

  class Mesh
  {
  public:
    std::vector<Cell> cells; // Morton-ordered leaves

    static Mesh create(const std::string &geometry, unsigned n_ref_global, unsigned n_ref_local);

    int  max_level() const;
    int  n_global_levels() const { return max_level() + 1; }
    bool uniform() const;

    // refine flagged leaves, then restore 2:1 balance (faces+edges+corners)
    void refine(const std::vector<char> &flags);
    void refine_global(unsigned times = 1);
    // one global coarsening step: every complete family of 8 leaves -> parent, then balance
    // (MGTransferGlobalCoarseningTools::create_geometric_coarsening_sequence, ref :2219-2224)
    Mesh coarsen_global() const;
    // local-smoothing level l of the refinement tree: ALL cells of refinement level l, active or not (SURVEY.md A.1,
    // construct_multigrid_hierarchy, ref :2044-2046).  The result does not cover the domain where coarser leaves live.
    Mesh level_mesh(int level) const;

    void rebuild_index();
    // leaf index of the cell (level,x,y,z); -1 if that cell is not a leaf
    int64_t leaf(int level, uint32_t x, uint32_t y, uint32_t z) const;
    // relation of the region (level,x,y,z) to the leaves:
    //   returns leaf index and sets rel = 0 (is a leaf), -k (contained in a leaf k levels coarser),
    //   +1 (subdivided further; returns -1)
    int64_t locate(int level, uint32_t x, uint32_t y, uint32_t z, int &rel) const;

    static uint64_t key(int level, uint32_t x, uint32_t y, uint32_t z)
    {
      return (uint64_t(level) << 57) | (uint64_t(x) << 38) | (uint64_t(y) << 19) | uint64_t(z);
    }
    void center(const Cell &c, double ctr[3]) const;

  private:
    void    balance();
    FlatMap index_;
  };
} // namespace mgb
