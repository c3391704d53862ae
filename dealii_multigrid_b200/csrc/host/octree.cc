#include "octree.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <new>
#include <stdexcept>

#include <sys/mman.h>

namespace mgb
{
  // ------------------------------------------------------------------ FlatMap
  void *huge_alloc(size_t bytes)
  {
    void *p = nullptr;
    if (bytes >= (size_t(4) << 20))
      {
        const size_t two_mb = size_t(2) << 20, rounded = (bytes + two_mb - 1) / two_mb * two_mb;
        if (posix_memalign(&p, two_mb, rounded) != 0)
          throw std::bad_alloc();
        madvise(p, rounded, MADV_HUGEPAGE); // advice only: without huge pages the table is merely slower
      }
    else
      {
        p = std::malloc(bytes ? bytes : 1);
        if (!p)
          throw std::bad_alloc();
      }
    return p;
  }
  void huge_free(void *p) { std::free(p); }

  uint64_t FlatMap::hash(uint64_t k)
  {
    k ^= k >> 33;
    k *= 0xff51afd7ed558ccdULL;
    k ^= k >> 33;
    k *= 0xc4ceb9fe1a85ec53ULL;
    k ^= k >> 33;
    return k;
  }

  void FlatMap::clear()
  {
    slots_.clear();
    n_    = 0;
    mask_ = 0;
  }

  void FlatMap::reserve(size_t n)
  {
    size_t cap = 16;
    while (cap < n + n / 2 + 2) // load factor <= 2/3 (linear probing, 16-byte slots)
      cap <<= 1;
    if (cap <= slots_.size())
      return;
    std::vector<Slot, HugeAllocator<Slot>> old;
    old.swap(slots_);
    slots_.assign(cap, Slot{~uint64_t(0), -1});
    mask_ = cap - 1;
    n_    = 0;
    for (const Slot &sl : old)
      if (sl.key != ~uint64_t(0))
        insert(sl.key, sl.val);
  }

  void FlatMap::grow() { reserve(std::max<size_t>(16, 2 * n_ + 8)); }

  void FlatMap::insert(uint64_t key, int64_t value)
  {
    if (slots_.empty() || 3 * (n_ + 1) > 2 * slots_.size())
      grow();
    size_t i = hash(key) & mask_;
    while (slots_[i].key != ~uint64_t(0) && slots_[i].key != key)
      i = (i + 1) & mask_;
    if (slots_[i].key != key)
      ++n_;
    slots_[i] = Slot{key, value};
  }

  int64_t FlatMap::find_or_insert(uint64_t key, int64_t value)
  {
    if (slots_.empty() || 3 * (n_ + 1) > 2 * slots_.size())
      grow();
    size_t i = hash(key) & mask_;
    while (slots_[i].key != ~uint64_t(0))
      {
        if (slots_[i].key == key)
          return slots_[i].val;
        i = (i + 1) & mask_;
      }
    slots_[i] = Slot{key, value};
    ++n_;
    return value;
  }

  int64_t FlatMap::find(uint64_t key) const
  {
    if (slots_.empty())
      return -1;
    size_t i = hash(key) & mask_;
    while (slots_[i].key != ~uint64_t(0))
      {
        if (slots_[i].key == key)
          return slots_[i].val;
        i = (i + 1) & mask_;
      }
    return -1;
  }

  // --------------------------------------------------------------------- Mesh
  int Mesh::max_level() const
  {
    int m = 0;
    for (const auto &c : cells)
      m = std::max<int>(m, c.level);
    return m;
  }

  bool Mesh::uniform() const
  {
    for (const auto &c : cells)
      if (c.level != cells[0].level)
        return false;
    return true;
  }

  void Mesh::rebuild_index()
  {
    index_.clear();
    index_.reserve(cells.size());
    for (size_t i = 0; i < cells.size(); ++i)
      index_.insert(key(cells[i].level, cells[i].x, cells[i].y, cells[i].z), (int64_t)i);
  }

  int64_t Mesh::leaf(int level, uint32_t x, uint32_t y, uint32_t z) const
  {
    return index_.find(key(level, x, y, z));
  }

  int64_t Mesh::locate(int level, uint32_t x, uint32_t y, uint32_t z, int &rel) const
  {
    for (int l = level; l >= 0; --l)
      {
        const int     s = level - l;
        const int64_t i = index_.find(key(l, x >> s, y >> s, z >> s));
        if (i >= 0)
          {
            rel = -s;
            return i;
          }
      }
    rel = 1;
    return -1;
  }

  void Mesh::center(const Cell &c, double ctr[3]) const
  {
    const double h = 2.0 / double(1u << c.level);
    ctr[0]         = -1.0 + h * (c.x + 0.5);
    ctr[1]         = -1.0 + h * (c.y + 0.5);
    ctr[2]         = -1.0 + h * (c.z + 0.5);
  }

  void Mesh::refine(const std::vector<char> &flags)
  {
    std::vector<char> f = flags;
    while (true)
      {
        bool any = false;
        for (char c : f)
          any |= (c != 0);
        if (!any)
          break;
        std::vector<Cell> out;
        out.reserve(cells.size() + 7 * std::count(f.begin(), f.end(), 1));
        for (size_t i = 0; i < cells.size(); ++i)
          {
            const Cell &c = cells[i];
            if (!f[i])
              out.push_back(c);
            else
              for (unsigned ch = 0; ch < 8; ++ch)
                out.push_back(Cell{2 * c.x + (ch & 1), 2 * c.y + ((ch >> 1) & 1), 2 * c.z + ((ch >> 2) & 1),
                                   uint8_t(c.level + 1)});
          }
        cells.swap(out);
        rebuild_index();
        // 2:1 balance across faces, edges and corners (p4est CONNECT_FULL): a leaf that is more than
        // one level coarser than any of the 26 same-level neighbour regions of a leaf must be refined
        f.assign(cells.size(), 0);
        if (uniform())
          break;
        for (size_t i = 0; i < cells.size(); ++i)
          {
            const Cell &c = cells[i];
            if (c.level < 2)
              continue;
            const int64_t N = int64_t(1) << c.level;
            for (int dz = -1; dz <= 1; ++dz)
              for (int dy = -1; dy <= 1; ++dy)
                for (int dx = -1; dx <= 1; ++dx)
                  {
                    if (!dx && !dy && !dz)
                      continue;
                    const int64_t nx = int64_t(c.x) + dx, ny = int64_t(c.y) + dy, nz = int64_t(c.z) + dz;
                    if (nx < 0 || ny < 0 || nz < 0 || nx >= N || ny >= N || nz >= N)
                      continue;
                    int           rel;
                    const int64_t j = locate(c.level, (uint32_t)nx, (uint32_t)ny, (uint32_t)nz, rel);
                    if (j >= 0 && rel < -1)
                      f[j] = 1;
                  }
          }
      }
  }

  void Mesh::refine_global(unsigned times)
  {
    for (unsigned t = 0; t < times; ++t)
      {
        std::vector<char> f(cells.size(), 1);
        refine(f);
      }
  }

  Mesh Mesh::coarsen_global() const
  {
    Mesh out;
    out.cells.reserve(cells.size() / 8 + 8);
    size_t i = 0;
    while (i < cells.size())
      {
        const Cell &c   = cells[i];
        bool        fam = false;
        if (c.level > 0 && (c.x & 1) == 0 && (c.y & 1) == 0 && (c.z & 1) == 0 && i + 8 <= cells.size())
          {
            fam = true;
            for (unsigned ch = 0; ch < 8 && fam; ++ch)
              {
                const Cell &s = cells[i + ch];
                fam           = s.level == c.level && s.x == c.x + (ch & 1) && s.y == c.y + ((ch >> 1) & 1) &&
                      s.z == c.z + ((ch >> 2) & 1);
              }
          }
        if (fam)
          {
            out.cells.push_back(Cell{c.x >> 1, c.y >> 1, c.z >> 1, uint8_t(c.level - 1)});
            i += 8;
          }
        else
          {
            out.cells.push_back(c);
            i += 1;
          }
      }
    out.rebuild_index();
    // restore 2:1 balance (may re-refine some of the freshly created parents)
    std::vector<char> f(out.cells.size(), 0);
    bool              any = false;
    if (!out.uniform())
      for (size_t k = 0; k < out.cells.size(); ++k)
        {
          const Cell &c = out.cells[k];
          if (c.level < 2)
            continue;
          const int64_t N = int64_t(1) << c.level;
          for (int dz = -1; dz <= 1; ++dz)
            for (int dy = -1; dy <= 1; ++dy)
              for (int dx = -1; dx <= 1; ++dx)
                {
                  if (!dx && !dy && !dz)
                    continue;
                  const int64_t nx = int64_t(c.x) + dx, ny = int64_t(c.y) + dy, nz = int64_t(c.z) + dz;
                  if (nx < 0 || ny < 0 || nz < 0 || nx >= N || ny >= N || nz >= N)
                    continue;
                  int           rel;
                  const int64_t j = out.locate(c.level, (uint32_t)nx, (uint32_t)ny, (uint32_t)nz, rel);
                  if (j >= 0 && rel < -1)
                    {
                      f[j] = 1;
                      any  = true;
                    }
                }
        }
    if (any)
      out.refine(f);
    return out;
  }

  Mesh Mesh::level_mesh(int level) const
  {
    // leaves are Morton-ordered, so the level-`level` ancestors of the leaves at least that fine appear in Morton order too
    Mesh out;
    for (const Cell &c : cells)
      {
        if (c.level < level)
          continue;
        const int  s = c.level - level;
        const Cell a{c.x >> s, c.y >> s, c.z >> s, uint8_t(level)};
        if (out.cells.empty() || out.cells.back().x != a.x || out.cells.back().y != a.y || out.cells.back().z != a.z)
          out.cells.push_back(a);
      }
    out.rebuild_index();
    return out;
  }

  Mesh Mesh::create(const std::string &geometry, unsigned n_ref_global, unsigned n_ref_local)
  {
    Mesh m;
    m.cells.push_back(Cell{0, 0, 0, 0});
    m.rebuild_index();

    const auto refine_where = [&m](auto &&pred) {
      std::vector<char> f(m.cells.size(), 0);
      for (size_t i = 0; i < m.cells.size(); ++i)
        f[i] = pred(m.cells[i]) ? 1 : 0;
      m.refine(f);
    };
    const auto center_in_negative_orthant = [&m](const Cell &c) {
      double ctr[3];
      m.center(c, ctr);
      return !(ctr[0] > 0.0 || ctr[1] > 0.0 || ctr[2] > 0.0);
    };
    const auto center_norm = [&m](const Cell &c) {
      double ctr[3];
      m.center(c, ctr);
      return std::sqrt(ctr[0] * ctr[0] + ctr[1] * ctr[1] + ctr[2] * ctr[2]);
    };

    if (geometry == "hypercube")
      {
        // ref multigrid_throughput.cc:2056-2060
        m.refine_global(n_ref_global);
      }
    else if (geometry == "quadrant_flexible")
      {
        // ref include/grid_generator.h:69-92
        m.refine_global(n_ref_global);
        for (unsigned i = 0; i < n_ref_local; ++i)
          refine_where(center_in_negative_orthant);
      }
    else if (geometry == "quadrant")
      {
        // ref include/grid_generator.h:34-65
        if (n_ref_global == 0)
          return m;
        m.refine_global(1);
        for (unsigned i = 1; i < n_ref_global; ++i)
          refine_where(center_in_negative_orthant);
        if ((unsigned)m.max_level() != n_ref_global)
          throw std::runtime_error("create_quadrant: n_global_levels-1 != n_refinements");
      }
    else if (geometry == "circle")
      {
        // ref include/grid_generator.h:3-30
        m.refine_global(std::min<unsigned>(n_ref_global, 3));
        const double r = 1.0 / (4.0 * 3.14159265358979323846);
        for (unsigned i = 3; i < n_ref_global; ++i)
          refine_where([&](const Cell &c) {
            const double h = 2.0 / double(1u << c.level);
            for (unsigned v = 0; v < 8; ++v)
              {
                const double px = -1.0 + h * (c.x + (v & 1)), py = -1.0 + h * (c.y + ((v >> 1) & 1)),
                             pz = -1.0 + h * (c.z + ((v >> 2) & 1));
                if (std::sqrt(px * px + py * py + pz * pz) < r)
                  return true;
              }
            return false;
          });
        if ((unsigned)m.max_level() != n_ref_global)
          throw std::runtime_error("create_circle: n_global_levels-1 != n_refinements");
      }
    else if (geometry == "annulus")
      {
        // ref include/grid_generator.h:96-140
        if (n_ref_global == 0)
          return m;
        for (int i = 0; i < int(n_ref_global) - 3; ++i)
          m.refine_global(1);
        if (n_ref_global >= 1)
          refine_where([&](const Cell &c) { return center_norm(c) < 0.55; });
        if (n_ref_global >= 2)
          refine_where([&](const Cell &c) {
            const double r = center_norm(c);
            return 0.3 <= r && r <= 0.43;
          });
        if (n_ref_global >= 3)
          refine_where([&](const Cell &c) {
            const double r = center_norm(c);
            return 0.335 <= r && r <= 0.39;
          });
      }
    else
      throw std::runtime_error("unknown GeometryType '" + geometry + "'");
    return m;
  }
} // namespace mgb
