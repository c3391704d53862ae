#include "chunks.h"

#include "dofmap.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <stdexcept>
#include <string>
#include <unordered_map>

namespace mgb
{
  // Chunk sizes are powers of two (whole octets of the Morton order) and chunk boundaries sit at multiples of the size in
  // the GLOBAL cell order: on a uniform mesh every chunk is then the same box (or one of two half boxes), sees its
  // neighbours in the same relative position, and all chunks of one kind share one table.
#ifndef MGB_CHUNK_NC4
#define MGB_CHUNK_NC4 32
#endif
#ifndef MGB_CHUNK_CAP4
#define MGB_CHUNK_CAP4 (MGB_CHUNK_NC4 == 32 ? 2880 : 1536)
#endif
  int chunk_cells(int p)
  {
    static const int nc[7] = {0, 256, 64, 64, MGB_CHUNK_NC4, 32, 16};
    return nc[p];
  }
  int chunk_warps(int p)
  {
    static const int w[7] = {0, 8, 7, 8, (MGB_CHUNK_NC4 + 5) / 6, 7, 4};
    return w[p];
  }
  int chunk_slot_capacity(int p)
  {
    static const int cap[7] = {0, 2112, 1792, 3072, MGB_CHUNK_CAP4, 5632, 5632};
    return cap[p];
  }

  void chunk_layout(int n, int bytes, int &ps, int &cs)
  {
    // must equal ChunkLayout<N, BYTES> of device/chunk.cuh (checked by its launcher): SlabLayout of device/cell_common.cuh
    // except for n = 5 doubles, where the tighter (27, 135) keeps three blocks resident per SM
    static const int t8[8][2] = {{0, 0}, {0, 0}, {5, 10}, {19, 57}, {17, 68}, {27, 135}, {37, 222}, {55, 385}};
    static const int t4[8][2] = {{0, 0}, {0, 0}, {5, 10}, {26, 105}, {17, 68}, {27, 153}, {37, 222}, {50, 369}};
    if (n < 2 || n > 7 || (bytes != 4 && bytes != 8))
      throw std::runtime_error("chunk_layout: unsupported size");
    ps = (bytes == 8 ? t8 : t4)[n][0];
    cs = (bytes == 8 ? t8 : t4)[n][1];
  }

  namespace
  {
    // open-addressing map u32 -> u32 with O(1) clear (generation stamps)
    struct LocalMap
    {
      std::vector<uint32_t> key, val, stamp;
      uint32_t              gen = 0, mask = 0;
      explicit LocalMap(uint32_t cap_pow2)
        : key(cap_pow2)
        , val(cap_pow2)
        , stamp(cap_pow2, 0)
        , mask(cap_pow2 - 1)
      {}
      void clear() { ++gen; }
      // returns pointer to the value; inserted = true when the key was new (value uninitialised)
      uint32_t *find_or_insert(uint32_t k, bool &inserted)
      {
        uint32_t h = (k * 2654435761u) & mask;
        while (true)
          {
            if (stamp[h] != gen)
              {
                stamp[h] = gen;
                key[h]   = k;
                inserted = true;
                return &val[h];
              }
            if (key[h] == k)
              {
                inserted = false;
                return &val[h];
              }
            h = (h + 1) & mask;
          }
      }
      const uint32_t *find(uint32_t k) const
      {
        uint32_t h = (k * 2654435761u) & mask;
        while (stamp[h] == gen)
          {
            if (key[h] == k)
              return &val[h];
            h = (h + 1) & mask;
          }
        return nullptr;
      }
    };

    // pool of deduplicated word blobs
    struct BlobPool
    {
      std::vector<uint32_t>                                &pool;
      std::unordered_map<uint64_t, std::vector<uint32_t>> index; // hash -> offsets
      uint32_t                                              n_unique = 0;
      explicit BlobPool(std::vector<uint32_t> &p)
        : pool(p)
      {}
      uint32_t add(const std::vector<uint32_t> &b)
      {
        uint64_t h = 1469598103934665603ull;
        for (uint32_t w : b)
          {
            h ^= w;
            h *= 1099511628211ull;
          }
        auto &cands = index[h];
        for (uint32_t off : cands)
          if (off + b.size() <= pool.size() && std::memcmp(&pool[off], b.data(), b.size() * 4) == 0)
            return off;
        // blobs start on 16-byte boundaries
        while (pool.size() % 4)
          pool.push_back(0);
        if (pool.size() + b.size() >= (size_t(1) << 32))
          throw std::runtime_error("chunk plan: table pool exceeds 2^32 words");
        const uint32_t off = (uint32_t)pool.size();
        pool.insert(pool.end(), b.begin(), b.end());
        cands.push_back(off);
        ++n_unique;
        return off;
      }
    };

    struct PullRun
    {
      uint32_t owner, contributor, owner_slot, rel, len;
    };
  } // namespace

  ChunkPlan build_chunk_plan(int p, long long n_cells, const uint32_t *etab, const uint16_t *mask, const uint8_t *level,
                             uint32_t n_owned, uint32_t n_ghost, int value_bytes, long long first_cell_global)
  {
    ChunkPlan plan;
    plan.degree      = p;
    plan.value_bytes = value_bytes;
    if (n_cells <= 0)
      return plan;
    const int n = p + 1;
    int       PS, CS;
    chunk_layout(n, value_bytes, PS, CS);
    const int NC = chunk_cells(p), CAP = chunk_slot_capacity(p);
    if (NC * CS >= 65536)
      throw std::runtime_error("chunk plan: cell storage positions exceed 16 bits");
    int esize[27];
    for (int e = 0; e < 27; ++e)
      {
        const int d = entity_dim(e);
        esize[e]    = d == 0 ? 1 : (d == 1 ? (p - 1) : (d == 2 ? (p - 1) * (p - 1) : (p - 1) * (p - 1) * (p - 1)));
      }
    const uint32_t shift_max = uint32_t(16 / value_bytes - 1);

    // ---- pass 0: the cell that touches an entity first, and per cell the lowest DoF it touches first
    std::vector<uint32_t> suffix_min(size_t(n_cells) + 1, n_owned);
    {
      std::vector<uint8_t> seen(n_owned, 0);
      for (long long c = 0; c < n_cells; ++c)
        {
          uint32_t lo = n_owned;
          for (int e = 0; e < 27; ++e)
            {
              const uint32_t t = etab[size_t(c) * 27 + e];
              if (esize[e] == 0 || (t & FLAG_BIT))
                continue;
              const uint32_t idx = t & INDEX_MASK;
              if (idx >= n_owned)
                {
                  if (idx + esize[e] > n_owned + n_ghost)
                    throw std::runtime_error("chunk plan: entity table index out of range");
                  continue;
                }
              if (!seen[idx])
                {
                  seen[idx] = 1;
                  lo        = std::min(lo, idx);
                }
            }
          suffix_min[c] = lo;
        }
      for (long long c = n_cells - 1; c >= 0; --c)
        suffix_min[c] = std::min(suffix_min[c], suffix_min[c + 1]);
      suffix_min[0] = 0; // leading DoFs nobody reads (Dirichlet) still belong to the first chunk
    }

    // ---- pass 1: chunk boundaries (cells and owned ranges)
    std::vector<uint32_t> cell_begin, owned_begin; // per chunk (+ end sentinel)
    LocalMap              lmap(8192);
    {
      long long c0 = 0;
      while (c0 < n_cells)
        {
          long long nc = std::min<long long>(NC - (first_cell_global + c0) % NC, n_cells - c0);
          while (true)
            {
              const uint32_t b0 = suffix_min[c0], b1 = suffix_min[c0 + nc];
              uint32_t       n_foreign = 0;
              lmap.clear();
              for (long long c = c0; c < c0 + nc; ++c)
                for (int e = 0; e < 27; ++e)
                  {
                    const uint32_t t = etab[size_t(c) * 27 + e];
                    if (esize[e] == 0 || (t & FLAG_BIT))
                      continue;
                    const uint32_t idx = t & INDEX_MASK;
                    if (idx >= b0 && idx < b1)
                      continue;
                    bool ins;
                    lmap.find_or_insert(idx, ins);
                    if (ins)
                      n_foreign += esize[e];
                  }
              if ((b1 - b0) + n_foreign + CHUNK_ZERO_SLOTS + shift_max <= uint32_t(CAP) && (b1 - b0) < 65536u)
                break;
              if (nc == 1)
                throw std::runtime_error("chunk plan: a single cell exceeds the slot capacity");
              nc = std::max<long long>(1, nc / 2);
            }
          cell_begin.push_back(uint32_t(c0));
          owned_begin.push_back(suffix_min[c0]);
          c0 += nc;
        }
      cell_begin.push_back(uint32_t(n_cells));
      owned_begin.push_back(n_owned);
    }
    const uint32_t K = uint32_t(cell_begin.size() - 1);
    plan.n_chunks    = K;
    plan.desc.assign(size_t(K) * CHUNK_DESC_WORDS, 0u);
    const auto owner_of = [&](uint32_t idx) -> uint32_t {
      // last chunk whose owned range starts at or before idx and is not empty there
      uint32_t k = uint32_t(std::upper_bound(owned_begin.begin(), owned_begin.begin() + K, idx) - owned_begin.begin()) - 1;
      return k;
    };

    // ---- pass 2: compute blobs
    BlobPool               blobs(plan.pool);
    std::vector<PullRun>   runs;
    std::vector<uint32_t>  src_deps_all;         // concatenated per chunk
    std::vector<uint32_t>  src_deps_off(K + 1, 0);
    std::vector<uint32_t>  blob;
    std::vector<uint32_t>  ent_cnt, ent_off, rec;
    std::vector<uint16_t>  ent_list;
    struct Foreign
    {
      uint32_t idx, len, type, dep, slot;
    };
    std::vector<Foreign>  foreign;
    std::unordered_map<uint64_t, std::vector<uint32_t>> key_cache; // hash of the cheap blob part -> blob offsets
    std::vector<uint32_t> dep_low, dep_late;
    uint64_t              side_off = 0;
    for (uint32_t k = 0; k < K; ++k)
      {
        const uint32_t c0 = cell_begin[k], nc = cell_begin[k + 1] - c0;
        const uint32_t b0 = owned_begin[k], b1 = owned_begin[k + 1], n_own = b1 - b0;
        foreign.clear();
        dep_low.clear();
        dep_late.clear();
        lmap.clear();
        bool hanging = false;
        for (uint32_t c = c0; c < c0 + nc; ++c)
          {
            hanging |= mask && (mask[c] >> 3) != 0;
            for (int e = 0; e < 27; ++e)
              {
                const uint32_t t = etab[size_t(c) * 27 + e];
                if (esize[e] == 0 || (t & FLAG_BIT))
                  continue;
                const uint32_t idx = t & INDEX_MASK;
                if (idx >= b0 && idx < b1)
                  continue;
                bool      ins;
                uint32_t *v = lmap.find_or_insert(idx, ins);
                if (!ins)
                  continue;
                *v = uint32_t(foreign.size());
                Foreign f{idx, uint32_t(esize[e]), 0u, 0u, 0u};
                if (idx >= n_owned)
                  f.type = 2;
                else
                  {
                    const uint32_t o = owner_of(idx);
                    if (o == k)
                      throw std::runtime_error("chunk plan: inconsistent owned range");
                    if (owner_of(idx + f.len - 1) != o)
                      throw std::runtime_error("chunk plan: an entity straddles two owned ranges");
                    f.type                  = o < k ? 0 : 1;
                    std::vector<uint32_t> &d = o < k ? dep_low : dep_late;
                    size_t                 q = std::find(d.begin(), d.end(), o) - d.begin();
                    if (q == d.size())
                      d.push_back(o);
                    f.dep = uint32_t(q);
                  }
                foreign.push_back(f);
              }
          }
        uint32_t n_f[3] = {0, 0, 0};
        for (const Foreign &f : foreign)
          n_f[f.type] += f.len;
        {
          uint32_t next[3] = {n_own, n_own + n_f[0], n_own + n_f[0] + n_f[1]};
          for (Foreign &f : foreign)
            {
              f.slot = next[f.type];
              next[f.type] += f.len;
            }
        }
        const uint32_t n_slots = n_own + n_f[0] + n_f[1] + n_f[2];
        const uint32_t zero    = n_slots;
        plan.max_slots         = std::max(plan.max_slots, n_slots);
        plan.n_foreign_low += n_f[0];
        plan.n_foreign_late += n_f[1];
        plan.n_foreign_ghost += n_f[2];
        if (dep_low.size() + dep_late.size() >= 65536)
          throw std::runtime_error("chunk plan: too many dependencies");

        // cheap part of the blob: header, cell records, sources of the foreign slots.  It determines the rest (the per-slot
        // contribution lists), so it doubles as the key of a cache: on a uniform mesh almost every chunk hits
        rec.assign(size_t(nc) * 32, 0u);
        for (uint32_t cl = 0; cl < nc; ++cl)
          {
            const uint32_t c = c0 + cl;
            for (int e = 0; e < 27; ++e)
              {
                const uint32_t t = etab[size_t(c) * 27 + e];
                if ((t & FLAG_BIT) != 0 || esize[e] == 0)
                  rec[size_t(cl) * 32 + e] = zero;
                else
                  {
                    const uint32_t idx       = t & INDEX_MASK;
                    rec[size_t(cl) * 32 + e] = (idx >= b0 && idx < b1) ? idx - b0 : foreign[*lmap.find(idx)].slot;
                  }
              }
            rec[size_t(cl) * 32 + 27] = uint32_t(mask ? mask[c] : 0) | (uint32_t(level[c]) << 16);
          }
        blob.assign(CB_HEADER_WORDS, 0u);
        blob[CB_N_CELLS]    = nc;
        blob[CB_N_OWNED]    = n_own;
        blob[CB_N_FP]       = n_f[0];
        blob[CB_N_FL]       = n_f[1];
        blob[CB_N_FG]       = n_f[2];
        blob[CB_N_DEP_LOW]  = uint32_t(dep_low.size());
        blob[CB_N_DEP_LATE] = uint32_t(dep_late.size());
        blob[CB_HANGING]    = hanging ? 1u : 0u;
        blob[CB_OFF_CELLREC] = uint32_t(blob.size());
        blob.insert(blob.end(), rec.begin(), rec.end());
        blob[CB_OFF_FSRC] = uint32_t(blob.size());
        {
          const size_t at = blob.size();
          blob.resize(at + n_f[0] + n_f[1], 0u);
          for (const Foreign &f : foreign)
            if (f.type < 2)
              {
                const uint32_t o   = f.type == 0 ? dep_low[f.dep] : dep_late[f.dep];
                const uint32_t ord = f.type == 0 ? f.dep : uint32_t(dep_low.size()) + f.dep;
                for (uint32_t j = 0; j < f.len; ++j)
                  blob[at + (f.slot - n_own) + j] = (ord << 16) | (f.idx + j - owned_begin[o]);
              }
        }
        uint32_t     blob_off = 0;
        const size_t key_size = blob.size();
        uint64_t     key_hash = 1469598103934665603ull;
        for (uint32_t w : blob)
          {
            key_hash ^= w;
            key_hash *= 1099511628211ull;
          }
        bool cached = false;
        {
          auto it = key_cache.find(key_hash);
          if (it != key_cache.end())
            for (uint32_t off : it->second)
              {
                // the cached blob starts with the same header/records/sources up to the two offsets not known yet
                const uint32_t *q = &plan.pool[off];
                bool            same = off + key_size <= plan.pool.size();
                for (size_t w = 0; same && w < key_size; ++w)
                  same = (w == CB_N_ROUNDS || w == CB_OFF_MERGE || w == CB_OFF_SLOTPOS || w == CB_TOTAL_WORDS) ? true : q[w] == blob[w];
                if (same)
                  {
                    blob_off = off;
                    cached   = true;
                    break;
                  }
              }
        }
        if (!cached)
          {
            // per-slot contribution lists: positions of the cell storage, in cell order
            ent_cnt.assign(n_slots + 1, 0u);
            for (int passno = 0; passno < 2; ++passno)
              {
                if (passno == 1)
                  {
                    ent_off.assign(n_slots + 1, 0u);
                    for (uint32_t s = 0; s < n_slots; ++s)
                      ent_off[s + 1] = ent_off[s] + ent_cnt[s];
                    ent_list.assign(ent_off[n_slots] + (ent_off[n_slots] & 1), 0);
                    std::fill(ent_cnt.begin(), ent_cnt.end(), 0u);
                  }
                for (uint32_t cl = 0; cl < nc; ++cl)
                  for (int kz = 0; kz < n; ++kz)
                    for (int j = 0; j < n; ++j)
                      for (int i = 0; i < n; ++i)
                        {
                          int e, off;
                          node_entity(p, i, j, kz, e, off);
                          if (rec[size_t(cl) * 32 + e] == zero)
                            continue;
                          const uint32_t s = rec[size_t(cl) * 32 + e] + uint32_t(off);
                          if (passno == 0)
                            ++ent_cnt[s];
                          else
                            ent_list[ent_off[s] + ent_cnt[s]++] = uint16_t(cl * CS + kz * PS + j * n + i);
                        }
              }
            for (uint32_t s = 0; s < n_slots; ++s)
              if (ent_cnt[s] > 255)
                throw std::runtime_error("chunk plan: more than 255 cell contributions to one DoF");
            if (ent_off[n_slots] >= (1u << 24))
              throw std::runtime_error("chunk plan: contribution list too long");
            // merge rounds: the copies p[0..m) of a DoF are summed as a binary tree, round r adds p[i + 2^r] to p[i] for
            // i = 0, 2^(r+1), ...; the sum ends up in p[0] (fixed order: reproducible)
            {
              uint32_t max_cnt = 1;
              for (uint32_t s = 0; s < n_slots; ++s)
                max_cnt = std::max(max_cnt, ent_cnt[s]);
              uint32_t n_rounds = 0;
              while ((1u << n_rounds) < max_cnt)
                ++n_rounds;
              std::vector<std::vector<uint32_t>> pairs(n_rounds);
              for (uint32_t s = 0; s < n_slots; ++s)
                for (uint32_t r = 0; (1u << r) < ent_cnt[s]; ++r)
                  for (uint32_t i = 0; i + (1u << r) < ent_cnt[s]; i += 2u << r)
                    pairs[r].push_back(uint32_t(ent_list[ent_off[s] + i]) | (uint32_t(ent_list[ent_off[s] + i + (1u << r)]) << 16));
              while (blob.size() % 4)
                blob.push_back(0u); // the epilogue tables [CB_OFF_MERGE, CB_TOTAL_WORDS) are fetched with one bulk copy
              blob[CB_N_ROUNDS]  = n_rounds;
              blob[CB_OFF_MERGE] = uint32_t(blob.size());
              uint32_t at = n_rounds + 1;
              for (uint32_t r = 0; r < n_rounds; ++r)
                {
                  blob.push_back(at);
                  at += uint32_t(pairs[r].size());
                }
              blob.push_back(at);
              for (uint32_t r = 0; r < n_rounds; ++r)
                blob.insert(blob.end(), pairs[r].begin(), pairs[r].end());
              blob[CB_OFF_SLOTPOS] = uint32_t(blob.size());
              for (uint32_t s = 0; s < n_slots; s += 2)
                {
                  const uint32_t p0 = ent_cnt[s] ? ent_list[ent_off[s]] : 0xffffu;
                  const uint32_t p1 = (s + 1 < n_slots && ent_cnt[s + 1]) ? ent_list[ent_off[s + 1]] : 0xffffu;
                  blob.push_back(p0 | (p1 << 16));
                }
            }
            while (blob.size() % 4)
              blob.push_back(0u);
            blob[CB_TOTAL_WORDS] = uint32_t(blob.size());
            blob_off             = blobs.add(blob);
            key_cache[key_hash].push_back(blob_off);
          }

        uint32_t *d       = &plan.desc[size_t(k) * CHUNK_DESC_WORDS];
        d[CD_BLOB]        = blob_off;
        d[CD_OWNED_BEGIN] = b0;
        d[CD_N_OWNED]     = n_own;
        if (side_off + n_f[0] >= (uint64_t(1) << 32))
          throw std::runtime_error("chunk plan: side buffer exceeds 2^32 entries");
        d[CD_SIDE_OFF]  = uint32_t(side_off);
        d[CD_GHOST_OFF] = uint32_t(plan.ghost_idx.size());
        src_deps_off[k] = uint32_t(src_deps_all.size());
        src_deps_all.insert(src_deps_all.end(), dep_low.begin(), dep_low.end());
        src_deps_all.insert(src_deps_all.end(), dep_late.begin(), dep_late.end());
        for (const Foreign &f : foreign)
          {
            if (f.type == 0)
              runs.push_back(PullRun{dep_low[f.dep], k, f.idx - owned_begin[dep_low[f.dep]], f.slot - n_own, f.len});
            else if (f.type == 2)
              for (uint32_t j = 0; j < f.len; ++j)
                plan.ghost_idx.push_back(f.idx + j);
          }
        side_off += n_f[0];
      }
    src_deps_off[K]      = uint32_t(src_deps_all.size());
    plan.side_size       = side_off;
    plan.n_compute_blobs = blobs.n_unique;

    // ---- pass 3: pull blobs (owner side of the side-buffer hand-over)
    std::vector<uint32_t> run_off(K + 2, 0);
    for (const PullRun &r : runs)
      ++run_off[r.owner + 1];
    for (uint32_t k = 0; k < K; ++k)
      run_off[k + 1] += run_off[k];
    std::vector<PullRun> sorted(runs.size());
    {
      std::vector<uint32_t> at(run_off.begin(), run_off.begin() + K);
      for (const PullRun &r : runs)
        sorted[at[r.owner]++] = r;
    }
    runs.clear();
    runs.shrink_to_fit();
    struct PullEntry
    {
      uint32_t slot, contributor, rel;
    };
    std::vector<PullEntry> entries;
    std::vector<uint32_t>  pull_deps;
    std::vector<uint32_t>  pull_deps_all, pull_deps_off(K + 1, 0);
    const uint32_t         before_pull = blobs.n_unique;
    for (uint32_t k = 0; k < K; ++k)
      {
        entries.clear();
        pull_deps.clear();
        for (uint32_t q = run_off[k]; q < run_off[k + 1]; ++q)
          for (uint32_t j = 0; j < sorted[q].len; ++j)
            entries.push_back(PullEntry{sorted[q].owner_slot + j, sorted[q].contributor, sorted[q].rel + j});
        std::sort(entries.begin(), entries.end(), [](const PullEntry &a, const PullEntry &b) {
          return a.slot != b.slot ? a.slot < b.slot : a.contributor < b.contributor;
        });
        blob.assign(PB_HEADER_WORDS, 0u);
        const uint32_t        n_own_k = plan.desc[size_t(k) * CHUNK_DESC_WORDS + CD_N_OWNED];
        std::vector<uint32_t> recw, listw;
        std::vector<uint16_t> opos(n_own_k + (n_own_k & 1), 0x7fffu);
        {
          const uint32_t *cb      = &plan.pool[plan.desc[size_t(k) * CHUNK_DESC_WORDS + CD_BLOB]];
          const uint32_t *slotpos = cb + cb[CB_OFF_SLOTPOS];
          for (uint32_t s = 0; s < n_own_k; ++s)
            {
              const uint32_t pos = (slotpos[s >> 1] >> ((s & 1) * 16)) & 0xffffu;
              if (pos != 0xffffu && pos >= 0x7fffu)
                throw std::runtime_error("chunk plan: cell storage position exceeds 15 bits");
              opos[s] = uint16_t(pos == 0xffffu ? 0x7fffu : pos);
            }
        }
        for (size_t q = 0; q < entries.size();)
          {
            size_t q1 = q;
            while (q1 < entries.size() && entries[q1].slot == entries[q].slot)
              ++q1;
            const uint32_t slot = entries[q].slot;
            if (slot >= n_own_k)
              throw std::runtime_error("chunk plan: pull target outside the owned range");
            opos[slot] |= 0x8000u;
            if (q1 - q >= 32 || slot >= 8192 || listw.size() >= 16384)
              throw std::runtime_error("chunk plan: pull record out of range");
            const uint32_t more_at = uint32_t(listw.size());
            uint32_t       first   = 0u;
            for (size_t t = q; t < q1; ++t)
              {
                size_t o = std::find(pull_deps.begin(), pull_deps.end(), entries[t].contributor) - pull_deps.begin();
                if (o == pull_deps.size())
                  pull_deps.push_back(entries[t].contributor);
                if (o >= 65536 || entries[t].rel >= 65536)
                  throw std::runtime_error("chunk plan: pull entry out of range");
                const uint32_t w = (uint32_t(o) << 16) | entries[t].rel;
                if (t == q)
                  first = w;
                else
                  listw.push_back(w);
              }
            recw.push_back(slot | (uint32_t(q1 - q) << 13) | (more_at << 18));
            recw.push_back(first);
            q = q1;
          }
        blob[PB_N_SLOTS]  = uint32_t(recw.size() / 2);
        blob[PB_N_DEPS]   = uint32_t(pull_deps.size());
        blob[PB_OFF_REC]  = uint32_t(blob.size());
        blob.insert(blob.end(), recw.begin(), recw.end());
        blob[PB_OFF_LIST] = uint32_t(blob.size());
        blob.insert(blob.end(), listw.begin(), listw.end());
        blob[PB_OFF_OPOS] = uint32_t(blob.size());
        for (size_t q = 0; q < opos.size(); q += 2)
          blob.push_back(uint32_t(opos[q]) | (uint32_t(opos[q + 1]) << 16));
        while (blob.size() % 4)
          blob.push_back(0u);
        blob[PB_TOTAL_WORDS] = uint32_t(blob.size());
        plan.desc[size_t(k) * CHUNK_DESC_WORDS + CD_PULL_BLOB] = blobs.add(blob);
        pull_deps_off[k] = uint32_t(pull_deps_all.size());
        pull_deps_all.insert(pull_deps_all.end(), pull_deps.begin(), pull_deps.end());
      }
    pull_deps_off[K]  = uint32_t(pull_deps_all.size());
    plan.n_pull_blobs = blobs.n_unique - before_pull;

    // dependencies: [src deps of all chunks | pull deps of all chunks]
    plan.deps = src_deps_all;
    plan.deps.insert(plan.deps.end(), pull_deps_all.begin(), pull_deps_all.end());
    for (uint32_t k = 0; k < K; ++k)
      {
        plan.desc[size_t(k) * CHUNK_DESC_WORDS + CD_SRC_DEPS]  = src_deps_off[k];
        plan.desc[size_t(k) * CHUNK_DESC_WORDS + CD_PULL_DEPS] = uint32_t(src_deps_all.size()) + pull_deps_off[k];
      }
    if (plan.deps.empty())
      plan.deps.push_back(0);
    if (plan.ghost_idx.empty())
      plan.ghost_idx.push_back(0);
    return plan;
  }

  // CPU walk through a plan exactly as device/chunk.cuh does it (chunks in descending order, DoF window, cell storage,
  // side buffers, pulls, late and ghost contributions), with the cell operator replaced by the diagonal map
  // y_e[q] = w(cell, q) x_e[q]; returns the largest deviation from the direct scatter-add over the entity table.
  double check_chunk_plan(const ChunkPlan &plan, long long n_cells, const uint32_t *etab, uint32_t n_owned, uint32_t n_ghost)
  {
    const int p = plan.degree, n = p + 1;
    int       PS, CS;
    chunk_layout(n, plan.value_bytes, PS, CS);
    const size_t        n_local = size_t(n_owned) + n_ghost;
    std::vector<double> src(n_local), ref(n_local, 0.0), out(n_local, 0.0), side(size_t(plan.side_size) + 1, 0.0);
    for (size_t i = 0; i < n_local; ++i)
      src[i] = 1.0 + double((i * 2654435761ull) % 1009) / 1009.0;
    const auto weight = [](uint32_t cell, int q) { return 1.0 + 0.001 * double((cell * 131u + uint32_t(q) * 7u) % 97u); };
    for (long long c = 0; c < n_cells; ++c)
      for (int k = 0; k < n; ++k)
        for (int j = 0; j < n; ++j)
          for (int i = 0; i < n; ++i)
            {
              int e, off;
              node_entity(p, i, j, k, e, off);
              const uint32_t t = etab[size_t(c) * 27 + e];
              if (t & FLAG_BIT)
                continue;
              const uint32_t g = (t & INDEX_MASK) + uint32_t(off);
              ref[g] += weight(uint32_t(c), (k * n + j) * n + i) * src[g];
            }
    // owned entries nobody touches are written as zero by the kernel; mark the output to see that every entry is written
    std::fill(out.begin(), out.begin() + n_owned, -777.0);
    std::vector<double>  xs, u_all(size_t(chunk_cells(p)) * CS);
    std::vector<uint8_t> stored(plan.n_chunks, 0), ready(plan.n_chunks, 0);
    uint32_t             cell0_of_chunk = 0;
    std::vector<uint32_t> first_cell(plan.n_chunks);
    for (uint32_t k = 0; k < plan.n_chunks; ++k)
      {
        first_cell[k] = cell0_of_chunk;
        cell0_of_chunk += plan.pool[plan.desc[size_t(k) * CHUNK_DESC_WORDS + CD_BLOB] + CB_N_CELLS];
      }
    if ((long long)cell0_of_chunk != n_cells)
      return 1e300;
    for (int c = int(plan.n_chunks) - 1; c >= 0; --c)
      {
        const uint32_t *d  = &plan.desc[size_t(c) * CHUNK_DESC_WORDS];
        const uint32_t *cb = &plan.pool[d[CD_BLOB]];
        const uint32_t  nc = cb[CB_N_CELLS], n_own = cb[CB_N_OWNED], n_fp = cb[CB_N_FP], n_fl = cb[CB_N_FL], n_fg = cb[CB_N_FG];
        const uint32_t  n_slots = n_own + n_fp + n_fl + n_fg, ob = d[CD_OWNED_BEGIN];
        const uint32_t *src_deps = &plan.deps[d[CD_SRC_DEPS]];
        const uint32_t *fsrc = cb + cb[CB_OFF_FSRC], *rec = cb + cb[CB_OFF_CELLREC], *merge = cb + cb[CB_OFF_MERGE];
        const uint32_t *slotpos = cb + cb[CB_OFF_SLOTPOS];
        if (n_own != d[CD_N_OWNED] || n_slots + CHUNK_ZERO_SLOTS > uint32_t(chunk_slot_capacity(p)))
          return 2e300;
        xs.assign(n_slots + CHUNK_ZERO_SLOTS, 0.0);
        for (uint32_t s = 0; s < n_own; ++s)
          xs[s] = src[ob + s];
        for (uint32_t t = 0; t < n_fp + n_fl; ++t)
          {
            const uint32_t dep = src_deps[fsrc[t] >> 16];
            xs[n_own + t]      = src[plan.desc[size_t(dep) * CHUNK_DESC_WORDS + CD_OWNED_BEGIN] + (fsrc[t] & 0xffffu)];
          }
        for (uint32_t t = 0; t < n_fg; ++t)
          xs[n_own + n_fp + n_fl + t] = src[plan.ghost_idx[d[CD_GHOST_OFF] + t]];
        std::fill(u_all.begin(), u_all.end(), 1e200); // poison: positions no cell writes must never be read
        for (uint32_t cl = 0; cl < nc; ++cl)
          for (int k = 0; k < n; ++k)
            for (int j = 0; j < n; ++j)
              for (int i = 0; i < n; ++i)
                {
                  int e, off;
                  node_entity(p, i, j, k, e, off);
                  u_all[size_t(cl) * CS + k * PS + j * n + i] =
                    weight(first_cell[c] + cl, (k * n + j) * n + i) * xs[rec[size_t(cl) * 32 + e] + uint32_t(off)];
                }
        for (uint32_t r = 0; r < cb[CB_N_ROUNDS]; ++r)
          {
            std::vector<uint8_t> touched(u_all.size(), 0);
            for (uint32_t q = merge[r]; q < merge[r + 1]; ++q)
              {
                const uint32_t dst = merge[q] & 0xffffu, src = merge[q] >> 16;
                if (touched[dst] || touched[src])
                  return 7e300; // a position twice in one round: the parallel merge would race
                touched[dst] = touched[src] = 1;
                u_all[dst] += u_all[src];
              }
          }
        const auto slot_sum = [&](uint32_t s) {
          const uint32_t pos = (slotpos[s >> 1] >> ((s & 1) * 16)) & 0xffffu;
          return pos == 0xffffu ? 0.0 : u_all[pos];
        };
        for (uint32_t t = 0; t < n_fp; ++t)
          side[size_t(d[CD_SIDE_OFF]) + t] = slot_sum(n_own + t);
        for (uint32_t t = 0; t < n_fg; ++t)
          out[plan.ghost_idx[d[CD_GHOST_OFF] + t]] += slot_sum(n_own + n_fp + n_fl + t);
        ready[c] = 1;
        const uint32_t *pb = &plan.pool[d[CD_PULL_BLOB]];
        const uint32_t *pull_deps = &plan.deps[d[CD_PULL_DEPS]], *prec = pb + pb[PB_OFF_REC], *plist = pb + pb[PB_OFF_LIST];
        for (uint32_t t = 0; t < pb[PB_N_DEPS]; ++t)
          if (pull_deps[t] <= uint32_t(c) || !ready[pull_deps[t]])
            return 3e300; // a dependency that was not started before this chunk
        const uint32_t *oposw = pb + pb[PB_OFF_OPOS];
        std::vector<double> pulled(n_own, 0.0);
        for (uint32_t q = 0; q < pb[PB_N_SLOTS]; ++q)
          {
            const uint32_t w0 = prec[2 * q], slot = w0 & 0x1fffu;
            double         v  = 0;
            for (uint32_t j = 0; j < ((w0 >> 13) & 31u); ++j)
              {
                const uint32_t e = j == 0 ? prec[2 * q + 1] : plist[(w0 >> 18) + j - 1];
                v += side[size_t(plan.desc[size_t(pull_deps[e >> 16]) * CHUNK_DESC_WORDS + CD_SIDE_OFF]) + (e & 0xffffu)];
              }
            if (!(((oposw[slot >> 1] >> ((slot & 1) * 16)) & 0x8000u)))
              return 6e300;
            pulled[slot] = v;
          }
        for (uint32_t s = 0; s < n_own; ++s)
          {
            const uint32_t w = (oposw[s >> 1] >> ((s & 1) * 16)) & 0xffffu, sp = (slotpos[s >> 1] >> ((s & 1) * 16)) & 0xffffu;
            if ((w & 0x7fffu) != (sp == 0xffffu ? 0x7fffu : sp) || (!(w & 0x8000u) && pulled[s] != 0.0))
              return 8e300;
          }
        for (uint32_t s = 0; s < n_own; ++s)
          out[ob + s] = slot_sum(s) + pulled[s];
        stored[c] = 1;
        for (uint32_t t = 0; t < n_fl; ++t)
          {
            const uint32_t w = fsrc[n_fp + t], dep = src_deps[w >> 16];
            if (dep <= uint32_t(c) || !stored[dep] || (w >> 16) < cb[CB_N_DEP_LOW])
              return 4e300;
            out[plan.desc[size_t(dep) * CHUNK_DESC_WORDS + CD_OWNED_BEGIN] + (w & 0xffffu)] += slot_sum(n_own + n_fp + t);
          }
        for (uint32_t t = 0; t < n_fp; ++t)
          if (src_deps[fsrc[t] >> 16] >= uint32_t(c))
            return 5e300;
      }
    double err = 0;
    for (size_t i = 0; i < n_local; ++i)
      err = std::max(err, std::fabs(out[i] - ref[i]) / std::max(1.0, std::fabs(ref[i])));
    return err;
  }
} // namespace mgb
