// Peer-memory backing of the vector partitioners: every rank allocates one segment holding, per partitioner, the
// flags and the double-buffered staging areas of the NVLink ghost exchange (device/halo.cu), publishes it with
// cudaIpc and maps the segments of all other ranks.  Collective; called once the partitioners of a problem exist
// (Problem::build, mgb_comm_enable_peer_exchange).  Partitioners that are not backed keep using the NCCL
// send/recv path, so a failure here (no peer access, more than HALO_MAX_PEERS ranks) only costs speed.
#include "comm.h"
#include "solver.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace mgb
{
  namespace
  {
    constexpr size_t HDR_BYTES = 1024; // 4 flag arrays [HALO_MAX_PEERS] u64 | 2 epochs | 4 tickets
    constexpr uint32_t NONE    = 0xffffffffu;

    struct Table // what the other ranks need to know about one of my partitioner blocks
    {
      uint64_t block_off;
      uint32_t n_ghost, n_import;
      uint32_t goff[HALO_MAX_PEERS]; // element offset in my ghost staging slot of the data owned by rank s
      uint32_t ioff[HALO_MAX_PEERS]; // element offset in my compress staging slot of the data sent by rank s
    };
    size_t align256(size_t x) { return (x + 255) / 256 * 256; }
    size_t ghost_bytes(const Table &t) { return align256(2 * size_t(t.n_ghost) * 8); }
    size_t block_bytes(const Table &t) { return HDR_BYTES + ghost_bytes(t) + align256(2 * size_t(t.n_import) * 8); }

    struct Segment
    {
      char               *base = nullptr;
      std::vector<char *> peer; // mapped bases, peer[me] = base
      int                 me = 0;
      ~Segment()
      {
        for (size_t r = 0; r < peer.size(); ++r)
          if (int(r) != me && peer[r])
            cudaIpcCloseMemHandle(peer[r]);
        if (base)
          cudaFree(base);
      }
    };

    // segments stay mapped until the communicator is finalised: a peer's last acknowledgement may still be in flight
    // when the objects of this rank are destroyed
    std::vector<std::shared_ptr<Segment>> &segments()
    {
      static std::vector<std::shared_ptr<Segment>> s;
      return s;
    }

    std::vector<DevicePartitioner *> &registry()
    {
      static std::vector<DevicePartitioner *> r;
      return r;
    }
  } // namespace

  void p2p_register(DevicePartitioner *p) { registry().push_back(p); }
  void p2p_unregister(DevicePartitioner *p)
  {
    auto &r = registry();
    r.erase(std::remove(r.begin(), r.end(), p), r.end());
  }

  bool p2p_back_partitioners()
  {
    Context &c = ctx();
    auto    &reg = registry();
    if (!c.comm || c.comm->size() < 2)
      {
        reg.clear();
        return false;
      }
    const char *env = std::getenv("MGB_P2P");
    const int   R = c.comm->size(), me = c.comm->rank();
    // On by default (MGB_P2P=0 keeps pack + grouped ncclSend/ncclRecv): the fused single-kernel exchange of device/halo.cu
    // costs 29-33 us per update_ghost_values + compress pair on 8 B200 against 46-67 us for NCCL, 19.6 against 21.7 ms per
    // solve at L=7 (profiles/r2n_*); the send/recv kernel pair of round 1 (15.5 us per exchange) was not faster than NCCL.
    // Without peer access between the GPUs the NCCL path stays in use.
    const bool  enabled = !(env && env[0] == '0') && R <= HALO_MAX_PEERS;
    // every rank must back the same number of partitioners (they are created in the same order everywhere)
    std::vector<long long> counts(R);
    const long long        mine = enabled ? (long long)reg.size() : -1;
    c.comm->allgather_host(&mine, counts.data(), sizeof(long long), c.stream);
    for (int r = 0; r < R; ++r)
      if (counts[r] != mine || mine <= 0)
        {
          reg.clear();
          return false;
        }
    const size_t       n = reg.size();
    std::vector<Table> tabs(n);
    size_t             total = 0;
    for (size_t k = 0; k < n; ++k)
      {
        const Partitioner &p = reg[k]->host;
        Table             &t = tabs[k];
        t.block_off = total;
        t.n_ghost   = p.n_ghost;
        t.n_import  = (uint32_t)p.import_indices.size();
        std::fill(t.goff, t.goff + HALO_MAX_PEERS, NONE);
        std::fill(t.ioff, t.ioff + HALO_MAX_PEERS, NONE);
        for (size_t j = 0; j < p.ghost_ranks.size(); ++j)
          t.goff[p.ghost_ranks[j]] = p.ghost_offsets[j];
        for (size_t i = 0; i < p.import_ranks.size(); ++i)
          t.ioff[p.import_ranks[i]] = p.import_offsets[i];
        total += block_bytes(t);
      }
    auto seg = std::make_shared<Segment>();
    seg->me  = me;
    seg->peer.assign(R, nullptr);
    int ok = 1;
    cudaIpcMemHandle_t handle;
    std::memset(&handle, 0, sizeof(handle));
    if (cudaMalloc((void **)&seg->base, std::max<size_t>(total, 256)) != cudaSuccess)
      ok = 0;
    if (ok && cudaMemset(seg->base, 0, std::max<size_t>(total, 256)) != cudaSuccess)
      ok = 0;
    if (ok && cudaIpcGetMemHandle(&handle, seg->base) != cudaSuccess)
      ok = 0;
    cudaDeviceSynchronize();
    std::vector<cudaIpcMemHandle_t> handles(R);
    c.comm->allgather_host(&handle, handles.data(), sizeof(handle), c.stream);
    std::vector<int> oks(R);
    c.comm->allgather_host(&ok, oks.data(), sizeof(int), c.stream);
    ok = *std::min_element(oks.begin(), oks.end());
    if (ok)
      for (int r = 0; r < R; ++r)
        {
          if (r == me)
            {
              seg->peer[r] = seg->base;
              continue;
            }
          void *ptr = nullptr;
          if (cudaIpcOpenMemHandle(&ptr, handles[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
            {
              cudaGetLastError();
              ok = 0;
              break;
            }
          seg->peer[r] = static_cast<char *>(ptr);
        }
    c.comm->allgather_host(&ok, oks.data(), sizeof(int), c.stream);
    ok = *std::min_element(oks.begin(), oks.end());
    std::vector<Table> all(n * size_t(R));
    c.comm->allgather_host(tabs.data(), all.data(), n * sizeof(Table), c.stream);
    if (!ok)
      {
        cudaGetLastError();
        reg.clear();
        return false;
      }
    const auto flag = [&](int rank, const Table &t, int array, int index) {
      return reinterpret_cast<unsigned long long *>(seg->peer[rank] + t.block_off) + array * HALO_MAX_PEERS + index;
    };
    for (size_t k = 0; k < n; ++k)
      {
        DevicePartitioner *dp = reg[k];
        const Partitioner &p  = dp->host;
        const Table       &t  = tabs[k];
        auto               b  = std::make_unique<P2PBacking>();
        b->segment            = seg;
        char *blk             = seg->base + t.block_off;
        auto *epochs          = reinterpret_cast<unsigned long long *>(blk + 4 * HALO_MAX_PEERS * 8);
        auto *tickets         = reinterpret_cast<unsigned *>(blk + 4 * HALO_MAX_PEERS * 8 + 16);
        HaloDir &g = b->ghost, &cp = b->compress;
        g.stage       = blk + HDR_BYTES;
        g.slot_bytes  = size_t(t.n_ghost) * 8;
        g.epoch       = epochs;
        g.ticket      = tickets;
        cp.stage      = blk + HDR_BYTES + ghost_bytes(t);
        cp.slot_bytes = size_t(t.n_import) * 8;
        cp.epoch      = epochs + 1;
        cp.ticket     = tickets + 2;
        // owner -> ghoster: I send to the ranks importing from me, I receive from the owners of my ghosts
        g.send.n = (int)p.import_ranks.size();
        for (int i = 0; i < g.send.n; ++i)
          {
            const int    q  = p.import_ranks[i];
            const Table &tq = all[size_t(q) * n + k];
            if (tq.goff[me] == NONE)
              throw std::runtime_error("p2p: inconsistent partitioners (import without matching ghost range)");
            g.send.off[i]               = p.import_offsets[i];
            g.send.remote_stage[i]      = seg->peer[q] + tq.block_off + HDR_BYTES + size_t(tq.goff[me]) * 8;
            g.send.remote_slot_bytes[i] = size_t(tq.n_ghost) * 8;
            g.send.remote_flag[i]       = flag(q, tq, 0, me); // q's ready_g[me]
            g.send.local_flag[i]        = flag(me, t, 1, q);  // my ack_g[q]
          }
        g.send.off[g.send.n] = t.n_import;
        g.recv.n             = (int)p.ghost_ranks.size();
        for (int j = 0; j < g.recv.n; ++j)
          {
            const int    r  = p.ghost_ranks[j];
            const Table &tr = all[size_t(r) * n + k];
            g.recv.off[j]         = p.ghost_offsets[j];
            g.recv.local_flag[j]  = flag(me, t, 0, r);  // my ready_g[r]
            g.recv.remote_flag[j] = flag(r, tr, 1, me); // r's ack_g[me]
          }
        g.recv.off[g.recv.n] = t.n_ghost;
        // ghoster -> owner (compress): I send my ghost tail to the owners, I receive from the ranks importing from me
        cp.send.n = (int)p.ghost_ranks.size();
        for (int j = 0; j < cp.send.n; ++j)
          {
            const int    r  = p.ghost_ranks[j];
            const Table &tr = all[size_t(r) * n + k];
            if (tr.ioff[me] == NONE)
              throw std::runtime_error("p2p: inconsistent partitioners (ghost range without matching import)");
            cp.send.off[j]               = p.ghost_offsets[j];
            cp.send.remote_stage[j]      = seg->peer[r] + tr.block_off + HDR_BYTES + ghost_bytes(tr) + size_t(tr.ioff[me]) * 8;
            cp.send.remote_slot_bytes[j] = size_t(tr.n_import) * 8;
            cp.send.remote_flag[j]       = flag(r, tr, 2, me); // r's ready_c[me]
            cp.send.local_flag[j]        = flag(me, t, 3, r);  // my ack_c[r]
          }
        cp.send.off[cp.send.n] = t.n_ghost;
        cp.recv.n              = (int)p.import_ranks.size();
        for (int i = 0; i < cp.recv.n; ++i)
          {
            const int    q  = p.import_ranks[i];
            const Table &tq = all[size_t(q) * n + k];
            cp.recv.off[i]         = p.import_offsets[i];
            cp.recv.local_flag[i]  = flag(me, t, 2, q);  // my ready_c[q]
            cp.recv.remote_flag[i] = flag(q, tq, 3, me); // q's ack_c[me]
          }
        cp.recv.off[cp.recv.n] = t.n_import;
        dp->p2p                = std::move(b);
      }
    segments().push_back(seg);
    if (me == 0 && std::getenv("MGB_VERBOSE"))
      std::fprintf(stderr, "mgb200: NVLink peer-memory ghost exchange enabled for %zu partitioners (%zu bytes of staging on rank 0)\n", n, total);
    reg.clear();
    return true;
  }

  void p2p_release_all()
  {
    Context &c = ctx();
    cudaDeviceSynchronize();
    if (c.comm && c.comm->size() > 1 && !segments().empty())
      {
        int              one = 1; // barrier: every rank has drained its exchange kernels
        std::vector<int> all(c.comm->size());
        c.comm->allgather_host(&one, all.data(), sizeof(int), c.stream);
      }
    segments().clear();
  }
} // namespace mgb
