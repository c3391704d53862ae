// Ghost exchange over NVLink peer memory: update_ghost_values / compress(add) of
// LinearAlgebra::distributed::Vector as two small kernels per exchange instead of a pack kernel,
// an NCCL send/recv group and an unpack kernel.
//
// Every rank maps the staging blocks of its peers (cudaIpc, host/p2p.cc).  One exchange of epoch e:
//   k_halo_send: waits until every destination has consumed epoch e-2 (same slot of the double-buffered staging
//                area), packs the values and stores them straight into the DESTINATION's staging block over
//                NVLink; the last thread block to finish publishes "ready = e" in the destination's memory
//                (st.release.sys after __threadfence_system).
//   k_halo_recv: spins (ld.acquire.sys) until every source has published e, copies the staged values into the
//                ghost tail (update_ghost_values) or adds them to the owned entries (compress), then acknowledges
//                "consumed = e" in the source's memory and advances the local epoch.
// The epoch counter lives in device memory, so the pair can be captured into a CUDA graph and replayed.
// On every GPU send(e) precedes recv(e) in stream order and send never waits for anything younger than e-2, so the
// protocol cannot deadlock.
#include "device.h"

#include <algorithm>

namespace mgb
{
  namespace
  {
    constexpr int HALO_BLOCK = 256;

    __device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p)
    {
      unsigned long long v;
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
      return v;
    }
    __device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v)
    {
      asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
    }

    // bounded spin: a peer that never answers (a rank died, mismatched call sequences) must not hang the GPU
    constexpr long long SPIN_LIMIT_CYCLES = 20000000000ll; // ~10 s
    template <class Pred> __device__ __forceinline__ void spin_until(Pred ready)
    {
      const long long t0 = clock64();
      while (!ready())
        if (clock64() - t0 > SPIN_LIMIT_CYCLES)
          __trap();
    }

    // peer index of element i in the concatenated per-peer ranges
    __device__ __forceinline__ int find_peer(const HaloSide &s, const uint32_t i)
    {
      int p = 0;
      while (p + 1 < s.n && i >= s.off[p + 1])
        ++p;
      return p;
    }

    // GHOST = true:  owner -> ghoster, values gathered by import_indices
    // GHOST = false: ghoster -> owner, the ghost tail is sent (and zeroed: compress ends with zero_out_ghost_values)
    template <typename T, bool GHOST>
    __global__ void __launch_bounds__(HALO_BLOCK)
      k_halo_send(const __grid_constant__ HaloDir hd, T *__restrict__ vec, const uint32_t n_owned, const uint32_t *__restrict__ import_indices)
    {
      const HaloSide          &s = hd.send;
      const unsigned long long e = *hd.epoch + 1;
      __shared__ bool          last;
      if (threadIdx.x < s.n) // while the destination still reads the slot of epoch e-2
        spin_until([&] { return ld_acquire_sys(s.local_flag[threadIdx.x]) + 2 >= e; });
      __syncthreads();
      const uint32_t total = s.off[s.n];
      for (uint32_t i = blockIdx.x * HALO_BLOCK + threadIdx.x; i < total; i += gridDim.x * HALO_BLOCK)
        {
          const int p = find_peer(s, i);
          T         v;
          if (GHOST)
            v = vec[import_indices[i]];
          else
            {
              v                = vec[n_owned + i];
              vec[n_owned + i] = T(0);
            }
          char *dst = s.remote_stage[p] + (e & 1) * s.remote_slot_bytes[p] + size_t(i - s.off[p]) * 8;
          *reinterpret_cast<T *>(dst) = v;
        }
      // the block's stores are ordered before thread 0's system-scope fence by the barrier (fences are cumulative)
      __syncthreads();
      if (threadIdx.x == 0)
        {
          __threadfence_system();
          last = atomicAdd(hd.ticket, 1u) == gridDim.x - 1;
        }
      __syncthreads();
      if (last)
        {
          __threadfence_system();
          if (threadIdx.x < s.n)
            st_release_sys(s.remote_flag[threadIdx.x], e);
          if (threadIdx.x == 0)
            *hd.ticket = 0;
        }
    }

    template <typename T, bool GHOST>
    __global__ void __launch_bounds__(HALO_BLOCK)
      k_halo_recv(const __grid_constant__ HaloDir hd, T *__restrict__ vec, const uint32_t n_owned, const uint32_t *__restrict__ import_indices)
    {
      const HaloSide          &r = hd.recv;
      const unsigned long long e = *hd.epoch + 1;
      __shared__ bool          last;
      if (threadIdx.x < r.n) // until the source has published this epoch
        spin_until([&] { return ld_acquire_sys(r.local_flag[threadIdx.x]) >= e; });
      __syncthreads();
      const uint32_t total = r.off[r.n];
      const char    *stage = hd.stage + (e & 1) * hd.slot_bytes;
      for (uint32_t i = blockIdx.x * HALO_BLOCK + threadIdx.x; i < total; i += gridDim.x * HALO_BLOCK)
        {
          const T v = __ldcg(reinterpret_cast<const T *>(stage + size_t(i) * 8)); // written by a peer: bypass L1
          if (GHOST)
            vec[n_owned + i] = v;
          else
            atomicAdd(vec + import_indices[i], v); // an owned entry may be imported by several ranks
        }
      __syncthreads();
      if (threadIdx.x == 0)
        {
          __threadfence();
          last = atomicAdd(hd.ticket + 1, 1u) == gridDim.x - 1;
        }
      __syncthreads();
      if (last)
        {
          if (threadIdx.x < r.n)
            st_release_sys(r.remote_flag[threadIdx.x], e);
          if (threadIdx.x == 0)
            {
              *hd.epoch     = e;
              hd.ticket[1] = 0;
            }
        }
    }

    __device__ __forceinline__ void st_relaxed_sys(unsigned long long *p, unsigned long long v)
    {
      asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
    }

    // The whole exchange in ONE kernel: pack + peer stores, one release per destination, poll for the sources' flags, unpack,
    // relaxed acknowledgement.  Against the send/recv pair this removes a kernel boundary, one of the two system-scope
    // fences of the send side and the fence of the acknowledging store (it publishes no data).  Every block first sends and
    // then receives, so the grid must be co-resident (a block that has not started cannot contribute to the flag the peer
    // waits for, while the resident blocks wait for the peer): the launcher caps it at one block per SM and only uses this
    // kernel when nothing else shares the GPU with it (not for the split cell loop on a second stream).
    template <typename T, bool GHOST>
    __global__ void __launch_bounds__(HALO_BLOCK)
      k_halo_exchange(const __grid_constant__ HaloDir hd, T *__restrict__ vec, const uint32_t n_owned, const uint32_t *__restrict__ import_indices)
    {
      const HaloSide          &s = hd.send, &r = hd.recv;
      const unsigned long long e = *hd.epoch + 1;
      __shared__ bool          last;
      // ---- send: the destinations must have consumed epoch e-2 (same slot of the double-buffered staging area)
      if (threadIdx.x < s.n)
        spin_until([&] { return ld_acquire_sys(s.local_flag[threadIdx.x]) + 2 >= e; });
      __syncthreads();
      const uint32_t n_send = s.off[s.n];
      for (uint32_t i = blockIdx.x * HALO_BLOCK + threadIdx.x; i < n_send; i += gridDim.x * HALO_BLOCK)
        {
          const int p = find_peer(s, i);
          T         v;
          if (GHOST)
            v = vec[import_indices[i]];
          else
            {
              v                = vec[n_owned + i];
              vec[n_owned + i] = T(0);
            }
          char *dst = s.remote_stage[p] + (e & 1) * s.remote_slot_bytes[p] + size_t(i - s.off[p]) * 8;
          *reinterpret_cast<T *>(dst) = v;
        }
      __syncthreads(); // orders the block's peer stores before the releasing store below (cumulativity)
      if (gridDim.x == 1)
        {
          if (threadIdx.x < s.n)
            st_release_sys(s.remote_flag[threadIdx.x], e);
        }
      else
        {
          if (threadIdx.x == 0)
            {
              __threadfence_system();
              last = atomicAdd(hd.ticket, 1u) == gridDim.x - 1;
            }
          __syncthreads();
          if (last)
            {
              if (threadIdx.x < s.n)
                st_release_sys(s.remote_flag[threadIdx.x], e);
              if (threadIdx.x == 0)
                *hd.ticket = 0;
            }
        }
      // ---- receive: until every source has published this epoch
      if (threadIdx.x < r.n)
        spin_until([&] { return ld_acquire_sys(r.local_flag[threadIdx.x]) >= e; });
      __syncthreads();
      const uint32_t n_recv = r.off[r.n];
      const char    *stage  = hd.stage + (e & 1) * hd.slot_bytes;
      for (uint32_t i = blockIdx.x * HALO_BLOCK + threadIdx.x; i < n_recv; i += gridDim.x * HALO_BLOCK)
        {
          const T v = __ldcg(reinterpret_cast<const T *>(stage + size_t(i) * 8)); // written by a peer: bypass L1
          if (GHOST)
            vec[n_owned + i] = v;
          else
            atomicAdd(vec + import_indices[i], v); // an owned entry may be imported by several ranks
        }
      __syncthreads(); // every load of the staging slot has returned its value
      if (gridDim.x > 1)
        {
          if (threadIdx.x == 0)
            {
              __threadfence();
              last = atomicAdd(hd.ticket + 1, 1u) == gridDim.x - 1;
            }
          __syncthreads();
          if (!last)
            return;
        }
      if (threadIdx.x < r.n)
        st_relaxed_sys(r.remote_flag[threadIdx.x], e); // "consumed": publishes no data, only has to follow the loads above
      if (threadIdx.x == 0)
        {
          *hd.epoch    = e;
          hd.ticket[1] = 0;
        }
    }

    inline unsigned halo_grid(uint32_t total)
    {
      const unsigned b = (total + HALO_BLOCK * 4 - 1) / (HALO_BLOCK * 4);
      return b < 1 ? 1u : (b > 296u ? 296u : b);
    }
  } // namespace

  template <typename T>
  void halo_exchange(const HaloDir &hd, bool ghost_direction, T *vec, uint32_t n_owned, const uint32_t *import_indices, cudaStream_t s,
                     bool fused)
  {
    const unsigned gs = halo_grid(hd.send.off[hd.send.n]), gr = halo_grid(hd.recv.off[hd.recv.n]);
    if (fused)
      {
        const unsigned g = std::min(148u, std::max(gs, gr)); // co-resident: at most one block per SM
        if (ghost_direction)
          k_halo_exchange<T, true><<<g, HALO_BLOCK, 0, s>>>(hd, vec, n_owned, import_indices);
        else
          k_halo_exchange<T, false><<<g, HALO_BLOCK, 0, s>>>(hd, vec, n_owned, import_indices);
        MGB_CUDA(cudaGetLastError());
        return;
      }
    if (ghost_direction)
      {
        k_halo_send<T, true><<<gs, HALO_BLOCK, 0, s>>>(hd, vec, n_owned, import_indices);
        k_halo_recv<T, true><<<gr, HALO_BLOCK, 0, s>>>(hd, vec, n_owned, import_indices);
      }
    else
      {
        k_halo_send<T, false><<<gs, HALO_BLOCK, 0, s>>>(hd, vec, n_owned, import_indices);
        k_halo_recv<T, false><<<gr, HALO_BLOCK, 0, s>>>(hd, vec, n_owned, import_indices);
      }
    MGB_CUDA(cudaGetLastError());
  }

  template void halo_exchange<double>(const HaloDir &, bool, double *, uint32_t, const uint32_t *, cudaStream_t, bool);
  template void halo_exchange<float>(const HaloDir &, bool, float *, uint32_t, const uint32_t *, cudaStream_t, bool);
} // namespace mgb
