// Chunked cell loop: the Cartesian Laplace cell kernel of operator.cuh (same fast-diagonalised sweeps, same hanging-node
// interpolation) reorganised around the chunk plan of host/chunks.h.
//
// One thread block takes chunks of consecutive cells from a ticket counter, highest chunk first.  Per chunk:
//   load     the chunk's owned DoF range by ONE bulk copy (cp.async.bulk -> UBLKCP, completion on an mbarrier), the
//            few foreign DoFs by ordinary loads, the cell records from the (deduplicated, L2-resident) table blob;
//   compute  every warp applies A_e to its cells out of the shared DoF window and leaves h A_e u_e in the cell storage;
//   hand-off sums of the foreign DoFs go to the chunk's side buffer (plain stores), then the chunk's flag is released;
//   finish   after the flags of the contributing chunks (all started earlier) are seen, their side-buffer entries are
//            added, every owned DoF is summed in a fixed order and written ONCE with plain coalesced stores, through the
//            epilogue: dst = A x, dst = b - A x, or the Chebyshev update x_new = (1+f1) x - f1 x_old + f2 D^-1 (b - A x).
// No atomics on the owned range, no zeroing of dst, no read-modify-write of dst; bitwise reproducible on meshes without
// "late" contributions (host/chunks.h).
#pragma once
#include "../host/chunks.h"
#include "operator.cuh"

namespace mgb
{
  constexpr int CHUNK_MAX_DEPS = 96;

  template <int P> struct ChunkCfg
  {
    static constexpr int n   = P + 1;
    static constexpr int CW  = 32 / n;
    static constexpr int NC  = P == 1 ? 256 : P <= 3 ? 64 : P <= 5 ? 32 : 16;                 // cells per chunk (whole octets)
    static constexpr int NB  = P == 1 ? 2 : 1;                                              // batches of CW cells per warp
    static constexpr int W   = (NC + CW * NB - 1) / (CW * NB);                              // warps per block
    static constexpr int CAP = P == 1 ? 2112 : P == 2 ? 1792 : P == 3 ? 3072 : P == 4 ? 2880 : 5632; // slots of the DoF window
    static constexpr int MINB = P == 4 ? 3 : 2;                                             // resident blocks per SM
    static constexpr int THREADS = W * 32;
  };

  template <int P, typename T> struct ChunkSmem
  {
    static constexpr int    CS      = SlabLayout<P + 1, sizeof(T)>::CS;
    static constexpr size_t u_bytes = (size_t(ChunkCfg<P>::NC) * CS * sizeof(T) + 15) / 16 * 16;
    static constexpr size_t x_bytes = (size_t(ChunkCfg<P>::CAP) * sizeof(T) + 15) / 16 * 16;
    static constexpr size_t r_bytes = size_t(ChunkCfg<P>::NC) * 32 * 4;
    static constexpr size_t d_bytes = size_t(CHUNK_MAX_DEPS) * 4 * 2;
    static constexpr size_t total   = u_bytes + x_bytes + r_bytes + d_bytes + 32;
  };

  __device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
  __device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t *p)
  {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
  }
  __device__ __forceinline__ void st_release_u32(uint32_t *p, uint32_t v)
  {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
  }
  // wait until *p == want; traps after ~4 s so that a broken dependency cannot hang the device
  __device__ __forceinline__ void spin_until(const uint32_t *p, const uint32_t want)
  {
    if (ld_acquire_u32(p) == want)
      return;
    const long long t0 = clock64();
    while (ld_acquire_u32(p) != want)
      {
        __nanosleep(64);
        if (clock64() - t0 > 8000000000ll)
          __trap();
      }
  }

  template <typename T> struct ChunkArgs
  {
    uint32_t        n_chunks;
    const uint32_t *desc, *pool, *deps, *ghost_idx;
    T              *side;
    uint32_t       *flags; // [n_chunks] ready | [n_chunks] stored
    uint32_t       *state; // ticket, epoch, finished blocks
    double          root_length;
    const T        *src;
    T              *out;
    // epilogue: 0: out = v; 1: out = b - v; 2: out = (1+f1) x + f2 dinv (b - v); 3: ... - f1 xold   (v = A x, x = src)
    int             mode;
    const T        *b, *dinv, *xold;
    T               f1, f2;
  };

  template <int P, typename T>
  __global__ void __launch_bounds__(ChunkCfg<P>::THREADS, ChunkCfg<P>::MINB) k_laplace_chunks(const ChunkArgs<T> a)
  {
    using Cfg = ChunkCfg<P>;
    constexpr int n = P + 1, n2 = n * n;
    constexpr int CW = Cfg::CW, NB = Cfg::NB, THREADS = Cfg::THREADS;
    constexpr int PS = SlabLayout<n, sizeof(T)>::PS, CS = SlabLayout<n, sizeof(T)>::CS;
    constexpr int NR = (CW * n2 + 31) / 32;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T        *u_all  = reinterpret_cast<T *>(smem_raw);
    T        *xs     = reinterpret_cast<T *>(smem_raw + ChunkSmem<P, T>::u_bytes);
    uint32_t *rec    = reinterpret_cast<uint32_t *>(smem_raw + ChunkSmem<P, T>::u_bytes + ChunkSmem<P, T>::x_bytes);
    uint32_t *s_dep  = rec + Cfg::NC * 32;          // owned_begin of the src deps
    uint32_t *s_side = s_dep + CHUNK_MAX_DEPS;      // side-buffer offset of the pull deps
    uint64_t *mbar   = reinterpret_cast<uint64_t *>(s_side + CHUNK_MAX_DEPS);
    int      *s_chunk = reinterpret_cast<int *>(mbar + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t epoch = *reinterpret_cast<volatile uint32_t *>(a.state + 1) + 1u;
    uint32_t *ready = a.flags, *stored = a.flags + a.n_chunks;
    if (tid == 0)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(mbar)) : "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    unsigned parity = 0;

    // static per-lane description of the columns of each round (operator.cuh)
    uint32_t col[NR];
#pragma unroll
    for (int r = 0; r < NR; ++r)
      {
        const int c  = r * 32 + lane;
        const int cl = c / n2, ij = c - cl * n2, i = ij % n, j = ij / n;
        const int tx = i == 0 ? 0 : (i == P ? 2 : 1), ty = j == 0 ? 0 : (j == P ? 2 : 1);
        const int sx = tx == 1 ? P - 1 : 1;
        col[r] = uint32_t(c < CW * n2 ? cl : 31) | uint32_t(tx + 3 * ty) << 5 |
                 uint32_t((tx == 1 ? i - 1 : 0) + (ty == 1 ? (j - 1) * sx : 0)) << 9 | uint32_t(sx * (ty == 1 ? P - 1 : 1)) << 14 |
                 uint32_t(c < CW * n2 ? cl * CS + ij : 0) << 19;
      }
    const auto col_cl = [&](const int r) { return int(col[r] & 31u); };
    const auto col_sm = [&](const int r) { return int(col[r] >> 19); };
    const int  scl = lane / n, st = lane - scl * n;
    const T    lk  = scl < CW ? FDM<P, T>::lam(st) : T(0);

    while (true)
      {
        __syncthreads(); // the previous chunk is finished: cell storage, DoF window and records are free
        if (tid == 0)
          *s_chunk = int(a.n_chunks) - 1 - int(atomicAdd(a.state, 1u));
        __syncthreads();
        const int c = *s_chunk;
        if (c < 0)
          break;
        const uint32_t *d  = a.desc + size_t(c) * CHUNK_DESC_WORDS;
        const uint32_t *cb = a.pool + __ldg(d + CD_BLOB);
        const uint32_t  n_cells = __ldg(cb + CB_N_CELLS), n_own = __ldg(cb + CB_N_OWNED), n_fp = __ldg(cb + CB_N_FP),
                       n_fl = __ldg(cb + CB_N_FL), n_fg = __ldg(cb + CB_N_FG), n_dep_low = __ldg(cb + CB_N_DEP_LOW),
                       n_dep_late = __ldg(cb + CB_N_DEP_LATE), hanging = __ldg(cb + CB_HANGING);
        const uint32_t  ob = __ldg(d + CD_OWNED_BEGIN), side_off = __ldg(d + CD_SIDE_OFF);
        const uint32_t *src_deps = a.deps + __ldg(d + CD_SRC_DEPS);
        const uint32_t  n_slots  = n_own + n_fp + n_fl + n_fg;
        // the bulk copy needs 16-byte aligned addresses on both sides: slot s lives at xs[shift + s]
        const uint32_t shift = uint32_t((reinterpret_cast<uintptr_t>(a.src + ob) & 15u) / sizeof(T));
        T             *xsp   = xs + shift;
        const uint32_t bulk_bytes = uint32_t((shift + n_own) * sizeof(T)) & ~15u;
        if (tid == 0 && bulk_bytes)
          {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(mbar)), "r"(bulk_bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(xs)),
                         "l"(a.src + ob - shift), "r"(bulk_bytes), "r"(smem_u32(mbar))
                         : "memory");
          }
        for (uint32_t t = tid; t < n_dep_low + n_dep_late; t += THREADS)
          s_dep[t] = __ldg(a.desc + size_t(__ldg(src_deps + t)) * CHUNK_DESC_WORDS + CD_OWNED_BEGIN);
        {
          const uint4 *g = reinterpret_cast<const uint4 *>(cb + __ldg(cb + CB_OFF_CELLREC));
          uint4       *s = reinterpret_cast<uint4 *>(rec);
          for (uint32_t t = tid; t < n_cells * 8; t += THREADS)
            s[t] = __ldg(g + t);
        }
        // tail of the owned range the 16-byte granular bulk copy leaves out
        for (uint32_t t = bulk_bytes / sizeof(T) + tid; t < shift + n_own; t += THREADS)
          xs[t] = a.src[ob - shift + t];
        for (uint32_t t = tid; t < CHUNK_ZERO_SLOTS; t += THREADS)
          xsp[n_slots + t] = T(0);
        __syncthreads();
        {
          const uint32_t *fsrc = cb + __ldg(cb + CB_OFF_FSRC);
          for (uint32_t t = tid; t < n_fp + n_fl; t += THREADS)
            {
              const uint32_t w = __ldg(fsrc + t);
              xsp[n_own + t]   = a.src[s_dep[w >> 16] + (w & 0xffffu)];
            }
          const uint32_t *gi = a.ghost_idx + __ldg(d + CD_GHOST_OFF);
          for (uint32_t t = tid; t < n_fg; t += THREADS)
            xsp[n_own + n_fp + n_fl + t] = a.src[__ldg(gi + t)];
        }
        if (bulk_bytes)
          {
            unsigned ok = 0;
            while (!ok)
              asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                           : "=r"(ok)
                           : "r"(smem_u32(mbar)), "r"(parity)
                           : "memory");
            parity ^= 1u;
          }
        __syncthreads();

        // ------------------------------------------------------------------ cells
#pragma unroll 1
        for (int batch = 0; batch < NB; ++batch)
          {
            const int first = (warp * NB + batch) * CW;
            const int ncw   = min(CW, int(n_cells) - first);
            if (ncw <= 0)
              continue;
            T              *u  = u_all + first * CS;
            const uint32_t *tb = rec + first * CTAB_STRIDE;
            const auto      column_slot = [&](const int r, const int k) -> uint32_t {
              const uint32_t e = tb[(col[r] & 31u) * CTAB_STRIDE + ((col[r] >> 5) & 15u) + (k == 0 ? 0 : (k == P ? 18 : 9))];
              return e + ((col[r] >> 9) & 31u) + ((k > 0 && k < P) ? (k - 1) * ((col[r] >> 14) & 31u) : 0);
            };
            bool warp_hn = false;
            if (hanging)
              {
                const unsigned mymask = lane < ncw ? (tb[lane * CTAB_STRIDE + 27] & 0xffffu) : 0u;
                warp_hn               = __any_sync(0xffffffffu, (mymask >> 3) != 0);
              }
            // gather from the DoF window in the column layout, z sweep, to the cell storage (round by round: the values of
            // one round are the only live registers)
#pragma unroll
            for (int r = 0; r < NR; ++r)
              if (col_cl(r) < ncw)
                {
                  T pre[n];
#pragma unroll
                  for (int k = 0; k < n; ++k)
                    pre[k] = xsp[column_slot(r, k)];
                  if (!warp_hn)
                    fdm_forward<P, T, 1>(pre);
#pragma unroll
                  for (int k = 0; k < n; ++k)
                    u[col_sm(r) + k * PS] = pre[k];
                }
            if (warp_hn)
              {
                // hanging-node interpolation on the nodal values (only warps that hold such cells), then the z sweep
                __syncwarp();
                hanging_node_sweeps_slab<P, T, false>(u, tb, ncw, lane);
#pragma unroll
                for (int r = 0; r < NR; ++r)
                  if (col_cl(r) < ncw)
                    {
                      T pre[n];
#pragma unroll
                      for (int k = 0; k < n; ++k)
                        pre[k] = u[col_sm(r) + k * PS];
                      fdm_forward<P, T, 1>(pre);
#pragma unroll
                      for (int k = 0; k < n; ++k)
                        u[col_sm(r) + k * PS] = pre[k];
                    }
              }
            __syncwarp();
            if (scl < ncw)
              {
                T  reg[n2];
                T *pl = u + scl * CS + st * PS;
#pragma unroll
                for (int r = 0; r < n2; ++r)
                  reg[r] = pl[r];
#pragma unroll
                for (int j = 0; j < n; ++j)
                  fdm_forward<P, T, 1>(reg + j * n);
#pragma unroll
                for (int i = 0; i < n; ++i)
                  fdm_forward<P, T, n>(reg + i);
#pragma unroll
                for (int j = 0; j < n; ++j)
#pragma unroll
                  for (int i = 0; i < n; ++i)
                    reg[j * n + i] *= (lk + FDM<P, T>::lam2(i, j));
#pragma unroll
                for (int i = 0; i < n; ++i)
                  fdm_backward<P, T, n>(reg + i);
#pragma unroll
                for (int j = 0; j < n; ++j)
                  fdm_backward<P, T, 1>(reg + j * n);
#pragma unroll
                for (int r = 0; r < n2; ++r)
                  pl[r] = reg[r];
              }
            __syncwarp();
            if (warp_hn)
              {
#pragma unroll
                for (int r = 0; r < NR; ++r)
                  if (col_cl(r) < ncw)
                    {
                      T res[n];
#pragma unroll
                      for (int k = 0; k < n; ++k)
                        res[k] = u[col_sm(r) + k * PS];
                      fdm_backward<P, T, 1>(res);
#pragma unroll
                      for (int k = 0; k < n; ++k)
                        u[col_sm(r) + k * PS] = res[k];
                    }
                __syncwarp();
                hanging_node_sweeps_slab<P, T, true>(u, tb, ncw, lane);
              }
#pragma unroll
            for (int r = 0; r < NR; ++r)
              if (col_cl(r) < ncw)
                {
                  T res[n];
#pragma unroll
                  for (int k = 0; k < n; ++k)
                    res[k] = u[col_sm(r) + k * PS];
                  if (!warp_hn)
                    fdm_backward<P, T, 1>(res);
                  const unsigned level = (tb[col_cl(r) * CTAB_STRIDE + 27] >> 16) & 0xffu;
                  const T        h     = T(a.root_length * __hiloint2double((1023 - (int)level) << 20, 0));
#pragma unroll
                  for (int k = 0; k < n; ++k)
                    u[col_sm(r) + k * PS] = h * res[k];
                }
          }
        __syncthreads();

        // ------------------------------------------------------------------ hand-off of the foreign sums
        const uint32_t *ent_hdr = cb + __ldg(cb + CB_OFF_ENT_HDR);
        const uint16_t *ent     = reinterpret_cast<const uint16_t *>(cb + __ldg(cb + CB_OFF_ENT));
        const auto      slot_sum = [&](const uint32_t s) -> T {
          const uint32_t hdr = __ldg(ent_hdr + s);
          const uint32_t cnt = hdr & 255u, off = hdr >> 8;
          T              v   = T(0);
          for (uint32_t j = 0; j < cnt; ++j)
            v += u_all[__ldg(ent + off + j)];
          return v;
        };
        for (uint32_t t = tid; t < n_fp; t += THREADS)
          __stcg(a.side + size_t(side_off) + t, slot_sum(n_own + t));
        if (n_fg)
          {
            const uint32_t *gi = a.ghost_idx + __ldg(d + CD_GHOST_OFF);
            for (uint32_t t = tid; t < n_fg; t += THREADS)
              atomic_add(a.out + __ldg(gi + t), slot_sum(n_own + n_fp + n_fl + t));
          }
        __threadfence();
        __syncthreads();
        if (tid == 0)
          st_release_u32(ready + c, epoch);

        // ------------------------------------------------------------------ contributions of the chunks started earlier
        const uint32_t *pb = a.pool + __ldg(d + CD_PULL_BLOB);
        const uint32_t  n_pull_slots = __ldg(pb + PB_N_SLOTS), n_pull_deps = __ldg(pb + PB_N_DEPS);
        if (n_pull_slots)
          {
            const uint32_t *pull_deps = a.deps + __ldg(d + CD_PULL_DEPS);
            for (uint32_t t = tid; t < n_pull_deps; t += THREADS)
              {
                const uint32_t dep = __ldg(pull_deps + t);
                spin_until(ready + dep, epoch);
                s_side[t] = __ldg(a.desc + size_t(dep) * CHUNK_DESC_WORDS + CD_SIDE_OFF);
              }
            __syncthreads();
            const uint32_t *prec = pb + __ldg(pb + PB_OFF_REC), *plist = pb + __ldg(pb + PB_OFF_LIST);
            for (uint32_t q = tid; q < n_pull_slots; q += THREADS)
              {
                const uint32_t w0 = __ldg(prec + 2 * q), w1 = __ldg(prec + 2 * q + 1);
                const uint32_t cnt = w0 >> 16;
                T              v   = T(0);
                for (uint32_t j = 0; j < cnt; ++j)
                  {
                    const uint32_t e = __ldg(plist + w1 + j);
                    v += __ldcg(a.side + size_t(s_side[e >> 16]) + (e & 0xffffu));
                  }
                u_all[w0 & 0xffffu] += v;
              }
            __syncthreads();
          }

        // ------------------------------------------------------------------ owned range: sum, epilogue, one plain store
        {
          T *out = a.out + ob;
          if (a.mode == 0)
            for (uint32_t s = tid; s < n_own; s += THREADS)
              out[s] = slot_sum(s);
          else if (a.mode == 1)
            for (uint32_t s = tid; s < n_own; s += THREADS)
              out[s] = a.b[ob + s] - slot_sum(s);
          else if (a.mode == 2)
            for (uint32_t s = tid; s < n_own; s += THREADS)
              out[s] = (T(1) + a.f1) * a.src[ob + s] + a.f2 * (a.dinv[ob + s] * (a.b[ob + s] - slot_sum(s)));
          else
            for (uint32_t s = tid; s < n_own; s += THREADS)
              out[s] = (T(1) + a.f1) * a.src[ob + s] - a.f1 * a.xold[ob + s] + a.f2 * (a.dinv[ob + s] * (a.b[ob + s] - slot_sum(s)));
        }
        __threadfence();
        __syncthreads();
        if (tid == 0)
          st_release_u32(stored + c, epoch);

        // ------------------------------------------------------------------ late contributions (owner has a higher index)
        if (n_fl)
          {
            for (uint32_t t = tid; t < n_dep_late; t += THREADS)
              spin_until(stored + __ldg(src_deps + n_dep_low + t), epoch);
            __syncthreads();
            const uint32_t *fsrc = cb + __ldg(cb + CB_OFF_FSRC);
            for (uint32_t t = tid; t < n_fl; t += THREADS)
              {
                const uint32_t w = __ldg(fsrc + n_fp + t);
                const uint32_t g = s_dep[w >> 16] + (w & 0xffffu);
                const T        v = slot_sum(n_own + n_fp + t);
                const T        coef = a.mode == 0 ? T(1) : (a.mode == 1 ? T(-1) : -a.f2 * a.dinv[g]);
                atomic_add(a.out + g, coef * v);
              }
          }
      }

    // the last block re-arms the ticket counter and advances the epoch for the next launch
    if (tid == 0)
      {
        __threadfence();
        if (atomicAdd(a.state + 2, 1u) == gridDim.x - 1)
          {
            a.state[0] = 0u;
            a.state[2] = 0u;
            __threadfence();
            a.state[1] = epoch;
          }
      }
  }

  template <int P, typename T> void launch_chunks(const ChunkArgs<T> &a, cudaStream_t s)
  {
    using Cfg = ChunkCfg<P>;
    if (a.n_chunks == 0)
      return;
    if (chunk_cells(P) != Cfg::NC || chunk_slot_capacity(P) != Cfg::CAP || chunk_warps(P) != Cfg::W)
      throw std::runtime_error("chunk configuration of host and device differ");
    const size_t smem = ChunkSmem<P, T>::total;
    static int   blocks_per_sm = 0, n_sm = 0;
    if (blocks_per_sm == 0)
      {
        int dev = 0;
        MGB_CUDA(cudaGetDevice(&dev));
        MGB_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
        MGB_CUDA(cudaFuncSetAttribute(k_laplace_chunks<P, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        MGB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, k_laplace_chunks<P, T>, Cfg::THREADS, smem));
        if (blocks_per_sm < 1)
          throw std::runtime_error("k_laplace_chunks does not fit on an SM");
      }
    const unsigned blocks = (unsigned)std::min<long long>(a.n_chunks, (long long)blocks_per_sm * n_sm);
    k_laplace_chunks<P, T><<<blocks, Cfg::THREADS, smem, s>>>(a);
    MGB_CUDA(cudaGetLastError());
  }
} // namespace mgb
