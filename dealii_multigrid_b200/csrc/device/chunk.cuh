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

  // MGB_CHUNK_ABLATE (experiments with tools/build_variants.sh; 0 = the shipped kernel):
  //   1: load phase only            (timing only, results wrong by construction)
  //   2: load + cell arithmetic     (timing only, results wrong by construction)
  //   3: FUNCTIONAL variant without the hand-off between chunks: after the merge rounds every slot's sum is added to the
  //      output with atomics (owned range coalesced); no flags, no side buffer, no waiting.  dst must be zero on entry
  //      (chunk_atomic_output() tells the host), the fused epilogues are not available.
#ifndef MGB_CHUNK_ABLATE
#define MGB_CHUNK_ABLATE 0
#endif

  // cell storage layout of the chunk kernel: SlabLayout, except that 5x5 planes of doubles are packed tighter (plane
  // stride 27, cell stride 135: the slab pattern stays conflict free, the column pattern gets a two-way conflict where a
  // half-warp crosses a cell boundary) so that three blocks fit on an SM
  template <int N, int BYTES> struct ChunkLayout : SlabLayout<N, BYTES>
  {};
  template <> struct ChunkLayout<5, 8>
  {
    static constexpr int PS = 27, CS = 135;
  };

  template <int P> struct ChunkCfg
  {
    static constexpr int n   = P + 1;
    static constexpr int CW  = 32 / n;
#ifndef MGB_CHUNK_NC4
#define MGB_CHUNK_NC4 32
#endif
#ifndef MGB_CHUNK_CAP4
#define MGB_CHUNK_CAP4 (MGB_CHUNK_NC4 == 32 ? 2880 : 1536)
#endif
    static constexpr int NC  = P == 1 ? 256 : P <= 3 ? 64 : P == 4 ? MGB_CHUNK_NC4 : P == 5 ? 32 : 16; // cells per chunk (whole octets)
    static constexpr int NB  = P == 1 ? 2 : 1;                                              // batches of CW cells per warp
    static constexpr int W   = (NC + CW * NB - 1) / (CW * NB);                              // warps per block
    static constexpr int CAP = P == 1 ? 2112 : P == 2 ? 1792 : P == 3 ? 3072 : P == 4 ? MGB_CHUNK_CAP4 : 5632; // slots of the DoF window
#ifndef MGB_CHUNK_MINB4
#define MGB_CHUNK_MINB4 3
#endif
    static constexpr int MINB = P == 4 ? MGB_CHUNK_MINB4 : 2;                               // resident blocks per SM
    static constexpr int THREADS = W * 32;
  };

  template <int P, typename T> struct ChunkSmem
  {
    static constexpr int    CS      = ChunkLayout<P + 1, sizeof(T)>::CS;
    static constexpr size_t u_bytes = (size_t(ChunkCfg<P>::NC) * CS * sizeof(T) + 15) / 16 * 16;
    static constexpr size_t x_bytes = (size_t(ChunkCfg<P>::CAP) * sizeof(T) + 15) / 16 * 16;
    static constexpr size_t r_bytes = size_t(ChunkCfg<P>::NC) * 32 * 4;
    static constexpr size_t d_bytes = size_t(CHUNK_MAX_DEPS) * 4 * 2;
    static constexpr size_t total   = u_bytes + x_bytes + r_bytes + d_bytes + 48;
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

  __device__ __forceinline__ void mbar_wait(uint64_t *bar, const unsigned parity)
  {
    unsigned ok = 0;
    while (!ok)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(ok)
                   : "r"(smem_u32(bar)), "r"(parity)
                   : "memory");
  }
  __device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, const uint32_t bytes, uint64_t *bar)
  {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
  }

  template <int P, typename T>
  __global__ void __launch_bounds__(ChunkCfg<P>::THREADS, ChunkCfg<P>::MINB) k_laplace_chunks(const ChunkArgs<T> a)
  {
    using Cfg = ChunkCfg<P>;
    constexpr int n = P + 1, n2 = n * n;
    constexpr int CW = Cfg::CW, NB = Cfg::NB, THREADS = Cfg::THREADS;
    constexpr int PS = ChunkLayout<n, sizeof(T)>::PS, CS = ChunkLayout<n, sizeof(T)>::CS;
    constexpr int NR = (CW * n2 + 31) / 32;
    // the epilogue tables are staged in the DoF window once every warp has gathered its cells (one batch per warp only)
    constexpr bool     STAGE       = NB == 1;
    constexpr uint32_t STAGE_WORDS = uint32_t(ChunkSmem<P, T>::x_bytes / 4);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T        *u_all  = reinterpret_cast<T *>(smem_raw);
    T        *xs     = reinterpret_cast<T *>(smem_raw + ChunkSmem<P, T>::u_bytes);
    uint32_t *rec    = reinterpret_cast<uint32_t *>(smem_raw + ChunkSmem<P, T>::u_bytes + ChunkSmem<P, T>::x_bytes);
    uint32_t *s_dep  = rec + Cfg::NC * 32;          // owned_begin of the src deps
    uint32_t *s_side = s_dep + CHUNK_MAX_DEPS;      // side-buffer offset of the pull deps
    uint64_t *mbar   = reinterpret_cast<uint64_t *>(s_side + CHUNK_MAX_DEPS); // [0]: DoF window + cell records, [1]: tables
    int      *s_chunk = reinterpret_cast<int *>(mbar + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t epoch = *reinterpret_cast<volatile uint32_t *>(a.state + 1) + 1u;
    uint32_t *ready = a.flags, *stored = a.flags + a.n_chunks;
    if (tid == 0)
      {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(mbar)) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(mbar + 1)) : "memory");
        *s_chunk = int(a.n_chunks) - 1 - int(atomicAdd(a.state, 1u));
      }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    unsigned parity0 = 0, parity1 = 0;

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
    int        prev_chunk = -1;

    while (true)
      {
        __syncthreads(); // the previous chunk is finished: cell storage, DoF window and records are free
        const int c = *s_chunk;
        if (tid == 0 && prev_chunk >= 0)
          st_release_u32(stored + prev_chunk, epoch); // its owned range is in memory (late contributions wait for this)
        if (c < 0)
          break;
        const uint32_t *d  = a.desc + size_t(c) * CHUNK_DESC_WORDS;
        const uint32_t *cb = a.pool + __ldg(d + CD_BLOB);
        const uint32_t *pb = a.pool + __ldg(d + CD_PULL_BLOB);
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
        if (tid == 0)
          {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(mbar)), "r"(bulk_bytes + n_cells * 128u)
                         : "memory");
            if (bulk_bytes)
              bulk_g2s(xs, a.src + ob - shift, bulk_bytes, mbar);
            bulk_g2s(rec, cb + CB_HEADER_WORDS, n_cells * 128u, mbar); // the cell records follow the header
          }
        for (uint32_t t = tid; t < n_dep_low + n_dep_late; t += THREADS)
          s_dep[t] = __ldg(a.desc + size_t(__ldg(src_deps + t)) * CHUNK_DESC_WORDS + CD_OWNED_BEGIN);
        // tail of the owned range the 16-byte granular bulk copy leaves out
        for (uint32_t t = bulk_bytes / sizeof(T) + tid; t < shift + n_own; t += THREADS)
          xs[t] = a.src[ob - shift + t];
        for (uint32_t t = tid; t < CHUNK_ZERO_SLOTS; t += THREADS)
          xsp[n_slots + t] = T(0);
        {
          // foreign DoFs: owner's range (from shared memory, hence the barrier) + offset; the first words are fetched
          // before the barrier, and FU loads per thread are in flight
          constexpr int   FU   = 3;
          const uint32_t *fsrc = cb + __ldg(cb + CB_OFF_FSRC);
          const uint32_t  n_f  = n_fp + n_fl;
          uint32_t        w[FU];
#pragma unroll
          for (int q = 0; q < FU; ++q)
            w[q] = tid + q * THREADS < n_f ? __ldg(fsrc + tid + q * THREADS) : 0u;
          __syncthreads();
          for (uint32_t base = tid; base < n_f; base += THREADS * FU)
            {
              if (base != uint32_t(tid))
#pragma unroll
                for (int q = 0; q < FU; ++q)
                  w[q] = base + q * THREADS < n_f ? __ldg(fsrc + base + q * THREADS) : 0u;
              T v[FU];
#pragma unroll
              for (int q = 0; q < FU; ++q)
                v[q] = base + q * THREADS < n_f ? a.src[s_dep[w[q] >> 16] + (w[q] & 0xffffu)] : T(0);
#pragma unroll
              for (int q = 0; q < FU; ++q)
                if (base + q * THREADS < n_f)
                  xsp[n_own + base + q * THREADS] = v[q];
            }
          const uint32_t *gi = a.ghost_idx + __ldg(d + CD_GHOST_OFF);
          for (uint32_t t = tid; t < n_fg; t += THREADS)
            xsp[n_own + n_fp + n_fl + t] = a.src[__ldg(gi + t)];
        }
        mbar_wait(mbar, parity0);
        parity0 ^= 1u;
        __syncthreads();

        // epilogue tables: [CB_OFF_MERGE, CB_TOTAL_WORDS) of the compute blob followed by the whole pull blob
        const uint32_t  off_merge = __ldg(cb + CB_OFF_MERGE), tc_words = __ldg(cb + CB_TOTAL_WORDS) - off_merge,
                       tp_words = __ldg(pb + PB_TOTAL_WORDS);
        const bool      staged = STAGE && tc_words + tp_words <= STAGE_WORDS;
        const uint32_t *tc = staged ? reinterpret_cast<const uint32_t *>(xs) : cb + off_merge;
        const uint32_t *tp = staged ? reinterpret_cast<const uint32_t *>(xs) + tc_words : pb;
        int             next_chunk = 0;

        // ------------------------------------------------------------------ cells
#pragma unroll 1
        for (int batch = 0; batch < (MGB_CHUNK_ABLATE == 1 ? 0 : NB); ++batch)
          {
            const int       first = (warp * NB + batch) * CW;
            const int       ncw   = max(0, min(CW, int(n_cells) - first));
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
            if (STAGE)
              {
                // every warp has read its DoFs: the window now receives the epilogue tables while the cells are computed
                __syncthreads();
                if (tid == 0)
                  {
                    next_chunk = int(a.n_chunks) - 1 - int(atomicAdd(a.state, 1u));
                    if (staged)
                      {
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(mbar + 1)),
                                     "r"((tc_words + tp_words) * 4u)
                                     : "memory");
                        bulk_g2s(xs, cb + off_merge, tc_words * 4u, mbar + 1);
                        bulk_g2s(reinterpret_cast<uint32_t *>(xs) + tc_words, pb, tp_words * 4u, mbar + 1);
                      }
                  }
              }
            if (warp_hn)
              {
                // hanging-node interpolation on the nodal values (only warps that hold such cells), then the z sweep
                __syncwarp();
                hanging_node_sweeps_slab<P, T, false, PS, CS>(u, tb, ncw, lane);
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
                hanging_node_sweeps_slab<P, T, true, PS, CS>(u, tb, ncw, lane);
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
        if ((!STAGE || MGB_CHUNK_ABLATE == 1) && tid == 0)
          next_chunk = int(a.n_chunks) - 1 - int(atomicAdd(a.state, 1u));
        __syncthreads();
        if (MGB_CHUNK_ABLATE == 1 || MGB_CHUNK_ABLATE == 2)
          {
            // timing ablation: nothing after the load (1) / the arithmetic (2); keep the arithmetic alive with a store that
            // never happens
            if (MGB_CHUNK_ABLATE == 2 && u_all[tid] == T(-1.2345e300))
              a.out[ob] = u_all[tid];
            if (staged)
              {
                mbar_wait(mbar + 1, parity1);
                parity1 ^= 1u;
              }
            if (tid == 0)
              *s_chunk = next_chunk;
            prev_chunk = -1;
            continue;
          }
        if (staged)
          {
            mbar_wait(mbar + 1, parity1);
            parity1 ^= 1u;
          }

        // ------------------------------------------------------------------ merge the copies of every DoF (binary tree)
        {
          const uint32_t n_rounds = __ldg(cb + CB_N_ROUNDS);
          for (uint32_t r = 0; r < n_rounds; ++r)
            {
              const uint32_t q0 = tc[r], q1 = tc[r + 1];
              for (uint32_t q = q0 + tid; q < q1; q += THREADS)
                {
                  const uint32_t w = tc[q];
                  u_all[w & 0xffffu] += u_all[w >> 16];
                }
              __syncthreads();
            }
        }
        // after the rounds the sum of slot s sits at position slotpos[s] of the cell storage
        const uint16_t *slotpos = reinterpret_cast<const uint16_t *>(tc + (__ldg(cb + CB_OFF_SLOTPOS) - off_merge));
        const auto      slot_value = [&](const uint32_t s) -> T {
          const uint32_t pos = slotpos[s];
          return pos != 0xffffu ? u_all[pos] : T(0);
        };
        if (MGB_CHUNK_ABLATE == 3)
          {
            // every slot's sum goes to the (zeroed) output with an atomic add: owned range coalesced, foreign ones scattered
            for (uint32_t s = tid; s < n_own; s += THREADS)
              {
                const uint32_t pos = slotpos[s];
                if (pos != 0xffffu)
                  atomic_add(a.out + ob + s, u_all[pos]);
              }
            const uint32_t *fsrc = cb + __ldg(cb + CB_OFF_FSRC);
            for (uint32_t t = tid; t < n_fp + n_fl; t += THREADS)
              {
                const uint32_t w = __ldg(fsrc + t);
                atomic_add(a.out + s_dep[w >> 16] + (w & 0xffffu), slot_value(n_own + t));
              }
            const uint32_t *gi = a.ghost_idx + __ldg(d + CD_GHOST_OFF);
            for (uint32_t t = tid; t < n_fg; t += THREADS)
              atomic_add(a.out + __ldg(gi + t), slot_value(n_own + n_fp + n_fl + t));
            if (tid == 0)
              *s_chunk = next_chunk;
            prev_chunk = -1;
            continue;
          }
        // ------------------------------------------------------------------ hand-off of the foreign sums
        for (uint32_t t = tid; t < n_fp; t += THREADS)
          __stcg(a.side + size_t(side_off) + t, slot_value(n_own + t));
        if (n_fg)
          {
            const uint32_t *gi = a.ghost_idx + __ldg(d + CD_GHOST_OFF);
            for (uint32_t t = tid; t < n_fg; t += THREADS)
              atomic_add(a.out + __ldg(gi + t), slot_value(n_own + n_fp + n_fl + t));
          }
        // one barrier orders the stores of all threads before thread 0's release (cumulativity of the release fence)
        __syncthreads();
        if (tid == 0)
          {
            st_release_u32(ready + c, epoch);
            *s_chunk = next_chunk;
          }

        // ------------------------------------------------------------------ owned range: epilogue, one plain store per DoF
        const uint32_t  n_pull_slots = __ldg(pb + PB_N_SLOTS), n_pull_deps = __ldg(pb + PB_N_DEPS);
        const uint16_t *opos = reinterpret_cast<const uint16_t *>(tp + __ldg(pb + PB_OFF_OPOS));
        T              *out = a.out + ob;
        const T        *xo = a.src + ob, *bo = a.b + ob, *dv = a.dinv + ob, *xl = a.xold + ob;
        const T         f1 = a.f1, f2 = a.f2;
        const int       mode = a.mode;
        const auto      finish = [&](const uint32_t s, const T v) {
          if (mode == 0)
            out[s] = v;
          else if (mode == 1)
            out[s] = bo[s] - v;
          else if (mode == 2)
            out[s] = (T(1) + f1) * xo[s] + f2 * (dv[s] * (bo[s] - v));
          else
            out[s] = (T(1) + f1) * xo[s] - f1 * xl[s] + f2 * (dv[s] * (bo[s] - v));
        };
        // first the DoFs no other chunk contributes to (three quarters), while the contributing chunks finish
#pragma unroll 2
        for (uint32_t s = tid; s < n_own; s += THREADS)
          {
            const uint32_t w = opos[s];
            if (!(w & 0x8000u))
              finish(s, w != 0x7fffu ? u_all[w] : T(0));
          }
        if (n_pull_slots)
          {
            const uint32_t *pull_deps = a.deps + __ldg(d + CD_PULL_DEPS);
            for (uint32_t t = tid; t < n_pull_deps; t += THREADS)
              {
                const uint32_t dep = __ldg(pull_deps + t);
                s_side[t]          = __ldg(a.desc + size_t(dep) * CHUNK_DESC_WORDS + CD_SIDE_OFF);
                spin_until(ready + dep, epoch);
              }
            __syncthreads();
            const uint2    *prec  = reinterpret_cast<const uint2 *>(tp + __ldg(pb + PB_OFF_REC));
            const uint32_t *plist = tp + __ldg(pb + PB_OFF_LIST);
            constexpr int   U     = 3;
            for (uint32_t base = tid; base < n_pull_slots; base += THREADS * U)
              {
                uint2 r[U];
                T     v[U];
#pragma unroll
                for (int q = 0; q < U; ++q)
                  {
                    r[q] = base + q * THREADS < n_pull_slots ? prec[base + q * THREADS] : make_uint2(0u, 0u);
                    v[q] = ((r[q].x >> 13) & 31u) ? __ldcg(a.side + size_t(s_side[r[q].y >> 16]) + (r[q].y & 0xffffu)) : T(0);
                  }
#pragma unroll
                for (int q = 0; q < U; ++q)
                  {
                    const uint32_t cnt = (r[q].x >> 13) & 31u, slot = r[q].x & 0x1fffu;
                    for (uint32_t j = 1; j < cnt; ++j)
                      {
                        const uint32_t e = plist[(r[q].x >> 18) + j - 1];
                        v[q] += __ldcg(a.side + size_t(s_side[e >> 16]) + (e & 0xffffu));
                      }
                    // own chunk first, then the other chunks in ascending order: the same order on every run
                    if (cnt)
                      {
                        const uint32_t w = opos[slot] & 0x7fffu;
                        finish(slot, (w != 0x7fffu ? u_all[w] : T(0)) + v[q]);
                      }
                  }
              }
          }

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
                const T        coef = a.mode == 0 ? T(1) : (a.mode == 1 ? T(-1) : -a.f2 * a.dinv[g]);
                atomic_add(a.out + g, coef * slot_value(n_own + n_fp + t));
              }
          }
        prev_chunk = c;
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
    {
      int ps = 0, cs = 0;
      chunk_layout(P + 1, (int)sizeof(T), ps, cs);
      if (chunk_cells(P) != Cfg::NC || chunk_slot_capacity(P) != Cfg::CAP || chunk_warps(P) != Cfg::W ||
          ps != ChunkLayout<P + 1, sizeof(T)>::PS || cs != ChunkLayout<P + 1, sizeof(T)>::CS)
        throw std::runtime_error("chunk configuration of host and device differ");
    }
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
