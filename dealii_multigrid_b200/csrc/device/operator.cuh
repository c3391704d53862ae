// Matrix-free Laplace cell kernels: the FEEvaluation gather_evaluate / submit_gradient /
// integrate_scatter pipeline of include/operator.h:461-472 for Cartesian cells.
//
// Math: on an axis-aligned cube of side h the cell matrix is
//   A_e = h (K x M x M  +  M x K x M  +  M x M x K)
// with the 1-D Gauss(n)-integrated mass/stiffness matrices of the unit interval.  It is evaluated in
// 7 one-dimensional sweeps
//   a = M_x u, b = K_x u;   c = M_y a, g = M_y b + K_y a;   r = M_z g + K_z c
// (identical to deal.II's 12-sweep collocation evaluation in exact arithmetic; differs at rounding level).
//
// Thread layout: n^2 threads per cell, CPB cells per block.  In every sweep a thread owns one complete
// 1-D line in registers (n x n dense mat-vec from constant memory); lines are exchanged between sweeps
// through shared memory.  Gather/scatter go through the 27-entry entity table; Dirichlet entities carry
// bit 31 (gather 0, no scatter); hanging faces/edges are interpolated in place (cell_common.cuh).
#pragma once
#include "cell_common.cuh"

#include <algorithm>
#include <stdexcept>

namespace mgb
{
  // number of cells per block for degree P
  template <int P> struct CellsPerBlock
  {
    static constexpr int value = P == 1 ? 64 : P == 2 ? 28 : P == 3 ? 16 : P == 4 ? 8 : P == 5 ? 7 : 4;
  };

  // apply A_e/h to the nodal values in sA (result in sA); sB is scratch.  Contains barriers.
  template <int P, typename T> __device__ __forceinline__ void laplace_sweeps(T *sA, T *sB, const int t, const bool active)
  {
    constexpr int n = P + 1, n2 = n * n;
    const double *Mm = c_fe[P].M;
    const double *Km = c_fe[P].K;
    const int     a = t % n, b = t / n;
    // ---- x: thread (j,k) = (a,b)
    if (active)
      {
        const int base = (b * n + a) * n;
        T         u[n];
#pragma unroll
        for (int i = 0; i < n; ++i)
          u[i] = sA[base + i];
#pragma unroll
        for (int i = 0; i < n; ++i)
          {
            T va = 0, vb = 0;
#pragma unroll
            for (int l = 0; l < n; ++l)
              {
                va += T(Mm[i * MAXN + l]) * u[l];
                vb += T(Km[i * MAXN + l]) * u[l];
              }
            sA[base + i] = va;
            sB[base + i] = vb;
          }
      }
    __syncthreads();
    // ---- y: thread (i,k) = (a,b)
    if (active)
      {
        const int base = b * n2 + a;
        T         ua[n], ub[n];
#pragma unroll
        for (int j = 0; j < n; ++j)
          {
            ua[j] = sA[base + j * n];
            ub[j] = sB[base + j * n];
          }
#pragma unroll
        for (int j = 0; j < n; ++j)
          {
            T vc = 0, vg = 0;
#pragma unroll
            for (int l = 0; l < n; ++l)
              {
                vc += T(Mm[j * MAXN + l]) * ua[l];
                vg += T(Mm[j * MAXN + l]) * ub[l] + T(Km[j * MAXN + l]) * ua[l];
              }
            sA[base + j * n] = vc;
            sB[base + j * n] = vg;
          }
      }
    __syncthreads();
    // ---- z: thread (i,j) = (a,b)
    if (active)
      {
        const int base = b * n + a;
        T         uc[n], ug[n];
#pragma unroll
        for (int k = 0; k < n; ++k)
          {
            uc[k] = sA[base + k * n2];
            ug[k] = sB[base + k * n2];
          }
#pragma unroll
        for (int k = 0; k < n; ++k)
          {
            T vr = 0;
#pragma unroll
            for (int l = 0; l < n; ++l)
              vr += T(Mm[k * MAXN + l]) * ug[l] + T(Km[k * MAXN + l]) * uc[l];
            sA[base + k * n2] = vr;
          }
      }
    __syncthreads();
  }

  // ------------------------------------------------------------------------------------------------------------
  // vmult cell kernel ("slab" layout, fast-diagonalised cell matrix)
  //
  //   A_e = h (Tm x Tm x Tm)^T diag(lam_i + lam_j + lam_k) (Tm x Tm x Tm)      (fe1d.h; 6 sweeps instead of 7/12)
  //
  // n = P+1 threads own one cell, floor(32/n) cells live in one warp and warps never talk to each other (only
  // __syncwarp after the prologue).  A thread first owns the (x,y) plane k of its cell in registers and does the x and y
  // sweeps there, the cell is transposed through shared memory so that the thread owns the (y,z) plane i for the
  // z sweep, the diagonal scaling and the transposed z sweep, and is transposed back for the transposed y and x
  // sweeps.  Every sweep is done in even-odd form (the modes are symmetric/antisymmetric).  Gather and scatter are
  // warp-cooperative and coalesced along the lexicographic node index through the 27-entry entity table.
  template <int P> struct SlabCfg
  {
    static constexpr int n   = P + 1;
    static constexpr int CW  = 32 / n;          // cells per warp
#ifndef MGB_SLAB_W
#define MGB_SLAB_W 4
#endif
#ifndef MGB_SLAB_MINB
#define MGB_SLAB_MINB 3
#endif
    static constexpr int W   = P <= 4 ? MGB_SLAB_W : 2;  // warps per block
    static constexpr int CPB = CW * W;          // cells per block
    static constexpr int MIN_BLOCKS = P <= 4 ? MGB_SLAB_MINB : 2; // resident blocks per SM the register allocation must allow
  };

  // warp-level version of hanging_node_sweeps for the padded slab layout; `u` = the warp's cells, tb = their 32-word records (word 27 = mask | level << 16)
  template <int P, typename T, bool TRANSPOSE, int PS = SlabLayout<P + 1, sizeof(T)>::PS, int CS = SlabLayout<P + 1, sizeof(T)>::CS>
  __device__ __forceinline__ void hanging_node_sweeps_slab(T *u, const uint32_t *tb, const int ncw, const int lane)
  {
    constexpr int n = P + 1, n2 = n * n;
#pragma unroll 1
    for (int dd = 0; dd < 3; ++dd)
      {
        const int d  = TRANSPOSE ? 2 - dd : dd;
        const int da = d == 0 ? 1 : 0, db = d == 2 ? 1 : 2;
        for (int w = lane; w < ncw * n2; w += 32)
          {
            const int      cl = w / n2, t = w % n2;
            const unsigned mask = tb[cl * 32 + 27] & 0xffffu;
            if (!(mask >> 3))
              continue;
            const int  a = t % n, b = t / n;
            const int  sa = (mask >> da) & 1, sb = (mask >> db) & 1, sd = (mask >> d) & 1;
            const bool fa = (mask >> (3 + da)) & 1, fb = (mask >> (3 + db)) & 1, ed = (mask >> (6 + d)) & 1;
            const bool on_a = (a == sa * P), on_b = (b == sb * P);
            if (!((fa && on_a) || (fb && on_b) || (ed && on_a && on_b)))
              continue;
            // element i of the line: d=0: (k=b, r=a*n+i); d=1: (k=b, r=i*n+a); d=2: (k=i, r=b*n+a)
            T        *base   = u + cl * CS + (d == 2 ? b * n + a : b * PS + (d == 0 ? a * n : a));
            const int stride = d == 0 ? 1 : (d == 1 ? n : PS);
            T         v[n], r[n];
#pragma unroll
            for (int i = 0; i < n; ++i)
              v[i] = base[i * stride];
            const double *S = c_fe[P].S[sd];
#pragma unroll
            for (int i = 0; i < n; ++i)
              {
                T acc = 0;
#pragma unroll
                for (int j = 0; j < n; ++j)
                  acc += T(TRANSPOSE ? S[j * MAXN + i] : S[i * MAXN + j]) * v[j];
                r[i] = acc;
              }
#pragma unroll
            for (int i = 0; i < n; ++i)
              base[i * stride] = r[i];
          }
        __syncwarp();
      }
  }

  // Persistent, software-pipelined kernel.  Every warp walks over its batches of CW cells (stride = all warps of the
  // grid).  Two thread layouts alternate inside a batch:
  //   column layout: lane <-> one (i,j) column of one cell (CW*n^2 columns, NR rounds of 32); the n values along z
  //                  live in registers.  Gather (coalesced along the lexicographic node index, straight into
  //                  registers), the z sweep, its transpose and the scatter (RED) happen here, without shared memory.
  //   slab layout:   lane <-> (cell, plane k); the n^2 values of the plane live in registers for the x and y sweeps,
  //                  the diagonal scaling and the transposed y and x sweeps.
  // The two transposes between the layouts are the only shared-memory traffic of the data path (4 n^3 words per cell).
  // Prefetching: the gather of batch b+1 is issued into registers as soon as the z sweep of batch b has consumed the
  // previous ones, and the cells' packed 32-word records (27 entity entries | mask+level word) of batch b+2 are
  // streamed into a 3-deep shared ring by 16-byte LDGSTS, so no warp waits for a load it has just issued.
  constexpr int CTAB_STRIDE = 32; // words per cell record

  template <int P, typename T> struct SlabSmem
  {
    static constexpr int n  = P + 1;
    static constexpr int CW = SlabCfg<P>::CW;
    static constexpr int CS = SlabLayout<n, sizeof(T)>::CS;
    // per warp: cell storage + 3 table buffers
    static constexpr size_t u_bytes            = (size_t(CW * CS) * sizeof(T) + 15) / 16 * 16;
    static constexpr size_t warp_bytes_aligned = u_bytes + 3 * size_t(CW) * CTAB_STRIDE * 4;
    static size_t           block_bytes(int warps) { return warps * warp_bytes_aligned; }
  };

  // MGB_SLAB_ABLATE (timing experiments with tools/build_variants.sh only -- the results are wrong by construction):
  //   1: no scatter (the atomic adds are predicated off by a value test that never holds)
  //   2: no gather (the column values are made up from the indices, no loads from src)
  //   3: no arithmetic and no shared-memory transposes (gathered values are scattered back directly)
#ifndef MGB_SLAB_ABLATE
#define MGB_SLAB_ABLATE 0
#endif
  // experiments with the side zeroing of k_laplace_apply<P, T, true>: kind of store (0 plain, 1 st.cs, 2 st.wt) and its place
  // in the batch (0 before the sweeps, 1 after the scatter)
#ifndef MGB_ZERO_STORE
#define MGB_ZERO_STORE 1
#endif
#ifndef MGB_ZERO_POS
#define MGB_ZERO_POS 0
#endif
  // ZERO: the kernel also writes zeros to a vector that takes no part in this product (the destination of the NEXT cell loop:
  // deal.II's cell_loop zeroes dst inside the loop as well, zero_dst_vector = true).  Batch b clears the b-th slice of
  // zero_per_batch 16-byte words; the kernel is bound by the L1 data pipe, so the stores ride along where a separate memset
  // costs a full pass at DRAM speed -- that was the idea; measured, a zeroing launch loses 0.19-0.29 ms against 0.17 ms for the
  // memset (profiles/r2v_zero_in_cell_loop.txt), so ChebyshevT only uses it with MGB_ZERO_IN_CELL_LOOP=1.
  template <int P, typename T, bool ZERO = false>
  __global__ void __launch_bounds__(SlabCfg<P>::W * 32, SlabCfg<P>::MIN_BLOCKS)
    k_laplace_apply(const uint32_t *__restrict__ ctab, const int any_hanging, const double root_length, const long long n_work,
                    const uint32_t *__restrict__ cell_list, const T *__restrict__ src, T *__restrict__ dst, T *zero_ptr = nullptr,
                    const long long zero_n = 0, const long long zero_per_batch = 0)
  {
    constexpr int n = P + 1, n2 = n * n;
    constexpr int CW = SlabCfg<P>::CW, W = SlabCfg<P>::W;
    constexpr int PS = SlabLayout<n, sizeof(T)>::PS, CS = SlabLayout<n, sizeof(T)>::CS;
    constexpr int NR = (CW * n2 + 31) / 32; // rounds of the column layout
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int      warp  = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char *wbase = smem_raw + warp * SlabSmem<P, T>::warp_bytes_aligned;
    T             *u     = reinterpret_cast<T *>(wbase);                                  // [CW*CS]
    uint32_t      *tbuf  = reinterpret_cast<uint32_t *>(wbase + SlabSmem<P, T>::u_bytes);          // [3][CW][32]

    const long long n_batches = (n_work + CW - 1) / CW;
    const long long stride    = (long long)gridDim.x * W;
    long long       b         = (long long)blockIdx.x * W + warp;
#ifdef MGB_SLAB_LOCKSTEP
    // experiment (-DMGB_SLAB_LOCKSTEP): the warps of a block walk through the unrolled loop body together and share its
    // instruction-cache lines (same trip count for all warps, one barrier per batch); see k_laplace_q1_slab
    const long long n_it = (n_batches - (long long)blockIdx.x * W + stride - 1) / stride;
#else
    if (b >= n_batches)
      return;
#endif

    // static per-lane description of the columns of each round, packed into one word:
    // cell (5 bits) | entity base tx+3ty (4) | offset inside the entity (5) | z stride inside the entity (5) | smem address (13)
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

    const auto issue_tables = [&](const long long batch, uint32_t *tb) {
      if (batch >= n_batches)
        return;
      const long long first = batch * CW;
      const int       ncw   = (int)min((long long)CW, n_work - first);
      for (int idx = lane; idx < ncw * (CTAB_STRIDE / 4); idx += 32)
        {
          const int       cl   = idx / (CTAB_STRIDE / 4), part = idx % (CTAB_STRIDE / 4);
          const long long cell = cell_list ? (long long)cell_list[first + cl] : first + cl;
          const unsigned  s    = (unsigned)__cvta_generic_to_shared(tb + cl * CTAB_STRIDE + part * 4);
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(ctab + cell * CTAB_STRIDE + part * 4) : "memory");
        }
    };
    // global index of node k of a column (bit 31 = Dirichlet)
    const auto column_index = [&](const uint32_t *tb, const int r, const int k) -> uint32_t {
      const uint32_t e = tb[(col[r] & 31u) * CTAB_STRIDE + ((col[r] >> 5) & 15u) + (k == 0 ? 0 : (k == P ? 18 : 9))];
      return e + ((col[r] >> 9) & 31u) + ((k > 0 && k < P) ? (k - 1) * ((col[r] >> 14) & 31u) : 0);
    };
    T          pre[NR][n];
    const auto issue_gather = [&](const long long batch, const uint32_t *tb) {
      const int ncw = batch < n_batches ? (int)min((long long)CW, n_work - batch * CW) : 0;
#pragma unroll
      for (int r = 0; r < NR; ++r)
#pragma unroll
        for (int k = 0; k < n; ++k)
          {
            pre[r][k] = T(0);
            if (col_cl(r) < ncw)
              {
                const uint32_t idx = column_index(tb, r, k);
#if MGB_SLAB_ABLATE == 2
                pre[r][k] = T(idx & 1023u) * T(1e-3);
#else
                if (!(idx & D_FLAG_BIT))
                  pre[r][k] = __ldg(src + idx);
#endif
              }
          }
    };

    // prologue
    issue_tables(b, tbuf);
    cp_async_wait_all();
    __syncwarp();
    issue_gather(b, tbuf);
    issue_tables(b + stride, tbuf + CW * CTAB_STRIDE);

    const int cl = lane / n, t = lane - cl * n;
    const T   lk = cl < CW ? FDM<P, T>::lam(t) : T(0);
#ifdef MGB_SLAB_LOCKSTEP
    for (int it = 0; it < (int)n_it; b += stride, ++it)
      {
        __syncthreads();
        if (b >= n_batches)
          continue;
#else
    int       it = 0;
    for (; b < n_batches; b += stride, ++it)
      {
#endif
        const uint32_t *tb  = tbuf + (it % 3) * (CW * CTAB_STRIDE);
        const int       ncw = (int)min((long long)CW, n_work - b * CW);
        const auto zero_slice = [&]() {
          constexpr long long EPW = 16 / sizeof(T); // elements per 16-byte word
          const long long     n16 = zero_n / EPW, z1 = min((b + 1) * zero_per_batch, n16);
          for (long long z = b * zero_per_batch + lane; z < z1; z += 32)
            {
#if MGB_ZERO_STORE == 1
              __stcs(reinterpret_cast<uint4 *>(zero_ptr) + z, make_uint4(0u, 0u, 0u, 0u)); // streaming: evict first
#elif MGB_ZERO_STORE == 2
              __stwt(reinterpret_cast<uint4 *>(zero_ptr) + z, make_uint4(0u, 0u, 0u, 0u)); // write through
#else
              reinterpret_cast<uint4 *>(zero_ptr)[z] = make_uint4(0u, 0u, 0u, 0u);
#endif
            }
          if (b == 0 && lane < int(zero_n - n16 * EPW))
            zero_ptr[n16 * EPW + lane] = T(0);
        };
        if (ZERO && MGB_ZERO_POS == 0)
          zero_slice();
        bool            warp_hn = false;
        if (any_hanging)
          {
            const unsigned mymask = lane < ncw ? (tb[lane * CTAB_STRIDE + 27] & 0xffffu) : 0u;
            warp_hn               = __any_sync(0xffffffffu, (mymask >> 3) != 0);
          }
        if (warp_hn)
          {
            // hanging-node interpolation on the nodal values (through shared memory; only warps that have such cells)
#pragma unroll
            for (int r = 0; r < NR; ++r)
              if (col_cl(r) < ncw)
#pragma unroll
                for (int k = 0; k < n; ++k)
                  u[col_sm(r) + k * PS] = pre[r][k];
            __syncwarp();
            hanging_node_sweeps_slab<P, T, false>(u, tb, ncw, lane);
#pragma unroll
            for (int r = 0; r < NR; ++r)
              if (col_cl(r) < ncw)
#pragma unroll
                for (int k = 0; k < n; ++k)
                  pre[r][k] = u[col_sm(r) + k * PS];
            __syncwarp();
          }
        // z sweep in the column layout, then to shared memory
#if MGB_SLAB_ABLATE == 3
        T keep[NR][n];
#pragma unroll
        for (int r = 0; r < NR; ++r)
#pragma unroll
          for (int k = 0; k < n; ++k)
            keep[r][k] = pre[r][k];
#else
#pragma unroll
        for (int r = 0; r < NR; ++r)
          if (col_cl(r) < ncw)
            {
              fdm_forward<P, T, 1>(pre[r]);
#pragma unroll
              for (int k = 0; k < n; ++k)
                u[col_sm(r) + k * PS] = pre[r][k];
            }
#endif
        // tables(b+stride) have landed (issued one iteration ago): start the gather of the next batch
        cp_async_wait_all();
        __syncwarp();
        issue_gather(b + stride, tbuf + ((it + 1) % 3) * (CW * CTAB_STRIDE));
        issue_tables(b + 2 * stride, tbuf + ((it + 2) % 3) * (CW * CTAB_STRIDE));

        if (MGB_SLAB_ABLATE != 3 && cl < ncw)
          {
            T  reg[n2];
            T *pl = u + cl * CS + t * PS; // plane k = t (z mode t): reg[j*n+i]
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
        // transposed z sweep in the column layout and scatter-add
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
#if MGB_SLAB_ABLATE == 3
#pragma unroll
              for (int k = 0; k < n; ++k)
                res[k] = keep[r][k];
#else
#pragma unroll
              for (int k = 0; k < n; ++k)
                res[k] = u[col_sm(r) + k * PS];
              if (!warp_hn)
                fdm_backward<P, T, 1>(res);
#endif
              const unsigned level = (tb[col_cl(r) * CTAB_STRIDE + 27] >> 16) & 0xffu;
              const T        h     = T(root_length * __hiloint2double((1023 - (int)level) << 20, 0));
#pragma unroll
              for (int k = 0; k < n; ++k)
                {
                  const uint32_t idx = column_index(tb, r, k);
#if MGB_SLAB_ABLATE == 1
                  if (h * res[k] == T(-1.2345e300)) // never: keeps the arithmetic alive without the atomics
#else
                  if (!(idx & D_FLAG_BIT))
#endif
                    atomic_add(dst + idx, h * res[k]);
                }
            }
        if (ZERO && MGB_ZERO_POS == 1)
          zero_slice();
        __syncwarp(); // u is rewritten by the next iteration's columns
      }
  }

  // local diagonal of C^T A_e C by applying the cell operator to unit vectors (MatrixFreeTools::compute_diagonal)
  template <int P, typename T>
  __global__ void __launch_bounds__(CellsPerBlock<P>::value *(P + 1) * (P + 1))
    k_laplace_diagonal(const uint32_t *__restrict__ etab, const uint16_t *__restrict__ maskv, const uint8_t *__restrict__ levelv,
                       const double root_length, const long long n_cells, T *__restrict__ diag)
  {
    constexpr int CPB = CellsPerBlock<P>::value;
    constexpr int n = P + 1, n2 = n * n, n3 = n2 * n;
    __shared__ T        sA[CPB][n3];
    __shared__ T        sB[CPB][n3];
    __shared__ T        sD[CPB][n3];
    __shared__ uint32_t stab[CPB][27];

    const int       cl     = threadIdx.x / n2;
    const int       t      = threadIdx.x % n2;
    const long long cell   = (long long)blockIdx.x * CPB + cl;
    const bool      active = cell < n_cells;
    unsigned        mask   = 0;
    if (active)
      {
        for (int e = t; e < 27; e += n2)
          stab[cl][e] = etab[cell * 27 + e];
        if (maskv)
          mask = maskv[cell];
      }
    const int block_hn = __syncthreads_or(active && (mask >> 3) != 0);

    if (!block_hn)
      {
        // closed form: diag(A_e)_{ijk} = h (K_ii M_jj M_kk + M_ii K_jj M_kk + M_ii M_jj K_kk)
        if (active)
          for (int q = t; q < n3; q += n2)
            {
              const int    i = q % n, j = (q / n) % n, k = q / n2;
              const double mi = c_fe[P].M[i * MAXN + i], mj = c_fe[P].M[j * MAXN + j], mk = c_fe[P].M[k * MAXN + k];
              const double ki = c_fe[P].K[i * MAXN + i], kj = c_fe[P].K[j * MAXN + j], kk = c_fe[P].K[k * MAXN + k];
              sD[cl][q]       = T(ki * mj * mk + mi * kj * mk + mi * mj * kk);
            }
      }
    else
      {
        for (int col = 0; col < n3; ++col)
          {
            if (active)
              for (int q = t; q < n3; q += n2)
                sA[cl][q] = (q == col) ? T(1) : T(0);
            __syncthreads();
            hanging_node_sweeps<P, T, false>(sA[cl], mask, t, active);
            laplace_sweeps<P, T>(sA[cl], sB[cl], t, active);
            hanging_node_sweeps<P, T, true>(sA[cl], mask, t, active);
            if (active && t == 0)
              sD[cl][col] = sA[cl][col];
            __syncthreads();
          }
      }
    __syncthreads();
    if (active)
      {
        const T h = T(root_length / double(1u << levelv[cell]));
        for (int q = t; q < n3; q += n2)
          {
            const int i = q % n, j = (q / n) % n, k = q / n2;
            int       e, off;
            node_to_entity<P>(i, j, k, e, off);
            const uint32_t tb = stab[cl][e];
            if (!(tb & D_FLAG_BIT))
              atomic_add(diag + (tb & D_INDEX_MASK) + off, h * sD[cl][q]);
          }
      }
  }

  template <int P, typename T>
  void launch_apply(const CellData &cd, const T *src, T *dst, const uint32_t *cell_list, long long n_list, cudaStream_t s)
  {
    constexpr int   CW = SlabCfg<P>::CW, W = SlabCfg<P>::W;
    const long long n_work = cell_list ? n_list : cd.n_cells;
    if (n_work <= 0)
      return;
    const size_t smem = SlabSmem<P, T>::block_bytes(W);
    static int   blocks_per_sm = 0, n_sm = 0; // per (P, T) instantiation
    if (blocks_per_sm == 0)
      {
        int dev = 0;
        MGB_CUDA(cudaGetDevice(&dev));
        MGB_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
        MGB_CUDA(cudaFuncSetAttribute(k_laplace_apply<P, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        MGB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, k_laplace_apply<P, T>, W * 32, smem));
        if (blocks_per_sm < 1)
          throw std::runtime_error("k_laplace_apply does not fit on an SM");
      }
    const long long n_batches = (n_work + CW - 1) / CW;
    const long long blocks    = std::min<long long>((n_batches + W - 1) / W, (long long)blocks_per_sm * n_sm);
    if (cd.zero_ptr && cd.zero_n > 0)
      {
        if (reinterpret_cast<uintptr_t>(cd.zero_ptr) & 15u)
          throw std::runtime_error("k_laplace_apply: the vector to be zeroed is not 16-byte aligned");
        static bool zero_attr = false; // the ZERO instantiation needs its own opt-in to the dynamic shared memory size
        if (!zero_attr)
          {
            MGB_CUDA(cudaFuncSetAttribute(k_laplace_apply<P, T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            zero_attr = true;
          }
        const long long n16 = cd.zero_n / (long long)(16 / sizeof(T));
        k_laplace_apply<P, T, true><<<(unsigned)blocks, W * 32, smem, s>>>(cd.ctab, cd.any_hanging, cd.root_length, n_work, cell_list, src, dst,
                                                                          static_cast<T *>(cd.zero_ptr), cd.zero_n,
                                                                          (n16 + n_batches - 1) / n_batches);
      }
    else
      k_laplace_apply<P, T><<<(unsigned)blocks, W * 32, smem, s>>>(cd.ctab, cd.any_hanging, cd.root_length, n_work, cell_list, src, dst);
    MGB_CUDA(cudaGetLastError());
  }

  template <int P, typename T> void launch_diagonal(const CellData &cd, T *diag, cudaStream_t s)
  {
    constexpr int CPB = CellsPerBlock<P>::value;
    if (cd.n_cells <= 0)
      return;
    const long long blocks = (cd.n_cells + CPB - 1) / CPB;
    k_laplace_diagonal<P, T><<<(unsigned)blocks, CPB *(P + 1) * (P + 1), 0, s>>>(cd.etab, cd.any_hanging ? cd.mask : nullptr, cd.level,
                                                                                cd.root_length, cd.n_cells, diag);
    MGB_CUDA(cudaGetLastError());
  }
} // namespace mgb
