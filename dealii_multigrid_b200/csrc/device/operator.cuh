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
    static constexpr int W   = P <= 4 ? 4 : 2;  // warps per block
    static constexpr int CPB = CW * W;          // cells per block
  };

  // warp-level version of hanging_node_sweeps for the padded slab layout; `u` = the warp's cells, smask = their masks
  template <int P, typename T, bool TRANSPOSE>
  __device__ __forceinline__ void hanging_node_sweeps_slab(T *u, const uint16_t *smask, const int ncw, const int lane)
  {
    constexpr int n = P + 1, n2 = n * n;
    constexpr int PS = SlabLayout<n, sizeof(T)>::PS, CS = SlabLayout<n, sizeof(T)>::CS;
#pragma unroll 1
    for (int dd = 0; dd < 3; ++dd)
      {
        const int d  = TRANSPOSE ? 2 - dd : dd;
        const int da = d == 0 ? 1 : 0, db = d == 2 ? 1 : 2;
        for (int w = lane; w < ncw * n2; w += 32)
          {
            const int      cl = w / n2, t = w % n2;
            const unsigned mask = smask[cl];
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

  template <int P, typename T>
  __global__ void __launch_bounds__(SlabCfg<P>::W * 32)
    k_laplace_apply(const uint32_t *__restrict__ etab, const uint16_t *__restrict__ maskv, const uint8_t *__restrict__ levelv,
                    const double root_length, const long long n_work, const uint32_t *__restrict__ cell_list,
                    const T *__restrict__ src, T *__restrict__ dst)
  {
    constexpr int n = P + 1, n2 = n * n, n3 = n2 * n;
    constexpr int CW = SlabCfg<P>::CW, W = SlabCfg<P>::W;
    constexpr int PS = SlabLayout<n, sizeof(T)>::PS, CS = SlabLayout<n, sizeof(T)>::CS;
    __shared__ T        sU[W][CW * CS];
    __shared__ uint32_t stab[W][CW][27];
    __shared__ T        sh[W][CW];
    __shared__ uint16_t smask[W][CW];
    __shared__ uint32_t slut[n3]; // node q -> entity (bits 24..), offset in entity (bits 16..23), smem address (bits 0..15)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int q = threadIdx.x; q < n3; q += W * 32)
      {
        const int i = q % n, j = (q / n) % n, k = q / n2;
        int       e, off;
        node_to_entity<P>(i, j, k, e, off);
        slut[q] = (uint32_t(e) << 24) | (uint32_t(off) << 16) | uint32_t(k * PS + j * n + i);
      }
    const long long first = ((long long)blockIdx.x * W + warp) * CW;
    const int       ncw   = (int)max(0ll, min((long long)CW, n_work - first));
    for (int idx = lane; idx < ncw * 27; idx += 32)
      {
        const int       cl   = idx / 27, e = idx - cl * 27;
        const long long cell = cell_list ? (long long)cell_list[first + cl] : first + cl;
        stab[warp][cl][e]    = etab[cell * 27 + e];
      }
    unsigned mymask = 0;
    if (lane < ncw)
      {
        const long long cell = cell_list ? (long long)cell_list[first + lane] : first + lane;
        if (maskv)
          mymask = maskv[cell];
        smask[warp][lane] = (uint16_t)mymask;
        sh[warp][lane]    = T(root_length / double(1u << levelv[cell]));
      }
    const bool warp_hn = __any_sync(0xffffffffu, (mymask >> 3) != 0);
    __syncthreads();

    T *u = sU[warp];
    // gather (Dirichlet entities read 0): asynchronous copies straight into shared memory, all in flight at once
    for (int cl = 0; cl < ncw; ++cl)
#pragma unroll
      for (int q = lane; q < n3; q += 32)
        {
          const uint32_t lut  = slut[q];
          const uint32_t tb   = stab[warp][cl][lut >> 24];
          const bool     flag = (tb & D_FLAG_BIT) != 0;
          cp_async_entry<T>(u + cl * CS + (lut & 0xffffu), src + (flag ? 0u : (tb & D_INDEX_MASK) + ((lut >> 16) & 0xffu)), !flag);
        }
    cp_async_wait_all();
    __syncwarp();
    if (warp_hn)
      hanging_node_sweeps_slab<P, T, false>(u, smask[warp], ncw, lane);

    const int  cl = lane / n, t = lane - cl * n;
    const bool active = cl < ncw;
    T          reg[n2];
    if (active)
      {
        T *pl = u + cl * CS + t * PS; // plane k = t: reg[j*n+i]
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
        for (int r = 0; r < n2; ++r)
          pl[r] = reg[r];
      }
    __syncwarp();
    if (active)
      {
        T *col = u + cl * CS + t; // plane i = t: reg[k*n+j] = element (i, j, k)
#pragma unroll
        for (int k = 0; k < n; ++k)
#pragma unroll
          for (int j = 0; j < n; ++j)
            reg[k * n + j] = col[k * PS + j * n];
#pragma unroll
        for (int j = 0; j < n; ++j)
          fdm_forward<P, T, n>(reg + j);
        const T li = FDM<P, T>::lam(t);
#pragma unroll
        for (int k = 0; k < n; ++k)
#pragma unroll
          for (int j = 0; j < n; ++j)
            reg[k * n + j] *= (li + FDM<P, T>::lam2(j, k));
#pragma unroll
        for (int j = 0; j < n; ++j)
          fdm_backward<P, T, n>(reg + j);
#pragma unroll
        for (int k = 0; k < n; ++k)
#pragma unroll
          for (int j = 0; j < n; ++j)
            col[k * PS + j * n] = reg[k * n + j];
      }
    __syncwarp();
    if (active)
      {
        T *pl = u + cl * CS + t * PS;
#pragma unroll
        for (int r = 0; r < n2; ++r)
          reg[r] = pl[r];
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
      hanging_node_sweeps_slab<P, T, true>(u, smask[warp], ncw, lane);

    // scatter-add
    for (int c2 = 0; c2 < ncw; ++c2)
      {
        const T h = sh[warp][c2];
#pragma unroll
        for (int q = lane; q < n3; q += 32)
          {
            const uint32_t lut = slut[q];
            const uint32_t tb  = stab[warp][c2][lut >> 24];
            if (!(tb & D_FLAG_BIT))
              atomic_add(dst + (tb & D_INDEX_MASK) + ((lut >> 16) & 0xffu), h * u[c2 * CS + (lut & 0xffffu)]);
          }
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
    constexpr int   CPB    = SlabCfg<P>::CPB;
    const long long n_work = cell_list ? n_list : cd.n_cells;
    if (n_work <= 0)
      return;
    const long long blocks = (n_work + CPB - 1) / CPB;
    k_laplace_apply<P, T><<<(unsigned)blocks, SlabCfg<P>::W * 32, 0, s>>>(cd.etab, cd.any_hanging ? cd.mask : nullptr, cd.level,
                                                                          cd.root_length, n_work, cell_list, src, dst);
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
