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

  template <int P, typename T>
  __global__ void __launch_bounds__(CellsPerBlock<P>::value *(P + 1) * (P + 1))
    k_laplace_apply(const uint32_t *__restrict__ etab, const uint16_t *__restrict__ maskv, const uint8_t *__restrict__ levelv,
                    const double root_length, const long long n_work, const uint32_t *__restrict__ cell_list,
                    const T *__restrict__ src, T *__restrict__ dst)
  {
    constexpr int CPB = CellsPerBlock<P>::value;
    constexpr int n = P + 1, n2 = n * n, n3 = n2 * n;
    __shared__ T        sA[CPB][n3];
    __shared__ T        sB[CPB][n3];
    __shared__ uint32_t sidx[CPB][n3];
    __shared__ uint32_t stab[CPB][27];

    const int       cl     = threadIdx.x / n2;
    const int       t      = threadIdx.x % n2;
    const long long w      = (long long)blockIdx.x * CPB + cl;
    const bool      active = w < n_work;
    const long long cell   = active ? (cell_list ? (long long)cell_list[w] : w) : 0;

    unsigned mask = 0;
    if (active)
      {
        for (int e = t; e < 27; e += n2)
          stab[cl][e] = etab[cell * 27 + e];
        if (maskv)
          mask = maskv[cell];
      }
    const int block_hn = __syncthreads_or(active && (mask >> 3) != 0);

    // gather
    if (active)
      for (int q = t; q < n3; q += n2)
        {
          const int i = q % n, j = (q / n) % n, k = q / n2;
          int       e, off;
          node_to_entity<P>(i, j, k, e, off);
          const uint32_t tb  = stab[cl][e];
          const uint32_t idx = (tb & D_INDEX_MASK) + off;
          sidx[cl][q]        = idx | (tb & D_FLAG_BIT);
          sA[cl][q]          = (tb & D_FLAG_BIT) ? T(0) : src[idx];
        }
    __syncthreads();
    if (block_hn)
      hanging_node_sweeps<P, T, false>(sA[cl], mask, t, active);

    laplace_sweeps<P, T>(sA[cl], sB[cl], t, active);

    if (block_hn)
      hanging_node_sweeps<P, T, true>(sA[cl], mask, t, active);

    if (active)
      {
        const T h = T(root_length / double(1u << levelv[cell]));
        for (int q = t; q < n3; q += n2)
          {
            const uint32_t idx = sidx[cl][q];
            if (!(idx & D_FLAG_BIT))
              atomic_add(dst + idx, h * sA[cl][q]);
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
    constexpr int   CPB    = CellsPerBlock<P>::value;
    const long long n_work = cell_list ? n_list : cd.n_cells;
    if (n_work <= 0)
      return;
    const long long blocks = (n_work + CPB - 1) / CPB;
    k_laplace_apply<P, T><<<(unsigned)blocks, CPB *(P + 1) * (P + 1), 0, s>>>(cd.etab, cd.any_hanging ? cd.mask : nullptr, cd.level,
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
