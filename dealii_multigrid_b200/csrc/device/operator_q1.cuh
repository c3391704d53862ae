// Laplace cell kernel for general (non-Cartesian) hexahedra with trilinear geometry (MappingQ1) -- SURVEY.md 8(f) rank 3:
// what Operator::do_cell_integral_* (include/operator.h:450-493) does through FEEvaluation on a non-Cartesian
// MatrixFree: values at the Gauss points (3 sweeps), reference gradient in collocation form (3 sweeps), per-point
// metric G = w det(J) J^-1 J^-T, and the transposed 6 sweeps.  The metric is recomputed from the 8 cell vertices at
// every quadrature point (192 B per cell instead of 48 n^3 B of stored metric terms), so the kernel's HBM traffic
// stays at the Cartesian kernel's 16 B/DoF + records.
//
// Layout: CPB cells per block, n^2 threads per cell, thread t = (a, b) owns the n points (a, b, 0..n-1); the x and y
// sweeps go through shared memory.  This is the straightforward formulation (the Cartesian kernel in operator.cuh is
// the tuned one); hanging-node constraints reuse hanging_node_sweeps.
#pragma once
#include "cell_common.cuh"
#include "operator.cuh"

namespace mgb
{
  struct FEConstQ
  {
    double Bq[MAXN * MAXN]; // Bq[q*MAXN+i] = l_i(x_q)
    double Dq[MAXN * MAXN]; // Dq[q*MAXN+r] = d/dx of the r-th Lagrange polynomial on the Gauss points at x_q
    double xq[MAXN], wq[MAXN];
  };
  __constant__ FEConstQ c_feq[MAX_DEGREE + 1];

  // cells per block: 4-5 n^3 shared arrays per cell must fit the 48 KB of static shared memory
  template <int P> struct Q1Cells
  {
    static constexpr int value = P == 1 ? 32 : P == 2 ? 16 : P == 3 ? 8 : P == 4 ? 8 : P == 5 ? 4 : 2;
  };

  // out(a,b,k) = sum_r Mat[.][.] in(...) along dimension DIM for the n points of thread (a,b); TRANS applies Mat^T.
  // in/out are lexicographic n^3 arrays in shared memory (in != out); no barrier inside.
  template <int P, typename T, int DIM, bool TRANS>
  __device__ __forceinline__ void sweep_q1(const double *Mat, const T *in, T *out, const int a, const int b)
  {
    constexpr int n = P + 1, n2 = n * n;
#pragma unroll
    for (int k = 0; k < n; ++k)
      {
        const int row = DIM == 0 ? a : (DIM == 1 ? b : k);
        T         acc = 0;
#pragma unroll
        for (int r = 0; r < n; ++r)
          {
            const T m = T(TRANS ? Mat[r * MAXN + row] : Mat[row * MAXN + r]);
            const T v = DIM == 0 ? in[k * n2 + b * n + r] : (DIM == 1 ? in[k * n2 + r * n + a] : in[r * n2 + b * n + a]);
            acc += m * v;
          }
        out[k * n2 + b * n + a] = acc;
      }
  }

  // sA (nodal values, n^3) <- A_e sA for the hexahedron with vertices xv[8][3] (lexicographic, x fastest).
  // sB, sC, sD: n^3 scratch each.  All threads of the block call it (barriers inside).
  template <int P, typename T>
  __device__ __forceinline__ void laplace_q1_cell(T *sA, T *sB, T *sC, T *sD, const double *xv, const int t, const bool active)
  {
    constexpr int n = P + 1, n2 = n * n;
    const int     a = t % n, b = t / n;
    const bool    on = active && t < n2;
    const double *B = c_feq[P].Bq, *D = c_feq[P].Dq;
    // values at the Gauss points: x, y, z
    if (on)
      sweep_q1<P, T, 0, false>(B, sA, sB, a, b);
    __syncthreads();
    if (on)
      sweep_q1<P, T, 1, false>(B, sB, sA, a, b);
    __syncthreads();
    if (on)
      sweep_q1<P, T, 2, false>(B, sA, sB, a, b); // sB = u at the Gauss points
    __syncthreads();
    T tx[n], ty[n], tz[n];
    if (on)
      {
        // reference gradient (collocation derivative) at this thread's points
#pragma unroll
        for (int k = 0; k < n; ++k)
          {
            T gx = 0, gy = 0, gz = 0;
#pragma unroll
            for (int r = 0; r < n; ++r)
              {
                gx += T(D[a * MAXN + r]) * sB[k * n2 + b * n + r];
                gy += T(D[b * MAXN + r]) * sB[k * n2 + r * n + a];
                gz += T(D[k * MAXN + r]) * sB[r * n2 + b * n + a];
              }
            // Jacobian of the trilinear map at (xq[a], xq[b], xq[k]): J[c][d] = d x_c / d xi_d
            const double xi = c_feq[P].xq[a], eta = c_feq[P].xq[b], zeta = c_feq[P].xq[k];
            const double lx[2] = {1.0 - xi, xi}, ly[2] = {1.0 - eta, eta}, lz[2] = {1.0 - zeta, zeta};
            double       J[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
#pragma unroll
            for (int v = 0; v < 8; ++v)
              {
                const int    vx = v & 1, vy = (v >> 1) & 1, vz = v >> 2;
                const double dx = (vx ? 1.0 : -1.0) * ly[vy] * lz[vz], dy = lx[vx] * (vy ? 1.0 : -1.0) * lz[vz],
                             dz = lx[vx] * ly[vy] * (vz ? 1.0 : -1.0);
#pragma unroll
                for (int c = 0; c < 3; ++c)
                  {
                    J[c][0] += xv[v * 3 + c] * dx;
                    J[c][1] += xv[v * 3 + c] * dy;
                    J[c][2] += xv[v * 3 + c] * dz;
                  }
              }
            // cofactors: inv(J) = C^T / det
            const double C00 = J[1][1] * J[2][2] - J[1][2] * J[2][1], C01 = J[1][2] * J[2][0] - J[1][0] * J[2][2],
                         C02 = J[1][0] * J[2][1] - J[1][1] * J[2][0];
            const double C10 = J[0][2] * J[2][1] - J[0][1] * J[2][2], C11 = J[0][0] * J[2][2] - J[0][2] * J[2][0],
                         C12 = J[0][1] * J[2][0] - J[0][0] * J[2][1];
            const double C20 = J[0][1] * J[1][2] - J[0][2] * J[1][1], C21 = J[0][2] * J[1][0] - J[0][0] * J[1][2],
                         C22 = J[0][0] * J[1][1] - J[0][1] * J[1][0];
            const double det = J[0][0] * C00 + J[0][1] * C01 + J[0][2] * C02;
            const double s   = c_feq[P].wq[a] * c_feq[P].wq[b] * c_feq[P].wq[k] / det;
            // physical gradient * det = C g (C[c][d] = cofactor of J[c][d]); flux = C^T (C g) * w / det
            const double g0 = double(gx), g1 = double(gy), g2 = double(gz);
            const double p0 = C00 * g0 + C01 * g1 + C02 * g2, p1 = C10 * g0 + C11 * g1 + C12 * g2,
                         p2 = C20 * g0 + C21 * g1 + C22 * g2;
            tx[k] = T(s * (C00 * p0 + C10 * p1 + C20 * p2));
            ty[k] = T(s * (C01 * p0 + C11 * p1 + C21 * p2));
            tz[k] = T(s * (C02 * p0 + C12 * p1 + C22 * p2));
          }
#pragma unroll
        for (int k = 0; k < n; ++k)
          {
            sC[k * n2 + b * n + a] = tx[k];
            sD[k * n2 + b * n + a] = ty[k];
          }
      }
    __syncthreads();
    if (on)
      {
        // transposed collocation derivative: v = Dx^T tx + Dy^T ty + Dz^T tz
#pragma unroll
        for (int k = 0; k < n; ++k)
          {
            T acc = 0;
#pragma unroll
            for (int r = 0; r < n; ++r)
              {
                acc += T(D[r * MAXN + a]) * sC[k * n2 + b * n + r];
                acc += T(D[r * MAXN + b]) * sD[k * n2 + r * n + a];
                acc += T(D[r * MAXN + k]) * tz[r];
              }
            sA[k * n2 + b * n + a] = acc;
          }
      }
    __syncthreads();
    // back to the nodal basis: z, y, x
    if (on)
      sweep_q1<P, T, 2, true>(B, sA, sB, a, b);
    __syncthreads();
    if (on)
      sweep_q1<P, T, 1, true>(B, sB, sA, a, b);
    __syncthreads();
    if (on)
      sweep_q1<P, T, 0, true>(B, sA, sB, a, b);
    __syncthreads();
    if (on)
#pragma unroll
      for (int k = 0; k < n; ++k)
        sA[k * n2 + b * n + a] = sB[k * n2 + b * n + a];
    __syncthreads();
  }

  // MODE 0: dst += A_cells src; MODE 1: diag += diag(C^T A_e C) by unit vectors (MatrixFreeTools::compute_diagonal)
  template <int P, typename T, int MODE>
  __global__ void __launch_bounds__(Q1Cells<P>::value *(P + 1) * (P + 1))
    k_laplace_q1(const uint32_t *__restrict__ etab, const uint16_t *__restrict__ maskv, const double *__restrict__ vertices,
                 const long long n_work, const uint32_t *__restrict__ cell_list, const T scale, const T *__restrict__ src,
                 T *__restrict__ dst)
  {
    constexpr int CPB = Q1Cells<P>::value;
    constexpr int n = P + 1, n2 = n * n, n3 = n2 * n;
    __shared__ T        sA[CPB][n3];
    __shared__ T        sB[CPB][n3];
    __shared__ T        sC[CPB][n3];
    __shared__ T        sD[CPB][n3];
    __shared__ T        sR[MODE == 1 ? CPB * n3 : 1]; // diagonal entries of the block's cells
    __shared__ uint32_t stab[CPB][27];
    __shared__ double   sxv[CPB][24];

    const int cl = threadIdx.x / n2;
    const int t  = threadIdx.x % n2;
    for (long long base = (long long)blockIdx.x * CPB; base < n_work; base += (long long)gridDim.x * CPB)
      {
        const long long w      = base + cl;
        const bool      active = w < n_work;
        const long long cell   = active ? (cell_list ? (long long)cell_list[w] : w) : 0;
        unsigned        mask   = 0;
        if (active)
          {
            for (int e = t; e < 27; e += n2)
              stab[cl][e] = etab[cell * 27 + e];
            for (int e = t; e < 24; e += n2)
              sxv[cl][e] = vertices[cell * 24 + e];
            if (maskv)
              mask = maskv[cell];
          }
        __syncthreads();
        if constexpr (MODE == 0)
          {
            if (active)
              for (int q = t; q < n3; q += n2)
                {
                  int e, off;
                  node_to_entity<P>(q % n, (q / n) % n, q / n2, e, off);
                  const uint32_t tb = stab[cl][e];
                  sA[cl][q]         = (tb & D_FLAG_BIT) ? T(0) : src[(tb & D_INDEX_MASK) + off];
                }
            __syncthreads();
            hanging_node_sweeps<P, T, false>(sA[cl], mask, t, active);
            laplace_q1_cell<P, T>(sA[cl], sB[cl], sC[cl], sD[cl], sxv[cl], t, active);
            hanging_node_sweeps<P, T, true>(sA[cl], mask, t, active);
            if (active)
              for (int q = t; q < n3; q += n2)
                {
                  int e, off;
                  node_to_entity<P>(q % n, (q / n) % n, q / n2, e, off);
                  const uint32_t tb = stab[cl][e];
                  if (!(tb & D_FLAG_BIT))
                    atomic_add(dst + (tb & D_INDEX_MASK) + off, scale * sA[cl][q]);
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
                laplace_q1_cell<P, T>(sA[cl], sB[cl], sC[cl], sD[cl], sxv[cl], t, active);
                hanging_node_sweeps<P, T, true>(sA[cl], mask, t, active);
                if (active && t == 0)
                  sR[cl * n3 + col] = sA[cl][col];
                __syncthreads();
              }
            if (active)
              for (int q = t; q < n3; q += n2)
                {
                  int e, off;
                  node_to_entity<P>(q % n, (q / n) % n, q / n2, e, off);
                  const uint32_t tb = stab[cl][e];
                  if (!(tb & D_FLAG_BIT))
                    atomic_add(dst + (tb & D_INDEX_MASK) + off, sR[cl * n3 + q]);
                }
          }
        __syncthreads(); // stab / sxv are rewritten by the next iteration
      }
  }

  template <int P, typename T, int MODE>
  void launch_q1(const CellData &cd, const T *src, T *dst, const uint32_t *cell_list, long long n_list, cudaStream_t s)
  {
    constexpr int   CPB    = Q1Cells<P>::value;
    const long long n_work = cell_list ? n_list : cd.n_cells;
    if (n_work <= 0)
      return;
    const long long blocks = std::min<long long>((n_work + CPB - 1) / CPB, 148LL * 16);
    k_laplace_q1<P, T, MODE><<<(unsigned)blocks, CPB *(P + 1) * (P + 1), 0, s>>>(cd.etab, cd.any_hanging ? cd.mask : nullptr, cd.vertices,
                                                                                 n_work, cell_list, T(cd.scale), src, dst);
    MGB_CUDA(cudaGetLastError());
  }
} // namespace mgb
