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

#include <cstdlib>
#include <string>

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

  // ------------------------------------------------------------------------------------------------------------
  // Tuned general-geometry kernel ("q1 slab"): the persistent-warp column/slab skeleton of k_laplace_apply
  // (operator.cuh) around the quadrature-point form of the operator.  With B = nodal basis at the Gauss points and
  // G = its derivative there (both n x n, 1-D),
  //   reference gradient  gx = G_x B_y B_z u,  gy = B_x G_y B_z u,  gz = B_x B_y G_z u
  //   flux at a point     t  = w C^T C g / det J          (C = cofactors of J, J from the trilinear vertex map)
  //   result              r  = B_z^T (G_x^T B_y^T tx + B_x^T G_y^T ty) + G_z^T (B_x^T B_y^T tz)
  // i.e. 16 one-dimensional sweeps, each in even-odd form.  Column layout (lane <-> one (i,j) column, z in registers):
  // gather, the two z sweeps (B_z u and G_z u into two shared arrays), their transposes and the scatter.  Slab layout
  // (lane <-> plane k of a cell, n^2 values in registers): the x/y sweeps and the metric.  The Jacobian at a point
  // comes from the 8 coefficient vectors of the trilinear map (Moebius transform of the 8 vertices, once per cell in
  // shared memory): 15 FMA per point instead of 8 x 15.  The metric is applied in a loop that is rolled over the
  // y index with the three gradient planes in shared memory (each thread only touches its own plane): the fully
  // unrolled register form has 6952 SASS instructions per batch and stalls on instruction fetch (ncu: no_instruction
  // 4 of 10 cycles per issue, profiles/r1s_ncu_q1_slab_L6.txt); the warps of a block also walk through the batches
  // in lockstep so that they share the fetched lines.  1/det by two Newton steps on the FP32 reciprocal.
  // Measured (hypercube L=6 p=4 FP64, smooth deformation): 2.15 ms for k_laplace_q1, 0.97 ms register form,
  // 0.50 ms this form (profiles/r1t_time_vmult_q1_variants.txt).  Degrees 1..4; higher degrees and the diagonal use
  // k_laplace_q1 above.
  // Build-time variants (measured in profiles/r1t_time_vmult_q1_variants.txt): the defaults are the rolled metric loop
  // through shared memory and block-lockstep batches; -DMGB_Q1_REG_METRIC / -DMGB_Q1_NO_LOCKSTEP restore the first form.
#if !defined(MGB_Q1_REG_METRIC) && !defined(MGB_Q1_SMEM_METRIC)
#define MGB_Q1_SMEM_METRIC
#endif
#if !defined(MGB_Q1_NO_LOCKSTEP) && !defined(MGB_Q1_LOCKSTEP)
#define MGB_Q1_LOCKSTEP
#endif
  struct FEConstQEO
  {
    // even-odd tables of 0: B, 1: G, 2: B^T, 3: G^T.  For A with A[n-1-q][n-1-l] = sigma A[q][l] (sigma = +1 for B, -1 for G),
    // row q < (n+1)/2:  m[q][l] = (A[q][l] + A[q][n-1-l])/2,  m[q][n-1-l] = (A[q][l] - A[q][n-1-l])/2  (l < n/2),
    //                   m[q][n/2] = A[q][n/2] (n odd)
    double m[4][MAXN * MAXN];
    float  mf[4][MAXN * MAXN];
  };
  __constant__ FEConstQEO c_feqeo[MAX_DEGREE + 1];

  template <int P, int MAT, typename T> struct QEO;
  template <int P, int MAT> struct QEO<P, MAT, double>
  {
    static __device__ __forceinline__ double at(int q, int l) { return c_feqeo[P].m[MAT][q * MAXN + l]; }
  };
  template <int P, int MAT> struct QEO<P, MAT, float>
  {
    static __device__ __forceinline__ float at(int q, int l) { return c_feqeo[P].mf[MAT][q * MAXN + l]; }
  };

  // v <- A v on a register line v[0], v[S], ..., v[(n-1) S]; MAT as in FEConstQEO
  template <int P, typename T, int S, int MAT> __device__ __forceinline__ void eo_sweep(T *v)
  {
    constexpr int  n = P + 1, no = n / 2, mid = n / 2;
    constexpr bool odd = (n % 2) == 1, antisym = (MAT & 1) != 0;
    T              e[no > 0 ? no : 1], o[no > 0 ? no : 1];
#pragma unroll
    for (int l = 0; l < no; ++l)
      {
        e[l] = v[l * S] + v[(n - 1 - l) * S];
        o[l] = v[l * S] - v[(n - 1 - l) * S];
      }
    const T vm = odd ? v[mid * S] : T(0);
#pragma unroll
    for (int q = 0; q < no; ++q)
      {
        T X = QEO<P, MAT, T>::at(q, 0) * e[0], Y = QEO<P, MAT, T>::at(q, n - 1) * o[0];
#pragma unroll
        for (int l = 1; l < no; ++l)
          {
            X += QEO<P, MAT, T>::at(q, l) * e[l];
            Y += QEO<P, MAT, T>::at(q, n - 1 - l) * o[l];
          }
        if (odd)
          X += QEO<P, MAT, T>::at(q, mid) * vm;
        v[q * S]           = X + Y;
        v[(n - 1 - q) * S] = antisym ? Y - X : X - Y;
      }
    if (odd)
      {
        T r;
        if (antisym)
          {
            r = QEO<P, MAT, T>::at(mid, n - 1) * o[0];
#pragma unroll
            for (int l = 1; l < no; ++l)
              r += QEO<P, MAT, T>::at(mid, n - 1 - l) * o[l];
          }
        else
          {
            r = QEO<P, MAT, T>::at(mid, mid) * vm;
#pragma unroll
            for (int l = 0; l < no; ++l)
              r += QEO<P, MAT, T>::at(mid, l) * e[l];
          }
        v[mid * S] = r;
      }
  }

  template <int P> struct Q1SlabCfg
  {
    static constexpr int n   = P + 1;
    static constexpr int CW  = 32 / n; // cells per warp
#ifndef MGB_Q1_W
#define MGB_Q1_W 4
#endif
#ifndef MGB_Q1_MINB
#define MGB_Q1_MINB 2
#endif
    static constexpr int W   = MGB_Q1_W; // warps per block
    static constexpr int MIN_BLOCKS = P == 1 ? 3 : MGB_Q1_MINB; // resident blocks per SM the register allocation must allow
  };
  template <int P, typename T> struct Q1SlabSmem
  {
    static constexpr int    n  = P + 1;
    static constexpr int    CW = Q1SlabCfg<P>::CW;
    static constexpr int    CS = SlabLayout<n, sizeof(T)>::CS;
    static constexpr size_t u_bytes    = (size_t(CW * CS) * sizeof(T) + 15) / 16 * 16;
    static constexpr size_t tab_bytes  = 2 * size_t(CW) * CTAB_STRIDE * 4;
#ifdef MGB_Q1_SMEM_METRIC
    static constexpr size_t g_bytes    = size_t(n * n) * 32 * sizeof(T); // gy of the slab phase, [n^2][lane]
#else
    static constexpr size_t g_bytes    = 0;
#endif
    static constexpr size_t warp_bytes = 2 * u_bytes + tab_bytes + size_t(CW) * 24 * sizeof(double) + g_bytes;
  };

  template <int P, typename T>
  __global__ void __launch_bounds__(Q1SlabCfg<P>::W * 32, Q1SlabCfg<P>::MIN_BLOCKS)
    k_laplace_q1_slab(const uint32_t *__restrict__ ctab, const int any_hanging, const double *__restrict__ vertices, const T scale,
                      const long long n_work, const uint32_t *__restrict__ cell_list, const T *__restrict__ src, T *__restrict__ dst)
  {
    constexpr int n = P + 1, n2 = n * n;
    constexpr int CW = Q1SlabCfg<P>::CW, W = Q1SlabCfg<P>::W;
    constexpr int PS = SlabLayout<n, sizeof(T)>::PS, CS = SlabLayout<n, sizeof(T)>::CS;
    constexpr int NR = (CW * n2 + 31) / 32; // rounds of the column layout
    using SM = Q1SlabSmem<P, T>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int      warp  = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char *wbase = smem_raw + warp * SM::warp_bytes;
    T             *u1    = reinterpret_cast<T *>(wbase);                                       // [CW*CS]  B_z u   / v1
    T             *u2    = reinterpret_cast<T *>(wbase + SM::u_bytes);                         // [CW*CS]  G_z u   / v2
    uint32_t      *tbuf  = reinterpret_cast<uint32_t *>(wbase + 2 * SM::u_bytes);              // [2][CW][32]
    double        *cf    = reinterpret_cast<double *>(wbase + 2 * SM::u_bytes + SM::tab_bytes); // [CW][8][3]
#ifdef MGB_Q1_SMEM_METRIC
    T *u3 = reinterpret_cast<T *>(wbase + 2 * SM::u_bytes + SM::tab_bytes + size_t(CW) * 24 * sizeof(double)) + lane; // [n^2][32]
#endif

    const long long n_batches = (n_work + CW - 1) / CW;
    const long long stride    = (long long)gridDim.x * W;
    long long       b         = (long long)blockIdx.x * W + warp;
#ifdef MGB_Q1_LOCKSTEP
    // the warps of a block walk through the (fully unrolled, > 100 KB) loop body together, so that they share the
    // instruction-cache lines: same trip count for all warps, one barrier per iteration
    const long long n_it = (n_batches - (long long)blockIdx.x * W + stride - 1) / stride;
#else
    if (b >= n_batches)
      return;
#endif

    // per-lane description of the columns of each round (as in k_laplace_apply):
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
    const auto column_index = [&](const uint32_t *tb, const int r, const int k) -> uint32_t {
      const uint32_t e = tb[(col[r] & 31u) * CTAB_STRIDE + ((col[r] >> 5) & 15u) + (k == 0 ? 0 : (k == P ? 18 : 9))];
      return e + ((col[r] >> 9) & 31u) + ((k > 0 && k < P) ? (k - 1) * ((col[r] >> 14) & 31u) : 0);
    };

    issue_tables(b, tbuf);
    const int    cl = lane / n, t = lane - cl * n; // slab layout: plane t of cell cl
    const double zeta = c_feq[P].xq[t < n ? t : 0];
#ifdef MGB_Q1_LOCKSTEP
    for (int it = 0; it < (int)n_it; b += stride, ++it)
      {
        __syncthreads();
        if (b >= n_batches)
          continue;
#else
    int          it = 0;
    for (; b < n_batches; b += stride, ++it)
      {
#endif
        uint32_t       *tb    = tbuf + (it & 1) * (CW * CTAB_STRIDE);
        const long long first = b * CW;
        const int       ncw   = (int)min((long long)CW, n_work - first);
        // vertices of the batch's cells (issued before the wait on the records)
        for (int idx = lane; idx < ncw * 24; idx += 32)
          {
            const int       c    = idx / 24;
            const long long cell = cell_list ? (long long)cell_list[first + c] : first + c;
            cf[idx]              = __ldg(vertices + cell * 24 + (idx - c * 24));
          }
        cp_async_wait_all();
        __syncwarp();
        issue_tables(b + stride, tbuf + ((it + 1) & 1) * (CW * CTAB_STRIDE));
        // gather (column layout)
        T pre[NR][n];
#pragma unroll
        for (int r = 0; r < NR; ++r)
#pragma unroll
          for (int k = 0; k < n; ++k)
            {
              pre[r][k] = T(0);
              if (col_cl(r) < ncw)
                {
                  const uint32_t idx = column_index(tb, r, k);
                  if (!(idx & D_FLAG_BIT))
                    pre[r][k] = __ldg(src + idx);
                }
            }
        // vertices -> coefficient vectors of the trilinear map, in place: entry s (bit 0: xi, 1: eta, 2: zeta) of component d
        for (int w = lane; w < ncw * 3; w += 32) // (cell, component) pairs: more than 32 of them at degree 1
          {
            double *c = cf + (w / 3) * 24 + (w % 3);
            double  x[8];
#pragma unroll
            for (int v = 0; v < 8; ++v)
              x[v] = c[v * 3];
#pragma unroll
            for (int bit = 1; bit < 8; bit <<= 1)
#pragma unroll
              for (int v = 0; v < 8; ++v)
                if (v & bit)
                  x[v] -= x[v ^ bit];
#pragma unroll
            for (int v = 0; v < 8; ++v)
              c[v * 3] = x[v];
          }
        bool warp_hn = false;
        if (any_hanging)
          {
            const unsigned mymask = lane < ncw ? (tb[lane * CTAB_STRIDE + 27] & 0xffffu) : 0u;
            warp_hn               = __any_sync(0xffffffffu, (mymask >> 3) != 0);
          }
        if (warp_hn)
          {
#pragma unroll
            for (int r = 0; r < NR; ++r)
              if (col_cl(r) < ncw)
#pragma unroll
                for (int k = 0; k < n; ++k)
                  u1[col_sm(r) + k * PS] = pre[r][k];
            __syncwarp();
            hanging_node_sweeps_slab<P, T, false>(u1, tb, ncw, lane);
#pragma unroll
            for (int r = 0; r < NR; ++r)
              if (col_cl(r) < ncw)
#pragma unroll
                for (int k = 0; k < n; ++k)
                  pre[r][k] = u1[col_sm(r) + k * PS];
            __syncwarp();
          }
        // z sweeps in the column layout: u1 = B_z u, u2 = G_z u
#pragma unroll
        for (int r = 0; r < NR; ++r)
          if (col_cl(r) < ncw)
            {
              T dz[n];
#pragma unroll
              for (int k = 0; k < n; ++k)
                dz[k] = pre[r][k];
              eo_sweep<P, T, 1, 0>(pre[r]);
              eo_sweep<P, T, 1, 1>(dz);
#pragma unroll
              for (int k = 0; k < n; ++k)
                {
                  u1[col_sm(r) + k * PS] = pre[r][k];
                  u2[col_sm(r) + k * PS] = dz[k];
                }
            }
        __syncwarp();
        if (cl < ncw)
          {
            T  ga[n2], gb[n2];
            T *p1 = u1 + cl * CS + t * PS, *p2 = u2 + cl * CS + t * PS; // plane k = t: reg[j*n+i]
            {
              // gz = B_y B_x (G_z u), kept in shared memory (this thread's own plane) while the metric is applied
              T gc[n2];
#pragma unroll
              for (int r = 0; r < n2; ++r)
                gc[r] = p2[r];
#pragma unroll
              for (int j = 0; j < n; ++j)
                eo_sweep<P, T, 1, 0>(gc + j * n);
#pragma unroll
              for (int i = 0; i < n; ++i)
                eo_sweep<P, T, n, 0>(gc + i);
#pragma unroll
              for (int r = 0; r < n2; ++r)
                p2[r] = gc[r];
            }
#pragma unroll
            for (int r = 0; r < n2; ++r)
              ga[r] = gb[r] = p1[r];
#pragma unroll
            for (int j = 0; j < n; ++j)
              {
                eo_sweep<P, T, 1, 1>(ga + j * n); // G_x
                eo_sweep<P, T, 1, 0>(gb + j * n); // B_x
              }
#pragma unroll
            for (int i = 0; i < n; ++i)
              {
                eo_sweep<P, T, n, 0>(ga + i); // gx = B_y G_x
                eo_sweep<P, T, n, 1>(gb + i); // gy = G_y B_x
              }
#ifdef MGB_Q1_SMEM_METRIC
#pragma unroll
            for (int r = 0; r < n2; ++r)
              {
                p1[r]      = ga[r];
                u3[r * 32] = gb[r];
              }
            // metric at the n^2 points of the plane
            // MT: arithmetic type of the metric (FP64; -DMGB_Q1_FLOAT_METRIC: FP32 in the float kernels, an experiment)
#ifdef MGB_Q1_FLOAT_METRIC
            using MT = T;
#else
            using MT = double;
#endif
            const double *c  = cf + cl * 24;
            const MT      wz = MT(c_feq[P].wq[t]);
            MT            A0[3], A1[3], B0[3];
#pragma unroll
            for (int d = 0; d < 3; ++d)
              {
                A0[d] = MT(c[1 * 3 + d]) + MT(c[5 * 3 + d]) * MT(zeta); // d x_d / d xi   = A0 + A1 eta
                A1[d] = MT(c[3 * 3 + d]) + MT(c[7 * 3 + d]) * MT(zeta);
                B0[d] = MT(c[2 * 3 + d]) + MT(c[6 * 3 + d]) * MT(zeta); // d x_d / d eta  = B0 + A1 xi
              }
#pragma unroll 1
            for (int j = 0; j < n; ++j) // rolled: the gradients of the plane live in shared memory
              {
                const MT eta = MT(c_feq[P].xq[j]);
                MT       J0[3], E0[3], F0[3];
#pragma unroll
                for (int d = 0; d < 3; ++d)
                  {
                    J0[d] = A0[d] + A1[d] * eta;
                    E0[d] = MT(c[4 * 3 + d]) + MT(c[6 * 3 + d]) * eta; // d x_d / d zeta = E0 + F0 xi
                    F0[d] = MT(c[5 * 3 + d]) + MT(c[7 * 3 + d]) * eta;
                  }
#pragma unroll
                for (int i = 0; i < n; ++i)
                  {
                    const MT xi = MT(c_feq[P].xq[i]);
                    MT       J1[3], J2[3];
#pragma unroll
                    for (int d = 0; d < 3; ++d)
                      {
                        J1[d] = B0[d] + A1[d] * xi;
                        J2[d] = E0[d] + F0[d] * xi;
                      }
                    // J[c][0] = J0[c], J[c][1] = J1[c], J[c][2] = J2[c]; cofactors C[c][d]
                    const MT C00 = J1[1] * J2[2] - J2[1] * J1[2], C01 = J2[1] * J0[2] - J0[1] * J2[2],
                             C02 = J0[1] * J1[2] - J1[1] * J0[2];
                    const MT C10 = J2[0] * J1[2] - J1[0] * J2[2], C11 = J0[0] * J2[2] - J2[0] * J0[2],
                             C12 = J1[0] * J0[2] - J0[0] * J1[2];
                    const MT C20 = J1[0] * J2[1] - J2[0] * J1[1], C21 = J2[0] * J0[1] - J0[0] * J2[1],
                             C22 = J0[0] * J1[1] - J1[0] * J0[1];
                    const MT det = J0[0] * C00 + J1[0] * C01 + J2[0] * C02;
#ifndef MGB_Q1_EXACT_DIV
                    MT rdet = MT(1.0f / float(det));     // two Newton steps: 6e-8 -> 4e-15 -> 1e-29 relative
                    rdet        = rdet * (MT(2) - det * rdet);
                    rdet        = rdet * (MT(2) - det * rdet);
                    const MT s = MT(c_feq[P].wq[i] * c_feq[P].wq[j]) * wz * rdet;
#else
                    const MT s = MT(c_feq[P].wq[i] * c_feq[P].wq[j]) * wz / det;
#endif
                    const int    r   = j * n + i;
                    const MT g0 = MT(p1[r]), g1 = MT(u3[r * 32]), g2 = MT(p2[r]);
                    const MT p0 = C00 * g0 + C01 * g1 + C02 * g2, p1v = C10 * g0 + C11 * g1 + C12 * g2,
                                 p2v = C20 * g0 + C21 * g1 + C22 * g2;
                    p1[r] = T(s * (C00 * p0 + C10 * p1v + C20 * p2v));
                    u3[r * 32] = T(s * (C01 * p0 + C11 * p1v + C21 * p2v));
                    p2[r] = T(s * (C02 * p0 + C12 * p1v + C22 * p2v));
                  }
              }
#pragma unroll
            for (int r = 0; r < n2; ++r)
              {
                ga[r] = p1[r];
                gb[r] = u3[r * 32];
              }
#else
            // metric at the n^2 points of the plane
            const double *c  = cf + cl * 24;
            const double  wz = c_feq[P].wq[t];
            double        A0[3], A1[3], B0[3];
#pragma unroll
            for (int d = 0; d < 3; ++d)
              {
                A0[d] = c[1 * 3 + d] + c[5 * 3 + d] * zeta; // d x_d / d xi   = A0 + A1 eta
                A1[d] = c[3 * 3 + d] + c[7 * 3 + d] * zeta;
                B0[d] = c[2 * 3 + d] + c[6 * 3 + d] * zeta; // d x_d / d eta  = B0 + A1 xi
              }
#pragma unroll
            for (int i = 0; i < n; ++i)
              {
                const double xi = c_feq[P].xq[i];
                double       J1[3], E[3], F[3];
#pragma unroll
                for (int d = 0; d < 3; ++d)
                  {
                    J1[d] = B0[d] + A1[d] * xi;
                    E[d]  = c[4 * 3 + d] + c[5 * 3 + d] * xi; // d x_d / d zeta = E + F eta
                    F[d]  = c[6 * 3 + d] + c[7 * 3 + d] * xi;
                  }
#pragma unroll
                for (int j = 0; j < n; ++j)
                  {
                    const double eta = c_feq[P].xq[j];
                    double       J0[3], J2[3];
#pragma unroll
                    for (int d = 0; d < 3; ++d)
                      {
                        J0[d] = A0[d] + A1[d] * eta;
                        J2[d] = E[d] + F[d] * eta;
                      }
                    // J[c][0] = J0[c], J[c][1] = J1[c], J[c][2] = J2[c]; cofactors C[c][d]
                    const double C00 = J1[1] * J2[2] - J2[1] * J1[2], C01 = J2[1] * J0[2] - J0[1] * J2[2],
                                 C02 = J0[1] * J1[2] - J1[1] * J0[2];
                    const double C10 = J2[0] * J1[2] - J1[0] * J2[2], C11 = J0[0] * J2[2] - J2[0] * J0[2],
                                 C12 = J1[0] * J0[2] - J0[0] * J1[2];
                    const double C20 = J1[0] * J2[1] - J2[0] * J1[1], C21 = J2[0] * J0[1] - J0[0] * J2[1],
                                 C22 = J0[0] * J1[1] - J1[0] * J0[1];
                    const double det = J0[0] * C00 + J1[0] * C01 + J2[0] * C02;
#ifndef MGB_Q1_EXACT_DIV
                    double rdet = double(1.0f / float(det)); // two Newton steps: 6e-8 -> 4e-15 -> 1e-29 relative
                    rdet        = rdet * (2.0 - det * rdet);
                    rdet        = rdet * (2.0 - det * rdet);
                    const double s = c_feq[P].wq[i] * c_feq[P].wq[j] * wz * rdet;
#else
                    const double s = c_feq[P].wq[i] * c_feq[P].wq[j] * wz / det;
#endif
                    const int    r   = j * n + i;
                    const double g0 = double(ga[r]), g1 = double(gb[r]), g2 = double(p2[r]);
                    const double p0 = C00 * g0 + C01 * g1 + C02 * g2, p1v = C10 * g0 + C11 * g1 + C12 * g2,
                                 p2v = C20 * g0 + C21 * g1 + C22 * g2;
                    ga[r] = T(s * (C00 * p0 + C10 * p1v + C20 * p2v));
                    gb[r] = T(s * (C01 * p0 + C11 * p1v + C21 * p2v));
                    p2[r] = T(s * (C02 * p0 + C12 * p1v + C22 * p2v));
                  }
              }
#endif
            // transposed y and x sweeps
#pragma unroll
            for (int i = 0; i < n; ++i)
              {
                eo_sweep<P, T, n, 2>(ga + i); // B_y^T tx
                eo_sweep<P, T, n, 3>(gb + i); // G_y^T ty
              }
#pragma unroll
            for (int j = 0; j < n; ++j)
              {
                eo_sweep<P, T, 1, 3>(ga + j * n); // G_x^T
                eo_sweep<P, T, 1, 2>(gb + j * n); // B_x^T
              }
#pragma unroll
            for (int r = 0; r < n2; ++r)
              p1[r] = ga[r] + gb[r];
            {
              T gc[n2];
#pragma unroll
              for (int r = 0; r < n2; ++r)
                gc[r] = p2[r];
#pragma unroll
              for (int i = 0; i < n; ++i)
                eo_sweep<P, T, n, 2>(gc + i); // B_y^T tz
#pragma unroll
              for (int j = 0; j < n; ++j)
                eo_sweep<P, T, 1, 2>(gc + j * n); // B_x^T
#pragma unroll
              for (int r = 0; r < n2; ++r)
                p2[r] = gc[r];
            }
          }
        __syncwarp();
        // transposed z sweeps in the column layout: r = B_z^T v1 + G_z^T v2, then scatter-add
        if (warp_hn)
          {
#pragma unroll
            for (int r = 0; r < NR; ++r)
              if (col_cl(r) < ncw)
                {
                  T r1[n], r2[n];
#pragma unroll
                  for (int k = 0; k < n; ++k)
                    {
                      r1[k] = u1[col_sm(r) + k * PS];
                      r2[k] = u2[col_sm(r) + k * PS];
                    }
                  eo_sweep<P, T, 1, 2>(r1);
                  eo_sweep<P, T, 1, 3>(r2);
#pragma unroll
                  for (int k = 0; k < n; ++k)
                    u1[col_sm(r) + k * PS] = r1[k] + r2[k];
                }
            __syncwarp();
            hanging_node_sweeps_slab<P, T, true>(u1, tb, ncw, lane);
          }
#pragma unroll
        for (int r = 0; r < NR; ++r)
          if (col_cl(r) < ncw)
            {
              T res[n];
#pragma unroll
              for (int k = 0; k < n; ++k)
                res[k] = u1[col_sm(r) + k * PS];
              if (!warp_hn)
                {
                  T r2[n];
#pragma unroll
                  for (int k = 0; k < n; ++k)
                    r2[k] = u2[col_sm(r) + k * PS];
                  eo_sweep<P, T, 1, 2>(res);
                  eo_sweep<P, T, 1, 3>(r2);
#pragma unroll
                  for (int k = 0; k < n; ++k)
                    res[k] += r2[k];
                }
#pragma unroll
              for (int k = 0; k < n; ++k)
                {
                  const uint32_t idx = column_index(tb, r, k);
                  if (!(idx & D_FLAG_BIT))
                    atomic_add(dst + idx, scale * res[k]);
                }
            }
        __syncwarp(); // u1/u2/cf/tb are rewritten by the next iteration
      }
  }

  // which general-geometry apply kernel: the slab kernel for degrees 1..4 unless MGB_Q1_KERNEL=basic
  inline bool q1_use_slab()
  {
    static const bool use = [] {
      const char *e = std::getenv("MGB_Q1_KERNEL");
      return !(e && std::string(e) == "basic");
    }();
    return use;
  }

  template <int P, typename T>
  void launch_q1_slab(const CellData &cd, const T *src, T *dst, const uint32_t *cell_list, long long n_list, cudaStream_t s)
  {
    constexpr int   CW = Q1SlabCfg<P>::CW, W = Q1SlabCfg<P>::W;
    const long long n_work = cell_list ? n_list : cd.n_cells;
    if (n_work <= 0)
      return;
    const size_t smem = Q1SlabSmem<P, T>::warp_bytes * W;
    static int   blocks_per_sm = 0, n_sm = 0; // per (P, T) instantiation
    if (blocks_per_sm == 0)
      {
        int dev = 0;
        MGB_CUDA(cudaGetDevice(&dev));
        MGB_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
        MGB_CUDA(cudaFuncSetAttribute(k_laplace_q1_slab<P, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        MGB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, k_laplace_q1_slab<P, T>, W * 32, smem));
        if (blocks_per_sm < 1)
          throw std::runtime_error("k_laplace_q1_slab does not fit on an SM");
      }
    const long long n_batches = (n_work + CW - 1) / CW;
    const long long blocks    = std::min<long long>((n_batches + W - 1) / W, (long long)blocks_per_sm * n_sm);
    k_laplace_q1_slab<P, T><<<(unsigned)blocks, W * 32, smem, s>>>(cd.ctab, cd.any_hanging, cd.vertices, T(cd.scale), n_work, cell_list, src,
                                                                   dst);
    MGB_CUDA(cudaGetLastError());
  }
} // namespace mgb
