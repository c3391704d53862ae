// Single-block coarse-grid solver: MGCoarseGridIterativeSolver<SolverCG, Operator, Identity | Chebyshev> at min_level
// (multigrid_throughput.cc:888-944) for coarse problems small enough to live in one thread block (the reference's
// geometries coarsen to ONE root cell: 125 DoFs at p=4).  The whole preconditioned CG -- SolverCG's recurrences
// (SURVEY.md A.6), ReductionControl's stopping rule and, for cg_with_chebyshev, PreconditionChebyshev::vmult -- runs in
// one kernel on the dense level matrix (assembled once from the matrix-free operator), with block-wide reductions in a
// fixed order.  It replaces hundreds of tiny launches and host synchronisations per V-cycle by one launch.
#include "device.h"

namespace mgb
{
  namespace
  {
    constexpr int CB = 256;

    __device__ inline double block_sum(double v, double *sh)
    {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1)
        v += __shfl_down_sync(0xffffffffu, v, o);
      __syncthreads(); // sh may still be read from the previous reduction
      if ((threadIdx.x & 31) == 0)
        sh[threadIdx.x >> 5] = v;
      __syncthreads();
      double s = 0;
#pragma unroll
      for (int w = 0; w < CB / 32; ++w)
        s += sh[w];
      return s;
    }

    // y = A x for the symmetric dense matrix A (n x n); x, y in shared memory
    template <typename T> __device__ inline void dense_matvec(const T *__restrict__ A, const int n, const T *x, T *y)
    {
      for (int i = threadIdx.x; i < n; i += CB)
        {
          T acc = 0;
          for (int j = 0; j < n; ++j)
            acc += A[size_t(j) * n + i] * x[j]; // symmetric: column access = coalesced row access
          y[i] = acc;
        }
      __syncthreads();
    }

    // PreconditionChebyshev::vmult (zero initial guess): x = p_k(D^-1 A) D^-1 b
    template <typename T>
    __device__ inline void chebyshev_vmult(const T *A, const int n, const T *dinv, const int degree, const double theta,
                                           const double delta, const T *b, T *x, T *xo, T *t)
    {
      for (int i = threadIdx.x; i < n; i += CB)
        x[i] = T(1.0 / theta) * (dinv[i] * b[i]);
      __syncthreads();
      if (degree < 2 || fabs(delta) < 1e-40)
        return;
      double rhok = delta / theta;
      const double sigma = theta / delta;
      for (int k = 0; k < degree - 1; ++k)
        {
          dense_matvec(A, n, x, t);
          const double rhokp = 1.0 / (2.0 * sigma - rhok);
          const T      f1 = T(rhokp * rhok), f2 = T(2.0 * rhokp / delta);
          rhok = rhokp;
          for (int i = threadIdx.x; i < n; i += CB)
            {
              const T xi  = x[i];
              const T xn  = (k == 0) ? (T(1) + f1) * xi + f2 * (dinv[i] * (b[i] - t[i]))
                                     : (T(1) + f1) * xi - f1 * xo[i] + f2 * (dinv[i] * (b[i] - t[i]));
              xo[i] = xi;
              x[i]  = xn;
            }
          __syncthreads();
        }
    }

    template <typename T>
    __global__ void __launch_bounds__(CB)
      k_coarse_cg(const T *__restrict__ A_glob, const int n, const int a_in_smem, const T *__restrict__ dinv_glob, const int cheb_degree,
                  const double theta, const double delta, const unsigned maxiter, const double abstol, const double reltol,
                  const T *__restrict__ b_glob, T *__restrict__ x_glob, int *__restrict__ n_iterations)
    {
      extern __shared__ __align__(16) unsigned char smem_raw[];
      __shared__ double red[CB / 32];
      T *x = reinterpret_cast<T *>(smem_raw), *g = x + n, *d = g + n, *h = d + n, *xo = h + n, *t = xo + n, *dinv = t + n, *b = dinv + n;
      T *As = b + n;
      const T *A = A_glob;
      if (a_in_smem)
        {
          for (size_t k = threadIdx.x; k < size_t(n) * n; k += CB)
            As[k] = A_glob[k];
          A = As;
        }
      const bool precond = cheb_degree > 0;
      for (int i = threadIdx.x; i < n; i += CB)
        {
          x[i]    = T(0);
          b[i]    = b_glob[i];
          g[i]    = -b[i];
          dinv[i] = precond ? dinv_glob[i] : T(1);
        }
      __syncthreads();
      double gg = 0;
      for (int i = threadIdx.x; i < n; i += CB)
        gg += double(g[i]) * double(g[i]);
      double         res         = sqrt(block_sum(gg, red));
      const double   reduced_tol = reltol * res;
      unsigned       it          = 0;
      double         gh          = 0;
      // ReductionControl::check(0, res)
      bool done = (res <= reduced_tol) || (res <= abstol);
      while (!done)
        {
          ++it;
          double gh_new;
          if (precond)
            {
              chebyshev_vmult(A, n, dinv, cheb_degree, theta, delta, g, h, xo, t);
              double v = 0;
              for (int i = threadIdx.x; i < n; i += CB)
                v += double(g[i]) * double(h[i]);
              gh_new = block_sum(v, red);
            }
          else
            gh_new = res * res;
          const T *z = precond ? h : g;
          if (it == 1)
            for (int i = threadIdx.x; i < n; i += CB)
              d[i] = -z[i];
          else
            {
              const T beta = T(gh_new / gh);
              for (int i = threadIdx.x; i < n; i += CB)
                d[i] = beta * d[i] - z[i];
            }
          gh = gh_new;
          __syncthreads();
          dense_matvec(A, n, d, h);
          double v = 0;
          for (int i = threadIdx.x; i < n; i += CB)
            v += double(d[i]) * double(h[i]);
          const T alpha = T(gh / block_sum(v, red));
          v             = 0;
          for (int i = threadIdx.x; i < n; i += CB)
            {
              x[i] += alpha * d[i];
              const T gi = g[i] + alpha * h[i];
              g[i]       = gi;
              v += double(gi) * double(gi);
            }
          res  = sqrt(fabs(block_sum(v, red)));
          done = (res <= reduced_tol) || (res <= abstol) || (it >= maxiter) || isnan(res);
        }
      for (int i = threadIdx.x; i < n; i += CB)
        x_glob[i] = x[i];
      if (threadIdx.x == 0)
        *n_iterations += int(it);
    }
  } // namespace

  size_t coarse_cg_smem_bytes(int n, size_t elem, bool matrix_in_smem)
  {
    return (8 * size_t(n) + (matrix_in_smem ? size_t(n) * n : 0)) * elem;
  }

  template <typename T>
  void coarse_cg_solve(const T *A, int n, const T *dinv, int cheb_degree, double theta, double delta, unsigned maxiter, double abstol,
                       double reltol, const T *b, T *x, int *n_iterations_dev, cudaStream_t s)
  {
    static int max_smem = 0;
    if (max_smem == 0)
      {
        int dev = 0;
        MGB_CUDA(cudaGetDevice(&dev));
        MGB_CUDA(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        MGB_CUDA(cudaFuncSetAttribute(k_coarse_cg<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem - 1024));
      }
    const bool   in_smem = coarse_cg_smem_bytes(n, sizeof(T), true) <= size_t(max_smem - 1024);
    const size_t smem    = coarse_cg_smem_bytes(n, sizeof(T), in_smem);
    k_coarse_cg<T><<<1, CB, smem, s>>>(A, n, in_smem ? 1 : 0, dinv, cheb_degree, theta, delta, maxiter, abstol, reltol, b, x,
                                       n_iterations_dev);
    MGB_CUDA(cudaGetLastError());
  }
  template void coarse_cg_solve<double>(const double *, int, const double *, int, double, double, unsigned, double, double, const double *,
                                        double *, int *, cudaStream_t);
  template void coarse_cg_solve<float>(const float *, int, const float *, int, double, double, unsigned, double, double, const float *,
                                       float *, int *, cudaStream_t);
} // namespace mgb
