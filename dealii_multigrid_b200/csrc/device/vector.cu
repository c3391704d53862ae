// BLAS-1, index gather/scatter and smoother vector updates for LinearAlgebra::distributed::Vector
// semantics.  All HBM-bound streaming kernels: 128-bit accesses where alignment allows,
// grid sized as a multiple of the SM count, deterministic two-stage reductions (fixed summation order
// for a fixed launch configuration, like deal.II's fixed pairwise order + MPI_Allreduce).
#include "device.h"

#include <cstdio>
#include <stdexcept>
#include <string>

namespace mgb
{
  void check_cuda(cudaError_t e, const char *what)
  {
    if (e != cudaSuccess)
      throw std::runtime_error(std::string("CUDA error in ") + what + ": " + cudaGetErrorString(e));
  }

  namespace
  {
    constexpr int BLOCK      = 256;
    constexpr int RED_BLOCKS = 148 * 8; // reduction grid: 8 CTAs per SM (full occupancy at 256 threads)

    // reduction grid: a function of n only (fixed summation order for a fixed vector length)
    inline int red_grid(size_t n)
    {
      const size_t b = (n + 2 * BLOCK - 1) / (2 * BLOCK);
      return int(b < 1 ? 1 : (b > size_t(RED_BLOCKS) ? size_t(RED_BLOCKS) : b));
    }

    inline int stream_grid(size_t n)
    {
      size_t b = (n + BLOCK - 1) / BLOCK;
      const size_t cap = 148 * 16;
      if (b > cap)
        b = cap;
      if (b < 1)
        b = 1;
      return int(b);
    }

    template <typename T, typename F> __global__ void k_map1(T *__restrict__ y, size_t n, F f)
    {
      for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        y[i] = f(y[i], i);
    }

    template <typename T> __global__ void k_set(T *__restrict__ y, T a, size_t n)
    {
      for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        y[i] = a;
    }
    template <typename T> __global__ void k_scale(T *__restrict__ y, T a, size_t n)
    {
      for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        y[i] *= a;
    }
    template <typename T> __global__ void k_add(T *__restrict__ y, T a, const T *__restrict__ x, size_t n)
    {
      for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        y[i] += a * x[i];
    }
    template <typename T> __global__ void k_sadd(T *__restrict__ y, T s, T a, const T *__restrict__ x, size_t n)
    {
      for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        y[i] = s * y[i] + a * x[i];
    }
    template <typename T> __global__ void k_equ(T *__restrict__ y, T a, const T *__restrict__ x, size_t n)
    {
      for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        y[i] = a * x[i];
    }
    template <typename T> __global__ void k_add_scalar(T *__restrict__ y, T a, size_t n)
    {
      for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        y[i] += a;
    }
    template <typename T, typename U> __global__ void k_convert(T *__restrict__ y, const U *__restrict__ x, size_t n)
    {
      for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        y[i] = T(x[i]);
    }

    // block reduction in fixed order; the last block to finish sums the per-block partials in index order
    __device__ inline void block_reduce_and_finish(double v, double *result, double *scratch, unsigned *ticket)
    {
      __shared__ double sh[BLOCK / 32];
      __shared__ bool   last;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1)
        v += __shfl_down_sync(0xffffffffu, v, o);
      if ((threadIdx.x & 31) == 0)
        sh[threadIdx.x >> 5] = v;
      __syncthreads();
      if (threadIdx.x == 0)
        {
          double s = 0;
          for (int w = 0; w < BLOCK / 32; ++w)
            s += sh[w];
          scratch[blockIdx.x] = s;
          __threadfence();
          const unsigned t = atomicAdd(ticket, 1u);
          last             = (t == gridDim.x - 1);
        }
      __syncthreads();
      if (last)
        {
          __threadfence();
          double s = 0;
          for (unsigned b = threadIdx.x; b < gridDim.x; b += BLOCK)
            s += scratch[b];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1)
            s += __shfl_down_sync(0xffffffffu, s, o);
          __syncthreads();
          if ((threadIdx.x & 31) == 0)
            sh[threadIdx.x >> 5] = s;
          __syncthreads();
          if (threadIdx.x == 0)
            {
              double r = 0;
              for (int w = 0; w < BLOCK / 32; ++w)
                r += sh[w];
              *result = r;
              *ticket = 0;
            }
        }
    }

    // The reduction loops are unrolled four times with independent accumulators (memory-level parallelism); the
    // summation order is a function of the launch configuration only, so results are reproducible run to run.
    template <typename T>
    __global__ void __launch_bounds__(BLOCK)
      k_dot(const T *__restrict__ x, const T *__restrict__ y, size_t n, double *result, double *scratch, unsigned *ticket)
    {
      const size_t stride = size_t(gridDim.x) * BLOCK;
      size_t       i      = blockIdx.x * size_t(BLOCK) + threadIdx.x;
      double       v0 = 0, v1 = 0, v2 = 0, v3 = 0;
      for (; i + 3 * stride < n; i += 4 * stride)
        {
          const T x0 = x[i], x1 = x[i + stride], x2 = x[i + 2 * stride], x3 = x[i + 3 * stride];
          const T y0 = y[i], y1 = y[i + stride], y2 = y[i + 2 * stride], y3 = y[i + 3 * stride];
          v0 += double(x0) * double(y0);
          v1 += double(x1) * double(y1);
          v2 += double(x2) * double(y2);
          v3 += double(x3) * double(y3);
        }
      for (; i < n; i += stride)
        v0 += double(x[i]) * double(y[i]);
      block_reduce_and_finish((v0 + v1) + (v2 + v3), result, scratch, ticket);
    }
    template <typename T>
    __global__ void __launch_bounds__(BLOCK) k_sum(const T *__restrict__ x, size_t n, double *result, double *scratch, unsigned *ticket)
    {
      double v = 0;
      for (size_t i = blockIdx.x * size_t(BLOCK) + threadIdx.x; i < n; i += size_t(gridDim.x) * BLOCK)
        v += double(x[i]);
      block_reduce_and_finish(v, result, scratch, ticket);
    }
    template <typename T>
    __global__ void __launch_bounds__(BLOCK)
      k_add_and_dot(T *__restrict__ g, T a, const T *__restrict__ h, size_t n, double *result, double *scratch, unsigned *ticket)
    {
      const size_t stride = size_t(gridDim.x) * BLOCK;
      size_t       i      = blockIdx.x * size_t(BLOCK) + threadIdx.x;
      double       v0 = 0, v1 = 0;
      for (; i + stride < n; i += 2 * stride)
        {
          const T g0 = g[i] + a * h[i], g1 = g[i + stride] + a * h[i + stride];
          g[i]          = g0;
          g[i + stride] = g1;
          v0 += double(g0) * double(g0);
          v1 += double(g1) * double(g1);
        }
      for (; i < n; i += stride)
        {
          const T gi = g[i] + a * h[i];
          g[i]       = gi;
          v0 += double(gi) * double(gi);
        }
      block_reduce_and_finish(v0 + v1, result, scratch, ticket);
    }

    // SolverCG's two vector updates of one iteration in one pass: x += a d;  g += a h;  result = g.g
    template <typename T>
    __global__ void __launch_bounds__(BLOCK) k_cg_update(T *__restrict__ x, T *__restrict__ g, T a, const T *__restrict__ d,
                                                         const T *__restrict__ h, size_t n, double *result, double *scratch, unsigned *ticket)
    {
      const size_t stride = size_t(gridDim.x) * BLOCK;
      size_t       i      = blockIdx.x * size_t(BLOCK) + threadIdx.x;
      double       v0 = 0, v1 = 0;
      for (; i + stride < n; i += 2 * stride)
        {
          const T d0 = d[i], d1 = d[i + stride], h0 = h[i], h1 = h[i + stride];
          const T x0 = x[i], x1 = x[i + stride], g0 = g[i] + a * h0, g1 = g[i + stride] + a * h1;
          x[i]          = x0 + a * d0;
          x[i + stride] = x1 + a * d1;
          g[i]          = g0;
          g[i + stride] = g1;
          v0 += double(g0) * double(g0);
          v1 += double(g1) * double(g1);
        }
      for (; i < n; i += stride)
        {
          x[i] += a * d[i];
          const T gi = g[i] + a * h[i];
          g[i]       = gi;
          v0 += double(gi) * double(gi);
        }
      block_reduce_and_finish(v0 + v1, result, scratch, ticket);
    }
    template <typename T>
    __global__ void k_idx_residual(T *__restrict__ dst, const T *__restrict__ a, const T *__restrict__ b, const uint32_t *__restrict__ idx, uint32_t n)
    {
      const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
      if (i < n)
        dst[idx[i]] = a[idx[i]] - b[idx[i]];
    }

    template <typename T> __global__ void k_idx_copy(T *__restrict__ dst, const T *__restrict__ src, const uint32_t *__restrict__ idx, uint32_t n)
    {
      const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
      if (i < n)
        dst[idx[i]] = src[idx[i]];
    }
    template <typename T> __global__ void k_idx_gather(T *__restrict__ buf, const T *__restrict__ src, const uint32_t *__restrict__ idx, uint32_t n)
    {
      const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
      if (i < n)
        buf[i] = src[idx[i]];
    }
    template <typename T> __global__ void k_idx_scatter(T *__restrict__ dst, const T *__restrict__ buf, const uint32_t *__restrict__ idx, uint32_t n)
    {
      const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
      if (i < n)
        dst[idx[i]] = buf[i];
    }
    template <typename T> __global__ void k_idx_scatter_add(T *__restrict__ dst, const T *__restrict__ buf, const uint32_t *__restrict__ idx, uint32_t n)
    {
      const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
      if (i < n)
        atomicAdd(dst + idx[i], buf[i]); // an owned entry may be imported by several ranks
    }
    template <typename T> __global__ void k_idx_set(T *__restrict__ dst, T a, const uint32_t *__restrict__ idx, uint32_t n)
    {
      const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
      if (i < n)
        dst[idx[i]] = a;
    }
    template <typename T, typename U>
    __global__ void k_idx_pair_copy(T *__restrict__ dst, const uint32_t *__restrict__ di, const U *__restrict__ src, const uint32_t *__restrict__ si, uint32_t n)
    {
      const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
      if (i < n)
        dst[di[i]] = T(src[si[i]]);
    }

    // out may alias x or x_old (the update is pointwise), hence no __restrict__ on those
    template <typename T, int IT>
    __global__ void k_cheb(T f1, T f2, const T *__restrict__ rhs, const T *__restrict__ dinv, const T *__restrict__ t, const T *x,
                           const T *x_old, T *out, size_t n)
    {
      for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        {
          if (IT == 0)
            out[i] = f2 * (dinv[i] * rhs[i]);
          else if (IT == 1)
            out[i] = (T(1) + f1) * x[i] + f2 * (dinv[i] * (rhs[i] - t[i]));
          else
            out[i] = (T(1) + f1) * x[i] - f1 * x_old[i] + f2 * (dinv[i] * (rhs[i] - t[i]));
        }
    }
    template <typename T> __global__ void k_invert_diag(T *__restrict__ d, size_t n)
    {
      for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        {
          const T v = d[i];
          d[i]      = (fabs(double(v)) > 1.0e-10) ? T(1) / v : T(1);
        }
    }
  } // namespace

  int reduction_scratch_size() { return RED_BLOCKS; }

#define LAUNCH_STREAM(kernel, n, ...)                                    \
  do                                                                     \
    {                                                                    \
      if ((n) > 0)                                                       \
        {                                                                \
          kernel<<<stream_grid(n), BLOCK, 0, s>>>(__VA_ARGS__);          \
          MGB_CUDA(cudaGetLastError());                                  \
        }                                                                \
    }                                                                    \
  while (0)

  template <typename T> void vec_set(T *x, T a, size_t n, cudaStream_t s)
  {
    if (a == T(0))
      {
        if (n)
          MGB_CUDA(cudaMemsetAsync(x, 0, n * sizeof(T), s));
      }
    else
      LAUNCH_STREAM(k_set<T>, n, x, a, n);
  }
  template <typename T> void vec_scale(T *x, T a, size_t n, cudaStream_t s) { LAUNCH_STREAM(k_scale<T>, n, x, a, n); }
  template <typename T> void vec_add(T *y, T a, const T *x, size_t n, cudaStream_t s) { LAUNCH_STREAM(k_add<T>, n, y, a, x, n); }
  template <typename T> void vec_sadd(T *y, T s_, T a, const T *x, size_t n, cudaStream_t s) { LAUNCH_STREAM(k_sadd<T>, n, y, s_, a, x, n); }
  template <typename T> void vec_equ(T *y, T a, const T *x, size_t n, cudaStream_t s) { LAUNCH_STREAM(k_equ<T>, n, y, a, x, n); }
  template <typename T> void vec_add_scalar(T *y, T a, size_t n, cudaStream_t s) { LAUNCH_STREAM(k_add_scalar<T>, n, y, a, n); }
  template <typename T, typename U> void vec_convert(T *y, const U *x, size_t n, cudaStream_t s)
  {
    LAUNCH_STREAM((k_convert<T, U>), n, y, x, n);
  }

  template <typename T>
  void vec_dot(const T *x, const T *y, size_t n, double *result_dev, double *scratch, unsigned *ticket, cudaStream_t s)
  {
    k_dot<T><<<red_grid(n), BLOCK, 0, s>>>(x, y, n, result_dev, scratch, ticket);
    MGB_CUDA(cudaGetLastError());
  }
  template <typename T> void vec_sum(const T *x, size_t n, double *result_dev, double *scratch, unsigned *ticket, cudaStream_t s)
  {
    k_sum<T><<<red_grid(n), BLOCK, 0, s>>>(x, n, result_dev, scratch, ticket);
    MGB_CUDA(cudaGetLastError());
  }
  template <typename T>
  void vec_add_and_dot(T *g, T a, const T *h, size_t n, double *result_dev, double *scratch, unsigned *ticket, cudaStream_t s)
  {
    k_add_and_dot<T><<<red_grid(n), BLOCK, 0, s>>>(g, a, h, n, result_dev, scratch, ticket);
    MGB_CUDA(cudaGetLastError());
  }

  template <typename T>
  void vec_cg_update(T *x, T *g, T a, const T *d, const T *h, size_t n, double *result_dev, double *scratch, unsigned *ticket, cudaStream_t s)
  {
    k_cg_update<T><<<red_grid(n), BLOCK, 0, s>>>(x, g, a, d, h, n, result_dev, scratch, ticket);
    MGB_CUDA(cudaGetLastError());
  }

#define LAUNCH_IDX(kernel, n, ...)                                         \
  do                                                                       \
    {                                                                      \
      if ((n) > 0)                                                         \
        {                                                                  \
          kernel<<<((n) + BLOCK - 1) / BLOCK, BLOCK, 0, s>>>(__VA_ARGS__); \
          MGB_CUDA(cudaGetLastError());                                    \
        }                                                                  \
    }                                                                      \
  while (0)

  template <typename T> void idx_copy(T *dst, const T *src, const uint32_t *idx, uint32_t n, cudaStream_t s) { LAUNCH_IDX(k_idx_copy<T>, n, dst, src, idx, n); }
  template <typename T> void idx_residual(T *dst, const T *a, const T *b, const uint32_t *idx, uint32_t n, cudaStream_t s) { LAUNCH_IDX(k_idx_residual<T>, n, dst, a, b, idx, n); }
  template <typename T> void idx_gather(T *buf, const T *src, const uint32_t *idx, uint32_t n, cudaStream_t s) { LAUNCH_IDX(k_idx_gather<T>, n, buf, src, idx, n); }
  template <typename T> void idx_scatter(T *dst, const T *buf, const uint32_t *idx, uint32_t n, cudaStream_t s) { LAUNCH_IDX(k_idx_scatter<T>, n, dst, buf, idx, n); }
  template <typename T> void idx_scatter_add(T *dst, const T *buf, const uint32_t *idx, uint32_t n, cudaStream_t s) { LAUNCH_IDX(k_idx_scatter_add<T>, n, dst, buf, idx, n); }
  template <typename T> void idx_set(T *dst, T a, const uint32_t *idx, uint32_t n, cudaStream_t s) { LAUNCH_IDX(k_idx_set<T>, n, dst, a, idx, n); }
  template <typename T, typename U>
  void idx_pair_copy(T *dst, const uint32_t *di, const U *src, const uint32_t *si, uint32_t n, cudaStream_t s)
  {
    LAUNCH_IDX((k_idx_pair_copy<T, U>), n, dst, di, src, si, n);
  }

  template <typename T>
  void cheb_update(int it, T f1, T f2, const T *rhs, const T *dinv, const T *t, const T *x, const T *x_old, T *out, size_t n,
                   cudaStream_t s)
  {
    if (it == 0)
      LAUNCH_STREAM((k_cheb<T, 0>), n, f1, f2, rhs, dinv, t, x, x_old, out, n);
    else if (it == 1)
      LAUNCH_STREAM((k_cheb<T, 1>), n, f1, f2, rhs, dinv, t, x, x_old, out, n);
    else
      LAUNCH_STREAM((k_cheb<T, 2>), n, f1, f2, rhs, dinv, t, x, x_old, out, n);
  }
  template <typename T> void invert_diagonal(T *d, size_t n, cudaStream_t s) { LAUNCH_STREAM(k_invert_diag<T>, n, d, n); }

  // explicit instantiations
#define INST(T)                                                                                                       \
  template void vec_set<T>(T *, T, size_t, cudaStream_t);                                                             \
  template void vec_scale<T>(T *, T, size_t, cudaStream_t);                                                           \
  template void vec_add<T>(T *, T, const T *, size_t, cudaStream_t);                                                  \
  template void vec_sadd<T>(T *, T, T, const T *, size_t, cudaStream_t);                                              \
  template void vec_equ<T>(T *, T, const T *, size_t, cudaStream_t);                                                  \
  template void vec_add_scalar<T>(T *, T, size_t, cudaStream_t);                                                      \
  template void vec_dot<T>(const T *, const T *, size_t, double *, double *, unsigned *, cudaStream_t);               \
  template void vec_sum<T>(const T *, size_t, double *, double *, unsigned *, cudaStream_t);                          \
  template void vec_add_and_dot<T>(T *, T, const T *, size_t, double *, double *, unsigned *, cudaStream_t);          \
  template void idx_copy<T>(T *, const T *, const uint32_t *, uint32_t, cudaStream_t);                                \
  template void idx_residual<T>(T *, const T *, const T *, const uint32_t *, uint32_t, cudaStream_t);                 \
  template void vec_cg_update<T>(T *, T *, T, const T *, const T *, size_t, double *, double *, unsigned *, cudaStream_t); \
  template void idx_gather<T>(T *, const T *, const uint32_t *, uint32_t, cudaStream_t);                              \
  template void idx_scatter<T>(T *, const T *, const uint32_t *, uint32_t, cudaStream_t);                             \
  template void idx_scatter_add<T>(T *, const T *, const uint32_t *, uint32_t, cudaStream_t);                         \
  template void idx_set<T>(T *, T, const uint32_t *, uint32_t, cudaStream_t);                                         \
  template void cheb_update<T>(int, T, T, const T *, const T *, const T *, const T *, const T *, T *, size_t, cudaStream_t); \
  template void invert_diagonal<T>(T *, size_t, cudaStream_t);
  INST(double)
  INST(float)
  template void vec_convert<double, float>(double *, const float *, size_t, cudaStream_t);
  template void vec_convert<float, double>(float *, const double *, size_t, cudaStream_t);
  template void vec_convert<double, double>(double *, const double *, size_t, cudaStream_t);
  template void vec_convert<float, float>(float *, const float *, size_t, cudaStream_t);
  template void idx_pair_copy<double, double>(double *, const uint32_t *, const double *, const uint32_t *, uint32_t, cudaStream_t);
  template void idx_pair_copy<float, double>(float *, const uint32_t *, const double *, const uint32_t *, uint32_t, cudaStream_t);
  template void idx_pair_copy<double, float>(double *, const uint32_t *, const float *, const uint32_t *, uint32_t, cudaStream_t);
  template void idx_pair_copy<float, float>(float *, const uint32_t *, const float *, const uint32_t *, uint32_t, cudaStream_t);
} // namespace mgb
