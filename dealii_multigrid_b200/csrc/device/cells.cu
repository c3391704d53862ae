// Translation unit of all cell-wise kernels (they share the constant-memory 1-D matrices).
#include "../host/fe1d.h"
#include "operator.cuh"
#include "operator_q1.cuh"
#include "chunk.cuh"
#include "transfer.cuh"

#include <mutex>
#include <stdexcept>

namespace mgb
{
  void upload_fe_constants(int degree)
  {
    static std::mutex mtx;
    static bool       done[MAX_DEGREE + 1] = {};
    if (degree < 1 || degree > MAX_DEGREE)
      throw std::runtime_error("degree out of range [1," + std::to_string(MAX_DEGREE) + "]");
    std::lock_guard<std::mutex> lock(mtx);
    if (done[degree])
      return;
    FE1D    fe(degree);
    FEConst h{};
    for (int i = 0; i < fe.n; ++i)
      for (int j = 0; j < fe.n; ++j)
        {
          h.M[i * MAXN + j]    = fe.M[i * fe.n + j];
          h.K[i * MAXN + j]    = fe.K[i * fe.n + j];
          h.S[0][i * MAXN + j] = fe.S[0][i * fe.n + j];
          h.S[1][i * MAXN + j] = fe.S[1][i * fe.n + j];
        }
    FEConstF hf{};
    for (int i = 0; i < fe.n; ++i)
      {
        h.lam[i] = fe.lambda[i];
        hf.lam[i] = (float)fe.lambda[i];
        for (int j = 0; j < fe.n; ++j)
          {
            h.Tm[i * MAXN + j]    = fe.Tm[i * fe.n + j];
            h.lam2[i * MAXN + j]  = fe.lambda[i] + fe.lambda[j];
            hf.Tm[i * MAXN + j]   = (float)fe.Tm[i * fe.n + j];
            hf.lam2[i * MAXN + j] = (float)(fe.lambda[i] + fe.lambda[j]);
          }
      }
    FEConstQ hq{};
    for (int q = 0; q < fe.n; ++q)
      {
        hq.xq[q] = fe.xq[q];
        hq.wq[q] = fe.wq[q];
        for (int i = 0; i < fe.n; ++i)
          {
            hq.Bq[q * MAXN + i] = fe.Bq[q * fe.n + i];
            hq.Dq[q * MAXN + i] = fe.Dq[q * fe.n + i];
          }
      }
    MGB_CUDA(cudaMemcpyToSymbol(c_feq, &hq, sizeof(FEConstQ), sizeof(FEConstQ) * degree));
    // even-odd tables of B, G = (d/dx of the nodal basis at the Gauss points) and their transposes (k_laplace_q1_slab)
    {
      const int           n = fe.n;
      std::vector<double> A[4] = {std::vector<double>(n * n), std::vector<double>(n * n), std::vector<double>(n * n),
                                  std::vector<double>(n * n)};
      for (int q = 0; q < n; ++q)
        for (int i = 0; i < n; ++i)
          {
            A[0][q * n + i] = fe.Bq[q * n + i];
            A[1][q * n + i] = lagrange_derivative(fe.nodes, i, fe.xq[q]);
            A[2][i * n + q] = A[0][q * n + i];
            A[3][i * n + q] = A[1][q * n + i];
          }
      static FEConstQEO he; // 3 KB: keep it off the stack
      he = FEConstQEO{};
      for (int mat = 0; mat < 4; ++mat)
        for (int q = 0; q < (n + 1) / 2; ++q)
          {
            for (int l = 0; l < n / 2; ++l)
              {
                const double a = A[mat][q * n + l], b = A[mat][q * n + n - 1 - l];
                he.m[mat][q * MAXN + l]         = 0.5 * (a + b);
                he.m[mat][q * MAXN + n - 1 - l] = 0.5 * (a - b);
              }
            if (n % 2 == 1)
              he.m[mat][q * MAXN + n / 2] = A[mat][q * n + n / 2];
            for (int l = 0; l < n; ++l)
              he.mf[mat][q * MAXN + l] = (float)he.m[mat][q * MAXN + l];
          }
      MGB_CUDA(cudaMemcpyToSymbol(c_feqeo, &he, sizeof(FEConstQEO), sizeof(FEConstQEO) * degree));
    }
    MGB_CUDA(cudaMemcpyToSymbol(c_fe, &h, sizeof(FEConst), sizeof(FEConst) * degree));
    MGB_CUDA(cudaMemcpyToSymbol(c_fef, &hf, sizeof(FEConstF), sizeof(FEConstF) * degree));
    done[degree] = true;
  }

#define DISPATCH_DEGREE(deg, CALL)                                   \
  switch (deg)                                                       \
    {                                                                \
      case 1: { constexpr int P = 1; CALL; break; }                  \
      case 2: { constexpr int P = 2; CALL; break; }                  \
      case 3: { constexpr int P = 3; CALL; break; }                  \
      case 4: { constexpr int P = 4; CALL; break; }                  \
      case 5: { constexpr int P = 5; CALL; break; }                  \
      case 6: { constexpr int P = 6; CALL; break; }                  \
      default: throw std::runtime_error("unsupported degree");      \
    }

  template <typename T>
  void cell_laplace_apply(const CellData &cd, const T *src, T *dst, const uint32_t *cell_list, long long n_list, cudaStream_t s)
  {
    if (cd.vertices)
      {
        if (cd.degree <= 4 && q1_use_slab())
          {
            switch (cd.degree)
              {
                case 1: launch_q1_slab<1, T>(cd, src, dst, cell_list, n_list, s); break;
                case 2: launch_q1_slab<2, T>(cd, src, dst, cell_list, n_list, s); break;
                case 3: launch_q1_slab<3, T>(cd, src, dst, cell_list, n_list, s); break;
                default: launch_q1_slab<4, T>(cd, src, dst, cell_list, n_list, s); break;
              }
            return;
          }
        DISPATCH_DEGREE(cd.degree, (launch_q1<P, T, 0>(cd, src, dst, cell_list, n_list, s)));
        return;
      }
    DISPATCH_DEGREE(cd.degree, (launch_apply<P, T>(cd, src, dst, cell_list, n_list, s)));
  }
  bool chunk_atomic_output() { return MGB_CHUNK_ABLATE == 3; }
  template <typename T>
  void cell_laplace_chunked(const CellData &cd, const ChunkDevice &ch, const T *x, T *out, const FusedEpilogue<T> &ep, cudaStream_t s)
  {
    if (cd.vertices)
      throw std::runtime_error("cell_laplace_chunked: Cartesian cells only");
    ChunkArgs<T> a;
    a.n_chunks    = ch.n_chunks;
    a.desc        = ch.desc;
    a.pool        = ch.pool;
    a.deps        = ch.deps;
    a.ghost_idx   = ch.ghost_idx;
    a.side        = static_cast<T *>(ch.side);
    a.flags       = ch.flags;
    a.state       = ch.state;
    a.root_length = cd.root_length;
    a.src         = x;
    a.out         = out;
    a.mode        = ep.mode;
    a.b           = ep.b;
    a.dinv        = ep.dinv;
    a.xold        = ep.xold;
    a.f1          = ep.f1;
    a.f2          = ep.f2;
    DISPATCH_DEGREE(cd.degree, (launch_chunks<P, T>(a, s)));
  }
  namespace
  {
    template <typename T>
    __global__ void k_idx_fused_identity(T *out, const T *x, const int mode, const T *dinv, const T f2, const uint32_t *idx, uint32_t n)
    {
      const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
      if (i < n)
        {
          const uint32_t g = idx[i];
          const T        coef = mode == 0 ? T(1) : (mode == 1 ? T(-1) : -f2 * dinv[g]);
          out[g] += coef * x[g];
        }
    }
  } // namespace
  template <typename T>
  void idx_fused_identity(T *out, const T *x, const FusedEpilogue<T> &ep, const uint32_t *idx, uint32_t n, cudaStream_t s)
  {
    if (n == 0)
      return;
    k_idx_fused_identity<T><<<(n + 255) / 256, 256, 0, s>>>(out, x, ep.mode, ep.dinv, ep.f2, idx, n);
    MGB_CUDA(cudaGetLastError());
  }
  template <typename T> void cell_laplace_diagonal(const CellData &cd, T *diag, cudaStream_t s)
  {
    if (cd.vertices)
      {
        DISPATCH_DEGREE(cd.degree, (launch_q1<P, T, 1>(cd, (const T *)nullptr, diag, nullptr, 0, s)));
        return;
      }
    DISPATCH_DEGREE(cd.degree, (launch_diagonal<P, T>(cd, diag, s)));
  }
  template <typename T> void transfer_prolongate_add(const TransferData &td, T *dst_fine, const T *src_coarse, cudaStream_t s)
  {
    launch_transfer<T, true>(td, dst_fine, src_coarse, s);
  }
  template <typename T> void transfer_restrict_add(const TransferData &td, T *dst_coarse, const T *src_fine, cudaStream_t s)
  {
    launch_transfer<T, false>(td, dst_coarse, src_fine, s);
  }

  template void cell_laplace_apply<double>(const CellData &, const double *, double *, const uint32_t *, long long, cudaStream_t);
  template void cell_laplace_apply<float>(const CellData &, const float *, float *, const uint32_t *, long long, cudaStream_t);
  template void cell_laplace_chunked<double>(const CellData &, const ChunkDevice &, const double *, double *, const FusedEpilogue<double> &, cudaStream_t);
  template void cell_laplace_chunked<float>(const CellData &, const ChunkDevice &, const float *, float *, const FusedEpilogue<float> &, cudaStream_t);
  template void idx_fused_identity<double>(double *, const double *, const FusedEpilogue<double> &, const uint32_t *, uint32_t, cudaStream_t);
  template void idx_fused_identity<float>(float *, const float *, const FusedEpilogue<float> &, const uint32_t *, uint32_t, cudaStream_t);
  template void cell_laplace_diagonal<double>(const CellData &, double *, cudaStream_t);
  template void cell_laplace_diagonal<float>(const CellData &, float *, cudaStream_t);
  template void transfer_prolongate_add<double>(const TransferData &, double *, const double *, cudaStream_t);
  template void transfer_prolongate_add<float>(const TransferData &, float *, const float *, cudaStream_t);
  template void transfer_restrict_add<double>(const TransferData &, double *, const double *, cudaStream_t);
  template void transfer_restrict_add<float>(const TransferData &, float *, const float *, cudaStream_t);
} // namespace mgb
