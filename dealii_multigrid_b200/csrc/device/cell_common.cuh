// Shared device helpers for the cell-wise kernels (operator.cu, transfer.cu).
#pragma once
#include "device.h"

namespace mgb
{
  constexpr uint32_t D_FLAG_BIT   = 0x80000000u;
  constexpr uint32_t D_INDEX_MASK = 0x7fffffffu;

  // 1-D matrices of FE_Q(p) on the unit interval, per degree (row-major, leading dimension MAXN)
  constexpr int MAXN = MAX_DEGREE + 1;
  struct FEConst
  {
    double M[MAXN * MAXN];
    double K[MAXN * MAXN];
    double S[2][MAXN * MAXN];
  };
  // defined once: this header is included by exactly one translation unit (cells.cu)
  __constant__ FEConst c_fe[MAX_DEGREE + 1];

  // lexicographic node (i,j,k) of FE_Q(P) -> entity (ternary) and offset inside the entity
  template <int P> __device__ __forceinline__ void node_to_entity(int i, int j, int k, int &e, int &off)
  {
    const int tx = i == 0 ? 0 : (i == P ? 2 : 1);
    const int ty = j == 0 ? 0 : (j == P ? 2 : 1);
    const int tz = k == 0 ? 0 : (k == P ? 2 : 1);
    e            = tx + 3 * ty + 9 * tz;
    int stride   = 1;
    off          = 0;
    if (tx == 1)
      {
        off += (i - 1);
        stride = P - 1;
      }
    if (ty == 1)
      {
        off += (j - 1) * stride;
        stride *= (P - 1);
      }
    if (tz == 1)
      off += (k - 1) * stride;
  }

  template <typename T> __device__ __forceinline__ void atomic_add(T *p, T v) { atomicAdd(p, v); }

  // In-place hanging-node interpolation of one cell's nodal values `u` (n^3, lexicographic, in shared
  // memory) -- SURVEY.md A.3 step 2 (TRANSPOSE=false, before evaluation) and step 6 (TRANSPOSE=true,
  // after integration).  Direction by direction, every line along d that lies in a constrained face
  // (normal != d) or is the constrained edge along d gets the 1-D subface matrix S[subcell_d] applied.
  // Called by the n^2 threads that own the cell (t = thread index within the cell); contains barriers:
  // ALL threads of the block must call it (cells without constraints pass mask with no face/edge bits).
  template <int P, typename T, bool TRANSPOSE>
  __device__ __forceinline__ void hanging_node_sweeps(T *u, const unsigned mask, const int t, const bool active)
  {
    constexpr int n = P + 1, n2 = n * n;
    const int     a = t % n, b = t / n;
#pragma unroll
    for (int dd = 0; dd < 3; ++dd)
      {
        const int d = TRANSPOSE ? 2 - dd : dd;
        // perpendicular dims in index order: d=0 -> (y,z); d=1 -> (x,z); d=2 -> (x,y)
        const int  da = d == 0 ? 1 : 0, db = d == 2 ? 1 : 2;
        const int  sa = (mask >> da) & 1, sb = (mask >> db) & 1, sd = (mask >> d) & 1;
        const bool fa = (mask >> (3 + da)) & 1, fb = (mask >> (3 + db)) & 1, ed = (mask >> (6 + d)) & 1;
        const bool on_a = (a == sa * P), on_b = (b == sb * P);
        const bool doit = active && t < n2 && ((fa && on_a) || (fb && on_b) || (ed && on_a && on_b));
        if (doit)
          {
            const int base   = d == 0 ? (b * n + a) * n : (d == 1 ? b * n2 + a : b * n + a);
            const int stride = d == 0 ? 1 : (d == 1 ? n : n2);
            T         v[n], r[n];
#pragma unroll
            for (int i = 0; i < n; ++i)
              v[i] = u[base + i * stride];
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
              u[base + i * stride] = r[i];
          }
        __syncthreads();
      }
  }
} // namespace mgb
