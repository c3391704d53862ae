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
    // fast diagonalisation (fe1d.h): K = Tm^T diag(lam) Tm, M = Tm^T Tm; lam2[j*MAXN+k] = lam[j]+lam[k]
    double Tm[MAXN * MAXN];
    double lam[MAXN];
    double lam2[MAXN * MAXN];
  };
  struct FEConstF // single-precision copy of the sweep data (MGNumberType=float kernels)
  {
    float Tm[MAXN * MAXN];
    float lam[MAXN];
    float lam2[MAXN * MAXN];
  };
  // defined once: this header is included by exactly one translation unit (cells.cu)
  __constant__ FEConst  c_fe[MAX_DEGREE + 1];
  __constant__ FEConstF c_fef[MAX_DEGREE + 1];

  template <int P, typename T> struct FDM;
  template <int P> struct FDM<P, double>
  {
    static __device__ __forceinline__ double Tm(int a, int l) { return c_fe[P].Tm[a * MAXN + l]; }
    static __device__ __forceinline__ double lam(int a) { return c_fe[P].lam[a]; }
    static __device__ __forceinline__ double lam2(int j, int k) { return c_fe[P].lam2[j * MAXN + k]; }
  };
  template <int P> struct FDM<P, float>
  {
    static __device__ __forceinline__ float Tm(int a, int l) { return c_fef[P].Tm[a * MAXN + l]; }
    static __device__ __forceinline__ float lam(int a) { return c_fef[P].lam[a]; }
    static __device__ __forceinline__ float lam2(int j, int k) { return c_fef[P].lam2[j * MAXN + k]; }
  };

  // 1-D sweeps on a register line v[0..n-1] (stride S inside a register array), even-odd form:
  // forward  v <- Tm v      (nodal values -> modes; symmetric modes first)
  // backward v <- Tm^T v    (modes -> nodal)
  template <int P, typename T, int S> __device__ __forceinline__ void fdm_forward(T *v)
  {
    constexpr int n = P + 1, ne = (n + 1) / 2, no = n / 2;
    T             e[ne], o[no > 0 ? no : 1];
#pragma unroll
    for (int l = 0; l < no; ++l)
      {
        e[l] = v[l * S] + v[(n - 1 - l) * S];
        o[l] = v[l * S] - v[(n - 1 - l) * S];
      }
    if (n % 2 == 1)
      e[ne - 1] = v[(n / 2) * S];
#pragma unroll
    for (int a = 0; a < ne; ++a)
      {
        T acc = FDM<P, T>::Tm(a, 0) * e[0];
#pragma unroll
        for (int l = 1; l < ne; ++l)
          acc += FDM<P, T>::Tm(a, l) * e[l];
        v[a * S] = acc;
      }
#pragma unroll
    for (int a = 0; a < no; ++a)
      {
        T acc = FDM<P, T>::Tm(ne + a, 0) * o[0];
#pragma unroll
        for (int l = 1; l < no; ++l)
          acc += FDM<P, T>::Tm(ne + a, l) * o[l];
        v[(ne + a) * S] = acc;
      }
  }
  template <int P, typename T, int S> __device__ __forceinline__ void fdm_backward(T *v)
  {
    constexpr int n = P + 1, ne = (n + 1) / 2, no = n / 2;
    T             e[ne], o[no > 0 ? no : 1];
#pragma unroll
    for (int l = 0; l < ne; ++l)
      {
        T acc = FDM<P, T>::Tm(0, l) * v[0];
#pragma unroll
        for (int a = 1; a < ne; ++a)
          acc += FDM<P, T>::Tm(a, l) * v[a * S];
        e[l] = acc;
      }
#pragma unroll
    for (int l = 0; l < no; ++l)
      {
        T acc = FDM<P, T>::Tm(ne, l) * v[ne * S];
#pragma unroll
        for (int a = 1; a < no; ++a)
          acc += FDM<P, T>::Tm(ne + a, l) * v[(ne + a) * S];
        o[l] = acc;
      }
#pragma unroll
    for (int l = 0; l < no; ++l)
      {
        v[l * S]           = e[l] + o[l];
        v[(n - 1 - l) * S] = e[l] - o[l];
      }
    if (n % 2 == 1)
      v[(n / 2) * S] = e[ne - 1];
  }

  // shared-memory layout of the slab kernels: one cell = n planes (k) of n*n values with plane stride PS and cell
  // stride CS, chosen so that the slab pattern, lane (cell, plane) -> cell*CS + plane*PS, and the column pattern,
  // lane c -> (c / n^2)*CS + c % n^2, are bank-conflict free (64-bit: per half-warp on 16 banks; 32-bit: per warp).
  template <int N, int BYTES> struct SlabLayout;
  template <> struct SlabLayout<2, 8> { static constexpr int PS = 5, CS = 10; };
  template <> struct SlabLayout<3, 8> { static constexpr int PS = 19, CS = 57; };
  template <> struct SlabLayout<4, 8> { static constexpr int PS = 17, CS = 68; };
  template <> struct SlabLayout<5, 8> { static constexpr int PS = 37, CS = 185; };
  template <> struct SlabLayout<6, 8> { static constexpr int PS = 37, CS = 222; };
  template <> struct SlabLayout<7, 8> { static constexpr int PS = 55, CS = 385; };
  template <> struct SlabLayout<2, 4> { static constexpr int PS = 5, CS = 10; };
  template <> struct SlabLayout<3, 4> { static constexpr int PS = 26, CS = 105; };
  template <> struct SlabLayout<4, 4> { static constexpr int PS = 17, CS = 68; };
  template <> struct SlabLayout<5, 4> { static constexpr int PS = 27, CS = 153; };
  template <> struct SlabLayout<6, 4> { static constexpr int PS = 37, CS = 222; };
  template <> struct SlabLayout<7, 4> { static constexpr int PS = 50, CS = 369; };

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

  // asynchronous global -> shared copy of one vector entry (LDGSTS); !valid zero-fills (Dirichlet entries)
  template <typename T> __device__ __forceinline__ void cp_async_entry(T *smem_dst, const T *gmem_src, const bool valid)
  {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    const unsigned b = valid ? (unsigned)sizeof(T) : 0u;
    if (sizeof(T) == 8)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(s), "l"(gmem_src), "r"(b) : "memory");
    else
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(s), "l"(gmem_src), "r"(b) : "memory");
  }
  __device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

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
