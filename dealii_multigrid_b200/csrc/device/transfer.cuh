// Cell-wise two-level transfer kernels: MGTwoLevelTransfer::prolongate_and_add / restrict_and_add
// (multigrid_throughput.cc:1600-1621).  One coarse cell per thread block:
//   prolongation: gather coarse DoFs through the coarse evaluation table (Dirichlet -> 0, hanging-node
//                 interpolation), 1-D embedding in x,y,z, multiply by the fine weights, scatter-add into
//                 the fine cells' own DoFs;
//   restriction : exact transpose.
// Geometric transfer: the fine patch is the coarse cell itself (identity) or its 8 children
// ((2p+1)^3 nodes, 1-D matrix P_h); polynomial transfer: same cell, 1-D matrix P_p.
// HBM-bound gather/scatter kernels; runtime degrees (no template explosion over (p_f, p_c) pairs).
#pragma once
#include "cell_common.cuh"

namespace mgb
{
  constexpr int TRANSFER_THREADS = 128;

  __device__ __forceinline__ void node_to_entity_rt(const int P, int i, int j, int k, int &e, int &off)
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

  // runtime-degree version of hanging_node_sweeps for one cell handled by the whole block
  template <typename T, bool TRANSPOSE> __device__ void hanging_node_sweeps_rt(T *u, const int P, const unsigned mask)
  {
    const int n = P + 1, n2 = n * n;
    for (int dd = 0; dd < 3; ++dd)
      {
        const int  d  = TRANSPOSE ? 2 - dd : dd;
        const int  da = d == 0 ? 1 : 0, db = d == 2 ? 1 : 2;
        const int  sa = (mask >> da) & 1, sb = (mask >> db) & 1, sd = (mask >> d) & 1;
        const bool fa = (mask >> (3 + da)) & 1, fb = (mask >> (3 + db)) & 1, ed = (mask >> (6 + d)) & 1;
        const double *S = c_fe[P].S[sd];
        for (int t = threadIdx.x; t < n2; t += blockDim.x)
          {
            const int  a = t % n, b = t / n;
            const bool on_a = (a == sa * P), on_b = (b == sb * P);
            if ((fa && on_a) || (fb && on_b) || (ed && on_a && on_b))
              {
                const int base   = d == 0 ? (b * n + a) * n : (d == 1 ? b * n2 + a : b * n + a);
                const int stride = d == 0 ? 1 : (d == 1 ? n : n2);
                T         v[MAXN], r[MAXN];
                for (int i = 0; i < n; ++i)
                  v[i] = u[base + i * stride];
                for (int i = 0; i < n; ++i)
                  {
                    T acc = 0;
                    for (int j = 0; j < n; ++j)
                      acc += T(TRANSPOSE ? S[j * MAXN + i] : S[i * MAXN + j]) * v[j];
                    r[i] = acc;
                  }
                for (int i = 0; i < n; ++i)
                  u[base + i * stride] = r[i];
              }
          }
        __syncthreads();
      }
  }

  // out[(c*NO + I)*..] 1-D product along one direction of a 3-index array.
  // in: [n2][n1][n0], apply along dimension `dim` a matrix Pm (NO x NI, row-major) or its transpose.
  template <typename T, bool TRANSPOSE>
  __device__ void sweep_rt(const T *in, T *out, const double *Pm, const int rows, const int cols, const int dim, const int n0,
                           const int n1, const int n2)
  {
    // input extent along dim is NI = TRANSPOSE ? rows : cols ; output extent NO = TRANSPOSE ? cols : rows
    const int NI = TRANSPOSE ? rows : cols, NO = TRANSPOSE ? cols : rows;
    const int o0 = dim == 0 ? NO : n0, o1 = dim == 1 ? NO : n1, o2 = dim == 2 ? NO : n2;
    const int total = o0 * o1 * o2;
    for (int q = threadIdx.x; q < total; q += blockDim.x)
      {
        const int i0 = q % o0, i1 = (q / o0) % o1, i2 = q / (o0 * o1);
        const int I  = dim == 0 ? i0 : (dim == 1 ? i1 : i2);
        T         acc = 0;
        for (int l = 0; l < NI; ++l)
          {
            const int j0 = dim == 0 ? l : i0, j1 = dim == 1 ? l : i1, j2 = dim == 2 ? l : i2;
            const T   m  = T(TRANSPOSE ? Pm[l * cols + I] : Pm[I * cols + l]);
            acc += m * in[(j2 * n1 + j1) * n0 + j0];
          }
        out[q] = acc;
      }
    __syncthreads();
  }

  template <typename T, bool PROLONGATE>
  __global__ void __launch_bounds__(TRANSFER_THREADS)
    k_transfer(const int PC, const int PF, const uint32_t *__restrict__ coarse_etab, const uint16_t *__restrict__ coarse_mask,
               const uint32_t *__restrict__ fine_own, const uint32_t *__restrict__ fine_first, const uint8_t *__restrict__ n_children,
               const uint8_t *__restrict__ weight_code, const double *__restrict__ P_h, const double *__restrict__ P_p,
               T *__restrict__ dst, const T *__restrict__ src)
  {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T *smem = reinterpret_cast<T *>(smem_raw);
    __shared__ uint32_t stab[27];

    const long long c    = blockIdx.x;
    const int       nc   = PC + 1;
    const int       nch  = n_children[c];
    const bool      hx   = nch == 8;                 // h-embedding into 8 children
    const bool      ident = (!hx) && PF == PC;       // same cell, same degree
    const int       NF   = hx ? 2 * PF + 1 : PF + 1; // 1-D size of the fine patch
    const double   *Pm   = hx ? P_h : P_p;           // NF x nc
    const unsigned  mask = coarse_mask ? coarse_mask[c] : 0u;
    const uint32_t  f0   = fine_first[c];

    T *bufC = smem;                    // nc^3
    T *buf1 = bufC + nc * nc * nc;     // NF*nc*nc
    T *buf2 = buf1 + NF * nc * nc;     // NF*NF*nc
    T *bufF = buf2 + NF * NF * nc;     // NF^3

    if (threadIdx.x < 27)
      stab[threadIdx.x] = coarse_etab[c * 27 + threadIdx.x];
    __syncthreads();

    // fine patch node -> (index, weight)
    const auto fine_node = [&](int q, uint32_t &idx, T &w) {
      const int I = q % NF, J = (q / NF) % NF, K = q / (NF * NF);
      int       cx = 0, cy = 0, cz = 0, i = I, j = J, k = K;
      if (hx)
        {
          cx = I > PF, cy = J > PF, cz = K > PF;
          i -= cx * PF, j -= cy * PF, k -= cz * PF;
        }
      int e, off;
      node_to_entity_rt(PF, i, j, k, e, off);
      const size_t   slot = (size_t(f0) + cx + 2 * cy + 4 * cz) * 27 + e;
      const unsigned code = weight_code[slot];
      idx                 = (fine_own[slot] & D_INDEX_MASK) + off;
      w                   = code ? T(1) / T(code) : T(0);
    };

    if (PROLONGATE)
      {
        for (int q = threadIdx.x; q < nc * nc * nc; q += blockDim.x)
          {
            const int i = q % nc, j = (q / nc) % nc, k = q / (nc * nc);
            int       e, off;
            node_to_entity_rt(PC, i, j, k, e, off);
            const uint32_t tb = stab[e];
            bufC[q]           = (tb & D_FLAG_BIT) ? T(0) : src[(tb & D_INDEX_MASK) + off];
          }
        __syncthreads();
        if (mask >> 3)
          hanging_node_sweeps_rt<T, false>(bufC, PC, mask);
        const T *fine = bufC;
        if (!ident)
          {
            sweep_rt<T, false>(bufC, buf1, Pm, NF, nc, 0, nc, nc, nc);
            sweep_rt<T, false>(buf1, buf2, Pm, NF, nc, 1, NF, nc, nc);
            sweep_rt<T, false>(buf2, bufF, Pm, NF, nc, 2, NF, NF, nc);
            fine = bufF;
          }
        for (int q = threadIdx.x; q < NF * NF * NF; q += blockDim.x)
          {
            uint32_t idx;
            T        w;
            fine_node(q, idx, w);
            if (w != T(0))
              atomic_add(dst + idx, w * fine[q]);
          }
      }
    else
      {
        T *fine = ident ? bufC : bufF;
        for (int q = threadIdx.x; q < NF * NF * NF; q += blockDim.x)
          {
            uint32_t idx;
            T        w;
            fine_node(q, idx, w);
            fine[q] = (w != T(0)) ? w * src[idx] : T(0);
          }
        __syncthreads();
        if (!ident)
          {
            sweep_rt<T, true>(bufF, buf2, Pm, NF, nc, 2, NF, NF, NF);
            sweep_rt<T, true>(buf2, buf1, Pm, NF, nc, 1, NF, NF, nc);
            sweep_rt<T, true>(buf1, bufC, Pm, NF, nc, 0, NF, nc, nc);
          }
        if (mask >> 3)
          hanging_node_sweeps_rt<T, true>(bufC, PC, mask);
        for (int q = threadIdx.x; q < nc * nc * nc; q += blockDim.x)
          {
            const int i = q % nc, j = (q / nc) % nc, k = q / (nc * nc);
            int       e, off;
            node_to_entity_rt(PC, i, j, k, e, off);
            const uint32_t tb = stab[e];
            if (!(tb & D_FLAG_BIT))
              atomic_add(dst + (tb & D_INDEX_MASK) + off, bufC[q]);
          }
      }
  }

  template <typename T, bool PROLONGATE> void launch_transfer(const TransferData &td, T *dst, const T *src, cudaStream_t s)
  {
    if (td.n_coarse_cells <= 0)
      return;
    const int    nc = td.degree_coarse + 1;
    const int    NF = std::max(2 * td.degree_fine + 1, td.degree_fine + 1);
    const size_t sh = sizeof(T) * size_t(nc * nc * nc + NF * nc * nc + NF * NF * nc + NF * NF * NF);
    k_transfer<T, PROLONGATE><<<(unsigned)td.n_coarse_cells, TRANSFER_THREADS, sh, s>>>(
      td.degree_coarse, td.degree_fine, td.coarse_etab, td.coarse_any_hanging ? td.coarse_mask : nullptr, td.fine_own, td.fine_first,
      td.n_children, td.weight_code, td.P_h, td.P_p, dst, src);
    MGB_CUDA(cudaGetLastError());
  }
} // namespace mgb
