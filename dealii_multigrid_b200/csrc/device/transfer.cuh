// Cell-wise two-level transfer kernels: MGTwoLevelTransfer::prolongate_and_add / restrict_and_add
// (multigrid_throughput.cc:1600-1621).
//   prolongation: gather the coarse cell's DoFs through its evaluation record (Dirichlet -> 0, hanging-node
//                 interpolation), 1-D embedding in x, y, z, multiply by the fine weights, add into the fine cells' own DoFs;
//   restriction : the exact transpose.
// Geometric (h) transfer: the fine patch of a coarse cell is the cell itself (identity) or its 8 Morton-contiguous
// children ((2p+1)^3 nodes, 1-D matrix P_h); polynomial (p) transfer: the same cell at a higher degree (P_p).
//
// One persistent warp per coarse cell (4 warps per block, no block-level synchronisation after the prologue).  The x and
// y embeddings run line-wise through shared memory; the z embedding runs in the column layout of the vmult kernel (lane
// <-> one (I,J) column of the fine patch, values along z in registers) so that the fine-side scatter/gather goes
// straight from/to registers, coalesced along the lexicographic fine node index.  The 1-D matrix is a kernel parameter
// (constant bank), degrees are template parameters.  HBM-bound: s (N_c + 2 N_f) bytes per prolongation.
#pragma once
#include "cell_common.cuh"

#include <algorithm>
#include <stdexcept>

namespace mgb
{
  constexpr int TRANSFER_WARPS = 4;
  // resident blocks per SM the register allocation must allow: the kernels wait on their global loads (ncu: 8.4 warps
  // stalled on long scoreboard per issue at 12 warps/SM), so occupancy pays more than the few spilled bytes cost --
  // measured on B200, hypercube L=7 p=4: prolongation 1.30 -> 1.01 ms, restriction 1.07 -> 0.84 ms from 3 to 5 blocks
  constexpr int transfer_min_blocks(int pf) { return pf <= 4 ? 5 : 2; }

  template <typename T> struct TransferMat
  {
    T v[(2 * MAX_DEGREE + 1) * MAXN]; // NF x nc row-major with leading dimension nc
  };

  template <int PC, int PF, bool HMODE, typename T> struct TransferCfg
  {
    static constexpr int nc  = PC + 1;
    static constexpr int NF  = HMODE ? 2 * PF + 1 : PF + 1;
    static constexpr int PSC = SlabLayout<nc, sizeof(T)>::PS; // plane stride of the coarse cell storage (hanging_node_sweeps_slab)
    static constexpr int n_C = nc * PSC, n_X = NF * nc * nc, n_XY = NF * NF * nc;
    static constexpr int NCH = HMODE ? 8 : 1;
    // per warp: values | fine weights (NCH*27 T) | coarse record (32 words) | fine own entries (NCH*27 words)
    static constexpr size_t val_bytes  = (size_t(n_C + n_X + n_XY) * sizeof(T) + 15) / 16 * 16;
    static constexpr size_t warp_bytes = (val_bytes + 32 * 4 + NCH * 27 * 4 + NCH * 27 * sizeof(T) + 15) / 16 * 16;
    static constexpr size_t lut_bytes  = (size_t(nc * nc * nc) * 4 + 15) / 16 * 16;
    static size_t           block_bytes() { return lut_bytes + TRANSFER_WARPS * warp_bytes; }
  };

  // fine patch coordinate along one axis -> (child bit, local node index, ternary entity digit)
  template <int PF, bool HMODE> __device__ __forceinline__ void fine_axis(const int I, int &c, int &i, int &t)
  {
    c = (HMODE && I > PF) ? 1 : 0;
    i = I - c * PF;
    t = i == 0 ? 0 : (i == PF ? 2 : 1);
  }

  template <int PC, int PF, bool HMODE, bool PROLONGATE, typename T>
  __global__ void __launch_bounds__(TRANSFER_WARPS * 32, transfer_min_blocks(PF))
    k_transfer(const long long n_coarse, const uint32_t *__restrict__ coarse_ctab, const int coarse_any_hanging,
               const uint32_t *__restrict__ fine_own, const uint32_t *__restrict__ fine_first, const uint8_t *__restrict__ n_children,
               const uint8_t *__restrict__ weight_code, const __grid_constant__ TransferMat<T> Pm, T *__restrict__ dst,
               const T *__restrict__ src)
  {
    using Cfg = TransferCfg<PC, PF, HMODE, T>;
    constexpr int nc = Cfg::nc, NF = Cfg::NF, PSC = Cfg::PSC, nc3 = nc * nc * nc;
    constexpr int NRC = (NF * NF + 31) / 32; // column rounds
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t      *slut  = reinterpret_cast<uint32_t *>(smem_raw); // coarse node q -> entity<<24 | offset<<16 | address in C
    const int      warp  = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char *wbase = smem_raw + Cfg::lut_bytes + warp * Cfg::warp_bytes;
    T             *C     = reinterpret_cast<T *>(wbase);
    T             *X     = C + Cfg::n_C;
    T             *XY    = X + Cfg::n_X;
    T             *sw    = reinterpret_cast<T *>(wbase + Cfg::val_bytes);
    uint32_t      *srec  = reinterpret_cast<uint32_t *>(sw + Cfg::NCH * 27);
    uint32_t      *sfo   = srec + 32;

    for (int q = threadIdx.x; q < nc3; q += TRANSFER_WARPS * 32)
      {
        const int i = q % nc, j = (q / nc) % nc, k = q / (nc * nc);
        int       e, off;
        node_to_entity<PC>(i, j, k, e, off);
        slut[q] = (uint32_t(e) << 24) | (uint32_t(off) << 16) | uint32_t(k * PSC + j * nc + i);
      }
    __syncthreads();

    // fine node (I,J,K) -> slot in the per-warp fine tables and offset inside the entity
    const auto fine_slot = [&](const int I, const int J, const int K, int &slot, int &off) {
      int cx, cy, cz, i, j, k, tx, ty, tz;
      fine_axis<PF, HMODE>(I, cx, i, tx);
      fine_axis<PF, HMODE>(J, cy, j, ty);
      fine_axis<PF, HMODE>(K, cz, k, tz);
      const int sx = tx == 1 ? PF - 1 : 1, sy = ty == 1 ? PF - 1 : 1;
      slot = (cx + 2 * cy + 4 * cz) * 27 + tx + 3 * ty + 9 * tz;
      off  = (tx == 1 ? i - 1 : 0) + (ty == 1 ? (j - 1) * sx : 0) + (tz == 1 ? (k - 1) * sx * sy : 0);
    };

    for (long long c = (long long)blockIdx.x * TRANSFER_WARPS + warp; c < n_coarse; c += (long long)gridDim.x * TRANSFER_WARPS)
      {
        // ---- tables
        const int nch = HMODE ? n_children[c] : 1;
#ifndef MGB_TRANSFER_NO_PREFETCH
        {
          // the warp's NEXT coarse cell: pull its record and fine tables into L2 now (no registers held, nothing waits), so
          // that the serial chain table -> gather -> compute of the next iteration starts from L2 hits instead of DRAM
          const long long cn = c + (long long)gridDim.x * TRANSFER_WARPS;
          if (cn < n_coarse)
            {
              const size_t   fn      = size_t(fine_first[cn]) * 27;
              const char    *ptrs[3] = {reinterpret_cast<const char *>(coarse_ctab + cn * 32), reinterpret_cast<const char *>(fine_own + fn),
                                        reinterpret_cast<const char *>(weight_code + fn)};
              constexpr int  lines0 = 1, lines1 = (Cfg::NCH * 27 * 4 + 127) / 128 + 1, lines2 = (Cfg::NCH * 27 + 127) / 128 + 1;
              const int      which  = lane < lines0 ? 0 : (lane < lines0 + lines1 ? 1 : 2);
              const int      idx    = lane - (which == 0 ? 0 : (which == 1 ? lines0 : lines0 + lines1));
              if (lane < lines0 + lines1 + lines2)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(ptrs[which] + size_t(idx) * 128));
            }
        }
#endif
        if (HMODE && nch == 0)
          continue; // local smoothing: parent without children on the fine level (warp-uniform)
        srec[lane]    = coarse_ctab[c * 32 + lane];
        {
          const size_t f0 = size_t(fine_first[c]) * 27;
          for (int idx = lane; idx < nch * 27; idx += 32)
            {
              sfo[idx]            = fine_own[f0 + idx] & D_INDEX_MASK;
              const unsigned code = weight_code[f0 + idx];
              sw[idx]             = code ? T(1) / T(code) : T(0);
            }
        }
        __syncwarp();
        const unsigned mask  = coarse_any_hanging ? (srec[27] & 0xffffu) : 0u;
        const bool     ident = HMODE && nch == 1;

        if (PROLONGATE)
          {
#pragma unroll
            for (int q = lane; q < nc3; q += 32)
              {
                const uint32_t lut = slut[q];
                const uint32_t tb  = srec[lut >> 24];
                C[lut & 0xffffu]   = (tb & D_FLAG_BIT) ? T(0) : src[(tb & D_INDEX_MASK) + ((lut >> 16) & 0xffu)];
              }
            __syncwarp();
            if (mask >> 3)
              hanging_node_sweeps_slab<PC, T, false>(C, srec, 1, lane);
            if (ident)
              {
                for (int q = lane; q < nc3; q += 32)
                  {
                    const uint32_t lut = slut[q];
                    const T        w   = sw[lut >> 24];
                    if (w != T(0))
                      atomic_add(dst + sfo[lut >> 24] + ((lut >> 16) & 0xffu), w * C[lut & 0xffffu]);
                  }
              }
            else
              {
                // x: lines (j,k)
                for (int ln = lane; ln < nc * nc; ln += 32)
                  {
                    const int j = ln % nc, k = ln / nc;
                    T         v[nc];
#pragma unroll
                    for (int l = 0; l < nc; ++l)
                      v[l] = C[k * PSC + j * nc + l];
#pragma unroll
                    for (int I = 0; I < NF; ++I)
                      {
                        T acc = Pm.v[I * nc] * v[0];
#pragma unroll
                        for (int l = 1; l < nc; ++l)
                          acc += Pm.v[I * nc + l] * v[l];
                        X[ln * NF + I] = acc;
                      }
                  }
                __syncwarp();
                // y: lines (I,k)
                for (int ln = lane; ln < NF * nc; ln += 32)
                  {
                    const int I = ln % NF, k = ln / NF;
                    T         v[nc];
#pragma unroll
                    for (int l = 0; l < nc; ++l)
                      v[l] = X[(k * nc + l) * NF + I];
#pragma unroll
                    for (int J = 0; J < NF; ++J)
                      {
                        T acc = Pm.v[J * nc] * v[0];
#pragma unroll
                        for (int l = 1; l < nc; ++l)
                          acc += Pm.v[J * nc + l] * v[l];
                        XY[(k * NF + J) * NF + I] = acc;
                      }
                  }
                __syncwarp();
                // z in the column layout + weighted scatter-add
#pragma unroll
                for (int r = 0; r < NRC; ++r)
                  {
                    const int col = r * 32 + lane;
                    if (col < NF * NF)
                      {
                        const int I = col % NF, J = col / NF;
                        T         v[nc];
#pragma unroll
                        for (int l = 0; l < nc; ++l)
                          v[l] = XY[l * NF * NF + col];
#pragma unroll
                        for (int K = 0; K < NF; ++K)
                          {
                            T acc = Pm.v[K * nc] * v[0];
#pragma unroll
                            for (int l = 1; l < nc; ++l)
                              acc += Pm.v[K * nc + l] * v[l];
                            int slot, off;
                            fine_slot(I, J, K, slot, off);
                            const T w = sw[slot];
                            if (w != T(0))
                              atomic_add(dst + sfo[slot] + off, w * acc);
                          }
                      }
                  }
              }
          }
        else
          {
            if (ident)
              {
                for (int q = lane; q < nc3; q += 32)
                  {
                    const uint32_t lut = slut[q];
                    const T        w   = sw[lut >> 24];
                    C[lut & 0xffffu]   = (w != T(0)) ? w * src[sfo[lut >> 24] + ((lut >> 16) & 0xffu)] : T(0);
                  }
              }
            else
              {
#pragma unroll
                for (int r = 0; r < NRC; ++r)
                  {
                    const int col = r * 32 + lane;
                    if (col < NF * NF)
                      {
                        const int I = col % NF, J = col / NF;
                        T         f[NF];
#pragma unroll
                        for (int K = 0; K < NF; ++K)
                          {
                            int slot, off;
                            fine_slot(I, J, K, slot, off);
                            const T w = sw[slot];
                            f[K]      = (w != T(0)) ? w * src[sfo[slot] + off] : T(0);
                          }
#pragma unroll
                        for (int l = 0; l < nc; ++l)
                          {
                            T acc = Pm.v[l] * f[0];
#pragma unroll
                            for (int K = 1; K < NF; ++K)
                              acc += Pm.v[K * nc + l] * f[K];
                            XY[l * NF * NF + col] = acc;
                          }
                      }
                  }
                __syncwarp();
                for (int ln = lane; ln < NF * nc; ln += 32)
                  {
                    const int I = ln % NF, k = ln / NF;
                    T         f[NF];
#pragma unroll
                    for (int J = 0; J < NF; ++J)
                      f[J] = XY[(k * NF + J) * NF + I];
#pragma unroll
                    for (int l = 0; l < nc; ++l)
                      {
                        T acc = Pm.v[l] * f[0];
#pragma unroll
                        for (int J = 1; J < NF; ++J)
                          acc += Pm.v[J * nc + l] * f[J];
                        X[(k * nc + l) * NF + I] = acc;
                      }
                  }
                __syncwarp();
                for (int ln = lane; ln < nc * nc; ln += 32)
                  {
                    const int j = ln % nc, k = ln / nc;
                    T         f[NF];
#pragma unroll
                    for (int I = 0; I < NF; ++I)
                      f[I] = X[ln * NF + I];
#pragma unroll
                    for (int l = 0; l < nc; ++l)
                      {
                        T acc = Pm.v[l] * f[0];
#pragma unroll
                        for (int I = 1; I < NF; ++I)
                          acc += Pm.v[I * nc + l] * f[I];
                        C[k * PSC + j * nc + l] = acc;
                      }
                  }
              }
            __syncwarp();
            if (mask >> 3)
              hanging_node_sweeps_slab<PC, T, true>(C, srec, 1, lane);
#pragma unroll
            for (int q = lane; q < nc3; q += 32)
              {
                const uint32_t lut = slut[q];
                const uint32_t tb  = srec[lut >> 24];
                if (!(tb & D_FLAG_BIT))
                  atomic_add(dst + (tb & D_INDEX_MASK) + ((lut >> 16) & 0xffu), C[lut & 0xffffu]);
              }
          }
        __syncwarp(); // tables and value buffers are rewritten by the next coarse cell
      }
  }

  template <int PC, int PF, bool HMODE, bool PROLONGATE, typename T>
  void launch_transfer_t(const TransferData &td, T *dst, const T *src, cudaStream_t s)
  {
    using Cfg = TransferCfg<PC, PF, HMODE, T>;
    const size_t smem = Cfg::block_bytes();
    static int   blocks_per_sm = 0, n_sm = 0; // per instantiation
    if (blocks_per_sm == 0)
      {
        int dev = 0;
        MGB_CUDA(cudaGetDevice(&dev));
        MGB_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
        MGB_CUDA(cudaFuncSetAttribute(k_transfer<PC, PF, HMODE, PROLONGATE, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        MGB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, k_transfer<PC, PF, HMODE, PROLONGATE, T>,
                                                               TRANSFER_WARPS * 32, smem));
        if (blocks_per_sm < 1)
          throw std::runtime_error("k_transfer does not fit on an SM");
      }
    TransferMat<T> Pm{};
    for (int I = 0; I < Cfg::NF; ++I)
      for (int l = 0; l < Cfg::nc; ++l)
        Pm.v[I * Cfg::nc + l] = T(td.P[I * Cfg::nc + l]);
    const long long blocks =
      std::min<long long>((td.n_coarse_cells + TRANSFER_WARPS - 1) / TRANSFER_WARPS, (long long)blocks_per_sm * n_sm);
    k_transfer<PC, PF, HMODE, PROLONGATE, T><<<(unsigned)blocks, TRANSFER_WARPS * 32, smem, s>>>(
      td.n_coarse_cells, td.coarse_ctab, td.coarse_any_hanging, td.fine_own, td.fine_first, td.n_children, td.weight_code, Pm, dst, src);
    MGB_CUDA(cudaGetLastError());
  }

  template <typename T, bool PROLONGATE> void launch_transfer(const TransferData &td, T *dst, const T *src, cudaStream_t s)
  {
    if (td.n_coarse_cells <= 0)
      return;
    const int pf = td.degree_fine, pc = td.degree_coarse;
#define MGB_TRANSFER_CASE(PC_, PF_, H_)                                   \
  if (pc == PC_ && pf == PF_ && td.h_mode == H_)                          \
    return launch_transfer_t<PC_, PF_, H_, PROLONGATE, T>(td, dst, src, s);
    MGB_TRANSFER_CASE(1, 1, true)
    MGB_TRANSFER_CASE(2, 2, true)
    MGB_TRANSFER_CASE(3, 3, true)
    MGB_TRANSFER_CASE(4, 4, true)
    MGB_TRANSFER_CASE(5, 5, true)
    MGB_TRANSFER_CASE(6, 6, true)
    // polynomial transfers of create_polynomial_coarsening_sequence(p, bisect) (multigrid_throughput.cc:1506-1510)
    MGB_TRANSFER_CASE(1, 2, false)
    MGB_TRANSFER_CASE(1, 3, false)
    MGB_TRANSFER_CASE(2, 4, false)
    MGB_TRANSFER_CASE(2, 5, false)
    MGB_TRANSFER_CASE(3, 6, false)
#undef MGB_TRANSFER_CASE
    throw std::runtime_error("two-level transfer: unsupported degree pair (fine " + std::to_string(pf) + ", coarse " + std::to_string(pc) +
                             (td.h_mode ? ", geometric)" : ", polynomial)"));
  }
} // namespace mgb
