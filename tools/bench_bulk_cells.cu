// bench_bulk_cells.cu -- feasibility test of the round-2 vmult gather/scatter on the REAL address
// spans of a mesh: every warp walks over batches of 6 cells like k_laplace_apply, but moves the cells' DoFs with 1-D bulk
// copies (cp.async.bulk global -> shared, mbarrier complete_tx, two batches in flight) and, in the second mode, adds a
// staged cell back with cp.reduce.async.bulk .add.f64.  Input: the span table written by tools/span_table.py.
//
// Prints the time of one pass over all cells for (a) gather only, (b) gather + reduce-add scatter, to be compared with the
// whole cell kernel (k_laplace_apply<4,double>: 0.173 ms for the 262 144 cells of L=6, 1.27 ms for L=7).  The gather result
// is checked (sum of the staged values of a vector of ones = number of staged DoFs), the scatter by the vector's sum.
//
// build (here):  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/bench_bulk_cells tools/bench_bulk_cells.cu
// run (GPU box): python tools/span_table.py --nref 6  (here, before the call; writes tools/spans_L6_p4.bin)
//                timeout 120 tools/bench_bulk_cells tools/spans_L6_p4.bin
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                                                     \
  do                                                                                                              \
    {                                                                                                             \
      cudaError_t e_ = (x);                                                                                       \
      if (e_ != cudaSuccess)                                                                                      \
        {                                                                                                         \
          std::fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, cudaGetErrorString(e_));                         \
          std::exit(1);                                                                                           \
        }                                                                                                         \
    }                                                                                                             \
  while (0)

constexpr int CW         = 6;   // cells per warp batch (p = 4: 32 / 5)
constexpr int STAGE_DOFS = 160; // staging doubles per cell (125 needed + rounding to 16-byte boundaries)
constexpr int MAX_SPANS  = 20;  // as written by span_table.py

__device__ __forceinline__ unsigned smem_addr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ bool mbar_wait(unsigned bar, unsigned parity)
{
  for (int spin = 0; spin < (1 << 22); ++spin)
    {
      unsigned ok;
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(ok)
                   : "r"(bar), "r"(parity)
                   : "memory");
      if (ok)
        return true;
    }
  return false;
}

// SCATTER = false: gather only; true: gather, then reduce-add the staged cells into dst
template <bool SCATTER>
__global__ void k_cells(const uint32_t *__restrict__ spans /* [n_cells][MAX_SPANS][3] */, const long long n_cells, const double *src,
                        double *dst, double *sums /* per block */, int *fail)
{
  extern __shared__ __align__(128) unsigned char smem[];
  const int  warp = threadIdx.x >> 5, lane = threadIdx.x & 31, W = blockDim.x >> 5;
  double    *stage = reinterpret_cast<double *>(smem) + size_t(warp) * 2 * CW * STAGE_DOFS; // [2][CW][STAGE_DOFS]
  __shared__ __align__(8) unsigned long long bars[32][2];
  __shared__ double                          block_sum[32];
  const unsigned bar0 = smem_addr(&bars[warp][0]), bar1 = smem_addr(&bars[warp][1]);
  if (lane == 0)
    {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0) : "memory");
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar1) : "memory");
    }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncwarp();

  const long long n_batches = (n_cells + CW - 1) / CW;
  const long long stride    = (long long)gridDim.x * W;
  long long       b         = (long long)blockIdx.x * W + warp;

  // issue the bulk copies of one batch into staging buffer `buf`; returns the staged DoFs of this lane's spans
  const auto issue = [&](const long long batch, const int buf) -> unsigned {
    unsigned my_dofs = 0;
    if (batch < n_batches)
      {
        const long long first = batch * CW;
        const int       ncw   = (int)min((long long)CW, n_cells - first);
        unsigned        desc[(CW * MAX_SPANS + 31) / 32][3];
#pragma unroll
        for (int r = 0; r < (CW * MAX_SPANS + 31) / 32; ++r)
          {
            const int w = r * 32 + lane, c = w / MAX_SPANS, k = w - c * MAX_SPANS;
            desc[r][1]  = 0;
            if (c < ncw)
              {
                const uint32_t *t = spans + ((first + c) * MAX_SPANS + k) * 3;
                desc[r][0] = t[0];
                desc[r][1] = t[1];
                desc[r][2] = t[2] + c * STAGE_DOFS;
                my_dofs += t[1];
              }
          }
        unsigned total = my_dofs;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
          total += __shfl_xor_sync(0xffffffffu, total, o);
        const unsigned bar = buf ? bar1 : bar0;
        if (lane == 0)
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(total * 8u) : "memory");
        __syncwarp();
#pragma unroll
        for (int r = 0; r < (CW * MAX_SPANS + 31) / 32; ++r)
          if (desc[r][1])
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                           smem_addr(stage + size_t(buf) * CW * STAGE_DOFS + desc[r][2])),
                         "l"(src + desc[r][0]), "r"(desc[r][1] * 8u), "r"(bar)
                         : "memory");
      }
    return my_dofs;
  };

  double   acc = 0;
  unsigned ph[2] = {0, 0};
  issue(b, 0);
  int it = 0;
  for (; b < n_batches; b += stride, ++it)
    {
      const int buf = it & 1;
      issue(b + stride, buf ^ 1); // the other buffer was consumed (and its reduce reads drained) in the previous iteration
      if (!mbar_wait(buf ? bar1 : bar0, ph[buf]))
        {
          atomicExch(fail, 1);
          return;
        }
      ph[buf] ^= 1;
      const long long first = b * CW;
      const int       ncw   = (int)min((long long)CW, n_cells - first);
      double         *sb    = stage + size_t(buf) * CW * STAGE_DOFS;
      // consume: what the column-layout read of the cell kernel would do (every staged value once)
      for (int c = 0; c < ncw; ++c)
        {
          const uint32_t *t = spans + ((first + c) * MAX_SPANS) * 3;
          unsigned        n = 0;
          for (int k = 0; k < MAX_SPANS; ++k)
            n += t[k * 3 + 1];
          for (unsigned i = lane; i < n; i += 32)
            acc += sb[c * STAGE_DOFS + i];
        }
      if (SCATTER)
        {
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
#pragma unroll
          for (int r = 0; r < (CW * MAX_SPANS + 31) / 32; ++r)
            {
              const int w = r * 32 + lane, c = w / MAX_SPANS, k = w - c * MAX_SPANS;
              if (c < ncw)
                {
                  const uint32_t *t = spans + ((first + c) * MAX_SPANS + k) * 3;
                  if (t[1])
                    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], %2;" ::"l"(dst + t[0]),
                                 "r"(smem_addr(sb + c * STAGE_DOFS + t[2])), "r"(t[1] * 8u)
                                 : "memory");
                }
            }
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); // staging buffer reusable
        }
      __syncwarp();
    }
  if (SCATTER)
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0)
    block_sum[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0)
    {
      double s = 0;
      for (int w = 0; w < W; ++w)
        s += block_sum[w];
      sums[blockIdx.x] = s;
    }
}

int main(int argc, char **argv)
{
  if (argc < 2)
    {
      std::fprintf(stderr, "usage: %s spans.bin [warps_per_block=4]\n", argv[0]);
      return 1;
    }
  FILE *f = std::fopen(argv[1], "rb");
  if (!f)
    {
      std::perror(argv[1]);
      return 1;
    }
  long long hdr[3];
  if (std::fread(hdr, sizeof(long long), 3, f) != 3 || hdr[1] != MAX_SPANS)
    {
      std::fprintf(stderr, "bad header\n");
      return 1;
    }
  const long long       n_cells = hdr[0], n_dofs = hdr[2];
  std::vector<uint32_t> h_spans(size_t(n_cells) * MAX_SPANS * 3);
  if (std::fread(h_spans.data(), 4, h_spans.size(), f) != h_spans.size())
    {
      std::fprintf(stderr, "short table\n");
      return 1;
    }
  std::fclose(f);
  double staged = 0;
  for (size_t i = 0; i < size_t(n_cells) * MAX_SPANS; ++i)
    staged += h_spans[i * 3 + 1];
  const int W = argc > 2 ? std::atoi(argv[2]) : 4;

  uint32_t *d_spans;
  double   *src, *dst, *sums;
  int      *fail;
  CK(cudaMalloc(&d_spans, h_spans.size() * 4));
  CK(cudaMemcpy(d_spans, h_spans.data(), h_spans.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&src, n_dofs * 8));
  CK(cudaMalloc(&dst, n_dofs * 8));
  CK(cudaMalloc(&fail, 4));
  CK(cudaMemset(fail, 0, 4));
  std::vector<double> ones(n_dofs, 1.0);
  CK(cudaMemcpy(src, ones.data(), n_dofs * 8, cudaMemcpyHostToDevice));
  int dev = 0, n_sm = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  const size_t smem = size_t(W) * 2 * CW * STAGE_DOFS * 8;
  CK(cudaFuncSetAttribute(k_cells<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(k_cells<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  std::printf("# %lld cells, %lld DoFs, %.1f staged DoFs per cell; %d warps per block, %zu bytes of staging per block\n", n_cells, n_dofs,
              staged / n_cells, W, smem);
  std::printf("# mode blocks/SM  ms/pass  Mcells/s  cycles/cell/SM(@1.965GHz)  check\n");
  for (int mode = 0; mode < 2; ++mode)
    for (int bps = 1; bps <= 227 * 1024 / (int)(smem + 1024) && bps <= 8; ++bps)
      {
        const int grid = n_sm * bps;
        CK(cudaMalloc(&sums, grid * 8));
        float best = 1e30f;
        for (int rep = 0; rep < 5; ++rep)
          {
            CK(cudaMemset(dst, 0, n_dofs * 8));
            CK(cudaEventRecord(e0));
            if (mode == 0)
              k_cells<false><<<grid, W * 32, smem>>>(d_spans, n_cells, src, dst, sums, fail);
            else
              k_cells<true><<<grid, W * 32, smem>>>(d_spans, n_cells, src, dst, sums, fail);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            CK(cudaGetLastError());
            float ms = 0;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (rep > 0 && ms < best)
              best = ms;
          }
        int h_fail = 0;
        CK(cudaMemcpy(&h_fail, fail, 4, cudaMemcpyDeviceToHost));
        if (h_fail)
          {
            std::printf("TIMEOUT (mbarrier wait)\n");
            return 2;
          }
        std::vector<double> h_sums(grid);
        CK(cudaMemcpy(h_sums.data(), sums, grid * 8, cudaMemcpyDeviceToHost));
        double got = 0;
        for (double v : h_sums)
          got += v;
        bool ok = got == staged;
        if (mode == 1)
          {
            std::vector<double> h_dst(n_dofs);
            CK(cudaMemcpy(h_dst.data(), dst, n_dofs * 8, cudaMemcpyDeviceToHost));
            double s = 0;
            for (double v : h_dst)
              s += v;
            ok = ok && s == staged;
          }
        std::printf("%s %d  %.4f  %.1f  %.1f  %s\n", mode == 0 ? "gather        " : "gather+scatter", bps, best, n_cells / best / 1e3,
                    best * 1e-3 * 1.965e9 * n_sm / n_cells, ok ? "ok" : "WRONG");
        CK(cudaFree(sums));
      }
  return 0;
}
