// bench_bulk.cu -- micro-benchmark for the round-2 vmult design: can the TMA unit move a cell's
// ~10 short address spans per direction fast enough?  Measures, per SM,
//   gather : cp.async.bulk.shared::cluster.global (1-D, mbarrier complete_tx) of S spans of B bytes per warp-iteration
//   scatter: cp.reduce.async.bulk.global.shared::cta .add.f64 of S spans of B bytes per warp-iteration
// from / to pseudo-random 16-byte-aligned offsets of a buffer larger than L2, and checks the reduce-add result.
// The vmult kernel at 0.45 of the HBM roofline needs one p=4 cell per ~96 cycles per SM, i.e. ~10 + 10 spans of
// 16..512 bytes (mean ~100 bytes) per 96 cycles per SM.
//
// build (here):  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/bench_bulk tools/bench_bulk.cu
// run (GPU box): timeout 120 tools/bench_bulk            (prints one line per span size; waits are bounded)
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

constexpr int WARPS     = 4;   // warps per block, each with its own staging area and mbarrier
constexpr int MAX_SPANS = 16;  // spans per warp-iteration
constexpr int MAX_BYTES = 512; // bytes per span

__device__ __forceinline__ unsigned smem_addr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint32_t hash32(uint32_t x)
{
  x ^= x >> 16;
  x *= 0x7feb352du;
  x ^= x >> 15;
  x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}

// bounded wait on an mbarrier phase; returns false on timeout
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

// MODE 0: gather (global -> shared), MODE 1: scatter reduce-add f64 (shared -> global)
template <int MODE>
__global__ void __launch_bounds__(WARPS * 32)
  k_bulk(double *buf, const uint32_t n_slots /* 16-byte slots in buf */, const int spans, const int bytes, const int iters, int *fail,
         double *sink)
{
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long bars[WARPS];
  const int      warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char *stage = smem + (size_t)warp * MAX_SPANS * MAX_BYTES;
  const unsigned bar   = smem_addr(&bars[warp]);
  if (lane == 0)
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
  if (MODE == 1)
    for (int i = lane; i < MAX_SPANS * MAX_BYTES / 8; i += 32)
      reinterpret_cast<double *>(stage)[i] = 1.0;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic-proxy writes (init, staging) before async-proxy use
  __syncwarp();
  const uint32_t wid      = (blockIdx.x * WARPS + warp);
  const uint32_t max_slot = n_slots - (uint32_t)(bytes / 16);
  double         acc      = 0;
  for (int it = 0; it < iters; ++it)
    {
      if (MODE == 0)
        {
          if (lane == 0)
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(spans * bytes) : "memory");
          __syncwarp();
          if (lane < spans)
            {
              const uint32_t slot = hash32(wid * 7919u + it * 31u + lane) % max_slot;
              const char    *src  = reinterpret_cast<const char *>(buf) + (size_t)slot * 16;
              asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             smem_addr(stage + lane * MAX_BYTES)),
                           "l"(src), "r"(bytes), "r"(bar)
                           : "memory");
            }
          if (!mbar_wait(bar, it & 1))
            {
              atomicExch(fail, 1);
              return;
            }
          if (lane < spans)
            acc += reinterpret_cast<const double *>(stage + lane * MAX_BYTES)[0];
          __syncwarp();
        }
      else
        {
          if (lane < spans)
            {
              const uint32_t slot = hash32(wid * 7919u + it * 31u + lane) % max_slot;
              char          *dst  = reinterpret_cast<char *>(buf) + (size_t)slot * 16;
              asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], %2;" ::"l"(dst),
                           "r"(smem_addr(stage + lane * MAX_BYTES)), "r"(bytes)
                           : "memory");
              asm volatile("cp.async.bulk.commit_group;" ::: "memory");
              // the staging area may be rewritten once the reads of the group are done (the data here never changes, the
              // wait is what a real kernel would pay)
              asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
          __syncwarp();
        }
    }
  if (MODE == 1 && lane < spans)
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  if (acc == 123.456)
    *sink = acc;
}

int main(int argc, char **argv)
{
  const int    iters = argc > 1 ? std::atoi(argv[1]) : 2000;
  const bool   quick = argc > 2; // any second argument: 256 MiB buffer, three span sizes, no host-side sum check
  const size_t bytes_buf = quick ? size_t(1) << 28 : size_t(1) << 30; // > L2 either way
  double      *buf;
  int         *fail;
  double      *sink;
  CK(cudaMalloc(&buf, bytes_buf));
  CK(cudaMalloc(&fail, sizeof(int)));
  CK(cudaMalloc(&sink, sizeof(double)));
  CK(cudaMemset(fail, 0, sizeof(int)));
  int dev = 0, n_sm = 0, khz = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
  const size_t smem = size_t(WARPS) * MAX_SPANS * MAX_BYTES;
  CK(cudaFuncSetAttribute(k_bulk<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(k_bulk<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  const uint32_t n_slots = (uint32_t)(bytes_buf / 16);
  std::printf("# %d SMs, %d kHz nominal; %d warps/block, spans of B bytes at random 16-byte-aligned offsets of a %zu MiB buffer\n", n_sm, khz,
              WARPS, bytes_buf >> 20);
  std::printf("# mode blocks/SM spans bytes  ms   Gspans/s  spans/cycle/SM  GB/s\n");
  for (int mode = 0; mode < 2; ++mode)
    for (int bps : {1, 2, 4})
      for (int spans : {10})
        for (int bytes : {16, 32, 64, 128, 256, 512})
          {
            if (quick && (bps != 2 || bytes == 16 || bytes == 64 || bytes == 256))
              continue;
            CK(cudaMemset(buf, 0, bytes_buf));
            const int grid = n_sm * bps;
            for (int rep = 0; rep < 2; ++rep) // first pass warms up
              {
                if (rep == 1)
                  CK(cudaMemset(buf, 0, bytes_buf));
                CK(cudaEventRecord(e0));
                if (mode == 0)
                  k_bulk<0><<<grid, WARPS * 32, smem>>>(buf, n_slots, spans, bytes, iters, fail, sink);
                else
                  k_bulk<1><<<grid, WARPS * 32, smem>>>(buf, n_slots, spans, bytes, iters, fail, sink);
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                CK(cudaGetLastError());
              }
            float ms = 0;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            int h_fail = 0;
            CK(cudaMemcpy(&h_fail, fail, sizeof(int), cudaMemcpyDeviceToHost));
            if (h_fail)
              {
                std::printf("TIMEOUT in mode %d (mbarrier wait)\n", mode);
                return 2;
              }
            const double n_spans = double(grid) * WARPS * spans * iters;
            std::printf("%s %d %d %d  %.3f  %.2f  %.3f  %.1f", mode == 0 ? "gather " : "scatter", bps, spans, bytes, ms, n_spans / ms / 1e6,
                        n_spans / (ms * 1e-3 * khz * 1e3) / n_sm, n_spans * bytes / ms / 1e6);
            if (mode == 1 && !quick)
              {
                // every reduce added 1.0 to bytes/8 doubles: the buffer must sum to n_spans * bytes / 8
                std::vector<double> h(bytes_buf / 8);
                CK(cudaMemcpy(h.data(), buf, bytes_buf, cudaMemcpyDeviceToHost));
                double sum = 0;
                for (double v : h)
                  sum += v;
                std::printf("  sum %s", sum == n_spans * bytes / 8 ? "ok" : "WRONG");
              }
            std::printf("\n");
          }
  return 0;
}
