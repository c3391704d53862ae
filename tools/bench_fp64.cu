// bench_fp64.cu -- measured FP64 peak of the device (DFMA throughput on all SMs), the second roof of the cell kernel
// (the reference's operation count makes the algorithm FP64-bound).  8 independent FMA chains per thread,
// 1024 threads per block, 2 blocks per SM; prints TFLOP/s (2 flop per FMA) for a few repetitions.
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(1024) k_dfma(double *out, const int iters, const double a, const double b)
{
  double x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
    x[i] = 1.0 + 1e-3 * (threadIdx.x + i);
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int r = 0; r < 16; ++r)
#pragma unroll
      for (int i = 0; i < 8; ++i)
        x[i] = fma(x[i], a, b);
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i)
    s += x[i];
  if (s == 12345.678)
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main()
{
  int dev = 0, n_sm = 0, clk = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, dev);
  double *out;
  cudaMalloc(&out, sizeof(double) * 1024 * n_sm * 2);
  const int   iters = 4000, blocks = n_sm * 2;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  k_dfma<<<blocks, 1024>>>(out, 100, 0.999999, 1e-7);
  cudaDeviceSynchronize();
  printf("# %d SMs, nominal %d kHz; %d blocks x 1024 threads, 8 chains, %d x 128 FMA per thread\n", n_sm, clk, blocks, iters);
  for (int rep = 0; rep < 5; ++rep)
    {
      cudaEventRecord(e0);
      k_dfma<<<blocks, 1024>>>(out, iters, 0.999999, 1e-7);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      const double flop = 2.0 * 128.0 * iters * 1024.0 * blocks;
      printf("rep %d: %.3f ms  %.2f TFLOP/s FP64 (%.1f FMA/cycle/SM at nominal clock)\n", rep, ms, flop / ms / 1e9,
             flop / 2.0 / (ms * 1e-3) / n_sm / (clk * 1e3));
    }
  if (cudaGetLastError() != cudaSuccess)
    return 1;
  return 0;
}
