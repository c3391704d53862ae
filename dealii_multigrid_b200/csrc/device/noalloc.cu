// mgb_vector_dot_no_alloc: the dot product of the library's BLAS-1 layer (device/vector.cu k_dot, reached through
// mgb_vector_dot) for callers that may not allocate -- code under CUDA graph capture, tight serving loops, harnesses that
// forbid cudaMalloc.  It has no counterpart in the reference.
//
// The existing reductions take their scratch (one partial sum per block, a ticket for last-block detection, the result
// scalar) from buffers that Context::ensure() obtains with cudaMalloc.  This variant keeps the scratch in a file-scope
// `static __device__` array of a stated, modest size instead:
//   * every GPU that loads the library holds the array (MGB_NO_ALLOC_SCRATCH_BYTES = 33 024 bytes) until the process exits;
//   * in HOST code the array's name is a host-side stand-in, not a device address, so the device address is looked up
//     once per device, with that device current, by cudaGetSymbolAddress (return code checked) and passed to the kernel;
//   * the entry point calls no cudaMalloc / cudaFree / allocating library routine and never touches Context::ensure();
//   * ALL calls share the array and it keeps the last call's contents: the ticket word, which must start at zero, is
//     cleared on the caller's stream before every launch; the partial sums need no clearing (every one that is read has been
//     written by the same launch).  Calls that overlap on the GPU (different streams, or several host threads on one
//     device) race on the array -- serialise them.
// One block reduces a fixed span of 4096 elements, so the summation order depends on n only (bitwise reproducible), and
// the 4096 partial sums of the array bound the vector length: n <= MGB_NO_ALLOC_MAX_N = 16 777 216.
#include "../../../include/mgb200.h"
#include "../host/solver.h"

#include <mutex>
#include <string>

namespace mgb
{
  std::string &capi_last_error(); // host/capi.cc (thread-local message behind mgb_last_error)

  namespace
  {
    constexpr int      NA_BLOCK        = 256;                 // threads per block
    constexpr int      NA_ITEMS        = 16;                  // elements per thread
    constexpr int      NA_SPAN         = NA_BLOCK * NA_ITEMS; // elements per block
    constexpr int      NA_MAX_PARTIALS = 4096;                // blocks per launch at most
    constexpr int      NA_TICKET_WORDS = 32;                  // the ticket lives in a 256-byte tail of its own
    constexpr uint64_t NA_MAX_N        = uint64_t(NA_SPAN) * NA_MAX_PARTIALS;
    static_assert(NA_MAX_N == MGB_NO_ALLOC_MAX_N, "include/mgb200.h states another maximum");
    static_assert((NA_MAX_PARTIALS + NA_TICKET_WORDS) * sizeof(double) == MGB_NO_ALLOC_SCRATCH_BYTES, "include/mgb200.h states another size");

    // [0, NA_MAX_PARTIALS): one partial sum per block; word NA_MAX_PARTIALS: the ticket (as unsigned)
    static __device__ __align__(256) double g_no_alloc_scratch[NA_MAX_PARTIALS + NA_TICKET_WORDS];

    // fixed-order tree over the block's values; result valid in thread 0
    __device__ __forceinline__ double block_sum(double v, double *sh)
    {
      sh[threadIdx.x] = v;
      __syncthreads();
#pragma unroll
      for (int s = NA_BLOCK / 2; s > 0; s >>= 1)
        {
          if (threadIdx.x < s)
            sh[threadIdx.x] += sh[threadIdx.x + s];
          __syncthreads();
        }
      return sh[0];
    }

    __global__ void __launch_bounds__(NA_BLOCK)
      k_dot_no_alloc(const double *__restrict__ x, const double *__restrict__ y, const unsigned long long n, double *scratch, double *result)
    {
      __shared__ double sh[NA_BLOCK];
      __shared__ bool   is_last;
      const unsigned long long base = (unsigned long long)blockIdx.x * NA_SPAN;
      double                   v    = 0;
#pragma unroll
      for (int k = 0; k < NA_ITEMS; ++k)
        {
          const unsigned long long i = base + (unsigned long long)k * NA_BLOCK + threadIdx.x;
          if (i < n)
            v += x[i] * y[i];
        }
      const double   total  = block_sum(v, sh);
      unsigned      *ticket = reinterpret_cast<unsigned *>(scratch + NA_MAX_PARTIALS);
      if (threadIdx.x == 0)
        {
          scratch[blockIdx.x] = total;
          __threadfence();
          is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
        }
      __syncthreads();
      if (!is_last)
        return;
      __threadfence();
      // the last block to arrive adds the partial sums, again in an order that depends on the grid size only
      double w = 0;
      for (unsigned b = threadIdx.x; b < gridDim.x; b += NA_BLOCK)
        w += *reinterpret_cast<volatile double *>(scratch + b);
      const double r = block_sum(w, sh);
      if (threadIdx.x == 0)
        {
          *result = r;
          *ticket = 0u; // leave the array ready for the next call (the host clears it as well)
        }
    }

    // device address of the scratch array, looked up once per device with that device current
    constexpr int MAX_DEVICES = 64;
    double       *g_scratch_addr[MAX_DEVICES] = {};
    std::mutex    g_scratch_mutex;

    double *scratch_address()
    {
      int dev = -1;
      MGB_CUDA(cudaGetDevice(&dev));
      if (dev < 0 || dev >= MAX_DEVICES)
        throw std::runtime_error("mgb_vector_dot_no_alloc: device ordinal " + std::to_string(dev) + " out of range");
      std::lock_guard<std::mutex> lock(g_scratch_mutex);
      if (!g_scratch_addr[dev])
        {
          void             *p = nullptr;
          const cudaError_t e = cudaGetSymbolAddress(&p, g_no_alloc_scratch);
          if (e != cudaSuccess || !p)
            throw std::runtime_error(std::string("mgb_vector_dot_no_alloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
          g_scratch_addr[dev] = static_cast<double *>(p);
        }
      return g_scratch_addr[dev];
    }
  } // namespace
} // namespace mgb

extern "C" int mgb_vector_dot_no_alloc(const double *x_dev, const double *y_dev, uint64_t n, double *result_dev, void *cuda_stream)
{
  using namespace mgb;
  try
    {
      if (n > NA_MAX_N)
        throw std::runtime_error("mgb_vector_dot_no_alloc: n = " + std::to_string((unsigned long long)n) + " exceeds the maximum of " +
                                 std::to_string((unsigned long long)NA_MAX_N) +
                                 " elements that the static scratch array covers (MGB_NO_ALLOC_MAX_N); use mgb_vector_dot");
      if (!result_dev || (n > 0 && (!x_dev || !y_dev)))
        throw std::runtime_error("mgb_vector_dot_no_alloc: null pointer (x_dev, y_dev and result_dev are device pointers of the caller)");
      double            *scratch = scratch_address();
      const cudaStream_t s       = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx().stream;
      // the ticket must start at zero whatever an earlier (possibly aborted) call left behind
      MGB_CUDA(cudaMemsetAsync(scratch + NA_MAX_PARTIALS, 0, NA_TICKET_WORDS * sizeof(double), s));
      const unsigned blocks = n == 0 ? 1u : unsigned((n + NA_SPAN - 1) / NA_SPAN);
      k_dot_no_alloc<<<blocks, NA_BLOCK, 0, s>>>(x_dev, y_dev, (unsigned long long)n, scratch, result_dev);
      MGB_CUDA(cudaGetLastError());
      ++ctx().launches;
      return 0;
    }
  catch (const std::exception &e)
    {
      capi_last_error() = e.what();
      return 1;
    }
  catch (...)
    {
      capi_last_error() = "unknown exception";
      return 2;
    }
}
