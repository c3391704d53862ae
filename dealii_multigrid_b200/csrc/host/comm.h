// Rank-to-rank communication used by the distributed vector (update_ghost_values / compress(add), dot
// products): the MPI calls of deal.II's LinearAlgebra::distributed::Vector become NCCL
// point-to-point groups and all-reduces on the compute stream.  One process per GPU.
#pragma once
#include <cstddef>
#include <cuda_runtime.h>
#include <vector>

namespace mgb
{
  struct Msg
  {
    int    peer;
    void  *ptr; // device pointer
    size_t bytes;
  };
  class Comm
  {
  public:
    virtual ~Comm() = default;
    virtual int  rank() const = 0;
    virtual int  size() const = 0;
    // all sends/recvs of one halo exchange, enqueued on `s` (ncclGroupStart/End)
    virtual void exchange(const std::vector<Msg> &sends, const std::vector<Msg> &recvs, cudaStream_t s) = 0;
    virtual void allreduce_sum(double *dev, int n, cudaStream_t s)                                       = 0;
    virtual void allreduce_max(double *dev, int n, cudaStream_t s)                                       = 0;
    // host-to-host all-gather of `bytes` per rank (setup only; synchronises the stream)
    virtual void allgather_host(const void *send, void *recv, size_t bytes, cudaStream_t s) = 0;
  };
  // NCCL-backed communicator; `unique_id` is the 128-byte ncclUniqueId broadcast by the launcher
  // (torch.distributed in bench.py / tests).  libnccl.so.2 is resolved with dlopen at run time.
  Comm *create_nccl_comm(const void *unique_id, int rank, int size, const char *libnccl_path);
  void  nccl_get_unique_id(void *out128, const char *libnccl_path);
} // namespace mgb
