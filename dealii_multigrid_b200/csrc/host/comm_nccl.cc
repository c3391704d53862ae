// NCCL communicator (one process per GPU; NVLink 5 / NVSwitch underneath).  libnccl.so.2 is resolved at
// run time with dlopen so the library itself has no link-time dependency on the torch-bundled NCCL.
#include "comm.h"

#include <cstring>
#include <dlfcn.h>
#include <stdexcept>
#include <string>

namespace mgb
{
  namespace
  {
    struct ncclUniqueId
    {
      char internal[128];
    };
    using ncclComm_t = void *;
    constexpr int ncclInt8 = 0, ncclFloat64 = 8, ncclSum = 0, ncclMax = 2;

    struct NcclApi
    {
      void *lib = nullptr;
      int (*GetUniqueId)(ncclUniqueId *)                                                               = nullptr;
      int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int)                                        = nullptr;
      int (*CommDestroy)(ncclComm_t)                                                                   = nullptr;
      int (*GroupStart)()                                                                              = nullptr;
      int (*GroupEnd)()                                                                                = nullptr;
      int (*Send)(const void *, size_t, int, int, ncclComm_t, cudaStream_t)                            = nullptr;
      int (*Recv)(void *, size_t, int, int, ncclComm_t, cudaStream_t)                                  = nullptr;
      int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t)               = nullptr;
      int (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t)                    = nullptr;
      const char *(*GetErrorString)(int)                                                               = nullptr;
    };

    NcclApi &api(const char *path)
    {
      static NcclApi a;
      if (a.lib)
        return a;
      const char *candidates[] = {path, "libnccl.so.2", "libnccl.so"};
      for (const char *c : candidates)
        {
          if (!c || !*c)
            continue;
          a.lib = dlopen(c, RTLD_NOW | RTLD_GLOBAL);
          if (a.lib)
            break;
        }
      if (!a.lib)
        throw std::runtime_error(std::string("cannot dlopen libnccl: ") + dlerror());
      const auto sym = [&](const char *name) {
        void *p = dlsym(a.lib, name);
        if (!p)
          throw std::runtime_error(std::string("libnccl: missing symbol ") + name);
        return p;
      };
      a.GetUniqueId    = reinterpret_cast<decltype(a.GetUniqueId)>(sym("ncclGetUniqueId"));
      a.CommInitRank   = reinterpret_cast<decltype(a.CommInitRank)>(sym("ncclCommInitRank"));
      a.CommDestroy    = reinterpret_cast<decltype(a.CommDestroy)>(sym("ncclCommDestroy"));
      a.GroupStart     = reinterpret_cast<decltype(a.GroupStart)>(sym("ncclGroupStart"));
      a.GroupEnd       = reinterpret_cast<decltype(a.GroupEnd)>(sym("ncclGroupEnd"));
      a.Send           = reinterpret_cast<decltype(a.Send)>(sym("ncclSend"));
      a.Recv           = reinterpret_cast<decltype(a.Recv)>(sym("ncclRecv"));
      a.AllReduce      = reinterpret_cast<decltype(a.AllReduce)>(sym("ncclAllReduce"));
      a.AllGather      = reinterpret_cast<decltype(a.AllGather)>(sym("ncclAllGather"));
      a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(sym("ncclGetErrorString"));
      return a;
    }

    void check(NcclApi &a, int r, const char *what)
    {
      if (r != 0)
        throw std::runtime_error(std::string("NCCL error in ") + what + ": " + (a.GetErrorString ? a.GetErrorString(r) : "?"));
    }

    class NcclComm : public Comm
    {
    public:
      NcclComm(NcclApi &a, const void *id, int rank, int size)
        : a_(a)
        , rank_(rank)
        , size_(size)
      {
        ncclUniqueId uid;
        std::memcpy(&uid, id, sizeof(uid));
        check(a_, a_.CommInitRank(&comm_, size, uid, rank), "ncclCommInitRank");
      }
      ~NcclComm() override
      {
        if (comm_)
          a_.CommDestroy(comm_);
      }
      int  rank() const override { return rank_; }
      int  size() const override { return size_; }
      void exchange(const std::vector<Msg> &sends, const std::vector<Msg> &recvs, cudaStream_t s) override
      {
        if (sends.empty() && recvs.empty())
          return;
        check(a_, a_.GroupStart(), "ncclGroupStart");
        for (const Msg &m : recvs)
          check(a_, a_.Recv(m.ptr, m.bytes, ncclInt8, m.peer, comm_, s), "ncclRecv");
        for (const Msg &m : sends)
          check(a_, a_.Send(m.ptr, m.bytes, ncclInt8, m.peer, comm_, s), "ncclSend");
        check(a_, a_.GroupEnd(), "ncclGroupEnd");
      }
      void allreduce_sum(double *dev, int n, cudaStream_t s) override
      {
        check(a_, a_.AllReduce(dev, dev, (size_t)n, ncclFloat64, ncclSum, comm_, s), "ncclAllReduce");
      }
      void allreduce_max(double *dev, int n, cudaStream_t s) override
      {
        check(a_, a_.AllReduce(dev, dev, (size_t)n, ncclFloat64, ncclMax, comm_, s), "ncclAllReduce");
      }

      void allgather_host(const void *send, void *recv, size_t bytes, cudaStream_t s) override
      {
        char *dev = nullptr;
        if (cudaMalloc((void **)&dev, bytes * size_t(size_ + 1)) != cudaSuccess)
          throw std::runtime_error("allgather_host: cudaMalloc failed");
        cudaMemcpyAsync(dev, send, bytes, cudaMemcpyHostToDevice, s);
        check(a_, a_.AllGather(dev, dev + bytes, bytes, ncclInt8, comm_, s), "ncclAllGather");
        cudaMemcpyAsync(recv, dev + bytes, bytes * size_t(size_), cudaMemcpyDeviceToHost, s);
        cudaStreamSynchronize(s);
        cudaFree(dev);
      }

    private:
      NcclApi   &a_;
      ncclComm_t comm_ = nullptr;
      int        rank_, size_;
    };
  } // namespace

  Comm *create_nccl_comm(const void *unique_id, int rank, int size, const char *libnccl_path)
  {
    return new NcclComm(api(libnccl_path), unique_id, rank, size);
  }
  void nccl_get_unique_id(void *out128, const char *libnccl_path)
  {
    NcclApi     &a = api(libnccl_path);
    ncclUniqueId id;
    check(a, a.GetUniqueId(&id), "ncclGetUniqueId");
    std::memcpy(out128, &id, sizeof(id));
  }
} // namespace mgb
