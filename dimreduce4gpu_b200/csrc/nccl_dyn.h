// Function table of the NCCL entry points the row-sharded fit uses (resolved with dlsym).
#pragma once
#include <cuda_runtime.h>
#include <cstddef>

namespace dr4g {

struct NcclUniqueId { char internal[128]; };  // == ncclUniqueId (nccl.h: NCCL_UNIQUE_ID_BYTES 128)
constexpr int kNcclFloat64 = 8;               // ncclFloat64 / ncclDouble
constexpr int kNcclSum = 0;                   // ncclSum

struct NcclApi {
  bool ok = false;
  int (*GetUniqueId)(NcclUniqueId*) = nullptr;
  int (*CommInitRank)(void** comm, int nranks, NcclUniqueId id, int rank) = nullptr;
  int (*CommDestroy)(void* comm) = nullptr;
  int (*AllReduce)(const void* send, void* recv, size_t count, int dtype, int op, void* comm,
                   cudaStream_t stream) = nullptr;
  int (*AllGather)(const void* send, void* recv, size_t sendcount, int dtype, void* comm, cudaStream_t stream) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};

}  // namespace dr4g
