// NCCL is bound at run time (dlopen) and only when a handle spans more than one rank: a single-GPU process never
// loads it, and a process that already holds another NCCL build (e.g. the one bundled with PyTorch) reuses that one
// instead of clashing on the libnccl.so.2 soname.
#pragma once
#include <dlfcn.h>
#include <nccl.h>  // types and enums only

#include "common.cuh"

namespace ufe {

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};

// nullptr + set_error() when libnccl cannot be loaded
NcclApi* nccl_api();

}  // namespace ufe
