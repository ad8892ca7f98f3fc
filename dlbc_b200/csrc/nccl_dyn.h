// nccl_dyn.h -- NCCL entry points resolved at run time (dlopen), so that libdlbc_cuda.so loads on a
// machine without NCCL and single-GPU contexts never touch it.  Only the point-to-point and
// all-reduce calls that replace the reference's MPI_Sendrecv halo exchange
// (src/dlbc/fields/parallel.d:70-103) and MPI_Allreduce diagnostics (lb/density.d:178) are bound.
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>  // types and enums only; no link-time dependency

namespace dlbc {

struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;

  // Returns nullptr on success, else a static description of what is missing.
  const char *load() {
    if (handle) return nullptr;
    // RTLD_NOLOAD first: reuse the copy the host process already has (e.g. the one torch bundles)
    handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!handle) handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!handle) handle = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!handle) return "libnccl.so.2 not found (needed for nranks > 1)";
#define DLBC_SYM(field, name)                                         \
  field = reinterpret_cast<decltype(field)>(dlsym(handle, name));     \
  if (!field) return "libnccl is missing " name;
    DLBC_SYM(GetUniqueId, "ncclGetUniqueId")
    DLBC_SYM(CommInitRank, "ncclCommInitRank")
    DLBC_SYM(CommDestroy, "ncclCommDestroy")
    DLBC_SYM(Send, "ncclSend")
    DLBC_SYM(Recv, "ncclRecv")
    DLBC_SYM(AllReduce, "ncclAllReduce")
    DLBC_SYM(GroupStart, "ncclGroupStart")
    DLBC_SYM(GroupEnd, "ncclGroupEnd")
    DLBC_SYM(GetErrorString, "ncclGetErrorString")
#undef DLBC_SYM
    return nullptr;
  }
};

inline NcclApi &nccl_api() {
  static NcclApi api;
  return api;
}

}  // namespace dlbc
