// NCCL bound at run time (dlopen), so that libsmcphys.so loads on machines without NCCL and uses the copy already in
// the process when there is one (torch ships its own libnccl.so.2; a Julia host would load the system one).
// Only the stable 2.x entry points the sharded filter needs.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

#include <string>

namespace smcphys {

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
  std::string err;

  bool load() {
    if (lib) return true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
      lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
      if (lib) break;
    }
    if (!lib) {
      err = std::string("dlopen(libnccl.so.2) failed: ") + dlerror();
      return false;
    }
#define SMC_NCCL_SYM(field, name)                                   \
  field = (decltype(field))dlsym(lib, name);                        \
  if (!field) {                                                     \
    err = std::string("NCCL symbol missing: ") + name;              \
    lib = nullptr;                                                  \
    return false;                                                   \
  }
    SMC_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
    SMC_NCCL_SYM(CommInitRank, "ncclCommInitRank")
    SMC_NCCL_SYM(CommInitAll, "ncclCommInitAll")
    SMC_NCCL_SYM(CommDestroy, "ncclCommDestroy")
    SMC_NCCL_SYM(AllReduce, "ncclAllReduce")
    SMC_NCCL_SYM(AllGather, "ncclAllGather")
    SMC_NCCL_SYM(GroupStart, "ncclGroupStart")
    SMC_NCCL_SYM(GroupEnd, "ncclGroupEnd")
    SMC_NCCL_SYM(GetErrorString, "ncclGetErrorString")
    SMC_NCCL_SYM(GetVersion, "ncclGetVersion")
#undef SMC_NCCL_SYM
    return true;
  }
};

inline NcclApi& nccl() {
  static NcclApi api;
  return api;
}

}  // namespace smcphys
