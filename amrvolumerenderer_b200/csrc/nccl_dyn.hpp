// nccl_dyn.hpp -- NCCL bound at run time (dlopen), so libamrvr_b200.so loads on a
// machine without NCCL and single-rank contexts never touch it.  When the host
// process already carries an NCCL (PyTorch's bundled libnccl.so.2) that copy is
// reused; otherwise the system library is opened.  AMRVR_NCCL_LIB overrides.
#pragma once

#include <dlfcn.h>
#include <nccl.h>

#include <cstdlib>
#include <string>

namespace amrvr {

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;

  // Returns an empty string on success, else why NCCL could not be bound.
  std::string load() {
    if (handle != nullptr) return {};
    const char* override_path = std::getenv("AMRVR_NCCL_LIB");
    const char* candidates[] = {override_path, "libnccl.so.2", "libnccl.so"};
    for (const char* name : candidates) {
      if (name == nullptr || *name == '\0') continue;
      handle = dlopen(name, RTLD_NOW | RTLD_NOLOAD);  // already in the process?
      if (handle == nullptr) handle = dlopen(name, RTLD_NOW | RTLD_LOCAL);
      if (handle != nullptr) break;
    }
    if (handle == nullptr) return std::string("cannot load NCCL: ") + dlerror();
    auto bind = [&](auto& fn, const char* symbol) {
      fn = reinterpret_cast<std::remove_reference_t<decltype(fn)>>(dlsym(handle, symbol));
      return fn != nullptr;
    };
    const bool ok = bind(GetUniqueId, "ncclGetUniqueId") && bind(CommInitRank, "ncclCommInitRank") &&
                    bind(CommDestroy, "ncclCommDestroy") && bind(Send, "ncclSend") &&
                    bind(Recv, "ncclRecv") && bind(AllGather, "ncclAllGather") &&
                    bind(AllReduce, "ncclAllReduce") &&
                    bind(GroupStart, "ncclGroupStart") && bind(GroupEnd, "ncclGroupEnd") &&
                    bind(GetErrorString, "ncclGetErrorString");
    if (!ok) {
      dlclose(handle);
      handle = nullptr;
      return "NCCL library lacks a required symbol";
    }
    return {};
  }
};

inline NcclApi& nccl_api() {
  static NcclApi api;
  return api;
}

}  // namespace amrvr
