// Minimal NCCL binding resolved at run time (dlopen), so that libbsnls.so has no link-time dependency on NCCL and
// uses the very library instance torch.distributed already loaded in the process (torch-bundled libnccl.so.2).
// Only what the landmark-sharded solve needs: unique id, communicator, all-reduce (sum / max of doubles).
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cstdlib>
#include <string>

namespace bsnls {

struct NcclUniqueId { char internal[128]; };   // ncclUniqueId (nccl.h: NCCL_UNIQUE_ID_BYTES = 128)
typedef struct ncclComm* NcclComm;
enum { kNcclSuccess = 0, kNcclSum = 0, kNcclMax = 2, kNcclFloat64 = 8 };

struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(NcclUniqueId*) = nullptr;
  int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
  int (*CommDestroy)(NcclComm) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  std::string error;

  bool Load() {
    if (lib) return true;
    const char* env = std::getenv("BSNLS_NCCL_LIB");
    // BSNLS_NCCL_LIB overrides; otherwise the soname: resolved to the instance already mapped into the process (the one
    // torch.distributed loaded) or through the loader's search path
    const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      if (!n) continue;
      lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (lib) break;
    }
    if (!lib) { error = std::string("cannot load NCCL: ") + dlerror(); return false; }
    auto sym = [&](const char* s) { void* p = dlsym(lib, s); if (!p) error = std::string("NCCL symbol missing: ") + s; return p; };
    GetUniqueId = reinterpret_cast<decltype(GetUniqueId)>(sym("ncclGetUniqueId"));
    CommInitRank = reinterpret_cast<decltype(CommInitRank)>(sym("ncclCommInitRank"));
    CommDestroy = reinterpret_cast<decltype(CommDestroy)>(sym("ncclCommDestroy"));
    AllReduce = reinterpret_cast<decltype(AllReduce)>(sym("ncclAllReduce"));
    GroupStart = reinterpret_cast<decltype(GroupStart)>(sym("ncclGroupStart"));
    GroupEnd = reinterpret_cast<decltype(GroupEnd)>(sym("ncclGroupEnd"));
    GetErrorString = reinterpret_cast<decltype(GetErrorString)>(sym("ncclGetErrorString"));
    if (!GetUniqueId || !CommInitRank || !CommDestroy || !AllReduce || !GroupStart || !GroupEnd || !GetErrorString) { dlclose(lib); lib = nullptr; return false; }
    return true;
  }
};

inline NcclApi& Nccl() { static NcclApi api; return api; }

}  // namespace bsnls
