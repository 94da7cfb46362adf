// NCCL, loaded at run time.  libdimreduce4gpu.so has no link-time NCCL dependency (single-GPU
// users and the reference's `ldd -r` check, ci/verify_native_so.sh, never need it); the row-sharded
// fit dlopen()s libnccl.so.2 on first use.  In a torch process that is the NCCL torch already
// loaded.  The reference has no collective at all.
#include <dlfcn.h>

#include <mutex>

#include "common.cuh"
#include "nccl_dyn.h"

namespace dr4g {

NcclApi* nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  static std::string err;
  std::call_once(once, [] {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* nm : names) {
      h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
      if (h) break;
    }
    if (!h) { err = std::string("cannot dlopen libnccl.so.2: ") + dlerror(); return; }
#define LOAD(field, sym)                                                      \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(h, sym));           \
  if (!api.field) { err = std::string("NCCL symbol missing: ") + sym; return; }
    LOAD(GetUniqueId, "ncclGetUniqueId");
    LOAD(CommInitRank, "ncclCommInitRank");
    LOAD(CommDestroy, "ncclCommDestroy");
    LOAD(AllReduce, "ncclAllReduce");
    LOAD(AllGather, "ncclAllGather");
    LOAD(GetErrorString, "ncclGetErrorString");
#undef LOAD
    api.ok = true;
  });
  if (!api.ok) fail("%s", err.c_str());
  return &api;
}

}  // namespace dr4g
