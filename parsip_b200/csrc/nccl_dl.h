// NCCL bound at run time: libpsgpu.so has no link-time dependency on it (single-GPU users never load it), and a process that
// already holds an NCCL (a torch.distributed application carries its own libnccl.so.2) keeps exactly one copy - the loader
// hands back the library that is already mapped under that SONAME. Only the entry points the z-slab exchange needs.
#ifndef PSGPU_NCCL_DL_H
#define PSGPU_NCCL_DL_H
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stddef.h>

namespace psgpu {

struct NcclApi {
    // ABI of nccl.h (2.x): opaque communicator, 128-byte unique id, int result (0 = success), enum data types
    typedef void* comm_t;
    struct unique_id { char internal[128]; };
    enum { kUint8 = 1, kUint32 = 3, kUint64 = 5, kFloat32 = 7 };
    int (*GetUniqueId)(unique_id*) = nullptr;
    int (*CommInitRank)(comm_t*, int, unique_id, int) = nullptr;
    int (*CommDestroy)(comm_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, comm_t, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, comm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, comm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
    void* handle = nullptr;

    bool load() {
        if (handle) return true;
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL | RTLD_NOLOAD);  // the copy the process already has, if any
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return false;
#define PS_SYM(field, name) *(void**)(&field) = dlsym(h, name); if (!field) return false
        PS_SYM(GetUniqueId, "ncclGetUniqueId");
        PS_SYM(CommInitRank, "ncclCommInitRank");
        PS_SYM(CommDestroy, "ncclCommDestroy");
        PS_SYM(AllGather, "ncclAllGather");
        PS_SYM(Send, "ncclSend");
        PS_SYM(Recv, "ncclRecv");
        PS_SYM(GroupStart, "ncclGroupStart");
        PS_SYM(GroupEnd, "ncclGroupEnd");
        PS_SYM(GetErrorString, "ncclGetErrorString");
        PS_SYM(GetVersion, "ncclGetVersion");
#undef PS_SYM
        handle = h;
        return true;
    }
};

inline NcclApi& nccl_api() { static NcclApi a; return a; }

}  // namespace psgpu
#endif
