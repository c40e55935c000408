// NCCL plumbing for the one collective of the step: an in-place fp32 sum all-reduce over the
// flat [gradients | loss | counts] buffer.  Replaces DDP's bucket all-reduce (reference
// train.py:46-52) and BaseTrainer.reduce_loss (base/base_trainer.py:165-174).  libnccl.so.2 is
// resolved at run time with dlopen (torch has it loaded already), so the library itself loads
// on machines without NCCL or a GPU.
#include "common.cuh"
#include <dlfcn.h>
#include <nccl.h>

namespace pdt {

struct nccl_api {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    // optional (NCCL >= 2.19): user-buffer registration
    ncclResult_t (*CommRegister)(const ncclComm_t, void*, size_t, void**) = nullptr;
    ncclResult_t (*CommDeregister)(const ncclComm_t, void*) = nullptr;
    ncclResult_t (*MemAlloc)(void**, size_t) = nullptr;
    ncclResult_t (*MemFree)(void*) = nullptr;
};

static nccl_api* api() {
    static nccl_api a;
    static bool tried = false;
    if (!tried) {
        tried = true;
        a.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (a.lib) {
            a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(a.lib, "ncclGetUniqueId");
            a.CommInitRank = (decltype(a.CommInitRank))dlsym(a.lib, "ncclCommInitRank");
            a.CommDestroy = (decltype(a.CommDestroy))dlsym(a.lib, "ncclCommDestroy");
            a.AllReduce = (decltype(a.AllReduce))dlsym(a.lib, "ncclAllReduce");
            a.GetErrorString = (decltype(a.GetErrorString))dlsym(a.lib, "ncclGetErrorString");
            a.CommRegister = (decltype(a.CommRegister))dlsym(a.lib, "ncclCommRegister");
            a.CommDeregister = (decltype(a.CommDeregister))dlsym(a.lib, "ncclCommDeregister");
            a.MemAlloc = (decltype(a.MemAlloc))dlsym(a.lib, "ncclMemAlloc");
            a.MemFree = (decltype(a.MemFree))dlsym(a.lib, "ncclMemFree");
        }
    }
    if (!a.lib || !a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllReduce || !a.GetErrorString) {
        set_error("libnccl.so.2 is not available: %s", a.lib ? "missing symbol" : dlerror());
        return nullptr;
    }
    return &a;
}

static int nccl_fail(nccl_api* a, const char* what, ncclResult_t r) {
    set_error("%s: %s", what, a->GetErrorString(r));
    return PDT_ENCCL;
}

}  // namespace pdt

using namespace pdt;

extern "C" int pdt_comm_unique_id(void* id128) {
    PDT_REQUIRE(id128, "null pointer");
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    nccl_api* a = api();
    if (!a) return PDT_ENCCL;
    const ncclResult_t r = a->GetUniqueId(reinterpret_cast<ncclUniqueId*>(id128));
    return r == ncclSuccess ? PDT_OK : nccl_fail(a, "ncclGetUniqueId", r);
}

extern "C" int pdt_comm_init(const void* id128, int rank, int world, void** comm) {
    PDT_REQUIRE(id128 && comm && world > 0 && rank >= 0 && rank < world, "bad arguments");
    nccl_api* a = api();
    if (!a) return PDT_ENCCL;
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    ncclComm_t c = nullptr;
    const ncclResult_t r = a->CommInitRank(&c, world, id, rank);
    if (r != ncclSuccess) return nccl_fail(a, "ncclCommInitRank", r);
    *comm = c;
    return PDT_OK;
}

extern "C" int pdt_comm_destroy(void* comm) {
    PDT_REQUIRE(comm, "null communicator");
    nccl_api* a = api();
    if (!a) return PDT_ENCCL;
    const ncclResult_t r = a->CommDestroy(reinterpret_cast<ncclComm_t>(comm));
    return r == ncclSuccess ? PDT_OK : nccl_fail(a, "ncclCommDestroy", r);
}

extern "C" int pdt_allreduce_flat(void* comm, float* buf, int64_t count, void* stream) {
    PDT_REQUIRE(comm && buf && count > 0, "bad arguments");
    nccl_api* a = api();
    if (!a) return PDT_ENCCL;
    const ncclResult_t r = a->AllReduce(buf, buf, (size_t)count, ncclFloat32, ncclSum,
                                        reinterpret_cast<ncclComm_t>(comm), (cudaStream_t)stream);
    return r == ncclSuccess ? PDT_OK : nccl_fail(a, "ncclAllReduce", r);
}

// User-buffer registration (the flat gradient buffer is one long-lived allocation, the case NCCL's registration is
// made for: with a buffer from ncclMemAlloc registered on the communicator, NCCL can reduce through NVLink SHARP
// (NVLS) / peer mappings straight out of and into it instead of staging through its own FIFO buffers).
extern "C" int pdt_comm_mem_alloc(int64_t bytes, void** dev_ptr) {
    PDT_REQUIRE(bytes > 0 && dev_ptr, "bad arguments");
    nccl_api* a = api();
    if (!a) return PDT_ENCCL;
    PDT_REQUIRE(a->MemAlloc != nullptr, "this NCCL has no ncclMemAlloc");
    const ncclResult_t r = a->MemAlloc(dev_ptr, (size_t)bytes);
    return r == ncclSuccess ? PDT_OK : nccl_fail(a, "ncclMemAlloc", r);
}

extern "C" int pdt_comm_mem_free(void* dev_ptr) {
    PDT_REQUIRE(dev_ptr, "null pointer");
    nccl_api* a = api();
    if (!a) return PDT_ENCCL;
    PDT_REQUIRE(a->MemFree != nullptr, "this NCCL has no ncclMemFree");
    const ncclResult_t r = a->MemFree(dev_ptr);
    return r == ncclSuccess ? PDT_OK : nccl_fail(a, "ncclMemFree", r);
}

extern "C" int pdt_comm_register(void* comm, void* buf, int64_t bytes, void** handle) {
    PDT_REQUIRE(comm && buf && bytes > 0 && handle, "bad arguments");
    nccl_api* a = api();
    if (!a) return PDT_ENCCL;
    PDT_REQUIRE(a->CommRegister != nullptr, "this NCCL has no ncclCommRegister");
    const ncclResult_t r = a->CommRegister(reinterpret_cast<ncclComm_t>(comm), buf, (size_t)bytes, handle);
    return r == ncclSuccess ? PDT_OK : nccl_fail(a, "ncclCommRegister", r);
}

extern "C" int pdt_comm_deregister(void* comm, void* handle) {
    PDT_REQUIRE(comm && handle, "bad arguments");
    nccl_api* a = api();
    if (!a) return PDT_ENCCL;
    PDT_REQUIRE(a->CommDeregister != nullptr, "this NCCL has no ncclCommDeregister");
    const ncclResult_t r = a->CommDeregister(reinterpret_cast<ncclComm_t>(comm), handle);
    return r == ncclSuccess ? PDT_OK : nccl_fail(a, "ncclCommDeregister", r);
}
