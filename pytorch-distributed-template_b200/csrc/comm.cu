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
