// NCCL communicator owned by the library (one per process), over NVLink 5 / NVSwitch.
//
// The reference reaches NCCL through c10d's ProcessGroupNCCL: the DDP reducer all-reduces
// each flat gradient bucket during backward (demo.py:107,109) and DDP's constructor
// broadcasts rank 0's parameters (demo.py:70,72; torch:nn/parallel/distributed.py:879-890).
// Here the collectives are issued from the C ABI on caller-provided streams, so the host
// layer can order them against its own kernels with events and no Python in between.
//
// NCCL is resolved at run time with dlopen/dlsym from the copy that PyTorch already
// loaded (site-packages/nvidia/nccl/lib/libnccl.so.2): two NCCL builds in one process
// would each own separate topology state, and the system header (2.27.3) does not match
// the bundled library (2.28.9).  Only the ABI-stable core entry points are used.
#include "common.cuh"
#include <dlfcn.h>
#include <cstring>

namespace {

// Minimal mirror of the public NCCL ABI (nccl.h): opaque comm, 128-byte id, enums.
struct NcclId { char bytes[128]; };
typedef void *NcclComm;
enum { kNcclFloat32 = 7, kNcclSum = 0, kNcclSuccess = 0 };

struct NcclApi {
    void *handle = nullptr;
    int (*GetUniqueId)(NcclId *) = nullptr;
    int (*CommInitRank)(NcclComm *, int, NcclId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*Broadcast)(const void *, void *, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int *) = nullptr;
};

NcclApi g_nccl;

template <typename F>
bool bind(F &slot, const char *name) {
    slot = reinterpret_cast<F>(dlsym(g_nccl.handle, name));
    return slot != nullptr;
}

int nccl_fail(int rc, const char *what) {
    return ddpb::fail(2000 + rc, "%s failed: %s", what, g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?");
}

int ensure_loaded() {
    if (g_nccl.handle) return 0;
    return ddpb_nccl_load(nullptr);
}

}  // namespace

#define DDPB_NCCL(call, what)                              \
    do {                                                   \
        int rc__ = (call);                                 \
        if (rc__ != kNcclSuccess) return nccl_fail(rc__, what); \
    } while (0)

DDPB_API int ddpb_nccl_load(const char *path_or_null) {
    if (g_nccl.handle) return 0;
    void *h = nullptr;
    if (path_or_null && *path_or_null) h = dlopen(path_or_null, RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);   // the copy torch loaded
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return ddpb::fail(3, "ddpb_nccl_load: cannot open libnccl.so.2: %s", dlerror());
    g_nccl.handle = h;
    bool ok = bind(g_nccl.GetUniqueId, "ncclGetUniqueId") && bind(g_nccl.CommInitRank, "ncclCommInitRank") &&
              bind(g_nccl.CommDestroy, "ncclCommDestroy") && bind(g_nccl.AllReduce, "ncclAllReduce") &&
              bind(g_nccl.Broadcast, "ncclBroadcast") && bind(g_nccl.GetErrorString, "ncclGetErrorString") &&
              bind(g_nccl.GetVersion, "ncclGetVersion");
    if (!ok) {
        g_nccl = NcclApi();
        return ddpb::fail(3, "ddpb_nccl_load: libnccl.so.2 lacks a required symbol");
    }
    return 0;
}

DDPB_API int ddpb_get_unique_id(char out[128]) {
    DDPB_REQUIRE(out != nullptr, "ddpb_get_unique_id: null output");
    if (int rc = ensure_loaded()) return rc;
    NcclId id;
    DDPB_NCCL(g_nccl.GetUniqueId(&id), "ncclGetUniqueId");
    std::memcpy(out, id.bytes, sizeof(id.bytes));
    return 0;
}

DDPB_API int ddpb_comm_create(const char id_bytes[128], int rank, int world, int device, void **comm_out) {
    DDPB_REQUIRE(id_bytes && comm_out && world >= 1 && rank >= 0 && rank < world, "ddpb_comm_create: bad arguments");
    if (int rc = ensure_loaded()) return rc;
    DDPB_CUDA(cudaSetDevice(device));
    NcclId id;
    std::memcpy(id.bytes, id_bytes, sizeof(id.bytes));
    NcclComm comm = nullptr;
    DDPB_NCCL(g_nccl.CommInitRank(&comm, world, id, rank), "ncclCommInitRank");
    *comm_out = comm;
    return 0;
}

DDPB_API int ddpb_comm_destroy(void *comm) {
    if (!comm) return 0;
    if (int rc = ensure_loaded()) return rc;
    DDPB_NCCL(g_nccl.CommDestroy((NcclComm)comm), "ncclCommDestroy");
    return 0;
}

DDPB_API int ddpb_broadcast_f32(void *comm, float *buf, int64_t n, int root, void *stream) {
    DDPB_REQUIRE(comm && buf && n >= 0, "ddpb_broadcast_f32: bad arguments");
    if (n == 0) return 0;
    DDPB_NCCL(g_nccl.Broadcast(buf, buf, (size_t)n, kNcclFloat32, root, (NcclComm)comm, (cudaStream_t)stream),
              "ncclBroadcast");
    return 0;
}

DDPB_API int ddpb_allreduce_sum_f32(void *comm, float *buf, int64_t n, void *stream) {
    DDPB_REQUIRE(comm && buf && n >= 0, "ddpb_allreduce_sum_f32: bad arguments");
    if (n == 0) return 0;
    DDPB_NCCL(g_nccl.AllReduce(buf, buf, (size_t)n, kNcclFloat32, kNcclSum, (NcclComm)comm, (cudaStream_t)stream),
              "ncclAllReduce");
    return 0;
}
