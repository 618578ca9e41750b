// comm.cu -- tensor-parallel collectives for the decode path (net-new: the reference is
// single-GPU, docs/scheduler.md:358).  One process per GPU, one NCCL communicator per
// process, collectives enqueued on the compute stream (graph-capturable).  NCCL is bound at
// run time with dlopen so the library has no link-time dependency on it (torch bundles its
// own copy; whichever libnccl.so.2 is already in the process is reused).
#include "common.cuh"
#include "../../include/pgk_b200.h"

#include <dlfcn.h>
#include <cstring>

namespace pgk {

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclInt8 = 0, ncclBfloat16 = 9, ncclSum = 0 };

static struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
} g_nccl;
static ncclComm_t g_comm = nullptr;
static int g_world = 1, g_rank = 0;

static int load_nccl() {
    if (g_nccl.lib) return 0;
    const char* names[] = {"libnccl.so.2", "libnccl.so", nullptr};
    for (int i = 0; names[i] && !g_nccl.lib; ++i) g_nccl.lib = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
    PGK_REQUIRE(g_nccl.lib != nullptr, "nccl_init failed: cannot dlopen libnccl.so.2 (%s)", dlerror());
#define PGK_SYM(field, name)                                                       \
    *(void**)(&g_nccl.field) = dlsym(g_nccl.lib, name);                            \
    PGK_REQUIRE(g_nccl.field != nullptr, "nccl_init failed: symbol %s missing", name)
    PGK_SYM(GetUniqueId, "ncclGetUniqueId");
    PGK_SYM(CommInitRank, "ncclCommInitRank");
    PGK_SYM(AllReduce, "ncclAllReduce");
    PGK_SYM(AllGather, "ncclAllGather");
    PGK_SYM(CommDestroy, "ncclCommDestroy");
    PGK_SYM(GetErrorString, "ncclGetErrorString");
#undef PGK_SYM
    return 0;
}

#define PGK_NCCL(call, what)                                                       \
    do {                                                                           \
        ncclResult_t _r = (call);                                                  \
        if (_r != 0) { set_error("%s failed: %s", what, g_nccl.GetErrorString(_r)); return 1; } \
    } while (0)

}  // namespace pgk

using namespace pgk;

extern "C" {

int pgk_nccl_unique_id(void* host_id128) {
    if (int rc = load_nccl()) return rc;
    ncclUniqueId id;
    PGK_NCCL(g_nccl.GetUniqueId(&id), "ncclGetUniqueId");
    memcpy(host_id128, &id, sizeof(id));
    return 0;
}
int pgk_nccl_init(const void* host_id128, int rank, int world) {
    if (int rc = load_nccl()) return rc;
    PGK_REQUIRE(g_comm == nullptr, "nccl_init failed: communicator already initialised");
    ncclUniqueId id;
    memcpy(&id, host_id128, sizeof(id));
    PGK_NCCL(g_nccl.CommInitRank(&g_comm, world, id, rank), "ncclCommInitRank");
    g_world = world;
    g_rank = rank;
    return 0;
}
int pgk_nccl_allreduce_bf16(void* buf, size_t count, void* stream) {
    PGK_REQUIRE(g_comm != nullptr, "nccl_allreduce failed: communicator not initialised");
    count_launch();
    PGK_NCCL(g_nccl.AllReduce(buf, buf, count, ncclBfloat16, ncclSum, g_comm, (cudaStream_t)stream), "ncclAllReduce");
    return 0;
}
int pgk_nccl_allgather(const void* send, void* recv, size_t bytes_per_rank, void* stream) {
    PGK_REQUIRE(g_comm != nullptr, "nccl_allgather failed: communicator not initialised");
    count_launch();
    PGK_NCCL(g_nccl.AllGather(send, recv, bytes_per_rank, ncclInt8, g_comm, (cudaStream_t)stream), "ncclAllGather");
    return 0;
}
int pgk_nccl_destroy(void) {
    if (g_comm) { g_nccl.CommDestroy(g_comm); g_comm = nullptr; }
    g_world = 1; g_rank = 0;
    return 0;
}

}  // extern "C"
