// runtime.cu -- device memory, streams, events, graph capture and error plumbing of the
// C ABI (include/pgk_b200.h).  Minimal B200 equivalent of the reference's native/core/
// (memory.cpp, stream.cpp, event.cpp, cuda_graph.cpp, device.cpp); no driver-API context
// juggling: one process = one GPU, CUDA runtime primary context.
#include "common.cuh"
#include "../../include/pgk_b200.h"

#include <atomic>
#include <cstdarg>
#include <cstring>
#include <mutex>

namespace pgk {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};
static std::atomic<int> g_pdl{0};
static int g_sm_count = 0;

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
    // same shape as the reference's "<fn> failed: <cudaGetErrorString>" (error.cuh:28-37)
    const char* base = strrchr(file, '/');
    set_error("%s failed: %s (%s:%d)", what, cudaGetErrorString(e), base ? base + 1 : file, line);
    cudaGetLastError();  // clear sticky-less errors
    return (int)e ? (int)e : 1;
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
bool pdl_enabled() { return g_pdl.load(std::memory_order_relaxed) != 0; }

int sm_count() {
    if (g_sm_count == 0) {
        int dev = 0, n = 0;
        if (cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
            g_sm_count = n;
        else
            g_sm_count = 148;
    }
    return g_sm_count;
}

}  // namespace pgk

using namespace pgk;

static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

extern "C" {

const char* pgk_last_error(void) { return g_err; }

int pgk_version(void) { return 100; }

int pgk_device_count(int* count) {
    PGK_CUDA(cudaGetDeviceCount(count));
    return 0;
}
int pgk_device_set(int device) {
    PGK_CUDA(cudaSetDevice(device));
    g_sm_count = 0;
    return 0;
}
int pgk_device_get(int* device) {
    PGK_CUDA(cudaGetDevice(device));
    return 0;
}
int pgk_device_props(int device, pgk_device_props_t* out) {
    cudaDeviceProp p;
    PGK_CUDA(cudaGetDeviceProperties(&p, device));
    memset(out, 0, sizeof(*out));
    strncpy(out->name, p.name, sizeof(out->name) - 1);
    out->sm_count = p.multiProcessorCount;
    out->cc_major = p.major;
    out->cc_minor = p.minor;
    out->total_mem = p.totalGlobalMem;
    out->l2_bytes = (size_t)p.l2CacheSize;
    out->max_smem_per_block = (int)p.sharedMemPerBlockOptin;
    return 0;
}
int pgk_device_sync(void) {
    PGK_CUDA(cudaDeviceSynchronize());
    return 0;
}
int pgk_mem_info(size_t* free_bytes, size_t* total_bytes) {
    PGK_CUDA(cudaMemGetInfo(free_bytes, total_bytes));
    return 0;
}

int pgk_malloc(void** ptr, size_t bytes) {
    *ptr = nullptr;
    if (bytes == 0) bytes = 1;
    PGK_CUDA(cudaMalloc(ptr, bytes));
    return 0;
}
int pgk_free(void* ptr) {
    if (ptr) {
        fp8_codes_forget(ptr);
        PGK_CUDA(cudaFree(ptr));
    }
    return 0;
}
int pgk_host_alloc(void** host_ptr, size_t bytes) {
    PGK_CUDA(cudaHostAlloc(host_ptr, bytes ? bytes : 1, cudaHostAllocDefault));
    return 0;
}
int pgk_host_free(void* host_ptr) {
    if (host_ptr) PGK_CUDA(cudaFreeHost(host_ptr));
    return 0;
}
int pgk_memset(void* ptr, int value, size_t bytes, void* stream) {
    PGK_CUDA(cudaMemsetAsync(ptr, value, bytes, S(stream)));
    return 0;
}
int pgk_memcpy_h2d(void* dst, const void* host_src, size_t bytes, void* stream) {
    PGK_CUDA(cudaMemcpyAsync(dst, host_src, bytes, cudaMemcpyHostToDevice, S(stream)));
    return 0;
}
int pgk_memcpy_d2h(void* host_dst, const void* src, size_t bytes, void* stream) {
    PGK_CUDA(cudaMemcpyAsync(host_dst, src, bytes, cudaMemcpyDeviceToHost, S(stream)));
    return 0;
}
int pgk_memcpy_d2d(void* dst, const void* src, size_t bytes, void* stream) {
    PGK_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, S(stream)));
    return 0;
}

int pgk_stream_create(void** stream) {
    cudaStream_t s;
    PGK_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    *stream = s;
    return 0;
}
int pgk_stream_destroy(void* stream) {
    PGK_CUDA(cudaStreamDestroy(S(stream)));
    return 0;
}
int pgk_stream_sync(void* stream) {
    PGK_CUDA(cudaStreamSynchronize(S(stream)));
    return 0;
}
int pgk_event_create(void** event) {
    cudaEvent_t e;
    PGK_CUDA(cudaEventCreate(&e));
    *event = e;
    return 0;
}
int pgk_event_destroy(void* event) {
    PGK_CUDA(cudaEventDestroy((cudaEvent_t)event));
    return 0;
}
int pgk_event_record(void* event, void* stream) {
    PGK_CUDA(cudaEventRecord((cudaEvent_t)event, S(stream)));
    return 0;
}
int pgk_event_sync(void* event) {
    PGK_CUDA(cudaEventSynchronize((cudaEvent_t)event));
    return 0;
}
int pgk_event_elapsed_ms(void* start, void* stop, float* ms) {
    PGK_CUDA(cudaEventElapsedTime(ms, (cudaEvent_t)start, (cudaEvent_t)stop));
    return 0;
}

int pgk_graph_begin(void* stream) {
    PGK_REQUIRE(stream != nullptr, "pgk_graph_begin failed: capture needs a non-default stream");
    PGK_CUDA(cudaStreamBeginCapture(S(stream), cudaStreamCaptureModeThreadLocal));
    return 0;
}
int pgk_graph_end(void* stream, void** graph_exec) {
    cudaGraph_t g = nullptr;
    PGK_CUDA(cudaStreamEndCapture(S(stream), &g));
    cudaGraphExec_t ge = nullptr;
    cudaError_t e = cudaGraphInstantiate(&ge, g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGraphInstantiate", __FILE__, __LINE__);
    *graph_exec = ge;
    return 0;
}
int pgk_graph_launch(void* graph_exec, void* stream) {
    PGK_CUDA(cudaGraphLaunch((cudaGraphExec_t)graph_exec, S(stream)));
    return 0;
}
int pgk_graph_destroy(void* graph_exec) {
    if (graph_exec) PGK_CUDA(cudaGraphExecDestroy((cudaGraphExec_t)graph_exec));
    return 0;
}

int pgk_set_pdl(int enabled) {
    g_pdl.store(enabled ? 1 : 0);
    return 0;
}
int64_t pgk_launch_count(void) { return g_launches.load(); }
void pgk_launch_count_reset(void) { g_launches.store(0); }

static void* g_flush_buf = nullptr;
static const size_t kFlushBytes = 256ull << 20;  // 2x the 126 MB L2
int pgk_flush_l2(void* stream) {
    if (!g_flush_buf) PGK_CUDA(cudaMalloc(&g_flush_buf, kFlushBytes));
    static int v = 0;
    PGK_CUDA(cudaMemsetAsync(g_flush_buf, (v++) & 0xff, kFlushBytes, S(stream)));
    return 0;
}

}  // extern "C"
