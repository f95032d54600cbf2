// api.cu -- device plumbing of the C ABI (include/b200_ggml.h) and the thread-local error string.
#include "b200_common.cuh"

#include <cstdarg>
#include <cstdio>
#include <cstring>

namespace b200 {

extern bool g_mmvq_pdl;
static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what) {
    set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
    return B200_ERR_CUDA;
}

}  // namespace b200

using namespace b200;

extern "C" {

const char *b200_last_error(void) { return g_err; }
const char *b200_version(void) { return "koboldcpp_b200 0.1 (sm_100a)"; }
void b200_set_pdl(int enabled) { g_mmvq_pdl = enabled != 0; }

int b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}
int b200_set_device(int device) { B200_CUDA_TRY(cudaSetDevice(device)); return B200_OK; }

int b200_device_info(int device, char *name, size_t name_len, size_t *free_bytes, size_t *total_bytes, int *sm_count, int *cc_major,
                     int *cc_minor) {
    cudaDeviceProp prop;
    B200_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (name && name_len) { strncpy(name, prop.name, name_len - 1); name[name_len - 1] = 0; }
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (cc_major) *cc_major = prop.major;
    if (cc_minor) *cc_minor = prop.minor;
    if (free_bytes || total_bytes) {
        int cur = 0;
        B200_CUDA_TRY(cudaGetDevice(&cur));
        B200_CUDA_TRY(cudaSetDevice(device));
        size_t f = 0, t = 0;
        cudaError_t e = cudaMemGetInfo(&f, &t);
        cudaSetDevice(cur);
        if (e != cudaSuccess) return cuda_fail(e, "cudaMemGetInfo");
        if (free_bytes) *free_bytes = f;
        if (total_bytes) *total_bytes = t;
    }
    return B200_OK;
}

int b200_malloc(void **ptr, size_t bytes) {
    if (!ptr) { set_error("b200_malloc: NULL out pointer"); return B200_ERR_INVALID; }
    cudaError_t e = cudaMalloc(ptr, bytes ? bytes : 1);
    if (e == cudaErrorMemoryAllocation) { cudaGetLastError(); *ptr = nullptr; set_error("b200_malloc: out of device memory (%zu bytes)", bytes); return B200_ERR_NOMEM; }
    if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    return B200_OK;
}
int b200_free(void *ptr) { B200_CUDA_TRY(cudaFree(ptr)); return B200_OK; }
int b200_malloc_host(void **ptr, size_t bytes) {
    if (!ptr) { set_error("b200_malloc_host: NULL out pointer"); return B200_ERR_INVALID; }
    B200_CUDA_TRY(cudaMallocHost(ptr, bytes ? bytes : 1));
    return B200_OK;
}
int b200_free_host(void *ptr) { B200_CUDA_TRY(cudaFreeHost(ptr)); return B200_OK; }
int b200_memcpy_h2d(void *dst, const void *src, size_t bytes, b200_stream_t stream) {
    B200_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return B200_OK;
}
int b200_memcpy_d2h(void *dst, const void *src, size_t bytes, b200_stream_t stream) {
    B200_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    return B200_OK;
}
int b200_memcpy_d2d(void *dst, const void *src, size_t bytes, b200_stream_t stream) {
    B200_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return B200_OK;
}
int b200_memset(void *dst, int value, size_t bytes, b200_stream_t stream) {
    B200_CUDA_TRY(cudaMemsetAsync(dst, value, bytes, (cudaStream_t)stream));
    return B200_OK;
}
int b200_stream_create(b200_stream_t *stream) {
    if (!stream) { set_error("b200_stream_create: NULL out pointer"); return B200_ERR_INVALID; }
    cudaStream_t s;
    B200_CUDA_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    *stream = (b200_stream_t)s;
    return B200_OK;
}
int b200_stream_destroy(b200_stream_t stream) { B200_CUDA_TRY(cudaStreamDestroy((cudaStream_t)stream)); return B200_OK; }
int b200_stream_synchronize(b200_stream_t stream) { B200_CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream)); return B200_OK; }
int b200_device_synchronize(void) { B200_CUDA_TRY(cudaDeviceSynchronize()); return B200_OK; }

int b200_event_create(void **event) {
    if (!event) { set_error("b200_event_create: NULL out pointer"); return B200_ERR_INVALID; }
    cudaEvent_t e;
    B200_CUDA_TRY(cudaEventCreate(&e));
    *event = (void *)e;
    return B200_OK;
}
int b200_event_destroy(void *event) { B200_CUDA_TRY(cudaEventDestroy((cudaEvent_t)event)); return B200_OK; }
int b200_event_record(void *event, b200_stream_t stream) { B200_CUDA_TRY(cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream)); return B200_OK; }
int b200_event_elapsed_ms(void *start, void *stop, float *ms) {
    B200_CUDA_TRY(cudaEventSynchronize((cudaEvent_t)stop));
    B200_CUDA_TRY(cudaEventElapsedTime(ms, (cudaEvent_t)start, (cudaEvent_t)stop));
    return B200_OK;
}

// stream capture -> executable graph, so a host in any language can replay a launch sequence without per-launch overhead
int b200_graph_begin_capture(b200_stream_t stream) {
    B200_CUDA_TRY(cudaStreamBeginCapture((cudaStream_t)stream, cudaStreamCaptureModeThreadLocal));
    return B200_OK;
}
int b200_graph_end_capture(b200_stream_t stream, void **graph_exec) {
    cudaGraph_t g = nullptr;
    B200_CUDA_TRY(cudaStreamEndCapture((cudaStream_t)stream, &g));
    cudaGraphExec_t e = nullptr;
    const cudaError_t err = cudaGraphInstantiate(&e, g, 0);
    cudaGraphDestroy(g);
    B200_CUDA_TRY(err);
    *graph_exec = e;
    return B200_OK;
}
int b200_graph_launch(void *graph_exec, b200_stream_t stream) { B200_CUDA_TRY(cudaGraphLaunch((cudaGraphExec_t)graph_exec, (cudaStream_t)stream)); return B200_OK; }
int b200_graph_destroy(void *graph_exec) { B200_CUDA_TRY(cudaGraphExecDestroy((cudaGraphExec_t)graph_exec)); return B200_OK; }

}  // extern "C"
