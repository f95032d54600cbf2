// backend_internal.h -- shared declarations of the ggml_backend glue (buffers, device/reg, graph executor).
#pragma once

#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/b200_ggml_backend.h"
#include "../../include/b200_ggml_ops.h"

// ---- dry run (B200_DRY_RUN=<n devices>): host-side validation of the drop-in boundary on a machine WITHOUT a GPU.
// Every CUDA runtime call of the backend glue (buffers, devices, events, streams, copies) is routed through these wrappers; with
// the flag set they emulate n devices in host memory, graph_compute plans the graph (fusion rules, supports_op coverage,
// launch count) and zero-fills the outputs instead of launching.  It exists so that the reference's own llama.cpp +
// ggml_backend_sched can drive the vtables in CI; NOTHING is computed (this is not a CPU fallback: results are zeros).
namespace b200 {
extern bool g_dry;
extern int g_dry_devices;
namespace dry {
inline int &cur() { static thread_local int d = 0; return d; }
inline cudaError_t SetDevice(int d) { if (!g_dry) return cudaSetDevice(d); cur() = d; return cudaSuccess; }
inline cudaError_t GetDevice(int *d) { if (!g_dry) return cudaGetDevice(d); *d = cur(); return cudaSuccess; }
inline cudaError_t GetDeviceCount(int *n) { if (!g_dry) return cudaGetDeviceCount(n); *n = g_dry_devices; return cudaSuccess; }
inline cudaError_t GetDeviceProperties(cudaDeviceProp *p, int d) {
    if (!g_dry) return cudaGetDeviceProperties(p, d);
    memset(p, 0, sizeof(*p)); snprintf(p->name, sizeof(p->name), "B200 (dry run, no device)"); return cudaSuccess;
}
inline cudaError_t MemGetInfo(size_t *f, size_t *t) { if (!g_dry) return cudaMemGetInfo(f, t); *f = (size_t)170 << 30; *t = (size_t)180 << 30; return cudaSuccess; }
inline cudaError_t Malloc(void **p, size_t n) {
    if (!g_dry) return cudaMalloc(p, n);
    return posix_memalign(p, 256, n ? n : 256) == 0 ? cudaSuccess : cudaErrorMemoryAllocation;
}
template <typename T> inline cudaError_t Malloc(T **p, size_t n) { return Malloc((void **)p, n); }
inline cudaError_t Free(void *p) { if (!g_dry) return cudaFree(p); free(p); return cudaSuccess; }
inline cudaError_t MallocHost(void **p, size_t n) { if (!g_dry) return cudaMallocHost(p, n); return Malloc(p, n); }
inline cudaError_t FreeHost(void *p) { if (!g_dry) return cudaFreeHost(p); free(p); return cudaSuccess; }
inline cudaError_t MemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind k, cudaStream_t st = 0) {
    if (!g_dry) return cudaMemcpyAsync(d, s, n, k, st);
    memmove(d, s, n); return cudaSuccess;
}
inline cudaError_t MemcpyPeerAsync(void *d, int dd, const void *s, int sd, size_t n, cudaStream_t st = 0) {
    if (!g_dry) return cudaMemcpyPeerAsync(d, dd, s, sd, n, st);
    memmove(d, s, n); return cudaSuccess;
}
inline cudaError_t Memcpy2DAsync(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h, cudaMemcpyKind k, cudaStream_t st = 0) {
    if (!g_dry) return cudaMemcpy2DAsync(d, dp, s, sp, w, h, k, st);
    for (size_t i = 0; i < h; i++) memmove((char *)d + i * dp, (const char *)s + i * sp, w);
    return cudaSuccess;
}
inline cudaError_t MemsetAsync(void *d, int v, size_t n, cudaStream_t st = 0) { if (!g_dry) return cudaMemsetAsync(d, v, n, st); memset(d, v, n); return cudaSuccess; }
inline cudaError_t Memset(void *d, int v, size_t n) { if (!g_dry) return cudaMemset(d, v, n); memset(d, v, n); return cudaSuccess; }
inline cudaError_t StreamSynchronize(cudaStream_t st) { return g_dry ? cudaSuccess : cudaStreamSynchronize(st); }
inline cudaError_t DeviceSynchronize() { return g_dry ? cudaSuccess : cudaDeviceSynchronize(); }
inline cudaError_t StreamCreateWithFlags(cudaStream_t *s, unsigned f) { if (!g_dry) return cudaStreamCreateWithFlags(s, f); *s = (cudaStream_t)0x10; return cudaSuccess; }
inline cudaError_t StreamDestroy(cudaStream_t s) { return g_dry ? cudaSuccess : cudaStreamDestroy(s); }
inline cudaError_t StreamWaitEvent(cudaStream_t s, cudaEvent_t e, unsigned f) { return g_dry ? cudaSuccess : cudaStreamWaitEvent(s, e, f); }
inline cudaError_t EventCreateWithFlags(cudaEvent_t *e, unsigned f) { if (!g_dry) return cudaEventCreateWithFlags(e, f); *e = (cudaEvent_t)0x20; return cudaSuccess; }
inline cudaError_t EventDestroy(cudaEvent_t e) { return g_dry ? cudaSuccess : cudaEventDestroy(e); }
inline cudaError_t EventRecord(cudaEvent_t e, cudaStream_t s) { return g_dry ? cudaSuccess : cudaEventRecord(e, s); }
inline cudaError_t EventSynchronize(cudaEvent_t e) { return g_dry ? cudaSuccess : cudaEventSynchronize(e); }
inline cudaError_t GetLastError() { return g_dry ? cudaSuccess : cudaGetLastError(); }
inline cudaError_t DeviceCanAccessPeer(int *c, int a, int b) { if (!g_dry) return cudaDeviceCanAccessPeer(c, a, b); *c = 1; return cudaSuccess; }
inline cudaError_t DeviceEnablePeerAccess(int d, unsigned f) { return g_dry ? cudaSuccess : cudaDeviceEnablePeerAccess(d, f); }
inline cudaError_t HostRegister(void *p, size_t n, unsigned f) { return g_dry ? cudaSuccess : cudaHostRegister(p, n, f); }
inline cudaError_t HostUnregister(void *p) { return g_dry ? cudaSuccess : cudaHostUnregister(p); }
}  // namespace dry
}  // namespace b200
// the glue files call the runtime through the wrappers (kernels and graph capture are never reached in a dry run)
#define cudaSetDevice ::b200::dry::SetDevice
#define cudaGetDevice ::b200::dry::GetDevice
#define cudaGetDeviceCount ::b200::dry::GetDeviceCount
#define cudaGetDeviceProperties ::b200::dry::GetDeviceProperties
#define cudaMemGetInfo ::b200::dry::MemGetInfo
#define cudaMalloc ::b200::dry::Malloc
#define cudaFree ::b200::dry::Free
#define cudaMallocHost ::b200::dry::MallocHost
#define cudaFreeHost ::b200::dry::FreeHost
#define cudaMemcpyAsync ::b200::dry::MemcpyAsync
#define cudaMemcpyPeerAsync ::b200::dry::MemcpyPeerAsync
#define cudaMemcpy2DAsync ::b200::dry::Memcpy2DAsync
#define cudaMemsetAsync ::b200::dry::MemsetAsync
#define cudaMemset ::b200::dry::Memset
#define cudaStreamSynchronize ::b200::dry::StreamSynchronize
#define cudaDeviceSynchronize ::b200::dry::DeviceSynchronize
#define cudaStreamCreateWithFlags ::b200::dry::StreamCreateWithFlags
#define cudaStreamDestroy ::b200::dry::StreamDestroy
#define cudaStreamWaitEvent ::b200::dry::StreamWaitEvent
#define cudaEventCreateWithFlags ::b200::dry::EventCreateWithFlags
#define cudaEventDestroy ::b200::dry::EventDestroy
#define cudaEventRecord ::b200::dry::EventRecord
#define cudaEventSynchronize ::b200::dry::EventSynchronize
#define cudaGetLastError ::b200::dry::GetLastError
#define cudaDeviceCanAccessPeer ::b200::dry::DeviceCanAccessPeer
#define cudaDeviceEnablePeerAccess ::b200::dry::DeviceEnablePeerAccess
#define cudaHostRegister ::b200::dry::HostRegister
#define cudaHostUnregister ::b200::dry::HostUnregister

namespace b200 {

// ---- the type traits the backend needs (ggml.c:740-1080 type_traits[]: blck_size, type_size) without linking ggml
struct TypeTraits { int blck; int size; };
inline TypeTraits type_traits(int t) {
    switch (t) {
        case GGML_TYPE_F32: return {1, 4};
        case GGML_TYPE_F16: return {1, 2};
        case GGML_TYPE_BF16: return {1, 2};
        case GGML_TYPE_Q4_0: return {32, 18};
        case GGML_TYPE_Q4_1: return {32, 20};
        case GGML_TYPE_Q5_0: return {32, 22};
        case GGML_TYPE_Q5_1: return {32, 24};
        case GGML_TYPE_Q8_0: return {32, 34};
        case GGML_TYPE_Q8_1: return {32, 36};
        case GGML_TYPE_Q2_K: return {256, 84};
        case GGML_TYPE_Q3_K: return {256, 110};
        case GGML_TYPE_Q4_K: return {256, 144};
        case GGML_TYPE_Q5_K: return {256, 176};
        case GGML_TYPE_Q6_K: return {256, 210};
        case GGML_TYPE_Q8_K: return {256, 292};
        case GGML_TYPE_IQ4_NL: return {32, 18};
        case GGML_TYPE_IQ4_XS: return {256, 136};
        case GGML_TYPE_I8: return {1, 1};
        case GGML_TYPE_I16: return {1, 2};
        case GGML_TYPE_I32: return {1, 4};
        case GGML_TYPE_I64: return {1, 8};
        case GGML_TYPE_F64: return {1, 8};
        default: return {0, 0};
    }
}
inline bool is_quantized(int t) { return type_traits(t).blck > 1; }
inline int64_t nelements(const ggml_tensor *t) { return t->ne[0] * t->ne[1] * t->ne[2] * t->ne[3]; }
inline int64_t nrows(const ggml_tensor *t) { return t->ne[1] * t->ne[2] * t->ne[3]; }
inline size_t row_size(int type, int64_t ne) { const TypeTraits tt = type_traits(type); return tt.blck ? (size_t)(ne / tt.blck) * tt.size : 0; }
// ggml_nbytes, ggml.c:3544-3564
inline size_t nbytes(const ggml_tensor *t) {
    const TypeTraits tt = type_traits(t->type);
    if (!tt.blck) return 0;
    size_t n;
    if (tt.blck == 1) {
        n = tt.size;
        for (int i = 0; i < GGML_MAX_DIMS; i++) n += (size_t)(t->ne[i] - 1) * t->nb[i];
    } else {
        n = (size_t)t->ne[0] * t->nb[0] / tt.blck;
        for (int i = 1; i < GGML_MAX_DIMS; i++) n += (size_t)(t->ne[i] - 1) * t->nb[i];
    }
    return n;
}
inline bool is_contiguous(const ggml_tensor *t) {
    const TypeTraits tt = type_traits(t->type);
    if (!tt.blck) return false;
    return t->nb[0] == (size_t)tt.size && t->nb[1] == t->nb[0] * (size_t)(t->ne[0] / tt.blck) && t->nb[2] == t->nb[1] * (size_t)t->ne[1] &&
           t->nb[3] == t->nb[2] * (size_t)t->ne[2];
}
inline bool same_shape(const ggml_tensor *a, const ggml_tensor *b) {
    return a->ne[0] == b->ne[0] && a->ne[1] == b->ne[1] && a->ne[2] == b->ne[2] && a->ne[3] == b->ne[3];
}
inline b200_tensor as_b200(const ggml_tensor *t) {
    b200_tensor d;
    d.data = t->data;
    d.type = (int32_t)t->type;
    d._pad = 0;
    for (int i = 0; i < 4; i++) { d.ne[i] = t->ne[i]; d.nb[i] = t->nb[i]; }
    return d;
}

// error convention of the reference backend: a CUDA failure is fatal (ggml_cuda_error -> GGML_ABORT, ggml-cuda.cu:63-73)
[[noreturn]] void backend_abort(const char *what, const char *detail);
#define B200_CHECK_CUDA(call)                                                          \
    do {                                                                               \
        cudaError_t e__ = (call);                                                      \
        if (e__ != cudaSuccess) ::b200::backend_abort(#call, cudaGetErrorString(e__)); \
    } while (0)
#define B200_CHECK_RC(call)                                                  \
    do {                                                                     \
        int rc__ = (call);                                                   \
        if (rc__ != 0) ::b200::backend_abort(#call, b200_last_error());      \
    } while (0)

// ---- contexts hanging off the vtable objects
struct DeviceCtx {
    int device;
    std::string name, description;
};
struct BuftCtx {
    int device;
    std::string name;
};
struct BufferCtx {
    int   device;
    void *base;
    size_t size;
};
struct SplitBuftCtx {
    float split[GGML_B200_MAX_DEVICES + 1];  // cumulative fractions, split[n_dev] = 1
    int   n_dev;
};
// tensor->extra of a tensor living in a split buffer: one row slice per device
struct SplitTensorExtra {
    void   *data[GGML_B200_MAX_DEVICES];
    int64_t row_low[GGML_B200_MAX_DEVICES], row_high[GGML_B200_MAX_DEVICES];
};

int device_count();
ggml_backend_dev_t device_get(int i);
bool buft_is_b200_device(ggml_backend_buffer_type_t buft);
bool buft_is_b200_split(ggml_backend_buffer_type_t buft);
bool buft_is_b200_host(ggml_backend_buffer_type_t buft);
bool buffer_is_b200_device(ggml_backend_buffer_t buf);

// graph executor (backend_graph.cu)
struct GraphExec;
GraphExec *graph_exec_new(int device);
void graph_exec_free(GraphExec *g);
cudaStream_t graph_exec_stream(GraphExec *g);
enum ggml_status graph_exec_compute(GraphExec *g, ggml_cgraph *cgraph);
bool op_supported(int device, const ggml_tensor *op);

extern bool g_use_graphs, g_use_fusion;
extern int64_t g_stat_launches, g_stat_replays, g_stat_captures;

}  // namespace b200
