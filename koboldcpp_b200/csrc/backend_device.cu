// backend_device.cu -- reg / device / backend (stream) objects of the B200 ggml backend.
// Contracts: /root/reference/ggml/src/ggml-backend-impl.h:86-216; the reference implementation being replaced is
// ggml-cuda.cu:2355-2453 (backend), :2886-3256 (device), :3260-3340 (reg).
#include <chrono>
#include "backend_internal.h"

namespace b200 {

bool g_use_graphs = true, g_use_fusion = true;
bool g_dry = false;
int g_dry_devices = 0;
int64_t g_stat_launches = 0, g_stat_replays = 0, g_stat_captures = 0;

int device_count() {
    static int n = -1;
    if (n < 0) {
        int c = 0;
        if (const char *e = getenv("B200_DRY_RUN")) {  // host-side validation without a GPU: see backend_internal.h
            g_dry_devices = atoi(e) > 0 ? atoi(e) : 1;
            g_dry = true;
            fprintf(stderr, "B200 backend: DRY RUN with %d emulated device(s): graphs are planned, nothing is computed\n", g_dry_devices);
        }
        if (cudaGetDeviceCount(&c) != cudaSuccess) { cudaGetLastError(); c = 0; }
        n = c > GGML_B200_MAX_DEVICES ? GGML_B200_MAX_DEVICES : c;
        if (getenv("B200_NO_GRAPHS")) g_use_graphs = false;
        if (getenv("B200_NO_FUSION")) g_use_fusion = false;
        // programmatic dependent launch: a GEMV's weight prefetch overlaps the tail of its producer (b200_ggml.h)
        if (!g_dry) b200_set_pdl(getenv("B200_NO_PDL") ? 0 : 1);
    }
    return n;
}

// ------------------------------------------------------------------------------------------------ backend (stream)
struct BackendCtx {
    int device;
    std::string name;
    GraphExec *exec;
};
static ggml_guid k_guid = {0xb2, 0x00, 0x5a, 0x10, 0x0a, 0x6b, 0x6f, 0x62, 0x6f, 0x6c, 0x64, 0x63, 0x70, 0x70, 0x00, 0x01};

static const char *be_get_name(ggml_backend_t b) { return ((BackendCtx *)b->context)->name.c_str(); }
static void be_free(ggml_backend_t b) {
    BackendCtx *c = (BackendCtx *)b->context;
    graph_exec_free(c->exec);
    delete c;
    delete b;
}
static ggml_backend_buffer_type_t be_default_buft(ggml_backend_t b) { return ggml_backend_b200_buffer_type(((BackendCtx *)b->context)->device); }
static void be_set_tensor_async(ggml_backend_t b, ggml_tensor *t, const void *data, size_t off, size_t size) {
    BackendCtx *c = (BackendCtx *)b->context;
    B200_CHECK_CUDA(cudaSetDevice(c->device));
    B200_CHECK_CUDA(cudaMemcpyAsync((char *)t->data + off, data, size, cudaMemcpyHostToDevice, graph_exec_stream(c->exec)));
}
static void be_get_tensor_async(ggml_backend_t b, const ggml_tensor *t, void *data, size_t off, size_t size) {
    BackendCtx *c = (BackendCtx *)b->context;
    B200_CHECK_CUDA(cudaSetDevice(c->device));
    B200_CHECK_CUDA(cudaMemcpyAsync(data, (const char *)t->data + off, size, cudaMemcpyDeviceToHost, graph_exec_stream(c->exec)));
}
static bool be_cpy_tensor_async(ggml_backend_t src_be, ggml_backend_t dst_be, const ggml_tensor *src, ggml_tensor *dst) {
    if (!ggml_backend_is_b200(src_be) || !ggml_backend_is_b200(dst_be)) return false;
    if (!src->buffer || !dst->buffer || !buffer_is_b200_device(src->buffer) || !buffer_is_b200_device(dst->buffer)) return false;
    BackendCtx *sc = (BackendCtx *)src_be->context, *dc = (BackendCtx *)dst_be->context;
    if (src_be == dst_be) {
        B200_CHECK_CUDA(cudaSetDevice(sc->device));
        B200_CHECK_CUDA(cudaMemcpyAsync(dst->data, src->data, nbytes(dst), cudaMemcpyDeviceToDevice, graph_exec_stream(sc->exec)));
        return true;
    }
    // copy on the source stream, then make the destination stream wait for it (ggml-cuda.cu:2392-2445)
    B200_CHECK_CUDA(cudaSetDevice(sc->device));
    B200_CHECK_CUDA(cudaMemcpyPeerAsync(dst->data, dc->device, src->data, sc->device, nbytes(dst), graph_exec_stream(sc->exec)));
    cudaEvent_t ev;
    B200_CHECK_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    B200_CHECK_CUDA(cudaEventRecord(ev, graph_exec_stream(sc->exec)));
    B200_CHECK_CUDA(cudaSetDevice(dc->device));
    B200_CHECK_CUDA(cudaStreamWaitEvent(graph_exec_stream(dc->exec), ev, 0));
    B200_CHECK_CUDA(cudaEventDestroy(ev));
    return true;
}
static void be_synchronize(ggml_backend_t b) {
    BackendCtx *c = (BackendCtx *)b->context;
    B200_CHECK_CUDA(cudaSetDevice(c->device));
    B200_CHECK_CUDA(cudaStreamSynchronize(graph_exec_stream(c->exec)));
}
static enum ggml_status be_graph_compute(ggml_backend_t b, ggml_cgraph *g) {
    BackendCtx *c = (BackendCtx *)b->context;
    B200_CHECK_CUDA(cudaSetDevice(c->device));
    return graph_exec_compute(c->exec, g);
}
static void be_event_record(ggml_backend_t b, ggml_backend_event_t e) {
    BackendCtx *c = (BackendCtx *)b->context;
    B200_CHECK_CUDA(cudaEventRecord((cudaEvent_t)e->context, graph_exec_stream(c->exec)));
}
static void be_event_wait(ggml_backend_t b, ggml_backend_event_t e) {
    BackendCtx *c = (BackendCtx *)b->context;
    B200_CHECK_CUDA(cudaStreamWaitEvent(graph_exec_stream(c->exec), (cudaEvent_t)e->context, 0));
}
static const ggml_backend_i k_backend_iface = {
    be_get_name, be_free, be_default_buft, be_set_tensor_async, be_get_tensor_async, be_cpy_tensor_async, be_synchronize,
    /* graph_plan_* */ nullptr, nullptr, nullptr, nullptr, be_graph_compute,
    /* supports_op, supports_buft, offload_op: moved to the device interface */ nullptr, nullptr, nullptr,
    be_event_record, be_event_wait,
};

// ------------------------------------------------------------------------------------------------ device
static const char *dev_get_name(ggml_backend_dev_t d) { return ((DeviceCtx *)d->context)->name.c_str(); }
static const char *dev_get_description(ggml_backend_dev_t d) { return ((DeviceCtx *)d->context)->description.c_str(); }
static void dev_get_memory(ggml_backend_dev_t d, size_t *free, size_t *total) {
    ggml_backend_b200_get_device_memory(((DeviceCtx *)d->context)->device, free, total);
}
static enum ggml_backend_dev_type dev_get_type(ggml_backend_dev_t) { return GGML_BACKEND_DEVICE_TYPE_GPU_FULL; }
static void dev_get_props(ggml_backend_dev_t d, ggml_backend_dev_props *p) {
    p->name = dev_get_name(d);
    p->description = dev_get_description(d);
    p->type = GGML_BACKEND_DEVICE_TYPE_GPU_FULL;
    dev_get_memory(d, &p->memory_free, &p->memory_total);
    // async=false: weights are uploaded with whole-tensor set_tensor (src/llama.cpp:5151, 5187); host buffers and events exist
    p->caps = ggml_backend_dev_caps{false, true, false, true};
}
static ggml_backend_t dev_init_backend(ggml_backend_dev_t d, const char *) { return ggml_backend_b200_init(((DeviceCtx *)d->context)->device); }
static ggml_backend_buffer_type_t dev_get_buft(ggml_backend_dev_t d) { return ggml_backend_b200_buffer_type(((DeviceCtx *)d->context)->device); }
static ggml_backend_buffer_type_t dev_get_host_buft(ggml_backend_dev_t) { return ggml_backend_b200_host_buffer_type(); }
// (B200_HOST_TIMING: the scheduler asks this for every node of every graph -- ggml-backend.cpp:1615-1975 -- so its cost is part of
// what the caller pays per token)
static const bool g_sup_timing = getenv("B200_HOST_TIMING") != nullptr;
static double g_sup_s = 0;
static long g_sup_calls = 0;
static void sup_report() { fprintf(stderr, "B200 host timing: supports_op %ld calls, %.1f ns per call, %.3f ms in total\n", g_sup_calls, 1e9 * g_sup_s / (double)(g_sup_calls ? g_sup_calls : 1), 1e3 * g_sup_s); }
static bool dev_supports_op(ggml_backend_dev_t d, const ggml_tensor *op) {
    if (!g_sup_timing) return op_supported(((DeviceCtx *)d->context)->device, op);
    if (g_sup_calls++ == 0) atexit(sup_report);
    const auto t0 = std::chrono::steady_clock::now();
    const bool r = op_supported(((DeviceCtx *)d->context)->device, op);
    g_sup_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return r;
}
static bool dev_supports_buft(ggml_backend_dev_t d, ggml_backend_buffer_type_t buft) {
    if (buft_is_b200_split(buft)) return true;
    if (buft_is_b200_device(buft)) return ((BuftCtx *)buft->context)->device == ((DeviceCtx *)d->context)->device;
    return false;
}
static bool dev_offload_op(ggml_backend_dev_t, const ggml_tensor *op) {
    // take ops whose weights live on the host only for real batches (ggml-cuda.cu:3201-3208)
    const int64_t batch = op->op == GGML_OP_MUL_MAT_ID ? op->ne[2] : op->ne[1];
    return batch >= 32 && op->op != GGML_OP_GET_ROWS;
}
static ggml_backend_event_t dev_event_new(ggml_backend_dev_t d) {
    B200_CHECK_CUDA(cudaSetDevice(((DeviceCtx *)d->context)->device));
    cudaEvent_t ev;
    B200_CHECK_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    return new ggml_backend_event{d, ev};
}
static void dev_event_free(ggml_backend_dev_t, ggml_backend_event_t e) {
    cudaEventDestroy((cudaEvent_t)e->context);
    delete e;
}
static void dev_event_sync(ggml_backend_dev_t, ggml_backend_event_t e) { B200_CHECK_CUDA(cudaEventSynchronize((cudaEvent_t)e->context)); }
static const ggml_backend_device_i k_device_iface = {
    dev_get_name, dev_get_description, dev_get_memory, dev_get_type, dev_get_props, dev_init_backend, dev_get_buft, dev_get_host_buft,
    /* buffer_from_host_ptr */ nullptr, dev_supports_op, dev_supports_buft, dev_offload_op, dev_event_new, dev_event_free, dev_event_sync,
};

// ------------------------------------------------------------------------------------------------ reg
static const char *reg_get_name(ggml_backend_reg_t) { return "B200"; }
static size_t reg_get_device_count(ggml_backend_reg_t) { return (size_t)device_count(); }
static ggml_backend_dev_t reg_get_device(ggml_backend_reg_t, size_t i) { return device_get((int)i); }
static void *reg_get_proc_address(ggml_backend_reg_t, const char *name) {
    if (strcmp(name, "ggml_backend_split_buffer_type") == 0) return (void *)ggml_backend_b200_split_buffer_type;
    return nullptr;
}
static const ggml_backend_reg_i k_reg_iface = {reg_get_name, reg_get_device_count, reg_get_device, reg_get_proc_address};

ggml_backend_dev_t device_get(int i) {
    static std::mutex mu;
    static ggml_backend_device devs[GGML_B200_MAX_DEVICES];
    static bool init = false;
    std::lock_guard<std::mutex> lock(mu);
    if (i < 0 || i >= device_count()) return nullptr;
    if (!init) {
        for (int d = 0; d < device_count(); d++) {
            cudaDeviceProp prop;
            B200_CHECK_CUDA(cudaGetDeviceProperties(&prop, d));
            devs[d] = ggml_backend_device{k_device_iface, ggml_backend_b200_reg(), new DeviceCtx{d, "B200_" + std::to_string(d), prop.name}};
        }
        init = true;
    }
    return &devs[i];
}

}  // namespace b200

using namespace b200;

extern "C" {

ggml_backend_reg_t ggml_backend_b200_reg(void) {
    static ggml_backend_reg reg = {k_reg_iface, nullptr};
    return &reg;
}

ggml_backend_t ggml_backend_b200_init(int device) {
    if (device < 0 || device >= device_count()) {
        fprintf(stderr, "ggml_backend_b200_init: invalid device %d\n", device);
        return nullptr;
    }
    B200_CHECK_CUDA(cudaSetDevice(device));
    BackendCtx *c = new BackendCtx{device, "B200_" + std::to_string(device), graph_exec_new(device)};
    return new ggml_backend{&k_guid, k_backend_iface, device_get(device), c};
}

bool ggml_backend_is_b200(ggml_backend_t backend) { return backend != nullptr && backend->guid == &k_guid; }
int ggml_backend_b200_get_device_count(void) { return device_count(); }

void ggml_backend_b200_get_device_description(int device, char *description, size_t description_size) {
    if (!description || !description_size) return;
    ggml_backend_dev_t d = device_get(device);
    snprintf(description, description_size, "%s", d ? ((DeviceCtx *)d->context)->description.c_str() : "");
}

void ggml_backend_b200_get_device_memory(int device, size_t *free, size_t *total) {
    int prev = 0;
    B200_CHECK_CUDA(cudaGetDevice(&prev));
    B200_CHECK_CUDA(cudaSetDevice(device));
    size_t f = 0, t = 0;
    B200_CHECK_CUDA(cudaMemGetInfo(&f, &t));
    cudaSetDevice(prev);
    if (free) *free = f;
    if (total) *total = t;
}

void ggml_backend_b200_set_graphs(int enabled) { device_count(); g_use_graphs = enabled != 0; }
void ggml_backend_b200_set_fusion(int enabled) { device_count(); g_use_fusion = enabled != 0; }
void ggml_backend_b200_get_stats(int64_t *launches, int64_t *replays, int64_t *captures) {
    if (launches) *launches = g_stat_launches;
    if (replays) *replays = g_stat_replays;
    if (captures) *captures = g_stat_captures;
}
void *ggml_backend_b200_get_stream(ggml_backend_t b) { return ggml_backend_is_b200(b) ? (void *)graph_exec_stream(((BackendCtx *)b->context)->exec) : nullptr; }
void ggml_backend_b200_reset_stats(void) { g_stat_launches = g_stat_replays = g_stat_captures = 0; }

// ---- Route 1: the public symbols of ggml/include/ggml-cuda.h:22-44 under their reference names, so the library can be
// linked in place of ggml-cuda.o with -DGGML_USE_CUDA and gpttype_adapter.cpp's row-split / mmq knobs work unmodified.
B200_API void ggml_cuda_set_mul_mat_q(bool) {}  // gpttype_adapter.cpp:1883 -- the quantised tensor-core GEMM is always on here
B200_API ggml_backend_t ggml_backend_cuda_init(int device) { return ggml_backend_b200_init(device); }
B200_API bool ggml_backend_is_cuda(ggml_backend_t b) { return ggml_backend_is_b200(b); }
B200_API ggml_backend_buffer_type_t ggml_backend_cuda_buffer_type(int device) { return ggml_backend_b200_buffer_type(device); }
B200_API ggml_backend_buffer_type_t ggml_backend_cuda_split_buffer_type(const float *ts) { return ggml_backend_b200_split_buffer_type(ts); }
B200_API ggml_backend_buffer_type_t ggml_backend_cuda_host_buffer_type(void) { return ggml_backend_b200_host_buffer_type(); }
B200_API int ggml_backend_cuda_get_device_count(void) { return device_count(); }
B200_API void ggml_backend_cuda_get_device_description(int d, char *s, size_t n) { ggml_backend_b200_get_device_description(d, s, n); }
B200_API void ggml_backend_cuda_get_device_memory(int d, size_t *f, size_t *t) { ggml_backend_b200_get_device_memory(d, f, t); }
B200_API bool ggml_backend_cuda_register_host_buffer(void *buffer, size_t size) {
    if (cudaHostRegister(buffer, size, cudaHostRegisterPortable | cudaHostRegisterReadOnly) != cudaSuccess) { cudaGetLastError(); return false; }
    return true;
}
B200_API void ggml_backend_cuda_unregister_host_buffer(void *buffer) { if (cudaHostUnregister(buffer) != cudaSuccess) cudaGetLastError(); }
B200_API ggml_backend_reg_t ggml_backend_cuda_reg(void) { return ggml_backend_b200_reg(); }

}  // extern "C"
