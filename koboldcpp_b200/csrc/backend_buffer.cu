// backend_buffer.cu -- buffer types and buffers of the B200 ggml backend (device, pinned host, row-split).
//
// Mirrors the contracts of /root/reference/ggml/src/ggml-cuda.cu:424-621 (device buffer + type), :623-953 (split buffer
// + type), :955-1025 (host buffer type) behind the vtables of ggml-backend-impl.h:15-65.  Not a port: weights keep
// their on-disk block layout (the GEMV streams raw blocks through the TMA engine), so set/get_tensor are plain copies
// at any offset and no row padding is needed -- the buffer only over-allocates a fixed tail so kernels may read up to
// the next 16-byte boundary.
#include "backend_internal.h"

namespace b200 {

static const size_t BUFFER_TAIL_SLACK = 256;
static const size_t TENSOR_ALIGNMENT = 128;  // ggml-cuda.cu:567-571

[[noreturn]] void backend_abort(const char *what, const char *detail) {
    fprintf(stderr, "B200 backend error: %s: %s\n", what, detail ? detail : "");
    fflush(stderr);
    abort();
}

struct DeviceGuard {  // cudaSetDevice for the scope, restore afterwards
    int prev = 0;
    explicit DeviceGuard(int dev) {
        B200_CHECK_CUDA(cudaGetDevice(&prev));
        if (prev != dev) B200_CHECK_CUDA(cudaSetDevice(dev));
        else prev = -1;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// ------------------------------------------------------------------------------------------------ device buffer
static const char *buffer_get_name(ggml_backend_buffer_t b) { return ((BuftCtx *)b->buft->context)->name.c_str(); }
static void buffer_free(ggml_backend_buffer_t b) {
    BufferCtx *c = (BufferCtx *)b->context;
    DeviceGuard g(c->device);
    cudaFree(c->base);
    delete c;
}
static void *buffer_get_base(ggml_backend_buffer_t b) { return ((BufferCtx *)b->context)->base; }
static void buffer_init_tensor(ggml_backend_buffer_t, ggml_tensor *) {}  // layout is the host layout: nothing to prepare
static void buffer_memset_tensor(ggml_backend_buffer_t b, ggml_tensor *t, uint8_t v, size_t off, size_t size) {
    BufferCtx *c = (BufferCtx *)b->context;
    DeviceGuard g(c->device);
    B200_CHECK_CUDA(cudaMemsetAsync((char *)t->data + off, v, size, cudaStreamPerThread));
    B200_CHECK_CUDA(cudaStreamSynchronize(cudaStreamPerThread));
}
static void buffer_set_tensor(ggml_backend_buffer_t b, ggml_tensor *t, const void *data, size_t off, size_t size) {
    BufferCtx *c = (BufferCtx *)b->context;
    DeviceGuard g(c->device);
    B200_CHECK_CUDA(cudaMemcpyAsync((char *)t->data + off, data, size, cudaMemcpyHostToDevice, cudaStreamPerThread));
    B200_CHECK_CUDA(cudaStreamSynchronize(cudaStreamPerThread));
}
static void buffer_get_tensor(ggml_backend_buffer_t b, const ggml_tensor *t, void *data, size_t off, size_t size) {
    BufferCtx *c = (BufferCtx *)b->context;
    DeviceGuard g(c->device);
    B200_CHECK_CUDA(cudaMemcpyAsync(data, (const char *)t->data + off, size, cudaMemcpyDeviceToHost, cudaStreamPerThread));
    B200_CHECK_CUDA(cudaStreamSynchronize(cudaStreamPerThread));
}
static bool buffer_cpy_tensor(ggml_backend_buffer_t b, const ggml_tensor *src, ggml_tensor *dst) {
    if (!src->buffer || !buffer_is_b200_device(src->buffer)) return false;  // the caller falls back to get+set
    BufferCtx *sc = (BufferCtx *)src->buffer->context, *dc = (BufferCtx *)b->context;
    DeviceGuard g(dc->device);
    if (sc->device == dc->device) {
        B200_CHECK_CUDA(cudaMemcpyAsync(dst->data, src->data, nbytes(src), cudaMemcpyDeviceToDevice, cudaStreamPerThread));
    } else {
        B200_CHECK_CUDA(cudaMemcpyPeerAsync(dst->data, dc->device, src->data, sc->device, nbytes(src), cudaStreamPerThread));
    }
    B200_CHECK_CUDA(cudaStreamSynchronize(cudaStreamPerThread));
    return true;
}
static void buffer_clear(ggml_backend_buffer_t b, uint8_t v) {
    BufferCtx *c = (BufferCtx *)b->context;
    DeviceGuard g(c->device);
    B200_CHECK_CUDA(cudaDeviceSynchronize());
    B200_CHECK_CUDA(cudaMemset(c->base, v, c->size));
    B200_CHECK_CUDA(cudaDeviceSynchronize());
}
static const ggml_backend_buffer_i k_buffer_iface = {
    buffer_get_name, buffer_free, buffer_get_base, buffer_init_tensor, buffer_memset_tensor, buffer_set_tensor,
    buffer_get_tensor, buffer_cpy_tensor, buffer_clear, /* reset */ nullptr,
};
bool buffer_is_b200_device(ggml_backend_buffer_t buf) { return buf && buf->iface.get_name == buffer_get_name; }

// ggml_backend_buffer_init (ggml-backend.cpp:75-89) allocates the handle with `new` and ggml_backend_buffer_free deletes it;
// we build the identical object ourselves so the library has no link-time dependency on the host's libggml.
static ggml_backend_buffer_t make_buffer(ggml_backend_buffer_type_t buft, const ggml_backend_buffer_i &iface, void *ctx, size_t size) {
    return new ggml_backend_buffer{iface, buft, ctx, size, GGML_BACKEND_BUFFER_USAGE_ANY};
}

// ------------------------------------------------------------------------------------------------ device buffer type
static const char *buft_get_name(ggml_backend_buffer_type_t t) { return ((BuftCtx *)t->context)->name.c_str(); }
static ggml_backend_buffer_t buft_alloc_buffer(ggml_backend_buffer_type_t t, size_t size) {
    BuftCtx *bc = (BuftCtx *)t->context;
    DeviceGuard g(bc->device);
    void *p = nullptr;
    size = size ? size : 1;
    cudaError_t e = cudaMalloc(&p, size + BUFFER_TAIL_SLACK);
    if (e != cudaSuccess) {  // nullptr on OOM: the caller handles it (ggml-cuda.cu:555-560)
        cudaGetLastError();
        fprintf(stderr, "B200 backend: allocating %.2f MiB on device %d failed: %s\n", size / 1048576.0, bc->device, cudaGetErrorString(e));
        return nullptr;
    }
    return make_buffer(t, k_buffer_iface, new BufferCtx{bc->device, p, size}, size);
}
static size_t buft_get_alignment(ggml_backend_buffer_type_t) { return TENSOR_ALIGNMENT; }
static size_t buft_get_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor *t) { return nbytes(t); }
static bool buft_is_host_false(ggml_backend_buffer_type_t) { return false; }
static const ggml_backend_buffer_type_i k_buft_iface = {
    buft_get_name, buft_alloc_buffer, buft_get_alignment, /* get_max_size */ nullptr, buft_get_alloc_size, buft_is_host_false,
};
bool buft_is_b200_device(ggml_backend_buffer_type_t buft) { return buft && buft->iface.get_name == buft_get_name; }

// ------------------------------------------------------------------------------------------------ pinned host buffer type
static const char *host_name(ggml_backend_buffer_type_t) { return "B200_Host"; }
static const char *host_buffer_name(ggml_backend_buffer_t) { return "B200_Host"; }
static void host_buffer_free(ggml_backend_buffer_t b) { cudaFreeHost(b->context); }
static void *host_buffer_base(ggml_backend_buffer_t b) { return b->context; }
static void host_memset(ggml_backend_buffer_t, ggml_tensor *t, uint8_t v, size_t off, size_t size) { memset((char *)t->data + off, v, size); }
static void host_set(ggml_backend_buffer_t, ggml_tensor *t, const void *d, size_t off, size_t size) { memcpy((char *)t->data + off, d, size); }
static void host_get(ggml_backend_buffer_t, const ggml_tensor *t, void *d, size_t off, size_t size) { memcpy(d, (const char *)t->data + off, size); }
static bool host_cpy(ggml_backend_buffer_t, const ggml_tensor *src, ggml_tensor *dst) {
    if (src->buffer && src->buffer->buft && src->buffer->buft->iface.is_host && src->buffer->buft->iface.is_host(src->buffer->buft)) {
        memcpy(dst->data, src->data, nbytes(src));
        return true;
    }
    return false;
}
static void host_clear(ggml_backend_buffer_t b, uint8_t v) { memset(b->context, v, b->size); }
static const ggml_backend_buffer_i k_host_buffer_iface = {
    host_buffer_name, host_buffer_free, host_buffer_base, nullptr, host_memset, host_set, host_get, host_cpy, host_clear, nullptr,
};
static ggml_backend_buffer_t host_alloc(ggml_backend_buffer_type_t t, size_t size) {
    void *p = nullptr;
    if (cudaMallocHost(&p, size ? size : 1) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;  // caller falls back to the CPU buffer type
    }
    return make_buffer(t, k_host_buffer_iface, p, size);
}
static size_t host_alignment(ggml_backend_buffer_type_t) { return 64; }
static bool host_is_host(ggml_backend_buffer_type_t) { return true; }
static const ggml_backend_buffer_type_i k_host_buft_iface = {host_name, host_alloc, host_alignment, nullptr, nullptr, host_is_host};
bool buft_is_b200_host(ggml_backend_buffer_type_t buft) { return buft && buft->iface.get_name == host_name; }

}  // namespace b200

using namespace b200;

extern "C" ggml_backend_buffer_type_t ggml_backend_b200_buffer_type(int device) {
    static std::mutex mu;
    static ggml_backend_buffer_type types[GGML_B200_MAX_DEVICES];
    static bool init = false;
    std::lock_guard<std::mutex> lock(mu);
    if (device < 0 || device >= device_count()) return nullptr;
    if (!init) {
        for (int i = 0; i < device_count() && i < GGML_B200_MAX_DEVICES; i++) {
            types[i] = ggml_backend_buffer_type{k_buft_iface, device_get(i), new BuftCtx{i, "B200_" + std::to_string(i)}};
        }
        init = true;
    }
    return &types[device];
}

extern "C" ggml_backend_buffer_type_t ggml_backend_b200_host_buffer_type(void) {
    static ggml_backend_buffer_type t = {k_host_buft_iface, device_count() > 0 ? device_get(0) : nullptr, nullptr};
    return &t;
}
