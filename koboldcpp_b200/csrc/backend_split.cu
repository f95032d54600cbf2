// backend_split.cu -- (a13) the row-split buffer type: weight rows sharded over the visible devices.
//
// Contract of the reference's split buffer (ggml-cuda.cu:623-953): get_base returns a dummy pointer, tensors keep their
// per-device row slices in tensor->extra, set/get_tensor take whole tensors only, no views, clear is a no-op.
// Row ranges come from cumulative tensor_split fractions rounded down to multiples of 128 rows (get_row_split, :638-651);
// a NULL / all-zero split divides rows in proportion to each device's free memory (:175-191).
#include "backend_internal.h"

namespace b200 {

static const int64_t SPLIT_ROW_ROUNDING = 128;

struct SplitBufferCtx {
    std::vector<SplitTensorExtra *> extras;
};

static void row_split(const SplitBuftCtx *s, int64_t nrows_total, int dev, int64_t *lo, int64_t *hi) {
    int64_t a = (int64_t)(nrows_total * (double)s->split[dev]);
    a -= a % SPLIT_ROW_ROUNDING;
    int64_t b;
    if (dev == s->n_dev - 1) b = nrows_total;
    else {
        b = (int64_t)(nrows_total * (double)s->split[dev + 1]);
        b -= b % SPLIT_ROW_ROUNDING;
    }
    *lo = a;
    *hi = b;
}

static const char *split_buffer_name(ggml_backend_buffer_t) { return "B200_Split"; }
static void split_buffer_free(ggml_backend_buffer_t b) {
    SplitBufferCtx *c = (SplitBufferCtx *)b->context;
    SplitBuftCtx *s = (SplitBuftCtx *)b->buft->context;
    for (SplitTensorExtra *e : c->extras) {
        for (int d = 0; d < s->n_dev; d++) {
            if (e->data[d]) { cudaSetDevice(d); cudaFree(e->data[d]); }
        }
        delete e;
    }
    delete c;
}
static void *split_buffer_base(ggml_backend_buffer_t) { return (void *)0x1000; }  // dummy, as in the reference (:699-704)

static void split_init_tensor(ggml_backend_buffer_t b, ggml_tensor *t) {
    if (t->view_src != nullptr) backend_abort("split buffer", "views of split tensors are not supported");
    SplitBufferCtx *c = (SplitBufferCtx *)b->context;
    SplitBuftCtx *s = (SplitBuftCtx *)b->buft->context;
    SplitTensorExtra *e = new SplitTensorExtra();
    memset(e, 0, sizeof(*e));
    c->extras.push_back(e);
    const int64_t nr = nrows(t);
    const size_t rb = row_size(t->type, t->ne[0]);
    for (int d = 0; d < s->n_dev; d++) {
        row_split(s, nr, d, &e->row_low[d], &e->row_high[d]);
        const int64_t n = e->row_high[d] - e->row_low[d];
        if (n <= 0) continue;
        B200_CHECK_CUDA(cudaSetDevice(d));
        B200_CHECK_CUDA(cudaMalloc(&e->data[d], (size_t)n * rb + 256));  // init_tensor cannot fail: abort like the reference (:734-738)
    }
    t->extra = e;
}
static void split_set_tensor(ggml_backend_buffer_t b, ggml_tensor *t, const void *data, size_t off, size_t size) {
    if (off != 0 || size != nbytes(t)) backend_abort("split buffer", "set_tensor must cover the whole tensor");
    SplitBuftCtx *s = (SplitBuftCtx *)b->buft->context;
    SplitTensorExtra *e = (SplitTensorExtra *)t->extra;
    const size_t rb = row_size(t->type, t->ne[0]);
    for (int d = 0; d < s->n_dev; d++) {
        const int64_t n = e->row_high[d] - e->row_low[d];
        if (n <= 0) continue;
        B200_CHECK_CUDA(cudaSetDevice(d));
        B200_CHECK_CUDA(cudaMemcpyAsync(e->data[d], (const char *)data + (size_t)e->row_low[d] * rb, (size_t)n * rb, cudaMemcpyHostToDevice, cudaStreamPerThread));
    }
    for (int d = 0; d < s->n_dev; d++) {
        B200_CHECK_CUDA(cudaSetDevice(d));
        B200_CHECK_CUDA(cudaStreamSynchronize(cudaStreamPerThread));
    }
}
static void split_get_tensor(ggml_backend_buffer_t b, const ggml_tensor *t, void *data, size_t off, size_t size) {
    if (off != 0 || size != nbytes(t)) backend_abort("split buffer", "get_tensor must cover the whole tensor");
    SplitBuftCtx *s = (SplitBuftCtx *)b->buft->context;
    SplitTensorExtra *e = (SplitTensorExtra *)t->extra;
    const size_t rb = row_size(t->type, t->ne[0]);
    for (int d = 0; d < s->n_dev; d++) {
        const int64_t n = e->row_high[d] - e->row_low[d];
        if (n <= 0) continue;
        B200_CHECK_CUDA(cudaSetDevice(d));
        B200_CHECK_CUDA(cudaMemcpyAsync((char *)data + (size_t)e->row_low[d] * rb, e->data[d], (size_t)n * rb, cudaMemcpyDeviceToHost, cudaStreamPerThread));
    }
    for (int d = 0; d < s->n_dev; d++) {
        B200_CHECK_CUDA(cudaSetDevice(d));
        B200_CHECK_CUDA(cudaStreamSynchronize(cudaStreamPerThread));
    }
}
static void split_clear(ggml_backend_buffer_t, uint8_t) {}
static const ggml_backend_buffer_i k_split_buffer_iface = {
    split_buffer_name, split_buffer_free, split_buffer_base, split_init_tensor, /* memset */ nullptr, split_set_tensor, split_get_tensor,
    /* cpy */ nullptr, split_clear, nullptr,
};

static const char *split_buft_name(ggml_backend_buffer_type_t) { return "B200_Split"; }
static ggml_backend_buffer_t split_alloc(ggml_backend_buffer_type_t t, size_t size) {
    // memory is allocated per tensor in init_tensor; the size only feeds the caller's accounting
    return new ggml_backend_buffer{k_split_buffer_iface, t, new SplitBufferCtx(), size, GGML_BACKEND_BUFFER_USAGE_ANY};
}
static size_t split_alignment(ggml_backend_buffer_type_t) { return 128; }
static size_t split_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor *t) { return nbytes(t); }
static bool split_is_host(ggml_backend_buffer_type_t) { return false; }
static const ggml_backend_buffer_type_i k_split_buft_iface = {split_buft_name, split_alloc, split_alignment, nullptr, split_alloc_size, split_is_host};
bool buft_is_b200_split(ggml_backend_buffer_type_t buft) { return buft && buft->iface.get_name == split_buft_name; }

}  // namespace b200

using namespace b200;

extern "C" ggml_backend_buffer_type_t ggml_backend_b200_split_buffer_type(const float *tensor_split) {
    static std::mutex mu;
    static std::vector<ggml_backend_buffer_type *> cache;
    std::lock_guard<std::mutex> lock(mu);
    const int n = device_count();
    if (n == 0) return nullptr;
    SplitBuftCtx s;
    memset(&s, 0, sizeof(s));
    s.n_dev = n;
    float frac[GGML_B200_MAX_DEVICES] = {0};
    float sum = 0;
    bool all_zero = true;
    if (tensor_split) {
        for (int i = 0; i < n; i++) all_zero = all_zero && tensor_split[i] == 0.0f;
    }
    for (int i = 0; i < n; i++) {
        if (!tensor_split || all_zero) {
            size_t f = 0, t = 0;
            ggml_backend_b200_get_device_memory(i, &f, &t);
            frac[i] = (float)t;
        } else {
            frac[i] = tensor_split[i];
        }
        sum += frac[i];
    }
    float acc = 0;
    for (int i = 0; i < n; i++) {
        s.split[i] = acc / sum;
        acc += frac[i];
    }
    s.split[n] = 1.0f;
    for (ggml_backend_buffer_type *t : cache) {
        if (memcmp(((SplitBuftCtx *)t->context)->split, s.split, sizeof(s.split)) == 0) return t;
    }
    ggml_backend_buffer_type *t = new ggml_backend_buffer_type{k_split_buft_iface, device_get(0), new SplitBuftCtx(s)};
    cache.push_back(t);
    return t;
}
