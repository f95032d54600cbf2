// host_llama.cu -- the caller side of the hot path: a llama / mixtral decoder that drives the B200 backend ONLY through
// the ggml_backend vtables (buffer type -> buffer -> set/get_tensor, backend -> graph_compute), the way llama.cpp does.
// It builds the per-batch graph node for node like the reference's build_llama (src/llama.cpp:10453-10617); the node
// constructors below fill struct ggml_tensor the way the ggml_* builders do (ggml.c: ggml_new_tensor_impl :3790-3870,
// ggml_view_impl :6130-6190, ggml_permute :6300-6350, ggml_rope_impl :6920-6975, ggml_flash_attn_ext :7130-7180, ...).
// No compute happens in this file.
#include <chrono>
#include <cmath>
#include <deque>
#include <map>
#include <unordered_set>

#include "../../include/b200_comm.h"
#include "../../include/b200_llama.h"
#include "backend_internal.h"

namespace b200 {
void set_error(const char *fmt, ...);
}
using namespace b200;

namespace {

const int KV_PAD = 256;  // llama_kv_cache_get_padding with flash attention, src/llama.cpp:4143-4146

// ------------------------------------------------------------------------------------------------ tensor arena
struct Arena {
    std::deque<ggml_tensor> pool;  // stable addresses
    size_t used = 0;
    void reset() { used = 0; }
    ggml_tensor *get() {
        if (used == pool.size()) pool.emplace_back();
        ggml_tensor *t = &pool[used++];
        memset(t, 0, sizeof(*t));
        return t;
    }
};

void set_shape(ggml_tensor *t, int type, int64_t ne0, int64_t ne1, int64_t ne2, int64_t ne3) {
    const TypeTraits tt = type_traits(type);
    t->type = (ggml_type)type;
    t->ne[0] = ne0; t->ne[1] = ne1; t->ne[2] = ne2; t->ne[3] = ne3;
    t->nb[0] = (size_t)tt.size;
    t->nb[1] = t->nb[0] * (size_t)(ne0 / tt.blck);
    t->nb[2] = t->nb[1] * (size_t)ne1;
    t->nb[3] = t->nb[2] * (size_t)ne2;
}

// bump allocator over one backend buffer (stands in for ggml-alloc: every graph gets the same addresses again)
struct Bump {
    ggml_backend_buffer_t buf = nullptr;
    char *base = nullptr;
    size_t size = 0, off = 0, align = 128;
    void *take(size_t bytes) {
        off = (off + align - 1) / align * align;
        if (off + bytes > size) return nullptr;
        void *p = base + off;
        off += bytes;
        return p;
    }
};

struct Builder {
    Arena *arena;
    Bump *bump;
    std::vector<ggml_tensor *> nodes;
    std::unordered_set<const ggml_tensor *> visited;
    bool oom = false;

    // ggml_build_forward_expand: post-order DFS over src[], left to right (ggml_visit_parents, ggml.c:19020-19070);
    // tensors with op NONE are leafs and not part of the node list
    void expand(ggml_tensor *t) {
        if (!t || !visited.insert(t).second) return;
        for (int i = 0; i < GGML_MAX_SRC; i++) expand(t->src[i]);
        if (t->op != GGML_OP_NONE) nodes.push_back(t);
    }

    ggml_tensor *alloc_result(int type, int64_t ne0, int64_t ne1 = 1, int64_t ne2 = 1, int64_t ne3 = 1) {
        ggml_tensor *t = arena->get();
        set_shape(t, type, ne0, ne1, ne2, ne3);
        t->data = bump->take(nbytes(t));
        if (!t->data) oom = true;
        t->buffer = bump->buf;
        return t;
    }
    ggml_tensor *op(ggml_tensor *t, int o, ggml_tensor *a, ggml_tensor *b = nullptr, ggml_tensor *c = nullptr, ggml_tensor *d = nullptr) {
        t->op = (ggml_op)o;
        t->src[0] = a; t->src[1] = b; t->src[2] = c; t->src[3] = d;
        return t;
    }
    // views share the data of their source (ggml_view_impl)
    ggml_tensor *view_of(ggml_tensor *a, int o, size_t offs) {
        ggml_tensor *t = arena->get();
        t->type = a->type;
        t->buffer = a->buffer;
        t->view_src = a->view_src ? a->view_src : a;
        t->view_offs = (a->view_src ? a->view_offs : 0) + offs;
        t->data = a->data ? (char *)a->data + offs : nullptr;
        t->op = (ggml_op)o;
        t->src[0] = a;
        return t;
    }
    ggml_tensor *reshape(ggml_tensor *a, int64_t ne0, int64_t ne1 = 1, int64_t ne2 = 1, int64_t ne3 = 1) {
        ggml_tensor *t = view_of(a, GGML_OP_RESHAPE, 0);
        set_shape(t, a->type, ne0, ne1, ne2, ne3);
        return t;
    }
    ggml_tensor *view_1d(ggml_tensor *a, int64_t ne0, size_t offs) {
        ggml_tensor *t = view_of(a, GGML_OP_VIEW, offs);
        set_shape(t, a->type, ne0, 1, 1, 1);
        memcpy(t->op_params, &offs, sizeof(offs));
        return t;
    }
    ggml_tensor *view_2d(ggml_tensor *a, int64_t ne0, int64_t ne1, size_t nb1, size_t offs) {
        ggml_tensor *t = view_of(a, GGML_OP_VIEW, offs);
        set_shape(t, a->type, ne0, ne1, 1, 1);
        t->nb[1] = nb1; t->nb[2] = nb1 * (size_t)ne1; t->nb[3] = t->nb[2];
        memcpy(t->op_params, &offs, sizeof(offs));
        return t;
    }
    ggml_tensor *view_3d(ggml_tensor *a, int64_t ne0, int64_t ne1, int64_t ne2, size_t nb1, size_t nb2, size_t offs) {
        ggml_tensor *t = view_of(a, GGML_OP_VIEW, offs);
        set_shape(t, a->type, ne0, ne1, ne2, 1);
        t->nb[1] = nb1; t->nb[2] = nb2; t->nb[3] = nb2 * (size_t)ne2;
        memcpy(t->op_params, &offs, sizeof(offs));
        return t;
    }
    ggml_tensor *permute(ggml_tensor *a, int ax0, int ax1, int ax2, int ax3) {
        ggml_tensor *t = view_of(a, GGML_OP_PERMUTE, 0);
        const int ax[4] = {ax0, ax1, ax2, ax3};
        for (int i = 0; i < 4; i++) { t->ne[ax[i]] = a->ne[i]; t->nb[ax[i]] = a->nb[i]; }
        for (int i = 0; i < 4; i++) t->op_params[i] = ax[i];
        return t;
    }
    ggml_tensor *rms_norm(ggml_tensor *a, float eps) {
        ggml_tensor *t = alloc_result(GGML_TYPE_F32, a->ne[0], a->ne[1], a->ne[2], a->ne[3]);
        memcpy(t->op_params, &eps, 4);
        return op(t, GGML_OP_RMS_NORM, a);
    }
    ggml_tensor *binary(int o, ggml_tensor *a, ggml_tensor *b) {
        return op(alloc_result(GGML_TYPE_F32, a->ne[0], a->ne[1], a->ne[2], a->ne[3]), o, a, b);
    }
    ggml_tensor *mul_mat(ggml_tensor *w, ggml_tensor *x) {
        return op(alloc_result(GGML_TYPE_F32, w->ne[1], x->ne[1], x->ne[2], x->ne[3]), GGML_OP_MUL_MAT, w, x);
    }
    // as [K, M, n_expert], b [K, n_b_rows, T], ids [n_used, T] -> [M, n_used, T]   (ggml.c:5693-5729)
    ggml_tensor *mul_mat_id(ggml_tensor *as, ggml_tensor *b, ggml_tensor *ids) {
        return op(alloc_result(GGML_TYPE_F32, as->ne[1], ids->ne[0], b->ne[2], 1), GGML_OP_MUL_MAT_ID, as, b, ids);
    }
    ggml_tensor *rope(ggml_tensor *a, ggml_tensor *pos, ggml_tensor *ff, int n_dims, int mode, int n_ctx_orig, float freq_base, float freq_scale) {
        ggml_tensor *t = alloc_result(GGML_TYPE_F32, a->ne[0], a->ne[1], a->ne[2], a->ne[3]);
        const float ext_factor = 0.0f, attn_factor = 1.0f, beta_fast = 32.0f, beta_slow = 1.0f;  // llama_context_default_params
        int32_t p[11] = {0, n_dims, mode, 0, n_ctx_orig};
        memcpy(p + 5, &freq_base, 4); memcpy(p + 6, &freq_scale, 4); memcpy(p + 7, &ext_factor, 4);
        memcpy(p + 8, &attn_factor, 4); memcpy(p + 9, &beta_fast, 4); memcpy(p + 10, &beta_slow, 4);
        memcpy(t->op_params, p, sizeof(p));
        return op(t, GGML_OP_ROPE, a, pos, ff);
    }
    // CPY(a, b): the result is a view of b (ggml_cpy_impl, ggml.c:5982-6005)
    ggml_tensor *cpy(ggml_tensor *a, ggml_tensor *b) {
        ggml_tensor *t = arena->get();
        *t = *b;
        t->view_src = b->view_src ? b->view_src : b;
        t->view_offs = b->view_src ? b->view_offs : 0;
        memset(t->op_params, 0, sizeof(t->op_params));
        memset(t->src, 0, sizeof(t->src));
        return op(t, GGML_OP_CPY, a, b);
    }
    ggml_tensor *cast(ggml_tensor *a, int type) {  // ggml_cast: CPY into a fresh tensor of `type`
        ggml_tensor *t = alloc_result(type, a->ne[0], a->ne[1], a->ne[2], a->ne[3]);
        return op(t, GGML_OP_CPY, a, t);
    }
    ggml_tensor *get_rows(ggml_tensor *a, ggml_tensor *ids) {
        return op(alloc_result(GGML_TYPE_F32, a->ne[0], ids->ne[0], ids->ne[1], 1), GGML_OP_GET_ROWS, a, ids);
    }
    ggml_tensor *flash_attn(ggml_tensor *q, ggml_tensor *k, ggml_tensor *v, ggml_tensor *mask, float scale) {
        // result [D, n_head, N] (ggml.c:7165: ne = { v->ne[0], q->ne[2], q->ne[1], q->ne[3] })
        ggml_tensor *t = alloc_result(GGML_TYPE_F32, v->ne[0], q->ne[2], q->ne[1], q->ne[3]);
        const float p[3] = {scale, 0.0f, 0.0f};
        memcpy(t->op_params, p, sizeof(p));
        return op(t, GGML_OP_FLASH_ATTN_EXT, q, k, v, mask);
    }
    ggml_tensor *silu(ggml_tensor *a) {
        ggml_tensor *t = alloc_result(GGML_TYPE_F32, a->ne[0], a->ne[1], a->ne[2], a->ne[3]);
        t->op_params[0] = GGML_UNARY_OP_SILU;
        return op(t, GGML_OP_UNARY, a);
    }
    ggml_tensor *soft_max(ggml_tensor *a) {
        ggml_tensor *t = alloc_result(GGML_TYPE_F32, a->ne[0], a->ne[1], a->ne[2], a->ne[3]);
        const float p[2] = {1.0f, 0.0f};
        memcpy(t->op_params, p, sizeof(p));
        return op(t, GGML_OP_SOFT_MAX, a);
    }
    ggml_tensor *top_k(ggml_tensor *a, int k) {  // ggml_top_k = argsort(desc) + view (ggml.c:7080-7095)
        ggml_tensor *s = alloc_result(GGML_TYPE_I32, a->ne[0], a->ne[1], a->ne[2], a->ne[3]);
        s->op_params[0] = GGML_SORT_ORDER_DESC;
        op(s, GGML_OP_ARGSORT, a);
        ggml_tensor *t = view_of(s, GGML_OP_VIEW, 0);
        set_shape(t, GGML_TYPE_I32, k, s->ne[1], s->ne[2], s->ne[3]);
        t->nb[1] = s->nb[1]; t->nb[2] = s->nb[2]; t->nb[3] = s->nb[3];
        return t;
    }
    // an exchange between the ranks of a tensor-parallel model (b200_comm.h): all-reduce in place, or all-gather along ne0
    ggml_tensor *collective(ggml_tensor *a, int kind, int world) {
        ggml_tensor *t;
        if (kind == B200_COLLECTIVE_ALL_REDUCE_SUM) {
            t = view_of(a, GGML_OP_MAP_CUSTOM1, 0);
            set_shape(t, a->type, a->ne[0], a->ne[1], a->ne[2], a->ne[3]);
        } else {
            t = alloc_result(GGML_TYPE_F32, a->ne[0] * world, a->ne[1], a->ne[2], a->ne[3]);
            op(t, GGML_OP_MAP_CUSTOM1, a);
        }
        t->op_params[0] = (int32_t)B200_COLLECTIVE_TAG;
        t->op_params[1] = kind;
        return t;
    }
    ggml_tensor *sum_rows(ggml_tensor *a) { return op(alloc_result(GGML_TYPE_F32, 1, a->ne[1], a->ne[2], a->ne[3]), GGML_OP_SUM_ROWS, a); }
};

}  // namespace

// ------------------------------------------------------------------------------------------------ model
struct LayerW {
    ggml_tensor *attn_norm = nullptr, *wq = nullptr, *wk = nullptr, *wv = nullptr, *wo = nullptr, *ffn_norm = nullptr;
    ggml_tensor *ffn_gate = nullptr, *ffn_up = nullptr, *ffn_down = nullptr;
    ggml_tensor *ffn_gate_inp = nullptr, *ffn_gate_exps = nullptr, *ffn_up_exps = nullptr, *ffn_down_exps = nullptr;
};

struct b200_llama {
    b200_llama_params hp;
    ggml_backend_t be = nullptr;
    ggml_backend_buffer_type_t buft = nullptr, split_buft = nullptr;
    std::map<std::string, ggml_tensor *> weights;
    std::deque<ggml_tensor> weight_pool;
    std::vector<ggml_backend_buffer_t> buffers;  // everything to free
    int64_t weight_bytes = 0, kv_bytes = 0;
    // resolved weights
    ggml_tensor *tok_embd = nullptr, *output_norm = nullptr, *output = nullptr, *rope_freqs = nullptr;
    std::vector<LayerW> layers;
    // KV cache (one tensor per layer, [n_embd_gqa * n_ctx], src/llama.cpp:3583-3584)
    std::vector<ggml_tensor *> k_l, v_l;
    // compute
    int32_t max_batch = 0;
    Bump bump;
    size_t region_inputs = 0, region_half = 0;  // layout of the compute buffer: [inputs | half A | half B | tail]
    ggml_tensor inp_tokens, inp_pos, inp_mask, inp_out_ids;
    // pinned staging of the per-call inputs: [tokens | positions | out ids | mask], reused every call behind `in_ev`
    unsigned char *pin = nullptr;
    size_t pin_bytes = 0;
    cudaEvent_t in_ev = nullptr;   // recorded behind the input copies of the last call
    bool in_ev_pending = false;
    // what the device-side KQ mask currently holds (single-token decode only moves one element per call)
    // the graph of the previous call, reusable when only the KV write position moved (single-token decode): per layer the two
    // destination views of the cache stores and the CPY nodes that alias them
    struct KvStore { ggml_tensor *view, *cpy; char *base; size_t row; };
    std::vector<KvStore> kv_stores;
    int32_t g_n_tokens = 0, g_n_outputs = 0;
    int64_t g_n_kv = -1;
    double t_build = 0, t_inputs = 0, t_compute = 0, t_sync = 0;  // host seconds per phase (B200_LLAMA_TIMING=1 prints them at free)
    int64_t t_calls = 0;
    int64_t mk_n_kv = -1, mk_rows = -1;
    int32_t mk_n_tokens = 0, mk_n_past = -1;
    Arena arena;
    ggml_cgraph graph;
    std::vector<ggml_tensor *> graph_nodes;
    ggml_tensor *logits_t = nullptr;
    int64_t last_launches = 0;
    bool finalized = false;
};

namespace {

#define LL_REQUIRE(cond, code, ...)      \
    do {                                 \
        if (!(cond)) {                   \
            b200::set_error(__VA_ARGS__); \
            return code;                 \
        }                                \
    } while (0)

bool is_matrix_weight(const std::string &name) {
    // what llama.cpp places in the split buffer type: the 2-D weight matrices of the layers and the output head
    // (llm_load_tensors uses buft_matrix for them, src/llama.cpp:7039-7062); norms, router and token_embd are not split
    if (name == "output.weight") return true;
    if (name.compare(0, 4, "blk.") != 0) return false;
    static const char *const k[] = {"attn_q.weight", "attn_k.weight", "attn_v.weight", "attn_output.weight", "ffn_gate.weight", "ffn_up.weight", "ffn_down.weight"};
    for (const char *s : k) {
        const size_t n = strlen(s);
        if (name.size() > n && name.compare(name.size() - n, n, s) == 0) return true;
    }
    return false;
}

ggml_tensor *alloc_in_buft(b200_llama *m, ggml_backend_buffer_type_t buft, ggml_tensor *t, int usage) {
    const size_t sz = buft->iface.get_alloc_size ? buft->iface.get_alloc_size(buft, t) : nbytes(t);
    ggml_backend_buffer_t buf = buft->iface.alloc_buffer(buft, sz);
    if (!buf) return nullptr;
    buf->usage = (ggml_backend_buffer_usage)usage;
    m->buffers.push_back(buf);
    t->buffer = buf;
    t->data = buf->iface.get_base(buf);
    if (buf->iface.init_tensor) buf->iface.init_tensor(buf, t);
    return t;
}

ggml_tensor *find_w(b200_llama *m, const std::string &name) {
    auto it = m->weights.find(name);
    return it == m->weights.end() ? nullptr : it->second;
}

}  // namespace

extern "C" int b200_llama_new(const b200_llama_params *p, b200_llama **out) {
    LL_REQUIRE(p && out, B200_ERR_INVALID, "b200_llama_new: NULL argument");
    LL_REQUIRE(p->n_embd > 0 && p->n_layer > 0 && p->n_head > 0 && p->n_head_kv > 0 && p->n_head % p->n_head_kv == 0 && p->n_vocab > 0 && p->n_ctx > 0,
               B200_ERR_INVALID, "b200_llama_new: bad hyper-parameters");
    LL_REQUIRE(p->n_ctx % KV_PAD == 0, B200_ERR_INVALID, "b200_llama_new: n_ctx must be a multiple of %d (flash-attention KV padding)", KV_PAD);
    LL_REQUIRE((p->type_k == B200_TYPE_F16 || p->type_k == B200_TYPE_Q8_0 || p->type_k == B200_TYPE_Q4_0) && p->type_v == p->type_k, B200_ERR_UNSUPPORTED,
               "b200_llama_new: KV cache types must be f16/f16, q8_0/q8_0 or q4_0/q4_0");
    LL_REQUIRE(ggml_backend_b200_get_device_count() > 0, B200_ERR_CUDA, "b200_llama_new: no CUDA device (there is no CPU fallback)");
    if (p->tp_world > 1) {
        LL_REQUIRE(p->tp_rank >= 0 && p->tp_rank < p->tp_world && b200_comm_world() == p->tp_world && b200_comm_rank() == p->tp_rank, B200_ERR_INVALID,
                   "b200_llama_new: tensor-parallel mode needs a communicator of %d ranks (b200_comm_init)", p->tp_world);
        LL_REQUIRE(p->n_head % p->tp_world == 0 && p->n_head_kv % p->tp_world == 0 && p->n_ff % p->tp_world == 0 && p->n_vocab % p->tp_world == 0 &&
                   (p->n_ff / p->tp_world) % 256 == 0, B200_ERR_UNSUPPORTED, "b200_llama_new: heads / ffn / vocabulary do not divide over %d ranks", p->tp_world);
        LL_REQUIRE(p->n_expert == 0 && p->n_split_devices <= 1, B200_ERR_UNSUPPORTED, "b200_llama_new: tensor-parallel mode is for dense models without the in-process split");
    }
    b200_llama *m = new b200_llama();
    m->hp = *p;
    if (m->hp.tp_world < 1) { m->hp.tp_world = 1; m->hp.tp_rank = 0; }
    if (m->hp.n_embd_head == 0) m->hp.n_embd_head = p->n_embd / p->n_head;
    if (m->hp.n_ctx_orig == 0) m->hp.n_ctx_orig = p->n_ctx;
    if (m->hp.rope_freq_base == 0.0f) m->hp.rope_freq_base = 10000.0f;
    if (m->hp.rope_freq_scale == 0.0f) m->hp.rope_freq_scale = 1.0f;
    if (m->hp.rms_eps == 0.0f) m->hp.rms_eps = 1e-5f;
    m->be = ggml_backend_b200_init(p->main_device);
    if (!m->be) { delete m; b200::set_error("b200_llama_new: device %d does not exist", p->main_device); return B200_ERR_INVALID; }
    m->buft = m->be->iface.get_default_buffer_type(m->be);
    if (p->n_split_devices > 1) {
        // found the way llama.cpp finds it: reg->get_proc_address("ggml_backend_split_buffer_type") (src/llama.cpp:3498-3510)
        ggml_backend_reg_t reg = ggml_backend_b200_reg();
        ggml_backend_split_buffer_type_t fn = (ggml_backend_split_buffer_type_t)reg->iface.get_proc_address(reg, "ggml_backend_split_buffer_type");
        float ts[16] = {0};
        bool any = false;
        for (int i = 0; i < 16; i++) { ts[i] = p->tensor_split[i]; any |= ts[i] != 0.0f; }
        if (!any) for (int i = 0; i < p->n_split_devices && i < 16; i++) ts[i] = 1.0f;
        m->split_buft = fn ? fn(ts) : nullptr;
        if (!m->split_buft) { b200_llama_free(m); b200::set_error("b200_llama_new: no split buffer type"); return B200_ERR_UNSUPPORTED; }
    }
    m->layers.resize((size_t)p->n_layer);
    memset(&m->graph, 0, sizeof(m->graph));
    *out = m;
    return B200_OK;
}

extern "C" void b200_llama_free(b200_llama *m) {
    if (!m) return;
    if (m->be) m->be->iface.synchronize(m->be);
    for (ggml_backend_buffer_t b : m->buffers) {
        if (b->iface.free_buffer) b->iface.free_buffer(b);
        delete b;  // ggml_backend_buffer_free: iface.free_buffer(buffer); delete buffer  (ggml-backend.cpp:95-104)
    }
    if (m->be) m->be->iface.free(m->be);
    if (getenv("B200_LLAMA_TIMING") && m->t_calls > 0) {
        const double k = 1e6 / (double)m->t_calls;
        fprintf(stderr, "b200_llama host time per decode call (us, %lld calls): build_graph %.1f  inputs %.1f  graph_compute %.1f  synchronize %.1f\n",
                (long long)m->t_calls, m->t_build * k, m->t_inputs * k, m->t_compute * k, m->t_sync * k);
    }
    if (m->pin) cudaFreeHost(m->pin);
    if (m->in_ev) cudaEventDestroy(m->in_ev);
    delete m;
}

extern "C" int b200_llama_set_tensor(b200_llama *m, const char *name, int type, int64_t ne0, int64_t ne1, int64_t ne2, const void *data) {
    LL_REQUIRE(m && name && data, B200_ERR_INVALID, "b200_llama_set_tensor: NULL argument");
    LL_REQUIRE(!m->finalized, B200_ERR_INVALID, "b200_llama_set_tensor: model already finalized");
    const TypeTraits tt = type_traits(type);
    LL_REQUIRE(tt.blck > 0 && ne0 > 0 && ne1 > 0 && ne2 > 0 && ne0 % tt.blck == 0, B200_ERR_INVALID, "b200_llama_set_tensor: bad type/shape for %s", name);
    LL_REQUIRE(m->weights.find(name) == m->weights.end(), B200_ERR_INVALID, "b200_llama_set_tensor: duplicate tensor %s", name);
    m->weight_pool.emplace_back();
    ggml_tensor *t = &m->weight_pool.back();
    memset(t, 0, sizeof(*t));
    set_shape(t, type, ne0, ne1, ne2, 1);
    snprintf(t->name, sizeof(t->name), "%s", name);
    ggml_backend_buffer_type_t buft = (m->split_buft && ne2 == 1 && is_matrix_weight(name)) ? m->split_buft : m->buft;
    if (!alloc_in_buft(m, buft, t, GGML_BACKEND_BUFFER_USAGE_WEIGHTS)) {
        m->weight_pool.pop_back();
        b200::set_error("b200_llama_set_tensor: out of device memory for %s (%zu bytes)", name, nbytes(t));
        return B200_ERR_NOMEM;
    }
    t->buffer->iface.set_tensor(t->buffer, t, data, 0, nbytes(t));
    m->weights[name] = t;
    m->weight_bytes += (int64_t)nbytes(t);
    return B200_OK;
}

extern "C" int b200_llama_clone_tensor(b200_llama *m, const char *name, const char *from) {
    LL_REQUIRE(m && name && from, B200_ERR_INVALID, "b200_llama_clone_tensor: NULL argument");
    ggml_tensor *src = find_w(m, from);
    LL_REQUIRE(src, B200_ERR_INVALID, "b200_llama_clone_tensor: no tensor %s", from);
    LL_REQUIRE(m->weights.find(name) == m->weights.end(), B200_ERR_INVALID, "b200_llama_clone_tensor: duplicate tensor %s", name);
    m->weight_pool.emplace_back();
    ggml_tensor *t = &m->weight_pool.back();
    memset(t, 0, sizeof(*t));
    set_shape(t, src->type, src->ne[0], src->ne[1], src->ne[2], 1);
    snprintf(t->name, sizeof(t->name), "%s", name);
    ggml_backend_buffer_type_t buft = src->buffer->buft;
    if (!alloc_in_buft(m, buft, t, GGML_BACKEND_BUFFER_USAGE_WEIGHTS)) {
        m->weight_pool.pop_back();
        b200::set_error("b200_llama_clone_tensor: out of device memory for %s", name);
        return B200_ERR_NOMEM;
    }
    // ggml_backend_tensor_copy: buffer cpy_tensor when the buffer offers it, else through the host (ggml-backend.cpp:390-420)
    if (!(t->buffer->iface.cpy_tensor && t->buffer->iface.cpy_tensor(t->buffer, src, t))) {
        std::vector<char> tmp(nbytes(src));
        src->buffer->iface.get_tensor(src->buffer, src, tmp.data(), 0, tmp.size());
        t->buffer->iface.set_tensor(t->buffer, t, tmp.data(), 0, tmp.size());
    }
    m->weights[name] = t;
    m->weight_bytes += (int64_t)nbytes(t);
    return B200_OK;
}

// ------------------------------------------------------------------------------------------------ finalize
extern "C" int b200_llama_finalize(b200_llama *m, int32_t max_batch) {
    LL_REQUIRE(m && max_batch > 0, B200_ERR_INVALID, "b200_llama_finalize: bad argument");
    LL_REQUIRE(!m->finalized, B200_ERR_INVALID, "b200_llama_finalize: called twice");
    const b200_llama_params &hp = m->hp;
    const bool moe = hp.n_expert > 0;
    m->tok_embd = find_w(m, "token_embd.weight");
    m->output_norm = find_w(m, "output_norm.weight");
    m->output = find_w(m, "output.weight");
    if (!m->output) m->output = m->tok_embd;  // tied embeddings (src/llama.cpp:7176-7182)
    m->rope_freqs = find_w(m, "rope_freqs.weight");
    LL_REQUIRE(m->tok_embd && m->output_norm && m->output, B200_ERR_INVALID, "b200_llama_finalize: token_embd / output_norm / output missing");
    for (int il = 0; il < hp.n_layer; il++) {
        LayerW &L = m->layers[(size_t)il];
        const std::string pre = "blk." + std::to_string(il) + ".";
        L.attn_norm = find_w(m, pre + "attn_norm.weight");
        L.wq = find_w(m, pre + "attn_q.weight"); L.wk = find_w(m, pre + "attn_k.weight");
        L.wv = find_w(m, pre + "attn_v.weight"); L.wo = find_w(m, pre + "attn_output.weight");
        L.ffn_norm = find_w(m, pre + "ffn_norm.weight");
        LL_REQUIRE(L.attn_norm && L.wq && L.wk && L.wv && L.wo && L.ffn_norm, B200_ERR_INVALID, "b200_llama_finalize: layer %d attention tensors missing", il);
        if (moe) {
            L.ffn_gate_inp = find_w(m, pre + "ffn_gate_inp.weight");
            L.ffn_gate_exps = find_w(m, pre + "ffn_gate_exps.weight");
            L.ffn_up_exps = find_w(m, pre + "ffn_up_exps.weight");
            L.ffn_down_exps = find_w(m, pre + "ffn_down_exps.weight");
            LL_REQUIRE(L.ffn_gate_inp && L.ffn_gate_exps && L.ffn_up_exps && L.ffn_down_exps, B200_ERR_INVALID, "b200_llama_finalize: layer %d expert tensors missing", il);
        } else {
            L.ffn_gate = find_w(m, pre + "ffn_gate.weight"); L.ffn_up = find_w(m, pre + "ffn_up.weight"); L.ffn_down = find_w(m, pre + "ffn_down.weight");
            LL_REQUIRE(L.ffn_gate && L.ffn_up && L.ffn_down, B200_ERR_INVALID, "b200_llama_finalize: layer %d ffn tensors missing", il);
        }
    }
    // KV cache: per layer one 1-D tensor of n_embd_gqa * n_ctx elements, zero-initialised (llama_kv_cache_init, src/llama.cpp:3534-3600)
    const int64_t n_embd_gqa = (int64_t)hp.n_embd_head * (hp.n_head_kv / hp.tp_world);  // this rank's KV heads
    for (int il = 0; il < hp.n_layer; il++) {
        for (int v = 0; v < 2; v++) {
            m->weight_pool.emplace_back();
            ggml_tensor *t = &m->weight_pool.back();
            memset(t, 0, sizeof(*t));
            set_shape(t, v ? hp.type_v : hp.type_k, n_embd_gqa * hp.n_ctx, 1, 1, 1);
            snprintf(t->name, sizeof(t->name), "cache_%c_l%d", v ? 'v' : 'k', il);
            if (!alloc_in_buft(m, m->buft, t, GGML_BACKEND_BUFFER_USAGE_ANY)) { b200::set_error("b200_llama_finalize: out of device memory for the KV cache"); return B200_ERR_NOMEM; }
            t->buffer->iface.clear(t->buffer, 0);
            (v ? m->v_l : m->k_l).push_back(t);
            m->kv_bytes += (int64_t)nbytes(t);
        }
    }
    // compute buffer: [inputs | half A | half B | tail]; layer il allocates from half il&1 (its input lives in the other half)
    m->max_batch = max_batch;
    const size_t B = (size_t)max_batch;
    const size_t mask_cols = (B + GGML_KQ_MASK_PAD - 1) / GGML_KQ_MASK_PAD * GGML_KQ_MASK_PAD;
    const size_t n_used = moe ? (size_t)hp.n_expert_used : 1;
    m->region_inputs = 4 * B * 3 + 4 * (size_t)hp.n_ctx * mask_cols + 2 * (size_t)hp.n_ctx * mask_cols + 4 * B * (size_t)hp.n_embd + 4096;
    m->region_half = 4 * B * (16 * (size_t)hp.n_embd + 6 * (size_t)hp.n_ff * n_used + 8 * (size_t)(hp.n_expert + 8)) + (1 << 16);
    const size_t tail = 4 * B * ((size_t)hp.n_vocab + 4 * (size_t)hp.n_embd) + (1 << 16);
    const size_t total = m->region_inputs + 2 * m->region_half + tail;
    ggml_backend_buffer_t cb = m->buft->iface.alloc_buffer(m->buft, total);
    LL_REQUIRE(cb, B200_ERR_NOMEM, "b200_llama_finalize: out of device memory for the compute buffer (%zu bytes)", total);
    cb->usage = GGML_BACKEND_BUFFER_USAGE_COMPUTE;
    m->buffers.push_back(cb);
    m->bump.buf = cb;
    m->bump.base = (char *)cb->iface.get_base(cb);
    m->bump.size = total;
    m->bump.align = m->buft->iface.get_alignment(m->buft);
    m->bump.off = 0;
    // graph inputs at fixed addresses (llama.cpp: lctx.inp_tokens / inp_pos / inp_KQ_mask / inp_out_ids, :16587-16720)
    auto mk_input = [&](ggml_tensor *t, int type, int64_t ne0, int64_t ne1, const char *name) {
        memset(t, 0, sizeof(*t));
        set_shape(t, type, ne0, ne1, 1, 1);
        t->buffer = cb;
        t->data = m->bump.take(nbytes(t));
        t->flags = GGML_TENSOR_FLAG_INPUT;
        snprintf(t->name, sizeof(t->name), "%s", name);
    };
    mk_input(&m->inp_tokens, GGML_TYPE_I32, max_batch, 1, "inp_tokens");
    mk_input(&m->inp_pos, GGML_TYPE_I32, max_batch, 1, "inp_pos");
    mk_input(&m->inp_out_ids, GGML_TYPE_I32, max_batch, 1, "inp_out_ids");
    mk_input(&m->inp_mask, GGML_TYPE_F32, hp.n_ctx, (int64_t)mask_cols, "KQ_mask");
    LL_REQUIRE(m->inp_mask.data, B200_ERR_NOMEM, "b200_llama_finalize: input region too small");
    m->finalized = true;
    return B200_OK;
}

extern "C" int b200_llama_kv_write(b200_llama *m, int32_t il, int32_t is_v, int32_t cell0, int32_t n, const void *data) {
    LL_REQUIRE(m && m->finalized && data, B200_ERR_INVALID, "b200_llama_kv_write: bad argument");
    LL_REQUIRE(il >= 0 && il < m->hp.n_layer && cell0 >= 0 && n >= 0 && cell0 + n <= m->hp.n_ctx, B200_ERR_INVALID, "b200_llama_kv_write: range");
    ggml_tensor *t = (is_v ? m->v_l : m->k_l)[(size_t)il];
    const size_t row = row_size(t->type, (int64_t)m->hp.n_embd_head * (m->hp.n_head_kv / m->hp.tp_world));
    t->buffer->iface.set_tensor(t->buffer, t, data, row * (size_t)cell0, row * (size_t)n);
    return B200_OK;
}

extern "C" int b200_llama_kv_read(b200_llama *m, int32_t il, int32_t is_v, int32_t cell0, int32_t n, void *data) {
    LL_REQUIRE(m && m->finalized && data, B200_ERR_INVALID, "b200_llama_kv_read: bad argument");
    LL_REQUIRE(il >= 0 && il < m->hp.n_layer && cell0 >= 0 && n >= 0 && cell0 + n <= m->hp.n_ctx, B200_ERR_INVALID, "b200_llama_kv_read: range");
    ggml_tensor *t = (is_v ? m->v_l : m->k_l)[(size_t)il];
    const size_t row = row_size(t->type, (int64_t)m->hp.n_embd_head * (m->hp.n_head_kv / m->hp.tp_world));
    m->be->iface.synchronize(m->be);
    t->buffer->iface.get_tensor(t->buffer, t, data, row * (size_t)cell0, row * (size_t)n);
    return B200_OK;
}

// ------------------------------------------------------------------------------------------------ graph (build_llama)
namespace {

int build_graph(b200_llama *m, int32_t n_tokens, int32_t n_past, int32_t n_outputs) {
    const b200_llama_params &hp = m->hp;
    const int W = hp.tp_world;  // tensor-parallel ranks: this rank's graph holds 1/W of the heads
    const int64_t D = hp.n_embd_head, n_head = hp.n_head / W, n_head_kv = hp.n_head_kv / W;
    const int64_t n_embd_gqa = D * n_head_kv;
    const int32_t kv_head = n_past;
    // kv_self.n = min(size, max(pad, GGML_PAD(cell_max, pad)))  (src/llama.cpp:17163-17171)
    int64_t n_kv = ((int64_t)(n_past + n_tokens) + KV_PAD - 1) / KV_PAD * KV_PAD;
    if (n_kv > hp.n_ctx) n_kv = hp.n_ctx;
    const size_t in_end = m->region_inputs, a0 = in_end, b0 = a0 + m->region_half, t0 = b0 + m->region_half;

    // Same token count, same n_kv, same outputs as the graph we still hold: every node is identical except the destination of the
    // 2 * n_layer cache stores, which moves with kv_head.  Patch those views instead of rebuilding ~900 nodes (the backend keys its
    // captured CUDA graph on everything BUT those destinations, so it replays the same executable either way).
    if (m->graph.n_nodes > 0 && m->g_n_tokens == n_tokens && m->g_n_kv == n_kv && m->g_n_outputs == n_outputs && !m->kv_stores.empty()) {
        for (const b200_llama::KvStore &ks : m->kv_stores) {
            const size_t offs = ks.row * (size_t)kv_head;
            ks.view->data = ks.base + offs;
            ks.view->view_offs = offs;
            memcpy(ks.view->op_params, &offs, sizeof(offs));
            ks.cpy->data = ks.view->data;
            ks.cpy->view_offs = offs;
        }
        return B200_OK;
    }
    m->kv_stores.clear();
    m->g_n_tokens = 0;

    m->arena.reset();
    Builder b;
    b.arena = &m->arena;
    b.bump = &m->bump;
    Bump &bp = m->bump;
    const size_t total = t0 + (bp.size > t0 ? 0 : 0);
    (void)total;
    const size_t buf_size = (size_t)bp.buf->size;
    auto region = [&](size_t lo, size_t hi) { bp.off = lo; bp.size = hi; };

    // per-graph shapes of the fixed-address inputs
    auto input_view = [&](const ggml_tensor &base, int64_t ne0, int64_t ne1) {
        ggml_tensor *t = m->arena.get();
        *t = base;
        set_shape(t, base.type, ne0, ne1, 1, 1);
        return t;
    };
    ggml_tensor *inp_tokens = input_view(m->inp_tokens, n_tokens, 1);
    ggml_tensor *inp_pos = input_view(m->inp_pos, n_tokens, 1);
    const int64_t mask_rows = ((int64_t)n_tokens + GGML_KQ_MASK_PAD - 1) / GGML_KQ_MASK_PAD * GGML_KQ_MASK_PAD;
    ggml_tensor *mask_f32 = input_view(m->inp_mask, n_kv, mask_rows);

    // everything allocated before the first layer lives in the tail of the inputs region
    size_t in_off = (size_t)((char *)m->inp_mask.data - bp.base) + (size_t)hp.n_ctx * ((size_t)m->inp_mask.ne[1]) * 4;
    region(in_off, in_end);
    ggml_tensor *inpL = b.get_rows(m->tok_embd, inp_tokens);             // llm_build_inp_embd, :9086-9120
    ggml_tensor *KQ_mask = b.cast(mask_f32, GGML_TYPE_F16);              // flash_attn: ggml_cast(inp_KQ_mask, F16), :10291-10298
    const float kq_scale = 1.0f / sqrtf((float)D);
    const int rope_mode = hp.rope_neox ? GGML_ROPE_TYPE_NEOX : 0;

    ggml_tensor *cur = nullptr;
    for (int il = 0; il < hp.n_layer; il++) {
        const LayerW &L = m->layers[(size_t)il];
        if (il & 1) region(b0, t0); else region(a0, b0);
        ggml_tensor *inpSA = inpL;
        cur = b.binary(GGML_OP_MUL, b.rms_norm(inpL, hp.rms_eps), L.attn_norm);  // llm_build_norm
        ggml_tensor *Qcur = b.mul_mat(L.wq, cur);
        ggml_tensor *Kcur = b.mul_mat(L.wk, cur);
        ggml_tensor *Vcur = b.mul_mat(L.wv, cur);
        Qcur = b.rope(b.reshape(Qcur, D, n_head, n_tokens), inp_pos, m->rope_freqs, (int)D, rope_mode, hp.n_ctx_orig, hp.rope_freq_base, hp.rope_freq_scale);
        Kcur = b.rope(b.reshape(Kcur, D, n_head_kv, n_tokens), inp_pos, m->rope_freqs, (int)D, rope_mode, hp.n_ctx_orig, hp.rope_freq_base, hp.rope_freq_scale);
        // llm_build_kv: expand q, k, v together, then the cache stores, then attention
        b.expand(Qcur); b.expand(Kcur); b.expand(Vcur);
        ggml_tensor *kc = m->k_l[(size_t)il], *vc = m->v_l[(size_t)il];
        ggml_tensor *k_view = b.view_1d(kc, (int64_t)n_tokens * n_embd_gqa, row_size(kc->type, n_embd_gqa) * (size_t)kv_head);
        ggml_tensor *k_cpy = b.cpy(Kcur, k_view);
        b.expand(k_cpy);
        ggml_tensor *v_view = b.view_1d(vc, (int64_t)n_tokens * n_embd_gqa, row_size(vc->type, n_embd_gqa) * (size_t)kv_head);
        ggml_tensor *v_cpy = b.cpy(Vcur, v_view);
        b.expand(v_cpy);
        m->kv_stores.push_back({k_view, k_cpy, (char *)kc->data, row_size(kc->type, n_embd_gqa)});
        m->kv_stores.push_back({v_view, v_cpy, (char *)vc->data, row_size(vc->type, n_embd_gqa)});
        // llm_build_kqv (flash-attention branch)
        ggml_tensor *q = b.permute(Qcur, 0, 2, 1, 3);
        ggml_tensor *k = b.view_3d(kc, D, n_kv, n_head_kv, row_size(kc->type, n_embd_gqa), row_size(kc->type, D), 0);
        ggml_tensor *v = b.view_3d(vc, D, n_kv, n_head_kv, row_size(vc->type, n_embd_gqa), row_size(vc->type, D), 0);
        cur = b.flash_attn(q, k, v, KQ_mask, kq_scale);
        cur = b.reshape(cur, D * n_head, n_tokens);
        b.expand(cur);
        cur = b.mul_mat(L.wo, cur);
        if (W > 1) cur = b.collective(cur, B200_COLLECTIVE_ALL_REDUCE_SUM, W);  // wo is split along K: partial sums
        if (il == hp.n_layer - 1 && n_outputs < n_tokens) {
            // skip computing output for unused tokens (:10541-10546)
            ggml_tensor *ids = input_view(m->inp_out_ids, n_outputs, 1);
            cur = b.get_rows(cur, ids);
            inpSA = b.get_rows(inpSA, ids);
        }
        ggml_tensor *ffn_inp = b.binary(GGML_OP_ADD, cur, inpSA);
        cur = b.binary(GGML_OP_MUL, b.rms_norm(ffn_inp, hp.rms_eps), L.ffn_norm);
        if (!L.ffn_gate_inp) {
            // llm_build_ffn, LLM_FFN_SILU + LLM_FFN_PAR (:9289-9414)
            ggml_tensor *tmp = b.mul_mat(L.ffn_up, cur);
            cur = b.mul_mat(L.ffn_gate, cur);
            cur = b.silu(cur);
            cur = b.binary(GGML_OP_MUL, cur, tmp);
            cur = b.mul_mat(L.ffn_down, cur);
            if (W > 1) cur = b.collective(cur, B200_COLLECTIVE_ALL_REDUCE_SUM, W);  // ffn_down is split along K
        } else {
            // llm_build_moe_ffn (:9416-9515), norm_w = true
            const int64_t n_expert = hp.n_expert, n_used = hp.n_expert_used, n_tok = cur->ne[1], n_embd = cur->ne[0];
            ggml_tensor *logits = b.mul_mat(L.ffn_gate_inp, cur);
            ggml_tensor *probs = b.soft_max(logits);
            ggml_tensor *sel = b.top_k(probs, (int)n_used);
            ggml_tensor *weights = b.get_rows(b.reshape(probs, 1, n_expert, n_tok), sel);
            weights = b.reshape(weights, n_used, n_tok);
            ggml_tensor *wsum = b.sum_rows(weights);
            weights = b.binary(GGML_OP_DIV, weights, wsum);
            weights = b.reshape(weights, 1, n_used, n_tok);
            cur = b.reshape(cur, n_embd, 1, n_tok);
            ggml_tensor *up = b.mul_mat_id(L.ffn_up_exps, cur, sel);
            ggml_tensor *gate = b.mul_mat_id(L.ffn_gate_exps, cur, sel);
            gate = b.silu(gate);
            ggml_tensor *par = b.binary(GGML_OP_MUL, up, gate);
            ggml_tensor *experts = b.mul_mat_id(L.ffn_down_exps, par, sel);
            experts = b.binary(GGML_OP_MUL, experts, weights);
            ggml_tensor *moe_out = nullptr;
            for (int64_t i = 0; i < n_used; i++) {
                ggml_tensor *ce = b.view_2d(experts, n_embd, n_tok, experts->nb[2], (size_t)i * experts->nb[1]);
                moe_out = i == 0 ? ce : b.binary(GGML_OP_ADD, moe_out, ce);
            }
            cur = moe_out;
        }
        cur = b.binary(GGML_OP_ADD, cur, ffn_inp);
        inpL = cur;
    }
    region(t0, buf_size);
    cur = b.binary(GGML_OP_MUL, b.rms_norm(inpL, hp.rms_eps), m->output_norm);
    cur = b.mul_mat(m->output, cur);
    if (W > 1) {  // the head is split by vocabulary rows
        LL_REQUIRE(n_outputs == 1, B200_ERR_UNSUPPORTED, "b200_llama_decode: tensor-parallel mode returns the logits of the last token only");
        cur = b.collective(cur, B200_COLLECTIVE_ALL_GATHER, W);
    }
    b.expand(cur);
    bp.size = buf_size;
    LL_REQUIRE(!b.oom, B200_ERR_NOMEM, "b200_llama_decode: compute buffer too small for this batch");
    m->logits_t = cur;
    m->graph_nodes.swap(b.nodes);
    memset(&m->graph, 0, sizeof(m->graph));
    m->graph.size = (int)m->graph_nodes.size();
    m->graph.n_nodes = (int)m->graph_nodes.size();
    m->graph.nodes = m->graph_nodes.data();
    m->g_n_tokens = n_tokens;
    m->g_n_kv = n_kv;
    m->g_n_outputs = n_outputs;
    return B200_OK;
}

int enqueue_decode(b200_llama *m, const int32_t *tokens, int32_t n_tokens, int32_t n_past, float *logits, int32_t logits_all) {
    LL_REQUIRE(m && m->finalized && tokens, B200_ERR_INVALID, "b200_llama_decode: bad argument");
    LL_REQUIRE(n_tokens > 0 && n_tokens <= m->max_batch, B200_ERR_INVALID, "b200_llama_decode: n_tokens %d outside (0, %d]", n_tokens, m->max_batch);
    LL_REQUIRE(n_past >= 0 && n_past + n_tokens <= m->hp.n_ctx, B200_ERR_INVALID, "b200_llama_decode: context overflow");
    for (int i = 0; i < n_tokens; i++) LL_REQUIRE(tokens[i] >= 0 && tokens[i] < m->hp.n_vocab, B200_ERR_INVALID, "b200_llama_decode: token id out of range");
    const int32_t n_outputs = logits_all ? n_tokens : 1;
    const auto tp0 = std::chrono::steady_clock::now();
    const int rc = build_graph(m, n_tokens, n_past, n_outputs);
    if (rc != B200_OK) return rc;
    const auto tp1 = std::chrono::steady_clock::now();
    ggml_backend_t be = m->be;
    // llama_set_inputs (:16587-16720): tokens, positions, causal mask (-inf above the diagonal and on padding rows), out ids
    const int64_t n_kv = m->graph_nodes.empty() ? 0 : ((int64_t)(n_past + n_tokens) + KV_PAD - 1) / KV_PAD * KV_PAD > m->hp.n_ctx
                                                          ? m->hp.n_ctx : ((int64_t)(n_past + n_tokens) + KV_PAD - 1) / KV_PAD * KV_PAD;
    const int64_t mask_rows = ((int64_t)n_tokens + GGML_KQ_MASK_PAD - 1) / GGML_KQ_MASK_PAD * GGML_KQ_MASK_PAD;
    // Staging is pinned (a pageable source makes the driver stage and serialise the copy) and reused: wait until the copies of
    // the previous call have left it.
    const size_t mask_floats = (size_t)(n_kv * mask_rows);
    const size_t off_pos = 4 * (size_t)m->max_batch, off_ids = 2 * off_pos, off_mask = (off_ids + 64 + 255) / 256 * 256;
    const size_t pin_need = off_mask + 4 * (size_t)m->hp.n_ctx * (size_t)((m->max_batch + GGML_KQ_MASK_PAD - 1) / GGML_KQ_MASK_PAD * GGML_KQ_MASK_PAD);
    if (m->pin_bytes < pin_need) {
        if (m->in_ev_pending) { cudaEventSynchronize(m->in_ev); m->in_ev_pending = false; }
        if (m->pin) cudaFreeHost(m->pin);
        m->pin = nullptr;
        m->pin_bytes = 0;
        LL_REQUIRE(cudaHostAlloc((void **)&m->pin, pin_need, cudaHostAllocDefault) == cudaSuccess, B200_ERR_NOMEM, "b200_llama_decode: cannot pin %zu bytes of input staging", pin_need);
        m->pin_bytes = pin_need;
        m->mk_n_kv = -1;
    }
    if (m->in_ev_pending) { cudaEventSynchronize(m->in_ev); m->in_ev_pending = false; }
    int32_t *tok_pin = (int32_t *)m->pin, *pos_pin = (int32_t *)(m->pin + off_pos), *ids_pin = (int32_t *)(m->pin + off_ids);
    float *mask_pin = (float *)(m->pin + off_mask);
    for (int i = 0; i < n_tokens; i++) { tok_pin[i] = tokens[i]; pos_pin[i] = n_past + i; }
    be->iface.set_tensor_async(be, &m->inp_tokens, tok_pin, 0, 4 * (size_t)n_tokens);
    // (created here: set_tensor_async has just made the backend's device current, and an event belongs to the device it was created on)
    if (!m->in_ev) LL_REQUIRE(cudaEventCreateWithFlags(&m->in_ev, cudaEventDisableTiming) == cudaSuccess, B200_ERR_CUDA, "b200_llama_decode: cudaEventCreate failed");
    be->iface.set_tensor_async(be, &m->inp_pos, pos_pin, 0, 4 * (size_t)n_tokens);
    // One more token behind the previous one with the same mask geometry: the device mask (which this decoder owns and nothing
    // else writes) differs from what it holds in ONE element -- cell n_past becomes visible.  Anything else: rebuild and send all.
    const bool one_more = n_tokens == 1 && m->mk_n_tokens == 1 && m->mk_n_kv == n_kv && m->mk_rows == mask_rows && n_past == m->mk_n_past + 1 && n_past < n_kv;
    if (one_more) {
        mask_pin[n_past] = 0.0f;
        be->iface.set_tensor_async(be, &m->inp_mask, mask_pin + n_past, 4 * (size_t)n_past, 4);
    } else {
        for (size_t i = 0; i < mask_floats; i++) mask_pin[i] = -INFINITY;
        for (int j = 0; j < n_tokens; j++) {
            float *row = mask_pin + (size_t)j * (size_t)n_kv;
            const int64_t lim = (int64_t)n_past + j;
            for (int64_t i = 0; i <= lim && i < n_kv; i++) row[i] = 0.0f;
        }
        be->iface.set_tensor_async(be, &m->inp_mask, mask_pin, 0, 4 * mask_floats);
    }
    m->mk_n_kv = n_kv; m->mk_rows = mask_rows; m->mk_n_tokens = n_tokens; m->mk_n_past = n_past;
    if (n_outputs < n_tokens) {
        ids_pin[0] = n_tokens - 1;
        be->iface.set_tensor_async(be, &m->inp_out_ids, ids_pin, 0, 4);
    }
    if (cudaEventRecord(m->in_ev, (cudaStream_t)ggml_backend_b200_get_stream(be)) == cudaSuccess) m->in_ev_pending = true;
    else { cudaGetLastError(); cudaStreamSynchronize((cudaStream_t)ggml_backend_b200_get_stream(be)); }  // no event: make the staging reusable the slow way
    const auto tp2 = std::chrono::steady_clock::now();
    int64_t l0 = 0, l1 = 0;
    ggml_backend_b200_get_stats(&l0, nullptr, nullptr);
    const ggml_status st = be->iface.graph_compute(be, &m->graph);
    const auto tp3 = std::chrono::steady_clock::now();
    m->t_build += std::chrono::duration<double>(tp1 - tp0).count();
    m->t_inputs += std::chrono::duration<double>(tp2 - tp1).count();
    m->t_compute += std::chrono::duration<double>(tp3 - tp2).count();
    m->t_calls++;
    ggml_backend_b200_get_stats(&l1, nullptr, nullptr);
    m->last_launches = l1 - l0;
    LL_REQUIRE(st == GGML_STATUS_SUCCESS, B200_ERR_CUDA, "b200_llama_decode: graph_compute failed (%d)", (int)st);
    if (logits) be->iface.get_tensor_async(be, m->logits_t, logits, 0, 4 * (size_t)m->hp.n_vocab * (size_t)n_outputs);
    return B200_OK;
}

}  // namespace

// Tensor parallel: an all-reduce whose peer never arrived gives up after its timeout and flags the window (comm.cu); the sums of
// that step are stale, so the step must not be reported as done.  One 4-byte read per synchronised step, tensor-parallel runs only.
static int peer_check(b200_llama *m) {
    if (m->hp.tp_world <= 1) return B200_OK;
    const int st = b200_comm_peer_status();
    LL_REQUIRE(st == 0, B200_ERR_CUDA, "b200_llama: a tensor-parallel peer did not arrive in time (peer status %d): the step's results are invalid", st);
    return B200_OK;
}

extern "C" int b200_llama_decode_async(b200_llama *m, const int32_t *tokens, int32_t n_tokens, int32_t n_past, float *logits, int32_t logits_all) {
    return enqueue_decode(m, tokens, n_tokens, n_past, logits, logits_all);
}
extern "C" int b200_llama_synchronize(b200_llama *m) {
    LL_REQUIRE(m && m->be, B200_ERR_INVALID, "b200_llama_synchronize: NULL model");
    m->be->iface.synchronize(m->be);
    return peer_check(m);
}
extern "C" int b200_llama_decode(b200_llama *m, const int32_t *tokens, int32_t n_tokens, int32_t n_past, float *logits, int32_t logits_all) {
    const int rc = enqueue_decode(m, tokens, n_tokens, n_past, logits, logits_all);
    if (rc != B200_OK) return rc;
    const auto ts0 = std::chrono::steady_clock::now();
    m->be->iface.synchronize(m->be);
    m->t_sync += std::chrono::duration<double>(std::chrono::steady_clock::now() - ts0).count();
    return peer_check(m);
}
extern "C" int b200_llama_replay(b200_llama *m) {
    LL_REQUIRE(m && m->finalized && m->graph.n_nodes > 0, B200_ERR_INVALID, "b200_llama_replay: nothing decoded yet");
    const ggml_status st = m->be->iface.graph_compute(m->be, &m->graph);
    LL_REQUIRE(st == GGML_STATUS_SUCCESS, B200_ERR_CUDA, "b200_llama_replay: graph_compute failed");
    return B200_OK;
}
extern "C" b200_stream_t b200_llama_stream(b200_llama *m) { return m && m->be ? ggml_backend_b200_get_stream(m->be) : nullptr; }
extern "C" int b200_llama_stats(b200_llama *m, int64_t *n_nodes, int64_t *n_kernels, int64_t *weight_bytes, int64_t *kv_bytes) {
    LL_REQUIRE(m, B200_ERR_INVALID, "b200_llama_stats: NULL model");
    if (n_nodes) *n_nodes = m->graph.n_nodes;
    if (n_kernels) *n_kernels = m->last_launches;
    if (weight_bytes) *weight_bytes = m->weight_bytes;
    if (kv_bytes) *kv_bytes = m->kv_bytes;
    return B200_OK;
}
