// backend_ops.cu -- per-node dispatch of the B200 ggml backend: supports_op (exact: graph_compute aborts on anything it
// said yes to and cannot run, like the reference, ggml-cuda.cu:2671-2675) and the node -> C-ABI kernel mapping.
// Replaces ggml_cuda_compute_forward / ggml_backend_cuda_device_supports_op (ggml-cuda.cu:2145-2340, 2959-3185).
#include "../../include/b200_comm.h"
#include "backend_exec.h"

namespace b200 {

static bool is_hot_weight_type(int t) {
    return t == GGML_TYPE_Q4_0 || t == GGML_TYPE_Q8_0 || t == GGML_TYPE_Q4_K || t == GGML_TYPE_Q5_K || t == GGML_TYPE_Q6_K ||
           t == GGML_TYPE_Q4_1 || t == GGML_TYPE_Q5_0 || t == GGML_TYPE_Q5_1 ||  // (N4: the legacy 32-value formats ...
           t == GGML_TYPE_Q2_K || t == GGML_TYPE_Q3_K ||                          //  ... the two small K-quants ...
           t == GGML_TYPE_IQ4_NL || t == GGML_TYPE_IQ4_XS;                        //  ... and the non-linear 4-bit formats)
}
static bool rows_contiguous_f32(const ggml_tensor *t) { return t->type == GGML_TYPE_F32 && t->nb[0] == 4; }
static bool in_split_buffer(const ggml_tensor *t) { return t->buffer && buft_is_b200_split(t->buffer->buft); }

// MUL_MAT with quantised weights: plain 2-D weight, activations [K, N] with contiguous 16-byte aligned rows
static bool mul_mat_q_ok(const ggml_tensor *w, const ggml_tensor *x, const ggml_tensor *dst) {
    if (!is_hot_weight_type(w->type) || x->type != GGML_TYPE_F32 || dst->type != GGML_TYPE_F32) return false;
    if (w->ne[2] != 1 || w->ne[3] != 1 || x->ne[2] != 1 || x->ne[3] != 1) return false;
    if (x->nb[0] != 4 || dst->nb[0] != 4 || (x->nb[1] & 15) || (dst->nb[1] & 3)) return false;
    if (w->nb[0] != (size_t)type_traits(w->type).size || w->nb[1] < row_size(w->type, w->ne[0])) return false;
    if ((w->type == GGML_TYPE_Q4_K || w->type == GGML_TYPE_Q5_K) && (w->nb[1] & 15)) return false;
    return w->ne[0] == x->ne[0] && dst->ne[0] == w->ne[1] && dst->ne[1] == x->ne[1];
}

// The 3-D / 4-D broadcast form (ggml.c:12486-12560; the reference loops i02 / i03 in ggml_cuda_op_mul_mat, ggml-cuda.cu:1358-1658):
// dst[:, :, i12, i13] = w[:, :, i12 / r2, i13 / r3] . x[:, :, i12, i13].  Every (i12, i13) slice must itself satisfy mul_mat_q_ok.
static bool mul_mat_q_bcast_ok(const ggml_tensor *w, const ggml_tensor *x, const ggml_tensor *dst) {
    if (!is_hot_weight_type(w->type) || x->type != GGML_TYPE_F32 || dst->type != GGML_TYPE_F32 || in_split_buffer(w)) return false;
    if (w->ne[2] < 1 || w->ne[3] < 1 || x->ne[2] % w->ne[2] || x->ne[3] % w->ne[3]) return false;
    if (dst->ne[2] != x->ne[2] || dst->ne[3] != x->ne[3] || x->ne[2] * x->ne[3] > 4096) return false;
    if (x->nb[0] != 4 || dst->nb[0] != 4 || ((x->nb[1] | x->nb[2] | x->nb[3]) & 15) || ((dst->nb[1] | dst->nb[2] | dst->nb[3]) & 3)) return false;
    if (w->nb[0] != (size_t)type_traits(w->type).size || w->nb[1] < row_size(w->type, w->ne[0])) return false;
    const size_t wal = (w->type == GGML_TYPE_Q4_K || w->type == GGML_TYPE_Q5_K) ? 15 : 1;  // slices keep the row alignment of the 2-D case
    if ((w->nb[1] | w->nb[2] | w->nb[3]) & wal) return false;
    return w->ne[0] == x->ne[0] && dst->ne[0] == w->ne[1] && dst->ne[1] == x->ne[1];
}

bool op_supported(int /*device*/, const ggml_tensor *op) {
    const ggml_tensor *s0 = op->src[0], *s1 = op->src[1];
    for (int i = 0; i < GGML_MAX_SRC; i++) {
        // split tensors can only be the weight of a MUL_MAT
        if (op->src[i] && in_split_buffer(op->src[i]) && !(op->op == GGML_OP_MUL_MAT && i == 0)) return false;
    }
    switch (op->op) {
        case GGML_OP_NONE: case GGML_OP_RESHAPE: case GGML_OP_VIEW: case GGML_OP_PERMUTE: case GGML_OP_TRANSPOSE:
            return true;
        case GGML_OP_MAP_CUSTOM1:  // a collective of the tensor-parallel decoder (b200_comm.h)
            if (!s0 || (uint32_t)op->op_params[0] != B200_COLLECTIVE_TAG || b200_comm_world() < 1) return false;
            if (s0->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32 || !is_contiguous(s0) || !is_contiguous(op)) return false;
            if (op->op_params[1] == B200_COLLECTIVE_ALL_REDUCE_SUM) return nelements(op) == nelements(s0);
            if (op->op_params[1] == B200_COLLECTIVE_ALL_GATHER) return nelements(op) == nelements(s0) * b200_comm_world();
            return false;
        case GGML_OP_MUL_MAT:
            if (!s0 || !s1) return false;
            if (is_quantized(s0->type)) return mul_mat_q_ok(s0, s1, op) || mul_mat_q_bcast_ok(s0, s1, op);
            return (s0->type == GGML_TYPE_F32 || s0->type == GGML_TYPE_F16) && s1->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 &&
                   s0->nb[0] == (size_t)type_traits(s0->type).size && s1->nb[0] == 4 && !in_split_buffer(s0);
        case GGML_OP_MUL_MAT_ID: {
            const ggml_tensor *ids = op->src[2];
            if (!s0 || !s1 || !ids) return false;
            if (!is_hot_weight_type(s0->type) || s1->type != GGML_TYPE_F32 || ids->type != GGML_TYPE_I32 || op->type != GGML_TYPE_F32) return false;
            if (s1->nb[0] != 4 || ((s1->nb[1] | s1->nb[2]) & 15) || op->nb[0] != 4 || ids->nb[0] != 4 || (ids->nb[1] & 3)) return false;
            if (s0->ne[3] != 1 || s1->ne[3] != 1) return false;
            if ((s0->type == GGML_TYPE_Q4_K || s0->type == GGML_TYPE_Q5_K) && ((s0->nb[1] | s0->nb[2]) & 15)) return false;
            if (ids->ne[0] * ids->ne[1] <= B200_GEMV_MAX_JOBS) return true;  // decode regime (T * n_used <= 16): expert GEMV
            // batches: device-side group-by-expert + one grouped tensor-core launch (b200_mul_mat_id_q)
            return s0->ne[0] % 256 == 0 && s0->ne[2] <= 256 && ids->ne[0] * ids->ne[1] <= 65535 && ids->ne[2] == 1 && s1->ne[2] == ids->ne[1] &&
                   (s1->ne[1] == 1 || s1->ne[1] == ids->ne[0]);
        }
        case GGML_OP_FLASH_ATTN_EXT: {
            const ggml_tensor *k = op->src[1], *v = op->src[2], *m = op->src[3];
            if (!s0 || !k || !v) return false;
            if (s0->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32 || s0->nb[0] != 4) return false;
            if (k->type != v->type || (k->type != GGML_TYPE_F16 && k->type != GGML_TYPE_Q8_0 && k->type != GGML_TYPE_Q4_0)) return false;
            {   // head sizes: 64 / 128 with either cache type; 80 / 96 / 112 / 256 with an f16 cache (ggml-cuda.cu:3162-3177)
                const int64_t D = s0->ne[0];
                const bool d_more = (D == 80 || D == 96 || D == 112 || D == 256) && k->type == GGML_TYPE_F16;
                if (D != 64 && D != 128 && !d_more) return false;
            }
            if (s0->ne[3] != 1 || k->ne[3] != 1 || v->ne[3] != 1 || k->ne[2] == 0 || s0->ne[2] % k->ne[2]) return false;
            if (m && (m->type != GGML_TYPE_F16 || m->nb[0] != 2)) return false;
            const size_t al = k->type == GGML_TYPE_F16 ? 15 : (s0->ne[0] == 128 ? 7 : 3);
            if ((k->nb[1] | k->nb[2] | v->nb[1] | v->nb[2]) & al) return false;
            return ((s0->nb[1] | s0->nb[2]) & 15) == 0 && is_contiguous(op);
        }
        case GGML_OP_RMS_NORM:
            return s0 && rows_contiguous_f32(s0) && rows_contiguous_f32(op);
        case GGML_OP_ADD: case GGML_OP_SUB: case GGML_OP_MUL: case GGML_OP_DIV:
            return s0 && s1 && s0->type == GGML_TYPE_F32 && s1->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32;
        case GGML_OP_SQR: case GGML_OP_SCALE: case GGML_OP_CLAMP:
            return s0 && s0->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && is_contiguous(s0) && is_contiguous(op);
        case GGML_OP_UNARY: {
            if (!s0 || s0->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32 || !is_contiguous(s0) || !is_contiguous(op)) return false;
            switch (op->op_params[0]) {
                case GGML_UNARY_OP_SILU: case GGML_UNARY_OP_GELU: case GGML_UNARY_OP_GELU_QUICK: case GGML_UNARY_OP_RELU:
                case GGML_UNARY_OP_TANH: case GGML_UNARY_OP_SIGMOID: return true;
                default: return false;
            }
        }
        case GGML_OP_ROPE: {
            if (!s0 || !s1 || s1->type != GGML_TYPE_I32 || (op->op_params[2] & ~GGML_ROPE_TYPE_NEOX) != 0 || (op->src[2] && op->src[2]->type != GGML_TYPE_F32)) return false;
            if (rows_contiguous_f32(s0) && rows_contiguous_f32(op)) return true;
            // f16: the in-place K-shift over an f16 KV cache (src/llama.cpp:10181-10186; reference rope.cu f16 path)
            return s0->type == GGML_TYPE_F16 && op->type == GGML_TYPE_F16 && s0->nb[0] == 2 && op->nb[0] == 2 && same_shape(s0, op) &&
                   ((s0->nb[1] | s0->nb[2] | s0->nb[3] | op->nb[1] | op->nb[2] | op->nb[3]) & 3) == 0;
        }
        case GGML_OP_CPY: case GGML_OP_DUP: case GGML_OP_CONT: {
            if (!s0) return false;
            const int a = s0->type, b = op->type;
            if (a == GGML_TYPE_F32 && (b == GGML_TYPE_Q8_0 || b == GGML_TYPE_Q4_0)) return s0->nb[0] == 4 && s0->ne[0] % 32 == 0 && op->ne[0] % 32 == 0 && ((s0->nb[1] | s0->nb[2] | s0->nb[3]) & 15) == 0;
            if ((a == GGML_TYPE_F32 || a == GGML_TYPE_F16) && (b == GGML_TYPE_F32 || b == GGML_TYPE_F16)) return true;
            if (a == GGML_TYPE_I32 && b == GGML_TYPE_I32) return true;
            // ggml_cast of a q8_0 K cache to f32 for the K-shift (src/llama.cpp:10165-10168; reference ggml-cuda.cu:3061)
            if (a == GGML_TYPE_Q8_0 && b == GGML_TYPE_F32) return s0->nb[0] == 34 && s0->ne[0] % 32 == 0 && op->nb[0] == 4 && ((s0->nb[1] | s0->nb[2] | s0->nb[3]) & 1) == 0;
            return a == b && is_contiguous(s0) && is_contiguous(op);
        }
        case GGML_OP_GET_ROWS:
            return s0 && s1 && s1->type == GGML_TYPE_I32 && op->type == GGML_TYPE_F32 &&
                   (s0->type == GGML_TYPE_F32 || s0->type == GGML_TYPE_F16 || is_hot_weight_type(s0->type));
        case GGML_OP_SOFT_MAX:
            return s0 && rows_contiguous_f32(s0) && rows_contiguous_f32(op) && s0->ne[0] <= 50 * 1024 &&
                   (!s1 || s1->type == GGML_TYPE_F16 || s1->type == GGML_TYPE_F32);
        case GGML_OP_ARGSORT:
            return s0 && s0->type == GGML_TYPE_F32 && is_contiguous(s0) && s0->ne[0] <= 1024;
        case GGML_OP_SUM_ROWS:
            return s0 && rows_contiguous_f32(s0);
        default:
            return false;
    }
}

// ------------------------------------------------------------------------------------------------ single-node execution
// ---- (a14) MUL_MAT with a row-split weight: every device computes its row slice, the main device gathers.
// Replaces the split branch of ggml_cuda_op_mul_mat (ggml-cuda.cu:1519-1657): activations are broadcast with peer copies,
// each device runs the same GEMV / tensor-core GEMM on its slice on its own stream, slices land directly in dst rows.
struct SplitPeer {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    float *x = nullptr; size_t x_bytes = 0;
    float *y = nullptr; size_t y_bytes = 0;
    void *ws = nullptr; size_t ws_bytes = 0;
};
static SplitPeer g_peer[GGML_B200_MAX_DEVICES];
static cudaEvent_t g_split_in_dev[64] = {};  // per main device: an event must be recorded on a stream of the device it was created on
static bool g_peer_enabled[GGML_B200_MAX_DEVICES][GGML_B200_MAX_DEVICES];

static void peer_prepare(int main_dev, int d, size_t xb, size_t yb, size_t wsb) {
    SplitPeer &p = g_peer[d];
    B200_CHECK_CUDA(cudaSetDevice(d));
    if (!p.stream) {
        B200_CHECK_CUDA(cudaStreamCreateWithFlags(&p.stream, cudaStreamNonBlocking));
        B200_CHECK_CUDA(cudaEventCreateWithFlags(&p.done, cudaEventDisableTiming));
    }
    if (!g_peer_enabled[d][main_dev]) {  // ggml_cuda_set_peer_access, ggml-cuda.cu:1286-1336
        int can = 0;
        B200_CHECK_CUDA(cudaDeviceCanAccessPeer(&can, d, main_dev));
        if (can) { cudaError_t e = cudaDeviceEnablePeerAccess(main_dev, 0); if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) backend_abort("cudaDeviceEnablePeerAccess", cudaGetErrorString(e)); cudaGetLastError(); }
        g_peer_enabled[d][main_dev] = true;
        B200_CHECK_CUDA(cudaSetDevice(main_dev));
        B200_CHECK_CUDA(cudaDeviceCanAccessPeer(&can, main_dev, d));
        if (can) { cudaError_t e = cudaDeviceEnablePeerAccess(d, 0); if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) backend_abort("cudaDeviceEnablePeerAccess", cudaGetErrorString(e)); cudaGetLastError(); }
        B200_CHECK_CUDA(cudaSetDevice(d));
    }
    auto grow = [&](void **ptr, size_t *have, size_t need) {
        if (need <= *have) return;
        B200_CHECK_CUDA(cudaStreamSynchronize(p.stream));
        if (*ptr) B200_CHECK_CUDA(cudaFree(*ptr));
        B200_CHECK_CUDA(cudaMalloc(ptr, need + 256));
        *have = need;
    };
    grow((void **)&p.x, &p.x_bytes, xb);
    grow((void **)&p.y, &p.y_bytes, yb);
    grow(&p.ws, &p.ws_bytes, wsb);
}

static void gemm_slice(int type, const void *W, size_t wrb, const float *X, int64_t xs, float *dst, int64_t ds, int64_t K, int64_t M, int64_t N,
                       void *ws, size_t wsb, cudaStream_t st, int64_t &launches) {
    if (N > 8 && ws) {
        const int rc = b200_mul_mat_q(type, W, wrb, X, xs, dst, ds, K, M, N, ws, wsb, st);
        if (rc == B200_OK) { launches += 2; return; }
        if (rc != B200_ERR_UNSUPPORTED) backend_abort("b200_mul_mat_q", b200_last_error());
    }
    for (int64_t c0 = 0; c0 < N; c0 += 8) {
        const int64_t nc = N - c0 < 8 ? N - c0 : 8;
        B200_CHECK_RC(b200_mul_mat_vec_q(type, W, wrb, X + c0 * xs, xs, dst + c0 * ds, ds, K, M, nc, st));
        launches++;
    }
}

static void run_mul_mat_split(ExecCtx &x, const ggml_tensor *n) {
    const ggml_tensor *w = n->src[0], *a = n->src[1];
    const SplitTensorExtra *e = (const SplitTensorExtra *)w->extra;
    const SplitBuftCtx *sc = (const SplitBuftCtx *)w->buffer->buft->context;
    const int64_t K = w->ne[0], N = a->ne[1];
    const int64_t xs = (int64_t)(a->nb[1] / 4), ds = (int64_t)(n->nb[1] / 4);
    const size_t wrb = row_size(w->type, K);
    if (x.device < 0 || x.device >= 64) backend_abort("run_mul_mat_split", "device ordinal out of range");
    cudaEvent_t &g_split_in = g_split_in_dev[x.device];
    if (!g_split_in) B200_CHECK_CUDA(cudaEventCreateWithFlags(&g_split_in, cudaEventDisableTiming));  // (the main device is current here)
    B200_CHECK_CUDA(cudaEventRecord(g_split_in, x.stream));  // activations are ready on the main device
    const size_t wsb_main = N > 8 ? b200_mul_mat_q_workspace_bytes(w->type, K, 1, N) : 0;
    for (int d = 0; d < sc->n_dev; d++) {
        const int64_t lo = e->row_low[d], nr = e->row_high[d] - lo;
        if (nr <= 0) continue;
        if (d == x.device) {
            gemm_slice(w->type, e->data[d], wrb, (const float *)a->data, xs, (float *)n->data + lo, ds, K, nr, N, wsb_main <= x.scratch_bytes ? x.scratch : nullptr,
                       x.scratch_bytes, x.stream, x.launches);
            continue;
        }
        peer_prepare(x.device, d, (size_t)N * K * 4, (size_t)N * nr * 4, wsb_main);
        SplitPeer &p = g_peer[d];
        B200_CHECK_CUDA(cudaStreamWaitEvent(p.stream, g_split_in, 0));
        // broadcast the activation rows (contiguous K floats each) and compute the slice on device d
        B200_CHECK_CUDA(cudaMemcpy2DAsync(p.x, (size_t)K * 4, a->data, a->nb[1], (size_t)K * 4, (size_t)N, cudaMemcpyDefault, p.stream));
        gemm_slice(w->type, e->data[d], wrb, p.x, K, p.y, nr, K, nr, N, p.ws, p.ws_bytes, p.stream, x.launches);
        // gather: slice rows [lo, lo+nr) of every column of dst on the main device
        B200_CHECK_CUDA(cudaMemcpy2DAsync((float *)n->data + lo, n->nb[1], p.y, (size_t)nr * 4, (size_t)nr * 4, (size_t)N, cudaMemcpyDefault, p.stream));
        B200_CHECK_CUDA(cudaEventRecord(p.done, p.stream));
    }
    B200_CHECK_CUDA(cudaSetDevice(x.device));
    for (int d = 0; d < sc->n_dev; d++) {
        if (d == x.device || e->row_high[d] - e->row_low[d] <= 0) continue;
        B200_CHECK_CUDA(cudaStreamWaitEvent(x.stream, g_peer[d].done, 0));
    }
}

static void run_mul_mat(ExecCtx &x, const ggml_tensor *n) {
    const ggml_tensor *w = n->src[0], *a = n->src[1];
    const void *xb_prev = x.xb_src;  // only the plain prefill GEMM below keeps the bf16 activations in the scratch meaningful
    x.xb_src = nullptr;
    if (in_split_buffer(w)) { run_mul_mat_split(x, n); return; }
    if (!is_quantized(w->type)) {
        b200_tensor tw = as_b200(w), ta = as_b200(a), td = as_b200(n);
        B200_CHECK_RC(b200_op_mul_mat_float(&tw, &ta, &td, x.stream));
        x.launches++;
        return;
    }
    const int64_t K = w->ne[0], M = w->ne[1], N = a->ne[1];
    const int64_t xs = (int64_t)(a->nb[1] / 4), ds = (int64_t)(n->nb[1] / 4);
    if (a->ne[2] != 1 || a->ne[3] != 1 || w->ne[2] != 1 || w->ne[3] != 1) {
        // broadcast over dims 2 and 3: one GEMM / GEMV per (i12, i13) slice, the weight slice shared by r2 x r3 of them
        const int64_t r2 = a->ne[2] / w->ne[2], r3 = a->ne[3] / w->ne[3];
        const size_t wsb = N > 8 ? b200_mul_mat_q_workspace_bytes(w->type, K, M, N) : 0;
        for (int64_t i13 = 0; i13 < a->ne[3]; i13++) {
            for (int64_t i12 = 0; i12 < a->ne[2]; i12++) {
                const char *ws = (const char *)w->data + (i12 / r2) * w->nb[2] + (i13 / r3) * w->nb[3];
                const float *xp = (const float *)((const char *)a->data + i12 * a->nb[2] + i13 * a->nb[3]);
                float *dp = (float *)((char *)n->data + i12 * n->nb[2] + i13 * n->nb[3]);
                gemm_slice(w->type, ws, w->nb[1], xp, xs, dp, ds, K, M, N, wsb <= x.scratch_bytes ? x.scratch : nullptr, x.scratch_bytes, x.stream, x.launches);
            }
        }
        return;
    }
    // N <= 8: the HBM-bound GEMV.  Larger batches: tensor-core GEMM when built, else the GEMV over column groups of 8
    // (every group re-streams the weights; correct, and only the stop-gap for prefill until b200_mul_mat_q lands)
    if (N > 8) {
        const size_t ws = b200_mul_mat_q_workspace_bytes(w->type, K, M, N);
        if (ws <= x.scratch_bytes) {
            // Q / K / V and gate / up read one activation: the first of the group converts it to bf16, the others find it in the scratch
            const bool have_xb = xb_prev && xb_prev == a->data && x.xb_k == K && x.xb_n == N && x.xb_stride == xs;
            const int rc = b200_mul_mat_q(w->type, w->data, w->nb[1], have_xb ? nullptr : (const float *)a->data, xs, (float *)n->data, ds, K, M, N, x.scratch,
                                          x.scratch_bytes, x.stream);
            if (rc == B200_OK) {
                x.launches++;
                x.xb_src = a->data; x.xb_k = K; x.xb_n = N; x.xb_stride = xs;
                return;
            }
            if (rc != B200_ERR_UNSUPPORTED) backend_abort("b200_mul_mat_q", b200_last_error());
        }
    }
    for (int64_t c0 = 0; c0 < N; c0 += 8) {
        const int64_t nc = N - c0 < 8 ? N - c0 : 8;
        B200_CHECK_RC(b200_mul_mat_vec_q(w->type, w->data, w->nb[1], (const float *)a->data + c0 * xs, xs, (float *)n->data + c0 * ds, ds, K, M, nc, x.stream));
        x.launches++;
    }
}

static void run_mul_mat_id(ExecCtx &x, const ggml_tensor *n) {
    const ggml_tensor *as = n->src[0], *b = n->src[1], *ids = n->src[2];
    // as [K, M, n_expert]; b [K, n_b_rows, T]; ids [n_used, T]; dst [M, n_used, T]   (ggml.c:5693-5729)
    if (ids->ne[0] * ids->ne[1] > B200_GEMV_MAX_JOBS) {
        const size_t need = b200_mul_mat_id_q_workspace_bytes(as->type, as->ne[0], as->ne[1], ids->ne[0], ids->ne[1], (int32_t)as->ne[2]);
        if (need > x.scratch_bytes) backend_abort("mul_mat_id workspace", "workspace smaller than planned");
        B200_CHECK_RC(b200_mul_mat_id_q(as->type, as->data, as->nb[1], as->nb[2], (int32_t)as->ne[2], (const int32_t *)ids->data, (int64_t)(ids->nb[1] / 4),
                                        (const float *)b->data, (int64_t)(b->nb[1] / 4), (int64_t)(b->nb[2] / 4), (float *)n->data, (int64_t)(n->nb[1] / 4),
                                        (int64_t)(n->nb[2] / 4), as->ne[0], as->ne[1], ids->ne[0], b->ne[1], ids->ne[1], x.scratch, x.scratch_bytes, x.stream));
        x.launches += 3;
        return;
    }
    B200_CHECK_RC(b200_mul_mat_id_vec_q(as->type, as->data, as->nb[1], as->nb[2], (int32_t)as->ne[2], (const int32_t *)ids->data, (int64_t)(ids->nb[1] / 4),
                                        (const float *)b->data, (int64_t)(b->nb[1] / 4), (int64_t)(b->nb[2] / 4), (float *)n->data, (int64_t)(n->nb[1] / 4),
                                        (int64_t)(n->nb[2] / 4), as->ne[0], as->ne[1], ids->ne[0], b->ne[1], ids->ne[1], x.stream));
    x.launches++;
}

static void run_flash_attn(ExecCtx &x, const ggml_tensor *n) {
    const ggml_tensor *q = n->src[0], *k = n->src[1], *v = n->src[2], *m = n->src[3];
    b200_fattn_args a;
    memset(&a, 0, sizeof(a));
    a.q = q->data; a.k = k->data; a.v = v->data; a.mask = m ? m->data : nullptr; a.dst = (float *)n->data;
    a.D = q->ne[0]; a.N = q->ne[1]; a.n_head = q->ne[2]; a.n_head_kv = k->ne[2]; a.n_kv = k->ne[1];
    a.type_k = k->type; a.type_v = v->type;
    a.q_nb1 = q->nb[1]; a.q_nb2 = q->nb[2]; a.k_nb1 = k->nb[1]; a.k_nb2 = k->nb[2]; a.v_nb1 = v->nb[1]; a.v_nb2 = v->nb[2];
    a.mask_nb1 = m ? m->nb[1] : 0;
    memcpy(&a.scale, (const float *)n->op_params + 0, 4);
    memcpy(&a.max_bias, (const float *)n->op_params + 1, 4);
    memcpy(&a.logit_softcap, (const float *)n->op_params + 2, 4);
    const size_t need = b200_flash_attn_workspace_bytes(&a);
    if (need > x.fa_ws_bytes) backend_abort("flash_attn workspace", "workspace smaller than planned");
    B200_CHECK_RC(b200_flash_attn_ext(&a, x.fa_ws, x.fa_ws_bytes, x.stream));
    x.launches++;
}

size_t node_fa_workspace(const ggml_tensor *n) {
    if (n->op != GGML_OP_FLASH_ATTN_EXT) return 0;
    const ggml_tensor *q = n->src[0], *k = n->src[1], *v = n->src[2];
    b200_fattn_args a;
    memset(&a, 0, sizeof(a));
    a.q = q->data; a.k = k->data; a.v = v->data; a.dst = (float *)n->data;
    a.D = q->ne[0]; a.N = q->ne[1]; a.n_head = q->ne[2]; a.n_head_kv = k->ne[2]; a.n_kv = k->ne[1];
    a.type_k = k->type; a.type_v = v->type;
    a.q_nb1 = q->nb[1]; a.q_nb2 = q->nb[2]; a.k_nb1 = k->nb[1]; a.k_nb2 = k->nb[2]; a.v_nb1 = v->nb[1]; a.v_nb2 = v->nb[2];
    return b200_flash_attn_workspace_bytes(&a);
}

size_t node_scratch_bytes(const ggml_tensor *n) {
    if (n->op == GGML_OP_MUL_MAT_ID && n->src[0] && n->src[2] && n->src[2]->ne[0] * n->src[2]->ne[1] > B200_GEMV_MAX_JOBS)
        return b200_mul_mat_id_q_workspace_bytes(n->src[0]->type, n->src[0]->ne[0], n->src[0]->ne[1], n->src[2]->ne[0], n->src[2]->ne[1], (int32_t)n->src[0]->ne[2]);
    if (n->op != GGML_OP_MUL_MAT || !n->src[0] || !n->src[1] || !is_quantized(n->src[0]->type) || n->src[1]->ne[1] <= 8) return 0;
    return b200_mul_mat_q_workspace_bytes(n->src[0]->type, n->src[0]->ne[0], n->src[0]->ne[1], n->src[1]->ne[1]);
}

static int unary_code(int ggml_unary) {
    switch (ggml_unary) {
        case GGML_UNARY_OP_SILU: return B200_UN_SILU;
        case GGML_UNARY_OP_GELU: return B200_UN_GELU;
        case GGML_UNARY_OP_GELU_QUICK: return B200_UN_GELU_QUICK;
        case GGML_UNARY_OP_RELU: return B200_UN_RELU;
        case GGML_UNARY_OP_TANH: return B200_UN_TANH;
        case GGML_UNARY_OP_SIGMOID: return B200_UN_SIGMOID;
    }
    return -1;
}

bool peer_all_reduce_ok(const ggml_tensor *n) {
    const ggml_tensor *s0 = n->src[0];
    return n->op == GGML_OP_MAP_CUSTOM1 && (uint32_t)n->op_params[0] == B200_COLLECTIVE_TAG && n->op_params[1] == B200_COLLECTIVE_ALL_REDUCE_SUM &&
           b200_comm_peer_ready() && s0 && s0->type == GGML_TYPE_F32 && n->type == GGML_TYPE_F32 && is_contiguous(s0) && is_contiguous(n) &&
           nelements(n) <= B200_COMM_PEER_MAX_FLOATS && nelements(n) % 4 == 0 && (((uintptr_t)s0->data | (uintptr_t)n->data) & 15) == 0;
}

void run_node(ExecCtx &x, const ggml_tensor *n, void *const *dst_slot) {
    const ggml_tensor *s0 = n->src[0], *s1 = n->src[1];
    switch (n->op) {
        case GGML_OP_NONE: case GGML_OP_RESHAPE: case GGML_OP_VIEW: case GGML_OP_PERMUTE: case GGML_OP_TRANSPOSE:
            return;
        case GGML_OP_MAP_CUSTOM1: {
            if (n->op_params[1] == B200_COLLECTIVE_ALL_REDUCE_SUM) {
                if (peer_all_reduce_ok(n)) {  // decode-size message: one kernel over NVLink peer memory (comm.cu)
                    B200_CHECK_RC(b200_all_reduce_sum_add_f32((const float *)s0->data, nullptr, (float *)n->data, nelements(n), x.stream));
                    break;
                }
                if (n->data != s0->data) B200_CHECK_CUDA(cudaMemcpyAsync(n->data, s0->data, nbytes(s0), cudaMemcpyDeviceToDevice, x.stream));
                B200_CHECK_RC(b200_all_reduce_sum_f32((float *)n->data, nelements(n), x.stream));
            } else {
                B200_CHECK_RC(b200_all_gather_f32((const float *)s0->data, (float *)n->data, nelements(s0), x.stream));
            }
            break;
        }
        case GGML_OP_MUL_MAT: run_mul_mat(x, n); return;
        case GGML_OP_MUL_MAT_ID: run_mul_mat_id(x, n); return;
        case GGML_OP_FLASH_ATTN_EXT: run_flash_attn(x, n); return;
        case GGML_OP_RMS_NORM: {
            float eps;
            memcpy(&eps, n->op_params, 4);
            b200_tensor a = as_b200(s0), d = as_b200(n);
            B200_CHECK_RC(b200_op_rms_norm(&a, nullptr, &d, eps, x.stream));
            break;
        }
        case GGML_OP_ADD: case GGML_OP_SUB: case GGML_OP_MUL: case GGML_OP_DIV: {
            const int op = n->op == GGML_OP_ADD ? B200_BIN_ADD : n->op == GGML_OP_SUB ? B200_BIN_SUB : n->op == GGML_OP_MUL ? B200_BIN_MUL : B200_BIN_DIV;
            b200_tensor a = as_b200(s0), b = as_b200(s1), d = as_b200(n);
            B200_CHECK_RC(b200_op_binary(op, &a, &b, &d, x.stream));
            break;
        }
        case GGML_OP_SQR: case GGML_OP_SCALE: case GGML_OP_CLAMP: case GGML_OP_UNARY: {
            float p0 = 0, p1 = 0;
            int op;
            if (n->op == GGML_OP_SQR) op = B200_UN_SQR;
            else if (n->op == GGML_OP_SCALE) { op = B200_UN_SCALE; memcpy(&p0, n->op_params, 4); }
            else if (n->op == GGML_OP_CLAMP) { op = B200_UN_CLAMP; memcpy(&p0, n->op_params, 4); memcpy(&p1, (const float *)n->op_params + 1, 4); }
            else op = unary_code(n->op_params[0]);
            b200_tensor a = as_b200(s0), d = as_b200(n);
            B200_CHECK_RC(b200_op_unary(op, &a, &d, p0, p1, x.stream));
            break;
        }
        case GGML_OP_ROPE: {
            b200_rope_params rp;
            rp.n_dims = n->op_params[1]; rp.mode = n->op_params[2]; rp.n_ctx_orig = n->op_params[4];
            memcpy(&rp.freq_base, n->op_params + 5, 4); memcpy(&rp.freq_scale, n->op_params + 6, 4); memcpy(&rp.ext_factor, n->op_params + 7, 4);
            memcpy(&rp.attn_factor, n->op_params + 8, 4); memcpy(&rp.beta_fast, n->op_params + 9, 4); memcpy(&rp.beta_slow, n->op_params + 10, 4);
            b200_tensor a = as_b200(s0), d = as_b200(n);
            B200_CHECK_RC(b200_op_rope(&a, (const int32_t *)s1->data, n->src[2] ? (const float *)n->src[2]->data : nullptr, &d, &rp, x.stream));
            break;
        }
        case GGML_OP_CPY: case GGML_OP_DUP: case GGML_OP_CONT: {
            // CPY(a, b): the result aliases b (ggml.c:5982-6005): node->data == src[1]->data
            b200_tensor a = as_b200(s0), d = as_b200(n);
            B200_CHECK_RC(b200_op_cpy_indirect(&a, &d, dst_slot, x.stream));
            break;
        }
        case GGML_OP_GET_ROWS: {
            b200_tensor a = as_b200(s0), i = as_b200(s1), d = as_b200(n);
            B200_CHECK_RC(b200_op_get_rows(&a, &i, &d, x.stream));
            break;
        }
        case GGML_OP_SOFT_MAX: {
            float scale = 1.0f, max_bias = 0.0f;
            memcpy(&scale, n->op_params, 4);
            memcpy(&max_bias, (const float *)n->op_params + 1, 4);
            b200_tensor a = as_b200(s0), d = as_b200(n), m;
            if (s1) m = as_b200(s1);
            B200_CHECK_RC(b200_op_soft_max(&a, s1 ? &m : nullptr, &d, scale, max_bias, x.stream));
            break;
        }
        case GGML_OP_ARGSORT: {
            b200_tensor a = as_b200(s0), d = as_b200(n);
            B200_CHECK_RC(b200_op_argsort(&a, &d, n->op_params[0], x.stream));
            break;
        }
        case GGML_OP_SUM_ROWS: {
            b200_tensor a = as_b200(s0), d = as_b200(n);
            B200_CHECK_RC(b200_op_sum_rows(&a, &d, x.stream));
            break;
        }
        default:
            backend_abort("graph_compute", "op not supported by the B200 backend (supports_op must have said so)");
    }
    x.launches++;
}

bool mul_mat_fusable(const ggml_tensor *n) {
    return n->op == GGML_OP_MUL_MAT && n->src[0] && n->src[1] && is_quantized(n->src[0]->type) && mul_mat_q_ok(n->src[0], n->src[1], n) &&
           n->src[1]->ne[1] <= 4 && !in_split_buffer(n->src[0]);
}

}  // namespace b200
