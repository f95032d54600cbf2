/* b200_ggml_ops.h -- C ABI of the hot-path-adjacent ops: everything else a llama / mixtral
 * graph contains, so that a whole decode graph stays on the device (no CPU fallback, CUDA-graph capturable).
 * Tensors are described the way ggml does: element counts ne[4] and BYTE strides nb[4] (ggml.h:584-619).
 * Each entry cites the reference CUDA kernel it replaces and the CPU function that defines its semantics
 * (paths under /root/reference/ggml/src/).
 */
#ifndef B200_GGML_OPS_H
#define B200_GGML_OPS_H

#include "b200_ggml.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    void   *data;
    int32_t type; /* enum b200_type / ggml_type */
    int32_t _pad;
    int64_t ne[4];
    size_t  nb[4];
} b200_tensor;

enum b200_binary_op { B200_BIN_ADD = 0, B200_BIN_SUB = 1, B200_BIN_MUL = 2, B200_BIN_DIV = 3 };
enum b200_unary_op {
    B200_UN_SILU = 0, B200_UN_GELU = 1, B200_UN_GELU_QUICK = 2, B200_UN_RELU = 3, B200_UN_TANH = 4, B200_UN_SIGMOID = 5,
    B200_UN_SQR = 6, B200_UN_SCALE = 7 /* p0 = factor */, B200_UN_CLAMP = 8 /* p0 = min, p1 = max */,
};

/* RMS_NORM (+ optional fused MUL by a per-column weight): y = x / sqrt(mean(x^2) + eps) [* w]
 * replaces rms_norm_f32, ggml-cuda/norm.cu:101-140; semantics ggml_compute_forward_rms_norm_f32, ggml.c:12060-12110.
 * x, dst: f32, rows contiguous (nb[0] == 4).  w: f32 [ne0] or NULL. */
B200_API int b200_op_rms_norm(const b200_tensor *x, const float *w, const b200_tensor *dst, float eps, b200_stream_t stream);

/* Fused RMS_NORM -> MUL(weight) pair as llama graphs emit it (src/llama.cpp:9130-9160 llm_build_norm): dst = norm(x) * w;
 * norm_out (optional, contiguous rows) also receives norm(x), the intermediate node's value. */
B200_API int b200_op_rms_norm_mul(const b200_tensor *x, const float *w, float *norm_out, const b200_tensor *dst, float eps, b200_stream_t stream);
/* Fused UNARY(SILU) -> MUL pair of the gated FFN (src/llama.cpp:9416-9515 llm_build_ffn, LLM_FFN_SILU + LLM_FFN_PAR):
 * dst = silu(gate) * up over n contiguous floats; act_out (optional) also receives silu(gate). */
B200_API int b200_op_silu_mul(const float *gate, const float *up, float *act_out, float *dst, int64_t n, b200_stream_t stream);

/* ADD / SUB / MUL / DIV with ggml broadcasting of src1 over src0 (ggml_can_repeat): f32 only.
 * replaces k_bin_bcast, ggml-cuda/binbcast.cu:26-120; semantics ggml_compute_forward_{add,sub,mul,div}_f32, ggml.c:10241-10900 */
B200_API int b200_op_binary(int op, const b200_tensor *a, const b200_tensor *b, const b200_tensor *dst, b200_stream_t stream);

/* elementwise f32 -> f32 on a contiguous tensor.  replaces ggml-cuda/unary.cu, scale.cu, clamp.cu */
B200_API int b200_op_unary(int op, const b200_tensor *x, const b200_tensor *dst, float p0, float p1, b200_stream_t stream);

/* ROPE (normal and NeoX ordering, YaRN extrapolation mix, optional frequency factors), f32.
 * replaces rope_norm / rope_neox, ggml-cuda/rope.cu:32-120; semantics ggml_compute_forward_rope_f32, ggml.c:14272-14400.
 * x: [ne0 = head dim, ne1 = heads, ne2 = tokens]; pos: i32 [ne2]; freq_factors: f32 [n_dims/2] or NULL. */
typedef struct {
    int32_t n_dims, mode, n_ctx_orig;
    float   freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow;
} b200_rope_params;
B200_API int b200_op_rope(const b200_tensor *x, const int32_t *pos, const float *freq_factors, const b200_tensor *dst,
                          const b200_rope_params *rp, b200_stream_t stream);

/* CPY / CONT / DUP: same element count, arbitrary strides; f32->f32, f32->f16, f16->f16, f16->f32, f32->q8_0 (the KV-cache
 * store; quantize_row_q8_0_ref arithmetic, ggml-quants.c:849-872).  For q8_0 destinations src rows must be contiguous.
 * replaces ggml-cuda/cpy.cu:34-440 */
B200_API int b200_op_cpy(const b200_tensor *src, const b200_tensor *dst, b200_stream_t stream);
/* Same, but the destination base pointer is read on the device from *dst_slot when the kernel runs (dst->data is ignored
 * when dst_slot != NULL).  Used for the KV-cache store inside a replayed CUDA graph: the write position moves every token
 * (reference: cudaGraphExecKernelNodeSetParams patching of the cpy nodes, ggml-cuda.cu:2593-2650). */
B200_API int b200_op_cpy_indirect(const b200_tensor *src, const b200_tensor *dst, void *const *dst_slot, b200_stream_t stream);

/* The four nodes llm_build_kv emits per layer in one launch (src/llama.cpp:9161-9203, 10504-10523): q_dst = ROPE(q),
 * k_dst = ROPE(k), cache_k[new rows] = CPY(k_dst), cache_v[new rows] = CPY(v).  q [D, n_head, T], k [D, n_head_kv, T],
 * v [D*n_head_kv, T]: contiguous f32.  kv_type: F16 or Q8_0; the T new cache rows are contiguous at k_cache / v_cache, or at
 * *k_slot / *v_slot read on the device (CUDA-graph replay with a moving write position).  NORM rope ordering only. */
B200_API int b200_op_rope_kv_store(const b200_tensor *q, const b200_tensor *q_dst, const b200_tensor *k, const b200_tensor *k_dst,
                                   const b200_tensor *v, const int32_t *pos, const float *freq_factors, const b200_rope_params *rp,
                                   int kv_type, void *k_cache, void *v_cache, void *const *k_slot, void *const *v_slot, b200_stream_t stream);

/* GET_ROWS: dst[i] = dequantise(src row ids[i]); src f32 / f16 / q4_0 / q8_0 / q4_K / q5_K / q6_K (K-quant token embeddings
 * stay on the device -- the reference falls back to the CPU for them, ggml-cuda.cu:3033-3047).
 * replaces k_get_rows / k_get_rows_float, ggml-cuda/getrows.cu:5-120 */
B200_API int b200_op_get_rows(const b200_tensor *src, const b200_tensor *ids, const b200_tensor *dst, b200_stream_t stream);

/* SOFT_MAX over ne0 with optional f16/f32 mask [ne0, >= ne1] broadcast over heads, scale and ALiBi max_bias.
 * replaces soft_max_f32, ggml-cuda/softmax.cu:15-120; semantics ggml_compute_forward_soft_max_f32, ggml.c:13900-14005 */
B200_API int b200_op_soft_max(const b200_tensor *x, const b200_tensor *mask, const b200_tensor *dst, float scale, float max_bias,
                              b200_stream_t stream);

/* ARGSORT of every row (ne0 <= 1024), i32 indices out; order: 0 ascending, 1 descending (ties keep index order).
 * replaces k_argsort_f32_i32, ggml-cuda/argsort.cu:11-100 */
/* The router of llm_build_moe_ffn (src/llama.cpp:9416-9470) in one launch, replacing SOFT_MAX -> ARGSORT(desc) + VIEW(top n_used) ->
 * GET_ROWS -> SUM_ROWS -> DIV: per token, probs = softmax(logits[n_expert]); ids[k] = index of the k-th largest prob (ties: lowest
 * index); w_gather[k] = probs[ids[k]]; wsum = sum_k w_gather[k]; w_out[k] = w_gather[k] / wsum.  ld_* are row strides in elements;
 * w_gather / w_out are [n_tok][n_used] contiguous (they may alias); n_expert <= 256, n_used <= 32. */
B200_API int b200_op_moe_router(const float *logits, int64_t ld_logits, float *probs, int64_t ld_probs, int32_t *ids, int64_t ld_ids, float *w_gather,
                                float *wsum, float *w_out, int32_t n_expert, int32_t n_used, int64_t n_tok, b200_stream_t stream);
B200_API int b200_op_argsort(const b200_tensor *x, const b200_tensor *dst, int order, b200_stream_t stream);

/* SUM_ROWS: dst[row] = sum over ne0.  replaces k_sum_rows_f32, ggml-cuda/sumrows.cu:3-30 */
B200_API int b200_op_sum_rows(const b200_tensor *x, const b200_tensor *dst, b200_stream_t stream);

/* MUL_MAT with an f32 or f16 weight matrix (MoE router, small dense layers): dst[n][m] = sum_k W[m][k] X[n][k], with ggml's
 * broadcast over dims 2 and 3.  A bandwidth-oriented warp-per-output kernel; the quantised hot path lives in b200_ggml.h.
 * replaces the cuBLAS / mul_mat_vec_nc paths for float weights, ggml-cuda.cu:1042-1284 */
B200_API int b200_op_mul_mat_float(const b200_tensor *w, const b200_tensor *x, const b200_tensor *dst, b200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif
