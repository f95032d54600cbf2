/* b200_ggml.h -- kernel-level C ABI of the B200-native GGML hot path (libkcpp_b200.so).
 *
 * Plain C: raw device pointers, sizes, strides and a CUDA stream handle.  No torch / C++ types.
 * Every entry point names the reference interface it replaces (paths under /root/reference/).
 * The ggml_backend-level drop-in boundary (reg / device / buffer type / buffer / backend vtables)
 * is declared in b200_ggml_backend.h and is built on top of these calls.
 *
 * Conventions
 *   - ggml shape notation: MUL_MAT(W, X): W = [K, M] (K contiguous, quantised along K),
 *     X = [K, N] f32, dst = [M, N] f32 (ggml/src/ggml.c:5664-5679).
 *   - All functions return B200_OK (0) or a b200_status error; they never abort the process and
 *     never allocate device memory behind the caller's back (a workspace is passed where needed).
 *     b200_last_error() returns a thread-local description of the last failure.
 *   - `stream` is a cudaStream_t (NULL = the legacy default stream).  All kernel entry points are
 *     asynchronous with respect to the host.
 *   - There is NO CPU fallback: on a machine without a CUDA device every compute entry point
 *     returns B200_ERR_CUDA.
 */
#ifndef B200_GGML_H
#define B200_GGML_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B200_API __attribute__((visibility("default")))
#else
#define B200_API
#endif

typedef void *b200_stream_t; /* cudaStream_t */

enum b200_status {
    B200_OK              = 0,
    B200_ERR_UNSUPPORTED = 1, /* valid ggml op, but a type/shape/stride this build has no kernel for */
    B200_ERR_INVALID     = 2, /* malformed arguments (NULL pointers, K not a multiple of the block size, ...) */
    B200_ERR_CUDA        = 3, /* CUDA runtime failure (see b200_last_error) */
    B200_ERR_NOMEM       = 4,
};

/* enum ggml_type values (ggml/include/ggml.h:361-398) */
enum b200_type {
    B200_TYPE_F32  = 0,
    B200_TYPE_F16  = 1,
    B200_TYPE_Q4_0 = 2,
    B200_TYPE_Q4_1 = 3,
    B200_TYPE_Q5_0 = 6,
    B200_TYPE_Q5_1 = 7,
    B200_TYPE_Q8_0 = 8,
    B200_TYPE_Q8_1 = 9,
    B200_TYPE_Q2_K = 10,
    B200_TYPE_Q3_K = 11,
    B200_TYPE_Q4_K = 12,
    B200_TYPE_Q5_K = 13,
    B200_TYPE_Q6_K = 14,
    B200_TYPE_IQ4_NL = 20,
    B200_TYPE_IQ4_XS = 23,
    B200_TYPE_I32  = 26,
};

B200_API const char *b200_last_error(void);
B200_API const char *b200_version(void);
/* Launch the kernels with the programmatic-dependent-launch attribute (griddepcontrol): a kernel's weight
 * prefetch then overlaps the tail of its producer.  Off by default; the ggml backend turns it on for its graphs. */
B200_API void b200_set_pdl(int enabled);

/* ---- device plumbing (so ctypes / cgo / JNI hosts need no CUDA binding of their own) ------------
 * replaces: ggml_backend_cuda_get_device_count / cudaMalloc / cudaMemcpyAsync call sites in
 * ggml/src/ggml-cuda.cu:3262-3270, 546-560, 471-485 */
B200_API int b200_device_count(void);
B200_API int b200_set_device(int device);
B200_API int b200_device_info(int device, char *name, size_t name_len, size_t *free_bytes, size_t *total_bytes,
                              int *sm_count, int *cc_major, int *cc_minor);
B200_API int b200_malloc(void **ptr, size_t bytes);
B200_API int b200_free(void *ptr);
B200_API int b200_malloc_host(void **ptr, size_t bytes); /* pinned */
B200_API int b200_free_host(void *ptr);
B200_API int b200_memcpy_h2d(void *dst, const void *src, size_t bytes, b200_stream_t stream);
B200_API int b200_memcpy_d2h(void *dst, const void *src, size_t bytes, b200_stream_t stream);
B200_API int b200_memcpy_d2d(void *dst, const void *src, size_t bytes, b200_stream_t stream);
B200_API int b200_memset(void *dst, int value, size_t bytes, b200_stream_t stream);
B200_API int b200_stream_create(b200_stream_t *stream);
B200_API int b200_stream_destroy(b200_stream_t stream);
B200_API int b200_stream_synchronize(b200_stream_t stream);
B200_API int b200_device_synchronize(void);
/* CUDA-event timing on `stream` (used by bench.py; torch.cuda.Event only sees torch's stream) */
B200_API int b200_event_create(void **event);
B200_API int b200_event_destroy(void *event);
B200_API int b200_event_record(void *event, b200_stream_t stream);
B200_API int b200_event_elapsed_ms(void *start, void *stop, float *ms); /* synchronises on `stop` */
/* Capture the launches issued on `stream` (a stream from b200_stream_create, not NULL) into an executable CUDA graph and
 * replay it: what ggml_backend_cuda_graph_compute does with cudaStreamBeginCapture / cudaGraphLaunch, ggml-cuda.cu:2700-2800 */
B200_API int b200_graph_begin_capture(b200_stream_t stream);
B200_API int b200_graph_end_capture(b200_stream_t stream, void **graph_exec);
B200_API int b200_graph_launch(void *graph_exec, b200_stream_t stream);
B200_API int b200_graph_destroy(void *graph_exec);

/* ---- (a2) block dequantisation ------------------------------------------------------------------
 * replaces: dequantize_block_* / to_fp32_cuda / to_fp16_cuda, ggml/src/ggml-cuda/convert.cu:7-445, 460-540
 * semantics: bit-exact with dequantize_row_{q4_0,q8_0,q4_K,q5_K,q6_K} (ggml/src/ggml-quants.c:1523, 1617,
 * 2556, 2764, 2978) for dst_type F32; for F16 the f32 result is rounded once (RNE).
 * n = number of values (multiple of the block size); src blocks contiguous. */
B200_API int b200_dequantize(int type, const void *src, void *dst, int dst_type, int64_t n, b200_stream_t stream);

/* ---- (a3) activation quantisation f32 -> block_q8_1 -----------------------------------------------
 * replaces: quantize_row_q8_1_cuda, ggml/src/ggml-cuda/quantize.cu:4-38, 129-141
 * semantics: bit-exact with the CPU definition quantize_row_q8_1_ref (ggml-quants.c:1160-1193):
 *   d = amax/127, id = d ? 1/d : 0, q = roundf(x*id), y.d = fp16(d), y.s = fp16(sum(q)*d).
 * x: N rows of K floats, row stride x_stride (in floats); y: N*K/32 contiguous 36-byte blocks. K % 32 == 0. */
B200_API int b200_quantize_q8_1(const float *x, int64_t x_stride, void *y, int64_t K, int64_t N, b200_stream_t stream);

/* ---- (a6) decode GEMV: dst[n][m] = sum_k W[m][k] * X[n][k], N <= 8 ---------------------------------
 * replaces: ggml_cuda_op_mul_mat_vec_q + mul_mat_vec_q<type,ncols_y> + the separate quantize_q8_1 launch,
 * ggml/src/ggml-cuda/mmvq.cu:50-202, 337-425; ggml-cuda.cu:1940-1941
 * W: M rows, row stride w_row_bytes (>= K/blck*block_bytes), type in {Q4_0,Q8_0,Q4_K,Q5_K,Q6_K};
 * X: f32, N rows of K (stride x_stride floats); dst: f32, N rows of M (stride dst_stride floats).
 * The activation is quantised to q8_1 inside the kernel (same arithmetic as b200_quantize_q8_1).
 * K must be a multiple of 256 for K-quants and of 32 for Q4_0/Q8_0.  W must be 16-byte aligned and
 * readable up to the next 16-byte boundary past its last row (true for any cudaMalloc'ed tensor). */
B200_API int b200_mul_mat_vec_q(int type, const void *W, size_t w_row_bytes, const float *X, int64_t x_stride,
                                float *dst, int64_t dst_stride, int64_t K, int64_t M, int64_t N, b200_stream_t stream);

/* Several GEMVs that share a launch (fused Q/K/V or gate/up): job j computes
 * dst_j[n][m] = W_j[m][:] . X_j[n][:] with its own M_j; all jobs share type, K and N.
 * When ids != NULL this is MUL_MAT_ID (next entry): W_j = as + ids[id_index_j] * expert_bytes, read on the
 * device -- no host synchronisation. */
typedef struct {
    const void    *W;
    const float   *X;
    float         *dst;
    int64_t        M;
    int32_t        id_index; /* index into ids[] or -1 */
    int32_t        _pad;     /* 0, or the weight type of THIS job when it differs from the launch's `type` (decode only: N == 1, one X,
                                contiguous rows; Q4_K+Q6_K, Q5_K+Q6_K, Q4_K+Q5_K: attn_q/attn_k + attn_v of the *_K_M mixes in one launch) */
} b200_gemv_job;
#define B200_GEMV_MAX_JOBS 16
B200_API int b200_mul_mat_vec_q_multi(int type, const b200_gemv_job *jobs, int n_jobs, size_t w_row_bytes,
                                      int64_t x_stride, int64_t dst_stride, int64_t K, int64_t N,
                                      const int32_t *ids, size_t expert_bytes, int32_t n_expert, b200_stream_t stream);

/* The same launch with the neighbouring graph nodes folded in (what llm_build_norm + llm_build_lora_mm + ggml_add emit around
 * every matmul of a layer, src/llama.cpp:9254-9287, 10479-10560):
 *   norm_weight != NULL: X is replaced by RMS_NORM(X, norm_eps) * norm_weight while it is quantised (X rows of K floats; the
 *                        normalised activations are only written when norm_out / norm_mul_out ask for them);
 *   residual[j] != NULL: dst_j = W_j . X + residual[j]   (same layout as dst_j; may alias it). */
typedef struct {
    const float *norm_weight; /* f32 [K], 16-byte aligned, or NULL */
    float        norm_eps;
    int32_t      silu_pair;   /* != 0: n_jobs == 2, N == 1, jobs[0] = gate, jobs[1] = up (same M, same X):
                                 jobs[0].dst[m] = silu(gate[m] . x) * (up[m] . x) -- UNARY(SILU) + MUL of llm_build_ffn
                                 (src/llama.cpp:9289-9414) folded into the epilogue; jobs[1].dst is not written */
    const float *residual[B200_GEMV_MAX_JOBS];
    /* The folded nodes remain graph nodes: another split of ggml_backend_sched may read them (the reference backend never
     * elides a node's output).  Each of these is optional (NULL = do not store) and costs one extra store per value:
     *   raw_dst[j]:   with residual[j]: the un-added product W_j . X (the MUL_MAT node), same layout as dst_j;
     *   norm_out / norm_mul_out: with norm_weight: RMS_NORM(X) and RMS_NORM(X) * norm_weight, N rows of K floats
     *                 (row stride K), stored by one CTA of the launch;
     *   gate_out / up_out / silu_out: with silu_pair: gate . x, up . x and silu(gate . x), M floats each.
     * A destination that aliases a later one in this list (ggml-alloc in-place reuse) must be passed as NULL. */
    float       *raw_dst[B200_GEMV_MAX_JOBS];
    float       *norm_out, *norm_mul_out;
    float       *gate_out, *up_out, *silu_out;
} b200_gemv_fuse;
B200_API int b200_mul_mat_vec_q_fused(int type, const b200_gemv_job *jobs, int n_jobs, size_t w_row_bytes, int64_t x_stride,
                                      int64_t dst_stride, int64_t K, int64_t N, const b200_gemv_fuse *fuse, b200_stream_t stream);

/* ---- (a9) MUL_MAT_ID (expert GEMV), ids read on the device ------------------------------------------
 * replaces: ggml_cuda_mul_mat_id incl. its cudaMemcpyAsync(ids)+cudaStreamSynchronize and the
 * k_copy_src1_to_contiguous / k_copy_dst_from_contiguous gather-scatter, ggml/src/ggml-cuda.cu:1954-2139
 * semantics (ggml.c:5693-5729): dst[t][e][:] = as[ids[t][e]] (M x K) . b[t][e % n_b_rows][:]
 *   as: n_expert matrices, expert stride expert_bytes, row stride w_row_bytes
 *   ids: i32, element (t, e) at ids[t*ids_stride + e];  b: f32 [T][n_b_rows][K] (strides in floats)
 *   dst: f32 [T][n_used][M] (strides in floats).  T*n_used <= 16 per call (decode regime).
 * An out-of-range expert id makes the affected rows NaN (the reference asserts; a device kernel cannot abort the host). */
B200_API int b200_mul_mat_id_vec_q(int type, const void *as, size_t w_row_bytes, size_t expert_bytes, int32_t n_expert,
                                   const int32_t *ids, int64_t ids_stride, const float *b, int64_t b_stride_row, int64_t b_stride_tok,
                                   float *dst, int64_t dst_stride_row, int64_t dst_stride_tok,
                                   int64_t K, int64_t M, int64_t n_used, int64_t n_b_rows, int64_t T, b200_stream_t stream);

/* MUL_MAT_ID for batches (prompt processing of a mixture-of-experts model), same tensor layout as b200_mul_mat_id_vec_q:
 * replaces the batch branch of ggml_cuda_mul_mat_id (ggml/src/ggml-cuda.cu:2066-2139: ids -> host + stream sync, then per expert
 * k_copy_src1_to_contiguous / mul_mat / k_copy_dst_from_contiguous, :1954-2001).  The (token, slot) pairs are grouped by expert
 * ON THE DEVICE (count, prefix sum, sorted row lists, tile table -- no host synchronisation, CUDA-graph capturable), the selected
 * activation rows are gathered as bf16, and ONE grouped tcgen05 launch computes every expert's tile list and scatters the f32
 * results to dst.  K must be a multiple of 256, n_expert <= 256, n_used * T <= 65535.  Rows whose id is outside [0, n_expert)
 * are skipped.  workspace: b200_mul_mat_id_q_workspace_bytes() bytes, 256-byte aligned. */
B200_API size_t b200_mul_mat_id_q_workspace_bytes(int type, int64_t K, int64_t M, int64_t n_used, int64_t T, int32_t n_expert);
B200_API int b200_mul_mat_id_q(int type, const void *as, size_t w_row_bytes, size_t expert_bytes, int32_t n_expert,
                               const int32_t *ids, int64_t ids_stride, const float *b, int64_t b_stride_row, int64_t b_stride_tok,
                               float *dst, int64_t dst_stride_row, int64_t dst_stride_tok, int64_t K, int64_t M, int64_t n_used,
                               int64_t n_b_rows, int64_t T, void *workspace, size_t workspace_bytes, b200_stream_t stream);

/* ---- (a7) prefill GEMM on tcgen05: dst[n][m] = sum_k W[m][k] * X[n][k], any N ------------------------
 * replaces: ggml_cuda_op_mul_mat_q + mul_mat_q<type,...> + stream-k fixup + quantize_mmq_q8_1 and the cuBLAS
 * dequantise-then-GEMM fallback, ggml/src/ggml-cuda/mmq.cuh:2495-2905, mmq.cu:3-151, ggml-cuda.cu:1186-1284
 * workspace: b200_mul_mat_q_workspace_bytes() bytes of device memory (may be NULL when that returns 0).
 * The call first converts X to bf16 at the front of the workspace.  X == NULL: the workspace still holds the converted activations
 * of the previous call with the same K and N (sibling matmuls over one activation: Q / K / V, gate / up) and they are used as is. */
B200_API size_t b200_mul_mat_q_workspace_bytes(int type, int64_t K, int64_t M, int64_t N);
B200_API int b200_mul_mat_q(int type, const void *W, size_t w_row_bytes, const float *X, int64_t x_stride,
                            float *dst, int64_t dst_stride, int64_t K, int64_t M, int64_t N,
                            void *workspace, size_t workspace_bytes, b200_stream_t stream);

/* ---- (a10-a12) FLASH_ATTN_EXT -----------------------------------------------------------------------
 * replaces: ggml_cuda_flash_attn_ext -> flash_attn_vec_ext_f16/f32 + flash_attn_combine_results (decode) and
 * flash_attn_ext_f16 (wmma, prefill) incl. the full-tensor dequantisation of a quantised KV cache,
 * ggml/src/ggml-cuda/fattn.cu:298-345, fattn-vec-f16.cuh:8-299, fattn-common.cuh:519-708, fattn-wmma-f16.cuh:13-506
 * layouts (ggml.h:1802-1817): q f32 [D, N, n_head] (strides in bytes), k/v type_k/type_v [D, n_kv, n_head_kv]
 * views into the cache (row stride nb1, head stride nb2, bytes), mask f16 [n_kv, >=N] or NULL (row stride bytes),
 * dst f32 [D, n_head, N] contiguous.  K and V types: F16, Q8_0 or Q4_0 (both the same).  D in {64, 128}; with an F16 cache also {80, 96, 112, 256}
 * (the head sizes of ggml-cuda.cu:3162-3177).  n_head % n_head_kv == 0. */
typedef struct {
    const void *q, *k, *v, *mask;
    float      *dst;
    int64_t     D, N, n_head, n_head_kv, n_kv;
    int32_t     type_k, type_v;
    size_t      q_nb1, q_nb2, k_nb1, k_nb2, v_nb1, v_nb2, mask_nb1;
    float       scale, max_bias, logit_softcap;
} b200_fattn_args;
B200_API size_t b200_flash_attn_workspace_bytes(const b200_fattn_args *a);
B200_API int b200_flash_attn_ext(const b200_fattn_args *a, void *workspace, size_t workspace_bytes, b200_stream_t stream);

/* b200_flash_attn_ext without a caller-owned workspace and without any allocation (no cudaMalloc / cudaFree, directly or
 * through a library): for callers under CUDA graph capture, serving loops and harnesses that forbid allocating.
 * No counterpart in the reference.
 * The split partials and arrival counters live in one static device array of B200_FATTN_STATIC_WS_BYTES (8 MiB) that every
 * GPU that loads the library holds until the process exits.  Same kernels, plan and results as b200_flash_attn_ext.
 * Shapes whose b200_flash_attn_workspace_bytes() exceeds the array return B200_ERR_UNSUPPORTED (b200_last_error names both
 * sizes).  ALL calls on a device share the array: calls that overlap on the GPU (different streams, or graph branches that
 * run concurrently) race -- serialise them or use b200_flash_attn_ext with separate workspaces. */
#define B200_FATTN_STATIC_WS_BYTES (8u << 20)
B200_API size_t b200_flash_attn_static_workspace_max(void);
B200_API int b200_flash_attn_ext_static(const b200_fattn_args *a, b200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_GGML_H */
