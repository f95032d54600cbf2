/* b200_llama.h -- C ABI of the host-side decoder that sits ABOVE the ggml_backend boundary (libkcpp_b200.so).
 *
 * This is the caller side of the hot path ("the callers either side of the path"): it does what
 * llama.cpp does for a llama / mixtral model -- builds the per-batch ggml graph node for node like build_llama
 * (/root/reference/src/llama.cpp:10453-10617, llm_build_kv :9648-9680, llm_build_kv_store :9161-9203, llm_build_kqv
 * :9517-9646, llm_build_ffn :9289-9414, llm_build_moe_ffn :9416-9515), places weights / KV cache / activations in
 * buffers obtained through the backend's buffer-type vtable, feeds the graph inputs with set_tensor, and calls the
 * backend's graph_compute -- the same five vtables llama.cpp would call (b200_ggml_backend.h).  It exists so that
 * bench.py and the tests can run whole-token decode through the reference-facing plugin interface on a machine that
 * has no reference checkout; under koboldcpp the caller is llama.cpp itself.
 *
 * All functions return 0 or a b200_status; b200_last_error() describes the failure.  Not thread safe per model.
 */
#ifndef B200_LLAMA_H
#define B200_LLAMA_H

#include "b200_ggml.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int32_t n_embd, n_layer, n_head, n_head_kv, n_embd_head, n_ff, n_vocab;
    int32_t n_expert, n_expert_used; /* 0 / 0 for dense models */
    int32_t n_ctx;                   /* KV cells */
    int32_t type_k, type_v;          /* B200_TYPE_F16 or B200_TYPE_Q8_0 (llama_context_params.type_k/type_v) */
    int32_t rope_neox;               /* 0: LLAMA_ROPE_TYPE_NORM (llama, mixtral) */
    int32_t n_ctx_orig;
    float   rms_eps, rope_freq_base, rope_freq_scale;
    int32_t main_device;             /* device that runs the graph */
    int32_t n_split_devices;         /* 0/1: everything on main_device; >1: weight matrices in the row-split buffer type */
    float   tensor_split[16];        /* all zero: equal split */
    /* tensor-parallel mode, one process per GPU (b200_comm.h; a communicator of tp_world ranks must exist): the caller uploads
     * this rank's SHARD of every matrix -- attn_q/k/v, ffn_gate/up and output split by output rows (heads r*n_head/W ..,
     * ffn columns, vocabulary rows), attn_output and ffn_down split along K (their ne0 is n_embd/W resp. n_ff/W) -- while
     * token_embd and the norm vectors are replicated.  The graph then holds n_head/W heads and n_head_kv/W KV heads per rank,
     * one all-reduce after attn_output and one after ffn_down, and an all-gather of the logits.  0 / 1 = off. */
    int32_t tp_rank, tp_world;
} b200_llama_params;

typedef struct b200_llama b200_llama; /* opaque */

B200_API int  b200_llama_new(const b200_llama_params *p, b200_llama **out);
B200_API void b200_llama_free(b200_llama *m);

/* Upload one weight tensor under its GGUF name ("token_embd.weight", "blk.3.attn_q.weight", "blk.3.ffn_gate_exps.weight",
 * "output_norm.weight", "output.weight", ...; names as written by convert_hf_to_gguf.py / gguf-py/gguf/constants.py).
 * data: ggml_nbytes of the tensor in host memory, GGUF layout (ne0 contiguous).  Goes through buft->alloc_buffer,
 * buffer->init_tensor and buffer->set_tensor exactly like llama_model_loader::load_all_data (src/llama.cpp:5024-5219). */
B200_API int b200_llama_set_tensor(b200_llama *m, const char *name, int type, int64_t ne0, int64_t ne1, int64_t ne2, const void *data);
/* Same tensor contents as an already uploaded tensor but its own device memory (synthetic models: one generated layer
 * replicated device-to-device instead of 80 host uploads). */
B200_API int b200_llama_clone_tensor(b200_llama *m, const char *name, const char *from);

/* Allocate KV cache + compute buffers once all weights are set (llama_new_context_with_model, src/llama.cpp:19507-19594). */
B200_API int b200_llama_finalize(b200_llama *m, int32_t max_batch);

/* Raw access to KV cells [cell0, cell0+n) of layer il (is_v: 0 = K, 1 = V): rows of n_embd_gqa values in type_k / type_v,
 * as llama_state_seq_set_data would write them (ggml_backend_tensor_set with an offset). */
B200_API int b200_llama_kv_write(b200_llama *m, int32_t il, int32_t is_v, int32_t cell0, int32_t n, const void *data);
B200_API int b200_llama_kv_read(b200_llama *m, int32_t il, int32_t is_v, int32_t cell0, int32_t n, void *data);

/* llama_decode for one sequence: tokens[n_tokens] at positions n_past .. n_past+n_tokens-1, stored in KV cells of the same
 * index.  logits: host buffer receiving n_vocab floats of the LAST token (logits_all == 0) or n_tokens*n_vocab (!= 0).
 * Synchronous on return (like llama_decode + llama_get_logits).  logits may be NULL (then nothing is copied back). */
B200_API int b200_llama_decode(b200_llama *m, const int32_t *tokens, int32_t n_tokens, int32_t n_past, float *logits, int32_t logits_all);

/* Asynchronous halves of the same call, for device-side timing: enqueue copies + graph, then wait. */
B200_API int b200_llama_decode_async(b200_llama *m, const int32_t *tokens, int32_t n_tokens, int32_t n_past, float *logits, int32_t logits_all);
B200_API int b200_llama_synchronize(b200_llama *m);
/* Replay the last decoded graph again with inputs already resident (no H2D / D2H): the HBM-resident leg of bench.py. */
B200_API int b200_llama_replay(b200_llama *m);
B200_API b200_stream_t b200_llama_stream(b200_llama *m); /* the backend's stream (for CUDA-event timing) */

/* Introspection for tests: number of graph nodes / kernels launched by the last decode, bytes of weights resident. */
B200_API int b200_llama_stats(b200_llama *m, int64_t *n_nodes, int64_t *n_kernels, int64_t *weight_bytes, int64_t *kv_bytes);

#ifdef __cplusplus
}
#endif
#endif
