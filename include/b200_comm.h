/* b200_comm.h -- the exchange step of the row-split (tensor-parallel) hot path: one process per GPU, NCCL over NVLink.
 *
 * Replaces the star-shaped peer copies of the reference's multi-GPU matmul loop (/root/reference/ggml/src/ggml-cuda.cu:1519-1657:
 * broadcast the quantised activations from the main GPU, gather output rows back, 2 peer memcpys + 2 event waits per device per
 * matmul from one host thread) with collectives that live INSIDE the captured decode graph:
 *   - weights that produce a layer's hidden state (attn_output, ffn_down) are split along K: every rank holds K/W columns and
 *     computes a partial [n_embd] vector; the partials are summed with one all-reduce (2 per layer instead of 7 exchanges);
 *   - all other matrices are split by output rows and feed rank-local work (heads, ffn columns), so they need no exchange;
 *   - the output head is split by vocabulary rows and its logits are all-gathered.
 * The communicator belongs to the process (one GPU per process).  NCCL is loaded at run time (libnccl.so.2: the copy PyTorch
 * ships, or the system one); single-GPU use of libkcpp_b200.so never touches it.
 */
#ifndef B200_COMM_H
#define B200_COMM_H

#include "b200_ggml.h"

#ifdef __cplusplus
extern "C" {
#endif

#define B200_COMM_ID_BYTES 128 /* sizeof(ncclUniqueId) */

/* rank 0 creates the id; the caller broadcasts the 128 bytes to the other ranks by any means (bench.py: torch.distributed) */
B200_API int b200_comm_unique_id(void *id_out);
/* collective over all ranks: ncclCommInitRank on the CURRENT device, then one warm-up all-reduce so that no lazy NCCL
 * allocation happens later inside a CUDA-graph capture */
B200_API int b200_comm_init(const void *id, int rank, int world);
B200_API int b200_comm_destroy(void);
B200_API int b200_comm_rank(void);  /* -1 when no communicator exists */
B200_API int b200_comm_world(void); /*  0 when no communicator exists */

/* in-place sum over ranks of n floats; capturable in a CUDA graph */
B200_API int b200_all_reduce_sum_f32(float *buf, int64_t n, b200_stream_t stream);
/* dst[r * n .. (r+1) * n) = src of rank r */
B200_API int b200_all_gather_f32(const float *src, float *dst, int64_t n, b200_stream_t stream);

/* ---- peer window: decode-size all-reduce in ONE kernel over NVLink peer memory (one process per GPU, CUDA IPC).
 * Replaces, for messages of up to B200_COMM_PEER_MAX_FLOATS floats, the NCCL call AND the residual ADD node behind it: every rank
 * pushes its partial vector into a window in each peer's HBM, exchanges per-block release/acquire flags, and adds the W partials in
 * rank order (+ residual), so all ranks hold the bit-identical result.  Set-up: every rank creates its window and obtains a
 * 64-byte IPC handle, the caller all-gathers the handles (rank order) by any means, every rank attaches.  All ranks must issue the
 * same sequence of calls (they do: same graph).  Without a window the backend uses the NCCL path above. */
#define B200_COMM_IPC_BYTES 64            /* sizeof(cudaIpcMemHandle_t) */
#define B200_COMM_PEER_MAX_FLOATS 16384
B200_API int b200_comm_peer_window_create(void *handle_out);      /* after b200_comm_init; 2..8 ranks */
B200_API int b200_comm_peer_window_attach(const void *handles);   /* world x 64 bytes */
B200_API int b200_comm_peer_window_destroy(void);                 /* also done by b200_comm_destroy */
B200_API int b200_comm_peer_ready(void);                          /* 1 when attached */
B200_API int b200_comm_peer_status(void);                         /* 0 ok; 1 a wait timed out (peer lost): results invalid */
/* dst = (sum over ranks, in rank order, of src) + residual (residual may be NULL; dst may alias src or residual);
 * n % 4 == 0, n <= B200_COMM_PEER_MAX_FLOATS, 16-byte aligned pointers; capturable in a CUDA graph */
B200_API int b200_all_reduce_sum_add_f32(const float *src, const float *residual, float *dst, int64_t n, b200_stream_t stream);

/* How a ggml graph asks the backend for an exchange: a GGML_OP_MAP_CUSTOM1 node (ggml.h: dst = f(src0)) whose
 * op_params[0] == B200_COLLECTIVE_TAG and op_params[1] is one of the kinds below.  llama.cpp never emits such a node; the
 * tensor-parallel host decoder (b200_llama.h, tp_world > 1) does. */
#define B200_COLLECTIVE_TAG 0x0b200c01
enum { B200_COLLECTIVE_ALL_REDUCE_SUM = 0, /* dst (same shape, may alias src0) = sum over ranks of src0 */
       B200_COLLECTIVE_ALL_GATHER = 1 /* dst [ne0 * world] = concatenation over ranks of src0 [ne0] */ };

#ifdef __cplusplus
}
#endif
#endif
