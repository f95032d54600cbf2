// backend_exec.h -- interface between the graph executor (backend_graph.cu) and the per-node dispatch (backend_ops.cu).
#pragma once

#include "backend_internal.h"

namespace b200 {

struct ExecCtx {
    int device;            // the device graph_compute runs on (main device of a row-split model)
    cudaStream_t stream;
    void  *scratch;        // workspace of the prefill GEMM (b200_mul_mat_q)
    size_t scratch_bytes;
    void  *fa_ws;          // split-K workspace of FLASH_ATTN_EXT; its ticket counters are zero between launches
    size_t fa_ws_bytes;
    int64_t launches;      // kernels launched through this context
    // the activations whose bf16 copy sits at the front of `scratch` (left there by the last prefill GEMM); valid only while the
    // steps in between were prefill GEMMs too: the next sibling matmul over the same activation skips the conversion
    const void *xb_src = nullptr;
    int64_t xb_k = 0, xb_n = 0, xb_stride = 0;
};

// Run one graph node.  dst_slot != NULL (CPY only): the destination base pointer is read on the device from *dst_slot.
void run_node(ExecCtx &x, const ggml_tensor *n, void *const *dst_slot);
size_t node_fa_workspace(const ggml_tensor *n);
// an all-reduce node small enough for the one-kernel peer-memory exchange (comm.cu), with a peer window attached
bool peer_all_reduce_ok(const ggml_tensor *n);
size_t node_scratch_bytes(const ggml_tensor *n);
// a quantised-weight MUL_MAT in the decode regime that may share a launch with its siblings
bool mul_mat_fusable(const ggml_tensor *n);

}  // namespace b200
