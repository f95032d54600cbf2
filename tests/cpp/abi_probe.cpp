// Prints the size / offsets / enum values of everything that crosses the ggml_backend boundary, once from the reference
// headers (-DUSE_REF -I ggml/include -I ggml/src) and once from include/b200_ggml_abi.h; tests/test_abi.py diffs the two.
#include <cstddef>
#include <cstdio>
#ifdef USE_REF
#include "ggml.h"
#include "ggml-backend.h"
#include "ggml-backend-impl.h"
#include "ggml-impl.h"
extern "C" void ggml_abort(const char *, int, const char *, ...) { __builtin_trap(); }  // referenced by inline helpers of ggml-impl.h
#else
#include "b200_ggml_abi.h"
#endif
#define SZ(T) printf("sizeof(" #T ")=%zu\n", sizeof(T))
#define OFF(T, f) printf("offsetof(" #T "," #f ")=%zu\n", offsetof(T, f))
#define VAL(e) printf(#e "=%d\n", (int)(e))
int main() {
    SZ(ggml_tensor); OFF(ggml_tensor, type); OFF(ggml_tensor, buffer); OFF(ggml_tensor, ne); OFF(ggml_tensor, nb); OFF(ggml_tensor, op);
    OFF(ggml_tensor, op_params); OFF(ggml_tensor, flags); OFF(ggml_tensor, src); OFF(ggml_tensor, view_src); OFF(ggml_tensor, view_offs);
    OFF(ggml_tensor, data); OFF(ggml_tensor, name); OFF(ggml_tensor, extra);
    SZ(ggml_cgraph); OFF(ggml_cgraph, n_nodes); OFF(ggml_cgraph, nodes); OFF(ggml_cgraph, leafs);
    SZ(ggml_backend_buffer_type_i); SZ(ggml_backend_buffer_type); SZ(ggml_backend_buffer_i); SZ(ggml_backend_buffer); OFF(ggml_backend_buffer, buft);
    OFF(ggml_backend_buffer, context); OFF(ggml_backend_buffer, size); OFF(ggml_backend_buffer, usage);
    SZ(ggml_backend_i); OFF(ggml_backend_i, graph_compute); OFF(ggml_backend_i, event_wait); SZ(ggml_backend); OFF(ggml_backend, device); OFF(ggml_backend, context);
    SZ(ggml_backend_device_i); OFF(ggml_backend_device_i, supports_op); SZ(ggml_backend_device); SZ(ggml_backend_reg_i); SZ(ggml_backend_reg);
    SZ(ggml_backend_dev_props); SZ(ggml_backend_event);
    VAL(GGML_TYPE_F32); VAL(GGML_TYPE_F16); VAL(GGML_TYPE_Q4_0); VAL(GGML_TYPE_Q8_0); VAL(GGML_TYPE_Q8_1); VAL(GGML_TYPE_Q4_K); VAL(GGML_TYPE_Q5_K);
    VAL(GGML_TYPE_Q6_K); VAL(GGML_TYPE_I32); VAL(GGML_TYPE_BF16);
    VAL(GGML_OP_ADD); VAL(GGML_OP_SUB); VAL(GGML_OP_MUL); VAL(GGML_OP_DIV); VAL(GGML_OP_SQR); VAL(GGML_OP_SUM_ROWS); VAL(GGML_OP_RMS_NORM);
    VAL(GGML_OP_MUL_MAT); VAL(GGML_OP_MUL_MAT_ID); VAL(GGML_OP_SCALE); VAL(GGML_OP_CPY); VAL(GGML_OP_CONT); VAL(GGML_OP_RESHAPE); VAL(GGML_OP_VIEW);
    VAL(GGML_OP_PERMUTE); VAL(GGML_OP_TRANSPOSE); VAL(GGML_OP_GET_ROWS); VAL(GGML_OP_SOFT_MAX); VAL(GGML_OP_ROPE); VAL(GGML_OP_CLAMP);
    VAL(GGML_OP_ARGSORT); VAL(GGML_OP_FLASH_ATTN_EXT); VAL(GGML_OP_UNARY); VAL(GGML_OP_DUP);
    VAL(GGML_UNARY_OP_SILU); VAL(GGML_UNARY_OP_GELU); VAL(GGML_UNARY_OP_GELU_QUICK); VAL(GGML_UNARY_OP_RELU); VAL(GGML_UNARY_OP_TANH); VAL(GGML_UNARY_OP_SIGMOID);
    VAL(GGML_BACKEND_DEVICE_TYPE_GPU_FULL); VAL(GGML_STATUS_SUCCESS); VAL(GGML_MAX_NAME); VAL(GGML_MAX_SRC); VAL(GGML_MAX_OP_PARAMS); VAL(GGML_KQ_MASK_PAD);
    return 0;
}
