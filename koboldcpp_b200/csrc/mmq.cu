// mmq.cu -- (a7) prefill GEMM over GGUF block formats on the 5th-generation tensor cores (tcgen05 + TMEM).
// Placeholder translation unit until the tcgen05 kernel lands: the entry points exist so the C ABI is complete
// and report B200_ERR_UNSUPPORTED (there is deliberately no CPU or cuBLAS fallback).
#include "b200_common.cuh"

using namespace b200;

extern "C" size_t b200_mul_mat_q_workspace_bytes(int, int64_t, int64_t, int64_t) { return 0; }

extern "C" int b200_mul_mat_q(int, const void *, size_t, const float *, int64_t, float *, int64_t, int64_t, int64_t, int64_t, void *,
                              size_t, b200_stream_t) {
    set_error("b200_mul_mat_q: tcgen05 prefill GEMM not built yet");
    return B200_ERR_UNSUPPORTED;
}
