// b200_flash_attn_ext_static: FLASH_ATTN_EXT for callers that cannot hand over a workspace they allocated (code under CUDA graph
// capture, serving loops and harnesses that forbid cudaMalloc).  The split partials and arrival counters live in ONE file-scope
// device array of B200_FATTN_STATIC_WS_BYTES; every GPU that loads the library holds it until the process exits.
//
// No counterpart in the reference.  The kernels are the ones b200_flash_attn_ext
// launches (same plan, same results); only where the scratch comes from differs.
#include <cuda_runtime.h>

#include <mutex>

#include "b200_common.cuh"
#include "b200_ggml.h"

namespace b200 {
size_t fa_ticket_bytes(const b200_fattn_args *a);  // fattn.cu
}
using namespace b200;

// In host code this name is the host-side stand-in of the symbol, NOT a device address: it is resolved per device below.
static __device__ __align__(256) unsigned char g_fa_static_ws[B200_FATTN_STATIC_WS_BYTES];

namespace {
constexpr int kMaxDevices = 64;
std::mutex     g_ws_mutex;
unsigned char *g_ws_ptr[kMaxDevices];  // device address of g_fa_static_ws on each device, resolved once with that device current

// device address of the scratch array on the CURRENT device
int static_ws(unsigned char **out) {
    int dev = -1;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        set_error("b200_flash_attn_ext_static: device ordinal %d out of range", dev);
        return B200_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lock(g_ws_mutex);
    if (!g_ws_ptr[dev]) {
        unsigned char *ptr = nullptr;
        B200_CUDA_TRY(cudaGetSymbolAddress((void **)&ptr, g_fa_static_ws));
        if (!ptr || ((uintptr_t)ptr & 255)) {
            set_error("b200_flash_attn_ext_static: scratch symbol resolved to %p", (void *)ptr);
            return B200_ERR_CUDA;
        }
        g_ws_ptr[dev] = ptr;
    }
    *out = g_ws_ptr[dev];
    return B200_OK;
}
}  // namespace

extern "C" size_t b200_flash_attn_static_workspace_max(void) { return B200_FATTN_STATIC_WS_BYTES; }

extern "C" int b200_flash_attn_ext_static(const b200_fattn_args *a, b200_stream_t stream) {
    if (!a) { set_error("b200_flash_attn_ext_static: NULL arguments"); return B200_ERR_INVALID; }
    const size_t need = b200_flash_attn_workspace_bytes(a);  // 0 for arguments b200_flash_attn_ext will refuse: it reports them
    if (need > B200_FATTN_STATIC_WS_BYTES) {
        set_error("b200_flash_attn_ext_static: this shape needs %zu bytes of scratch, the static array holds %zu "
                  "(use b200_flash_attn_ext with a caller-owned workspace)", need, (size_t)B200_FATTN_STATIC_WS_BYTES);
        return B200_ERR_UNSUPPORTED;
    }
    unsigned char *ws = nullptr;
    const int rc = static_ws(&ws);
    if (rc != B200_OK) return rc;
    // The array keeps the previous call's contents.  The partials are fully written before they are read; the arrival counters
    // must start at zero (a finished launch leaves them zero, an aborted one may not): clear them in stream order.
    const size_t tickets = fa_ticket_bytes(a);
    if (need && tickets) B200_CUDA_TRY(cudaMemsetAsync(ws, 0, tickets, (cudaStream_t)stream));
    return b200_flash_attn_ext(a, ws, B200_FATTN_STATIC_WS_BYTES, stream);
}
