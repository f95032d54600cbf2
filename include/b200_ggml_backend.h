/* b200_ggml_backend.h -- the drop-in boundary: a ggml backend ("B200") behind the reference's own ggml_backend
 * interface (reg / device / buffer type / buffer / backend vtables, /root/reference/ggml/src/ggml-backend-impl.h).
 *
 * Two ways to attach:
 *   Route 2 (run time):  ggml_backend_register(ggml_backend_b200_reg());   before any llama_* call
 *   Route 1 (link time): link libkcpp_b200.so in place of ggml-cuda.o with -DGGML_USE_CUDA; the library also exports
 *                        the 12 public symbols of ggml/include/ggml-cuda.h:22-44 under their reference names.
 *
 * Each function cites the reference entry point it replaces (paths under /root/reference/ggml/).
 */
#ifndef B200_GGML_BACKEND_H
#define B200_GGML_BACKEND_H

#ifdef B200_USE_GGML_HEADERS
#include "ggml-backend-impl.h"
#include "ggml.h"
#else
#include "b200_ggml_abi.h"
#endif
#include "b200_ggml.h"

#ifdef __cplusplus
extern "C" {
#endif

#define GGML_B200_MAX_DEVICES 16 /* GGML_CUDA_MAX_DEVICES, include/ggml-cuda.h:20 */

/* replaces ggml_backend_cuda_reg, src/ggml-cuda.cu:3302-3340 */
B200_API ggml_backend_reg_t ggml_backend_b200_reg(void);
/* replaces ggml_backend_cuda_init, src/ggml-cuda.cu:3342-3362 (NULL when the device does not exist) */
B200_API ggml_backend_t ggml_backend_b200_init(int device);
B200_API bool ggml_backend_is_b200(ggml_backend_t backend); /* ggml_backend_is_cuda, :2830 */
/* device buffer type; replaces ggml_backend_cuda_buffer_type, src/ggml-cuda.cu:597-621 */
B200_API ggml_backend_buffer_type_t ggml_backend_b200_buffer_type(int device);
/* row-split buffer type: weight rows sharded over the devices by cumulative tensor_split[16] fractions (all-zero / NULL:
 * split by free memory); replaces ggml_backend_cuda_split_buffer_type, src/ggml-cuda.cu:918-953.
 * Also answered by reg->iface.get_proc_address(reg, "ggml_backend_split_buffer_type"). */
B200_API ggml_backend_buffer_type_t ggml_backend_b200_split_buffer_type(const float *tensor_split);
/* pinned host memory; replaces ggml_backend_cuda_host_buffer_type, src/ggml-cuda.cu:1003-1025 */
B200_API ggml_backend_buffer_type_t ggml_backend_b200_host_buffer_type(void);
B200_API int  ggml_backend_b200_get_device_count(void);
B200_API void ggml_backend_b200_get_device_description(int device, char *description, size_t description_size);
B200_API void ggml_backend_b200_get_device_memory(int device, size_t *free, size_t *total);

/* knobs of the B200 backend that the reference does not have */
B200_API void ggml_backend_b200_set_graphs(int enabled);  /* whole-graph CUDA-graph replay (default on; env B200_NO_GRAPHS=1) */
B200_API void ggml_backend_b200_set_fusion(int enabled);  /* node fusion in graph_compute (default on; env B200_NO_FUSION=1)   */
/* counters for tests / bench: kernels launched (or replayed) by graph_compute since the last reset */
B200_API void ggml_backend_b200_get_stats(int64_t *n_kernel_launches, int64_t *n_graph_replays, int64_t *n_graph_captures);
B200_API void ggml_backend_b200_reset_stats(void);
/* the backend's CUDA stream (cudaStream_t), for CUDA-event timing around graph_compute */
B200_API void *ggml_backend_b200_get_stream(ggml_backend_t backend);

#ifdef __cplusplus
}
#endif
#endif
