// comm.cu -- the exchange step of the tensor-parallel hot path (include/b200_comm.h): a process-wide NCCL communicator,
// loaded at run time so that single-GPU users of the library never need NCCL.
//
// Replaces the peer-copy star of ggml_cuda_op_mul_mat's split branch (/root/reference/ggml/src/ggml-cuda.cu:1519-1657,
// ggml_cuda_Memcpy2DPeerAsync :1338-1356) with all-reduce / all-gather calls that are captured into the decode CUDA graph.
#include <dlfcn.h>

#include <cstring>

#include "../../include/b200_comm.h"
#include "b200_common.cuh"

namespace b200 {

// the slice of nccl.h this file needs (stable ABI since NCCL 2.0)
typedef struct { char internal[B200_COMM_ID_BYTES]; } nccl_id_t;
typedef void *nccl_comm_t;
enum { NCCL_FLOAT32 = 7, NCCL_SUM = 0 };

struct NcclApi {
    void *handle = nullptr;
    int (*GetUniqueId)(nccl_id_t *) = nullptr;
    int (*CommInitRank)(nccl_comm_t *, int, nccl_id_t, int) = nullptr;
    int (*CommDestroy)(nccl_comm_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, nccl_comm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};
static NcclApi g_nccl;
static nccl_comm_t g_comm = nullptr;
static int g_rank = -1, g_world = 0;

static int nccl_load() {
    if (g_nccl.handle) return B200_OK;
    // RTLD_NOLOAD first: when PyTorch is in the process its bundled NCCL is already mapped under this soname
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW);
    if (!h) { set_error("b200_comm: cannot load libnccl.so.2 (%s)", dlerror()); return B200_ERR_UNSUPPORTED; }
    NcclApi a;
    a.handle = h;
    a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(h, "ncclGetUniqueId");
    a.CommInitRank = (decltype(a.CommInitRank))dlsym(h, "ncclCommInitRank");
    a.CommDestroy = (decltype(a.CommDestroy))dlsym(h, "ncclCommDestroy");
    a.AllReduce = (decltype(a.AllReduce))dlsym(h, "ncclAllReduce");
    a.AllGather = (decltype(a.AllGather))dlsym(h, "ncclAllGather");
    a.GetErrorString = (decltype(a.GetErrorString))dlsym(h, "ncclGetErrorString");
    if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllReduce || !a.AllGather || !a.GetErrorString) {
        set_error("b200_comm: libnccl.so.2 lacks a required symbol");
        return B200_ERR_UNSUPPORTED;
    }
    g_nccl = a;
    return B200_OK;
}

static int nccl_fail(int rc, const char *what) {
    set_error("NCCL error %d (%s) at %s", rc, g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?", what);
    return B200_ERR_CUDA;
}
#define B200_NCCL_TRY(call)                                  \
    do {                                                     \
        const int rc__ = (call);                             \
        if (rc__ != 0) return ::b200::nccl_fail(rc__, #call); \
    } while (0)

}  // namespace b200

using namespace b200;

extern "C" int b200_comm_unique_id(void *id_out) {
    if (!id_out) { set_error("b200_comm_unique_id: NULL argument"); return B200_ERR_INVALID; }
    int rc = nccl_load();
    if (rc) return rc;
    nccl_id_t id;
    B200_NCCL_TRY(g_nccl.GetUniqueId(&id));
    memcpy(id_out, &id, sizeof(id));
    return B200_OK;
}

extern "C" int b200_comm_init(const void *id, int rank, int world) {
    if (!id || world < 1 || rank < 0 || rank >= world) { set_error("b200_comm_init: bad arguments"); return B200_ERR_INVALID; }
    if (g_comm) { set_error("b200_comm_init: a communicator already exists"); return B200_ERR_INVALID; }
    int rc = nccl_load();
    if (rc) return rc;
    nccl_id_t nid;
    memcpy(&nid, id, sizeof(nid));
    B200_NCCL_TRY(g_nccl.CommInitRank(&g_comm, world, nid, rank));
    g_rank = rank;
    g_world = world;
    // warm-up: channels, proxy threads and staging buffers are set up here, not inside a stream capture
    float *tmp = nullptr;
    B200_CUDA_TRY(cudaMalloc(&tmp, 4096 * sizeof(float) * (size_t)(world + 1)));
    B200_CUDA_TRY(cudaMemset(tmp, 0, 4096 * sizeof(float) * (size_t)(world + 1)));
    B200_NCCL_TRY(g_nccl.AllReduce(tmp, tmp, 4096, NCCL_FLOAT32, NCCL_SUM, g_comm, nullptr));
    B200_NCCL_TRY(g_nccl.AllGather(tmp, tmp + 4096, 4096, NCCL_FLOAT32, g_comm, nullptr));
    B200_CUDA_TRY(cudaDeviceSynchronize());
    B200_CUDA_TRY(cudaFree(tmp));
    return B200_OK;
}

extern "C" int b200_comm_destroy(void) {
    if (g_comm) {
        cudaDeviceSynchronize();
        g_nccl.CommDestroy(g_comm);
        g_comm = nullptr;
    }
    g_rank = -1;
    g_world = 0;
    return B200_OK;
}

extern "C" int b200_comm_rank(void) { return g_rank; }
extern "C" int b200_comm_world(void) { return g_world; }

extern "C" int b200_all_reduce_sum_f32(float *buf, int64_t n, b200_stream_t stream) {
    if (!g_comm) { set_error("b200_all_reduce_sum_f32: no communicator (b200_comm_init)"); return B200_ERR_INVALID; }
    if (!buf || n < 0) { set_error("b200_all_reduce_sum_f32: bad arguments"); return B200_ERR_INVALID; }
    if (n == 0) return B200_OK;
    B200_NCCL_TRY(g_nccl.AllReduce(buf, buf, (size_t)n, NCCL_FLOAT32, NCCL_SUM, g_comm, (cudaStream_t)stream));
    return B200_OK;
}

extern "C" int b200_all_gather_f32(const float *src, float *dst, int64_t n, b200_stream_t stream) {
    if (!g_comm) { set_error("b200_all_gather_f32: no communicator (b200_comm_init)"); return B200_ERR_INVALID; }
    if (!src || !dst || n < 0) { set_error("b200_all_gather_f32: bad arguments"); return B200_ERR_INVALID; }
    if (n == 0) return B200_OK;
    B200_NCCL_TRY(g_nccl.AllGather(src, dst, (size_t)n, NCCL_FLOAT32, g_comm, (cudaStream_t)stream));
    return B200_OK;
}
