// comm.cu -- the exchange step of the tensor-parallel hot path (include/b200_comm.h): a process-wide NCCL communicator,
// loaded at run time so that single-GPU users of the library never need NCCL.
//
// Replaces the peer-copy star of ggml_cuda_op_mul_mat's split branch (/root/reference/ggml/src/ggml-cuda.cu:1519-1657,
// ggml_cuda_Memcpy2DPeerAsync :1338-1356) with all-reduce / all-gather calls that are captured into the decode CUDA graph.
#include <dlfcn.h>

#include <cstring>

#include "../../include/b200_comm.h"
#include "b200_common.cuh"
#include <cstdlib>

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
static bool g_pw_created = false, g_pw_ready = false;  // peer window (below)
extern bool g_mmvq_pdl;                                 // mmvq.cu: programmatic dependent launch is on
static void pw_release();

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
    if (g_pw_created) { cudaDeviceSynchronize(); pw_release(); }
    if (g_comm) {
        cudaDeviceSynchronize();
        g_nccl.CommDestroy(g_comm);
        g_comm = nullptr;
    }
    g_rank = -1;
    g_world = 0;
    return B200_OK;
}

// ------------------------------------------------------------------------------------------------ peer window (NVLink one-shot all-reduce)
// Decode-size all-reduces (n_embd floats: 16-32 KB) are latency-bound: NCCL costs ~10 us per call here (launch + its own protocol)
// and the residual ADD behind it is another launch.  Here every rank owns a "window" in its HBM that all peers map through CUDA
// IPC (one process per GPU, NVLink/NVSwitch peer access).  One kernel per rank, no fence and no separate flag: every 8-byte word
// PUSHED into a peer's window carries 4 bytes of payload and the 4-byte number of the call (the idea of NCCL's LL protocol), so a
// receiver simply polls the words it needs until they show the current call number -- one NVLink traversal end to end.  A
// 16-byte store holds two such self-validating halves.  Each thread then adds the W partials of its 4 elements in RANK ORDER
// (+ residual): every rank computes the bit-identical sum.  Two alternating slots make a trailing barrier unnecessary: a rank
// can only reach call e+2 (which rewrites slot e & 1) after it received every peer's data of call e+1, which a peer sends
// only after it finished reading call e.
namespace b200 {

constexpr int    PW_MAX_WORLD = 8;
constexpr int    PW_THREADS   = 128;                       // 4 floats per thread
constexpr int    PW_BLOCK_F   = PW_THREADS * 4;            // floats per block
constexpr int    PW_MAX_N     = B200_COMM_PEER_MAX_FLOATS;
constexpr int    PW_MAX_BLK   = PW_MAX_N / PW_BLOCK_F;
constexpr size_t PW_BYTES     = (size_t)2 * PW_MAX_WORLD * PW_MAX_N * 8;  // [slot][source rank][element] x {payload, call number}

struct PeerWin {
    unsigned char *win[PW_MAX_WORLD];  // win[r]: rank r's window as mapped into this process (win[rank] is the local allocation)
    unsigned *ctr;                     // local: calls completed per block
    unsigned *status;                  // local: set to 1 when a wait timed out (a peer died): results are then invalid
    int rank, world;
};
static PeerWin g_pw = {};

__device__ __forceinline__ void st_ll(unsigned char *p, float a, float b, unsigned e) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %2};" ::"l"(p), "r"(__float_as_uint(a)), "r"(e), "r"(__float_as_uint(b)) : "memory");
}
__device__ __forceinline__ uint4 ld_ll(const unsigned char *p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long pw_now() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

__global__ void __launch_bounds__(PW_THREADS) all_reduce_push_kernel(const PeerWin w, const float *__restrict__ src, const float *__restrict__ residual,
                                                                     float *__restrict__ dst, int n) {
    const int b = blockIdx.x, t = threadIdx.x;
    const int i = (b * PW_THREADS + t) * 4;
    const bool active = i < n;            // n % 4 == 0 (checked by the host)
    pdl_wait();                           // src, and the call counter of the previous exchange, are written by kernels in front of us
    pdl_trigger();
    const unsigned epoch = w.ctr[b] + 1;  // the number of this call: identical on every rank (same graph, same order)
    const int slot = (int)(epoch & 1u);
    if (active) {
        const float4 v = *reinterpret_cast<const float4 *>(src + i);
        const size_t off = ((size_t)(slot * PW_MAX_WORLD + w.rank) * PW_MAX_N + (size_t)i) * 8;
#pragma unroll
        for (int p = 0; p < PW_MAX_WORLD; p++) {
            if (p < w.world && p != w.rank) {
                st_ll(w.win[p] + off, v.x, v.y, epoch);
                st_ll(w.win[p] + off + 16, v.z, v.w, epoch);
            }
        }
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        const unsigned long long t0 = pw_now();
#pragma unroll
        for (int r = 0; r < PW_MAX_WORLD; r++) {
            if (r >= w.world) continue;
            float4 x = v;
            if (r != w.rank) {
                const unsigned char *q = w.win[w.rank] + ((size_t)(slot * PW_MAX_WORLD + r) * PW_MAX_N + (size_t)i) * 8;
                uint4 lo = ld_ll(q), hi = ld_ll(q + 16);
                int spins = 0;
                while (lo.y != epoch || lo.w != epoch || hi.y != epoch || hi.w != epoch) {
                    if ((++spins & 1023) == 0 && pw_now() - t0 > 2000000000ull) { *w.status = 1; break; }  // 2 s: a peer is gone; do not hang
                    lo = ld_ll(q);
                    hi = ld_ll(q + 16);
                }
                x = make_float4(__uint_as_float(lo.x), __uint_as_float(lo.z), __uint_as_float(hi.x), __uint_as_float(hi.z));
            }
            if (r == 0) acc = x;
            else { acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w; }
        }
        if (residual != nullptr) {
            const float4 x = *reinterpret_cast<const float4 *>(residual + i);
            acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
        }
        *reinterpret_cast<float4 *>(dst + i) = acc;
    }
    __syncthreads();
    if (t == 0) w.ctr[b] = epoch;
}

static void pw_release() {
    if (g_pw_ready) {
        for (int r = 0; r < g_pw.world; r++)
            if (r != g_pw.rank && g_pw.win[r]) cudaIpcCloseMemHandle(g_pw.win[r]);
    }
    if (g_pw_created) {
        if (g_pw.win[g_pw.rank]) cudaFree(g_pw.win[g_pw.rank]);
        if (g_pw.ctr) cudaFree(g_pw.ctr);
    }
    g_pw = PeerWin();
    g_pw_created = g_pw_ready = false;
}

}  // namespace b200

extern "C" int b200_comm_peer_window_create(void *handle_out) {
    if (!handle_out) { set_error("b200_comm_peer_window_create: NULL argument"); return B200_ERR_INVALID; }
    if (!g_comm || g_world < 2 || g_world > PW_MAX_WORLD) { set_error("b200_comm_peer_window_create: needs a communicator of 2..%d ranks", PW_MAX_WORLD); return B200_ERR_UNSUPPORTED; }
    if (g_pw_created) { set_error("b200_comm_peer_window_create: a window already exists"); return B200_ERR_INVALID; }
    static_assert(sizeof(cudaIpcMemHandle_t) == B200_COMM_IPC_BYTES, "cudaIpcMemHandle_t size");
    unsigned char *win = nullptr;
    unsigned *ctr = nullptr;
    B200_CUDA_TRY(cudaMalloc(&win, PW_BYTES));
    B200_CUDA_TRY(cudaMemset(win, 0, PW_BYTES));  // call number 0 never occurs
    B200_CUDA_TRY(cudaMalloc(&ctr, (PW_MAX_BLK + 1) * sizeof(unsigned)));
    B200_CUDA_TRY(cudaMemset(ctr, 0, (PW_MAX_BLK + 1) * sizeof(unsigned)));
    B200_CUDA_TRY(cudaDeviceSynchronize());
    cudaIpcMemHandle_t h;
    const cudaError_t e = cudaIpcGetMemHandle(&h, win);
    if (e != cudaSuccess) { cudaFree(win); cudaFree(ctr); return cuda_fail(e, "cudaIpcGetMemHandle"); }
    memcpy(handle_out, &h, sizeof(h));
    g_pw = PeerWin();
    g_pw.rank = g_rank;
    g_pw.world = g_world;
    g_pw.win[g_rank] = win;
    g_pw.ctr = ctr;
    g_pw.status = ctr + PW_MAX_BLK;
    g_pw_created = true;
    return B200_OK;
}

extern "C" int b200_comm_peer_window_attach(const void *handles) {
    if (!handles || !g_pw_created || g_pw_ready) { set_error("b200_comm_peer_window_attach: create the window first (once)"); return B200_ERR_INVALID; }
    for (int r = 0; r < g_world; r++) {
        if (r == g_rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, (const unsigned char *)handles + (size_t)r * B200_COMM_IPC_BYTES, sizeof(h));
        void *p = nullptr;
        const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            for (int q = 0; q < r; q++) if (q != g_rank && g_pw.win[q]) { cudaIpcCloseMemHandle(g_pw.win[q]); g_pw.win[q] = nullptr; }
            return cuda_fail(e, "cudaIpcOpenMemHandle");
        }
        g_pw.win[r] = (unsigned char *)p;
    }
    g_pw_ready = true;
    return B200_OK;
}

extern "C" int b200_comm_peer_window_destroy(void) {
    if (g_pw_created) cudaDeviceSynchronize();
    pw_release();
    return B200_OK;
}

extern "C" int b200_comm_peer_ready(void) { return g_pw_ready ? 1 : 0; }

extern "C" int b200_comm_peer_status(void) {
    if (!g_pw_created) return 0;
    unsigned v = 0;
    if (cudaMemcpy(&v, g_pw.status, sizeof(v), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    return (int)v;
}

extern "C" int b200_all_reduce_sum_add_f32(const float *src, const float *residual, float *dst, int64_t n, b200_stream_t stream) {
    if (!g_pw_ready) { set_error("b200_all_reduce_sum_add_f32: no peer window (b200_comm_peer_window_attach)"); return B200_ERR_INVALID; }
    if (!src || !dst || n < 0) { set_error("b200_all_reduce_sum_add_f32: bad arguments"); return B200_ERR_INVALID; }
    if (n == 0) return B200_OK;
    if (n > PW_MAX_N || (n & 3) || ((uintptr_t)src & 15) || ((uintptr_t)dst & 15) || ((uintptr_t)residual & 15)) {
        set_error("b200_all_reduce_sum_add_f32: n=%lld must be a multiple of 4 up to %d with 16-byte aligned buffers", (long long)n, PW_MAX_N);
        return B200_ERR_UNSUPPORTED;
    }
    const int grid = (int)((n + PW_BLOCK_F - 1) / PW_BLOCK_F);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(PW_THREADS);
    cfg.stream = (cudaStream_t)stream;
    // The exchange kernel uses no shared memory but ASKS for 150 KB: its CTAs then never share an SM with a CTA of the GEMVs around
    // it.  Under programmatic dependent launch its 16 small CTAs otherwise become resident next to the (register-lean) decode GEMV
    // CTAs and poll for their peers' data there while the next GEMV's prologue runs on the same SM: Llama-3-70B over 8 GPUs 186 ->
    // 241 tok/s (profiles/r02_tp70_variants_n8.log).  B200_AR_SMEM overrides (0: none).
    static const int ar_smem = getenv("B200_AR_SMEM") ? atoi(getenv("B200_AR_SMEM")) : 150000;
    if (ar_smem > 0) {
        static bool attr_set_dev[16] = {};
        int cur_dev = 0;
        B200_CUDA_TRY(cudaGetDevice(&cur_dev));
        if (!attr_set_dev[cur_dev & 15]) {
            B200_CUDA_TRY(cudaFuncSetAttribute(all_reduce_push_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ar_smem));
            attr_set_dev[cur_dev & 15] = true;
        }
        cfg.dynamicSmemBytes = (size_t)ar_smem;
    }
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = g_mmvq_pdl ? 1 : 0;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, all_reduce_push_kernel, g_pw, src, residual, dst, (int)n));
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
