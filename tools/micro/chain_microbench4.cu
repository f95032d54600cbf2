// Developer micro-benchmark 4: what does a chain of DEPENDENT weight-streaming phases (a decode token's GEMVs) cost under three
// execution schemes?  Every phase streams `bytes` of distinct weights through TMA rings, needs a 16 KB "activation" vector that the
// previous phase wrote (all CTAs contribute), and writes its own slice of the next vector.
//   V0  one full-SM kernel per phase (512 threads, ~190 KB ring), launched back to back in a CUDA graph, optional PDL
//   V1  one HALF-SM kernel per phase (256 threads, ~95 KB ring, 2 CTAs/SM): under PDL phase k+1 is co-resident with phase k and
//       fills its ring before griddepcontrol.wait
//   V2  one persistent kernel for the whole chain: grid barrier (counter) between phases, rings run ahead across the barrier
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o chain_microbench4 chain_microbench4.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
constexpr int SEG = 4096;
constexpr int MAXP = 64;

struct Phases {
    const char *w[MAXP];
    unsigned    nseg[MAXP];   // segments per warp
    int         n;
};

__device__ __forceinline__ void mbar_wait(unsigned b, unsigned ph) {
    asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(b), "r"(ph) : "memory");
}
__device__ __forceinline__ unsigned consume(const unsigned char *st, int lane, int work) {
    const uint4 *p = reinterpret_cast<const uint4 *>(st) + lane;
    unsigned a = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) { const uint4 v = p[k * 32]; a += __dp4a(v.x, v.y, v.z) + v.w; }
    for (int k = 0; k < work; k++) a = a * 1664525u + 1013904223u;
    return a;
}

// ---- V0 / V1: one kernel per phase.  NW warps, NS stages.
template <int NW, int NS, int MINB>
__global__ void __launch_bounds__(NW * 32, MINB) phase_kernel(const char *w, unsigned nseg, const float *xin, float *xout, int work, int early) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ unsigned long long bars[NW][NS];
    __shared__ float xs[4096];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const char *src = w + ((size_t)blockIdx.x * NW + warp) * (size_t)nseg * SEG;
    unsigned char *ring = smem + (size_t)warp * NS * SEG;
    if (lane == 0) {
        for (int s = 0; s < NS; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[warp][s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    auto issue = [&](unsigned i) {
        if (lane == 0) {
            const unsigned b = smem_u32(&bars[warp][i % NS]);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(SEG) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(ring + (i % NS) * SEG)),
                         "l"(src + (size_t)i * SEG), "r"(SEG), "r"(b) : "memory");
        }
    };
    for (unsigned i = 0; i < NS && i < nseg; i++) issue(i);
    if (early) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    // the activation vector of this phase (16 KB, written by the previous phase)
    for (int i = threadIdx.x; i < 1024; i += NW * 32) reinterpret_cast<float4 *>(xs)[i] = __ldcg(reinterpret_cast<const float4 *>(xin) + i);
    __syncthreads();
    unsigned acc = __float_as_uint(xs[threadIdx.x]);
    for (unsigned i = 0; i < nseg; i++) {
        const unsigned b = smem_u32(&bars[warp][i % NS]), ph = (i / NS) & 1;
        mbar_wait(b, ph);
        acc += consume(ring + (i % NS) * SEG, lane, work);
        __syncwarp();
        if (lane == 0) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        if (i + NS < nseg) issue(i + NS);
    }
    if (!early) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    // every CTA writes its slice of the next vector
    const int per = 4096 / gridDim.x + 1;
    if (threadIdx.x < per) { const int o = blockIdx.x * per + threadIdx.x; if (o < 4096) xout[o] = (float)(acc & 0xff) * 1e-3f; }
}

// ---- V2: persistent chain
template <int NW, int NS>
__global__ void __launch_bounds__(NW * 32, 1) chain_kernel(const __grid_constant__ Phases P, float *xbuf0, float *xbuf1, unsigned *counter, int work, int warp_arrive) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ unsigned long long bars[NW][NS];
    __shared__ float xs[4096];
    __shared__ unsigned cta_count;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char *ring = smem + (size_t)warp * NS * SEG;
    if (lane == 0) {
        for (int s = 0; s < NS; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[warp][s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x == 0) cta_count = 0;
    __syncthreads();
    // producer cursor
    int ip = 0; unsigned is = 0, islot = 0;
    auto issue = [&]() {
        if (lane == 0) {
            const char *src = P.w[ip] + (((size_t)blockIdx.x * NW + warp) * (size_t)P.nseg[ip] + is) * SEG;
            const unsigned b = smem_u32(&bars[warp][islot]);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(SEG) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(ring + islot * SEG)),
                         "l"(src), "r"(SEG), "r"(b) : "memory");
        }
        islot = islot + 1 == NS ? 0 : islot + 1;
        if (++is == P.nseg[ip]) { is = 0; ip++; }
    };
    for (int s = 0; s < NS && ip < P.n; s++) issue();
    unsigned cslot = 0, cph = 0;
    for (int p = 0; p < P.n; p++) {
        const float *xin = (p & 1) ? xbuf1 : xbuf0;
        float *xout = (p & 1) ? xbuf0 : xbuf1;
        if (p > 0) {
            // grid barrier p: all CTAs have written their slice of xin
            if (threadIdx.x == 0) {
                const unsigned target = (unsigned)p * gridDim.x;
                unsigned v;
                do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory"); } while (v < target);
            }
            __syncthreads();
        }
        for (int i = threadIdx.x; i < 1024; i += NW * 32) reinterpret_cast<float4 *>(xs)[i] = __ldcg(reinterpret_cast<const float4 *>(xin) + i);
        __syncthreads();
        unsigned acc = __float_as_uint(xs[threadIdx.x]);
        const unsigned nseg = P.nseg[p];
        for (unsigned i = 0; i < nseg; i++) {
            mbar_wait(smem_u32(&bars[warp][cslot]), cph);
            acc += consume(ring + cslot * SEG, lane, work);
            __syncwarp();
            if (lane == 0) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            cslot = cslot + 1 == NS ? 0 : cslot + 1;
            cph ^= (cslot == 0);
            if (ip < P.n) issue();
        }
        const int per = 4096 / gridDim.x + 1;
        if (warp_arrive) {
            // each warp writes part of the slice, then the last warp of the CTA to finish arrives for the CTA
            if (lane < 2) { const int o = blockIdx.x * per + warp * 2 + lane; if (o < 4096 && warp * 2 + lane < per) xout[o] = (float)(acc & 0xff) * 1e-3f; }
            __syncwarp();
            if (lane == 0) {
                __threadfence();
                const unsigned c = atomicAdd(&cta_count, 1u);
                if (c == (unsigned)(p + 1) * NW - 1) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
            }
        } else {
            if (threadIdx.x < per) { const int o = blockIdx.x * per + threadIdx.x; if (o < 4096) xout[o] = (float)(acc & 0xff) * 1e-3f; }
            __syncthreads();
            if (threadIdx.x == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
        }
    }
}

__global__ void flushk(const uint4 *p, size_t n16, uint4 *sink) {
    uint4 a = make_uint4(0, 0, 0, 0);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) { const uint4 v = __ldcg(p + i); a.x ^= v.x; a.y ^= v.y; }
    if ((a.x ^ a.y) == 0x12345678u) *sink = a;
}

template <int NW, int NS, int MINB>
static double run_per_phase(const std::vector<size_t> &sizes, char *buf, float *x0, float *x1, bool pdl, int early, int work, int grid, uint4 *sink, char *flushbuf) {
    const size_t smem = (size_t)NW * NS * SEG;
    CK(cudaFuncSetAttribute(phase_kernel<NW, NS, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaStream_t st; CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    cudaGraph_t g; cudaGraphExec_t ge;
    CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    size_t off = 0;
    for (size_t k = 0; k < sizes.size(); k++) {
        const unsigned nseg = (unsigned)(sizes[k] / grid / NW / SEG);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid); cfg.blockDim = dim3(NW * 32); cfg.dynamicSmemBytes = smem; cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
        const char *w = buf + off;
        const float *xin = (k & 1) ? x1 : x0; float *xout = (k & 1) ? x0 : x1;
        CK(cudaLaunchKernelEx(&cfg, phase_kernel<NW, NS, MINB>, w, nseg, xin, xout, work, early));
        off += sizes[k];
    }
    CK(cudaStreamEndCapture(st, &g));
    CK(cudaGraphInstantiate(&ge, g, 0));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e9;
    for (int rep = 0; rep < 4; rep++) {
        flushk<<<148 * 4, 512, 0, st>>>((const uint4 *)flushbuf, (400u << 20) / 16, sink);
        CK(cudaEventRecord(e0, st)); CK(cudaGraphLaunch(ge, st)); CK(cudaEventRecord(e1, st)); CK(cudaStreamSynchronize(st));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
    }
    CK(cudaGraphExecDestroy(ge)); CK(cudaGraphDestroy(g)); CK(cudaStreamDestroy(st));
    return best * 1e3;
}

template <int NW, int NS>
static double run_chain(const std::vector<size_t> &sizes, char *buf, float *x0, float *x1, unsigned *counter, int work, int warp_arrive, int grid, uint4 *sink, char *flushbuf) {
    const size_t smem = (size_t)NW * NS * SEG;
    CK(cudaFuncSetAttribute(chain_kernel<NW, NS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    Phases P = {}; P.n = (int)sizes.size();
    size_t off = 0;
    for (size_t k = 0; k < sizes.size(); k++) { P.w[k] = buf + off; P.nseg[k] = (unsigned)(sizes[k] / grid / NW / SEG); off += sizes[k]; }
    cudaStream_t st; CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e9;
    for (int rep = 0; rep < 4; rep++) {
        flushk<<<148 * 4, 512, 0, st>>>((const uint4 *)flushbuf, (400u << 20) / 16, sink);
        CK(cudaMemsetAsync(counter, 0, 4, st));
        CK(cudaEventRecord(e0, st));
        chain_kernel<NW, NS><<<grid, NW * 32, smem, st>>>(P, x0, x1, counter, work, warp_arrive);
        CK(cudaEventRecord(e1, st)); CK(cudaStreamSynchronize(st));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
    }
    CK(cudaStreamDestroy(st));
    return best * 1e3;
}

int main() {
    const size_t big = (size_t)3 << 30;
    char *buf; float *x0, *x1; unsigned *counter; uint4 *sink;
    CK(cudaMalloc(&buf, big)); CK(cudaMalloc(&x0, 16384)); CK(cudaMalloc(&x1, 16384)); CK(cudaMalloc(&counter, 64)); CK(cudaMalloc(&sink, 64));
    CK(cudaMemset(buf, 1, big)); CK(cudaMemset(x0, 0, 16384)); CK(cudaMemset(x1, 0, 16384));
    char *flushbuf = buf + ((size_t)2 << 30);
    const int grid = 148;
    struct Mix { const char *name; std::vector<size_t> one; };
    std::vector<Mix> mixes = {
        {"7B layer (28,9.4,50.6,25.3 MB)", {28u << 20, 9856u << 10, 51814u << 10, 25907u << 10}},
        {"uniform 9.4 MB", {9856u << 10, 9856u << 10, 9856u << 10, 9856u << 10}},
        {"uniform 28 MB", {28u << 20, 28u << 20, 28u << 20, 28u << 20}},
        {"TP8 slices (3.5,1.2,6.3,3.2 MB)", {3584u << 10, 1232u << 10, 6476u << 10, 3238u << 10}},
    };
    for (auto &m : mixes) {
        std::vector<size_t> sizes;
        size_t total = 0;
        for (int rep = 0; rep < 10; rep++) for (size_t s : m.one) { const size_t q = (size_t)grid * 16 * SEG; const size_t r = s / q * q; sizes.push_back(r); total += r; }
        printf("== %s: %zu phases, %.1f MB total, ideal at 6.46 TB/s: %.1f us\n", m.name, sizes.size(), total / 1048576.0, total / 6.46e6);
        for (int work : {0, 150}) {
            double us;
            us = run_per_phase<16, 3, 1>(sizes, buf, x0, x1, false, 0, work, grid, sink, flushbuf); printf("  work %3d  V0 full-SM, no PDL            : %8.1f us  %6.0f GB/s\n", work, us, total / us / 1e3);
            us = run_per_phase<16, 3, 1>(sizes, buf, x0, x1, true, 0, work, grid, sink, flushbuf);  printf("  work %3d  V0 full-SM, PDL (late trigger) : %8.1f us  %6.0f GB/s\n", work, us, total / us / 1e3);
            us = run_per_phase<16, 3, 1>(sizes, buf, x0, x1, true, 1, work, grid, sink, flushbuf);  printf("  work %3d  V0 full-SM, PDL (early trigger): %8.1f us  %6.0f GB/s\n", work, us, total / us / 1e3);
            us = run_per_phase<8, 3, 2>(sizes, buf, x0, x1, true, 1, work, grid, sink, flushbuf);   printf("  work %3d  V1 half-SM 8w x 3 x 4K, PDL    : %8.1f us  %6.0f GB/s\n", work, us, total / us / 1e3);
            us = run_per_phase<8, 3, 2>(sizes, buf, x0, x1, false, 0, work, grid, sink, flushbuf);  printf("  work %3d  V1 half-SM 8w x 3 x 4K, no PDL : %8.1f us  %6.0f GB/s\n", work, us, total / us / 1e3);
            us = run_per_phase<8, 3, 2>(sizes, buf, x0, x1, true, 1, work, 2 * grid, sink, flushbuf); printf("  work %3d  V1 half-SM, grid 296, PDL      : %8.1f us  %6.0f GB/s\n", work, us, total / us / 1e3);
            us = run_per_phase<4, 6, 2>(sizes, buf, x0, x1, true, 1, work, grid, sink, flushbuf);  printf("  work %3d  V1 4w x 6 x 4K (96 KB), PDL    : %8.1f us  %6.0f GB/s\n", work, us, total / us / 1e3);
            us = run_chain<16, 3>(sizes, buf, x0, x1, counter, work, 0, grid, sink, flushbuf);      printf("  work %3d  V2 persistent, CTA arrive      : %8.1f us  %6.0f GB/s\n", work, us, total / us / 1e3);
            us = run_chain<16, 3>(sizes, buf, x0, x1, counter, work, 1, grid, sink, flushbuf);      printf("  work %3d  V2 persistent, warp arrive     : %8.1f us  %6.0f GB/s\n", work, us, total / us / 1e3);
        }
    }
    return 0;
}
