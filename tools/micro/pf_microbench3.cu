// Developer micro-benchmark 3: warming L2 with REAL loads that carry an L2 prefetch-size hint (one 4-byte load per 128 / 256 bytes),
// and a persistent "L2 warmer" kernel running CONCURRENTLY with a chain of dependent streaming kernels, paced by a progress word.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
constexpr int NW = 16, NS = 3, SEG = 4096;

__global__ void __launch_bounds__(512, 1) consumer(const char *base, size_t bytes, unsigned long long *tstamp, unsigned long long *progress, unsigned long long pos) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ unsigned long long bars[NW][NS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { if (tstamp) tstamp[2 * blockIdx.x] = gtime(); if (progress && blockIdx.x == 0) *(volatile unsigned long long *)progress = pos; }
    const size_t per_cta = bytes / gridDim.x / SEG * SEG, per_warp = per_cta / NW / SEG * SEG;
    const char *src = base + (size_t)blockIdx.x * per_cta + (size_t)warp * per_warp;
    const int nseg = (int)(per_warp / SEG);
    unsigned char *ring = smem + (size_t)warp * NS * SEG;
    if (lane == 0) {
        for (int s = 0; s < NS; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[warp][s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    auto issue = [&](int i) {
        if (lane == 0) {
            const unsigned b = smem_u32(&bars[warp][i % NS]);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(SEG) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(ring + (i % NS) * SEG)),
                         "l"(src + (size_t)i * SEG), "r"(SEG), "r"(b) : "memory");
        }
    };
    for (int i = 0; i < NS && i < nseg; i++) issue(i);
    unsigned acc = 0;
    for (int i = 0; i < nseg; i++) {
        const unsigned b = smem_u32(&bars[warp][i % NS]), ph = (i / NS) & 1;
        asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(b), "r"(ph) : "memory");
        acc += reinterpret_cast<const unsigned *>(ring + (i % NS) * SEG)[lane];
        __syncwarp();
        if (lane == 0) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        if (i + NS < nseg) issue(i + NS);
    }
    if (acc == 0x12345u && tstamp) tstamp[0] = 0;
    __syncthreads();
    if (threadIdx.x == 0 && tstamp) tstamp[2 * blockIdx.x + 1] = gtime();
}
template <int MODE>  // 0: plain u32 per 128 B; 1: L2::128B per 128 B; 2: L2::256B per 256 B; 3: plain u32 per 256 B
__global__ void touch(const char *base, size_t bytes, unsigned *sink) {
    constexpr size_t STEP = (MODE == 2 || MODE == 3) ? 256 : 128;
    unsigned a = 0;
    for (size_t o = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * STEP; o < bytes; o += (size_t)gridDim.x * blockDim.x * STEP) {
        unsigned v;
        if (MODE == 0 || MODE == 3) asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(base + o));
        else if (MODE == 1) asm volatile("ld.global.cg.L2::128B.u32 %0, [%1];" : "=r"(v) : "l"(base + o));
        else asm volatile("ld.global.cg.L2::256B.u32 %0, [%1];" : "=r"(v) : "l"(base + o));
        a ^= v;
    }
    if (a == 0x1234567u) *sink = a;
}
// persistent warmer: CTA b handles 64 KB chunks c = b, b+G, ... of the position space [0, total); lane: 8 loads x 256 B apart
__global__ void __launch_bounds__(32, 1) warmer(const char *base, unsigned long long total, const unsigned long long *progress, unsigned long long window,
                                                unsigned long long min_ahead, unsigned *sink, unsigned long long *stats) {
    constexpr unsigned long long CH = 65536;
    const int lane = threadIdx.x;
    unsigned acc = 0;
    unsigned long long issued = 0, skipped = 0, t_last = gtime(), last_prog = 0;
    for (unsigned long long pos = blockIdx.x * CH; pos < total; pos += (unsigned long long)gridDim.x * CH) {
        unsigned long long prog;
        for (;;) {
            prog = *(const volatile unsigned long long *)progress;
            if (pos < prog + window) break;
            const unsigned long long now = gtime();
            if (prog != last_prog) { last_prog = prog; t_last = now; }
            if (now - t_last > 200000ull) { pos = total; break; }  // 200 us without progress: give up
            __nanosleep(100);
        }
        if (pos >= total) break;
        if (pos < prog + min_ahead) { skipped++; continue; }
        const char *p = base + pos + lane * 256;
        unsigned v[8];
#pragma unroll
        for (int k = 0; k < 8; k++) asm volatile("ld.global.cg.L2::256B.u32 %0, [%1];" : "=r"(v[k]) : "l"(p + k * 8192));
#pragma unroll
        for (int k = 0; k < 8; k++) acc ^= v[k];
        issued++;
    }
    if (acc == 0x1234567u) *sink = acc;
    if (lane == 0 && stats) { atomicAdd(stats, issued); atomicAdd(stats + 1, skipped); }
}
__global__ void spin(long long cycles) { const long long t0 = clock64(); while (clock64() - t0 < cycles) {} }
__global__ void flushk(const uint4 *p, size_t n16, uint4 *sink) {
    uint4 a = make_uint4(0, 0, 0, 0);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) { const uint4 v = __ldcg(p + i); a.x ^= v.x; a.y ^= v.y; }
    if ((a.x ^ a.y) == 0x12345678u) *sink = a;
}

int main() {
    const size_t big = (size_t)3 << 30;
    char *buf; unsigned long long *ts, *progress, *stats; uint4 *sink;
    CK(cudaMalloc(&buf, big)); CK(cudaMalloc(&ts, 148 * 16)); CK(cudaMalloc(&sink, 64)); CK(cudaMalloc(&progress, 64)); CK(cudaMalloc(&stats, 64));
    CK(cudaMemset(buf, 1, big));
    CK(cudaFuncSetAttribute(consumer, cudaFuncAttributeMaxDynamicSharedMemorySize, NW * NS * SEG));
    if (getenv("CARVE")) {
        CK(cudaFuncSetAttribute(warmer, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        CK(cudaFuncSetAttribute(spin, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        CK(cudaFuncSetAttribute(consumer, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        printf("carve-out: max shared on every kernel\n");
    }
    char *flushbuf = buf + ((size_t)2 << 30);
    auto flush = [&]() { flushk<<<148 * 4, 512>>>((const uint4 *)flushbuf, (400u << 20) / 16, sink); CK(cudaDeviceSynchronize()); };
    auto run_consumer = [&](size_t bytes) {
        consumer<<<148, 512, NW * NS * SEG>>>(buf, bytes, ts, nullptr, 0);
        CK(cudaDeviceSynchronize());
        unsigned long long h[296]; CK(cudaMemcpy(h, ts, sizeof(h), cudaMemcpyDeviceToHost));
        unsigned long long t0 = ~0ull, t1 = 0;
        for (int i = 0; i < 148; i++) { if (h[2 * i] < t0) t0 = h[2 * i]; if (h[2 * i + 1] > t1) t1 = h[2 * i + 1]; }
        return (double)(t1 - t0) / 1e3;
    };
    printf("%-44s %6s %10s %10s\n", "variant", "MB", "cons_us", "GB/s");
    for (int mb : {28}) {
        const size_t bytes = (size_t)mb << 20;
        double us;
        flush(); us = run_consumer(bytes); printf("%-44s %6d %10.1f %10.0f\n", "cold", mb, us, bytes / us / 1e3);
        us = run_consumer(bytes); printf("%-44s %6d %10.1f %10.0f\n", "warm", mb, us, bytes / us / 1e3);
        flush(); touch<0><<<148 * 2, 512>>>(buf, bytes, (unsigned *)sink); spin<<<1, 32>>>(40 * 1900); us = run_consumer(bytes); printf("%-44s %6d %10.1f %10.0f\n", "touch u32 / 128B plain", mb, us, bytes / us / 1e3);
        flush(); touch<1><<<148 * 2, 512>>>(buf, bytes, (unsigned *)sink); spin<<<1, 32>>>(40 * 1900); us = run_consumer(bytes); printf("%-44s %6d %10.1f %10.0f\n", "touch u32 / 128B L2::128B", mb, us, bytes / us / 1e3);
        flush(); touch<2><<<148 * 2, 512>>>(buf, bytes, (unsigned *)sink); spin<<<1, 32>>>(40 * 1900); us = run_consumer(bytes); printf("%-44s %6d %10.1f %10.0f\n", "touch u32 / 256B L2::256B", mb, us, bytes / us / 1e3);
        flush(); touch<3><<<148 * 2, 512>>>(buf, bytes, (unsigned *)sink); spin<<<1, 32>>>(40 * 1900); us = run_consumer(bytes); printf("%-44s %6d %10.1f %10.0f\n", "touch u32 / 256B plain", mb, us, bytes / us / 1e3);
        // time of the touch kernels themselves (events, includes launch overhead)
        cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); float ms;
        flush(); CK(cudaEventRecord(e0)); touch<2><<<148 * 2, 512>>>(buf, bytes, (unsigned *)sink); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("%-44s %6d %10.1f %10.0f\n", "  (touch L2::256B kernel itself, cold)", mb, ms * 1e3, bytes / ms / 1e6);
    }
    // ---- chain with serial gaps, with and without a concurrent warmer
    cudaStream_t sa, sb; CK(cudaStreamCreateWithFlags(&sa, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&sb, cudaStreamNonBlocking));
    cudaEvent_t e0, e1, ef; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreateWithFlags(&ef, cudaEventDisableTiming));
    for (int mb : {9, 28, 48}) {
        for (int gap_us : {0, 3, 6}) {
            const int NK = 40;
            const size_t bytes = (size_t)mb << 20;
            for (int with_warmer = 0; with_warmer < 2; with_warmer++) {
                float best = 1e9;
                for (int rep = 0; rep < 3; rep++) {
                    flush();
                    CK(cudaMemset(progress, 0, 8)); CK(cudaMemset(stats, 0, 16));
                    CK(cudaDeviceSynchronize());
                    CK(cudaEventRecord(e0, sa));
                    if (with_warmer) {
                        CK(cudaEventRecord(ef, sa)); CK(cudaStreamWaitEvent(sb, ef, 0));
                        warmer<<<148, 32, 0, sb>>>(buf, (unsigned long long)NK * bytes, progress, 48ull << 20, 2ull << 20, (unsigned *)sink, stats);
                    }
                    for (int k = 0; k < NK; k++) {
                        if (gap_us) spin<<<1, 32, 0, sa>>>((long long)gap_us * 1900);
                        consumer<<<148, 512, NW * NS * SEG, sa>>>(buf + (size_t)k * bytes, bytes, nullptr, progress, (unsigned long long)k * bytes);
                    }
                    CK(cudaEventRecord(e1, sa));
                    CK(cudaDeviceSynchronize());
                    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
                    if (ms < best) best = ms;
                }
                unsigned long long st[2]; CK(cudaMemcpy(st, stats, 16, cudaMemcpyDeviceToHost));
                printf("chain %2d x %2d MB, gap %d us, warmer %d: %8.1f us total, %6.2f us/kernel, %6.0f GB/s effective (warmer chunks issued %llu skipped %llu)\n", NK, mb, gap_us,
                       with_warmer, best * 1e3, best * 1e3 / NK, (double)NK * bytes / best / 1e6, st[0], st[1]);
            }
        }
    }
    return 0;
}
