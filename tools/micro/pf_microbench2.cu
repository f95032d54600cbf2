// Developer micro-benchmark 2: is data pulled into L2 by a prefetch kernel actually HIT by a TMA-bulk streaming consumer?
// consumer: 148 CTAs x 16 warps, every warp streams its contiguous share through a 4-stage smem ring with cp.async.bulk (like mmvq).
// variants: cold | warm (same consumer ran just before) | bulk prefetch, same CTA->bytes mapping | bulk prefetch, CTA mapping shifted by 74 |
//           line prefetch (prefetch.global.L2) | plain LDG touch kernel (reads 1 word per 128B line)
// Timing: globaltimer inside the consumer (min start / max end over CTAs).
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
constexpr int NW = 16, NS = 3, SEG = 4096;

__global__ void __launch_bounds__(512, 1) consumer(const char *base, size_t bytes, unsigned long long *tstamp, unsigned *smids) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ unsigned long long bars[NW][NS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { tstamp[2 * blockIdx.x] = gtime(); unsigned s; asm("mov.u32 %0, %%smid;" : "=r"(s)); smids[blockIdx.x] = s; }
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
    if (acc == 0x12345u) tstamp[0] = 0;
    __syncthreads();
    if (threadIdx.x == 0) tstamp[2 * blockIdx.x + 1] = gtime();
}
__global__ void pf_bulk(const char *base, size_t bytes, int shift) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane) return;
    const size_t per_cta = bytes / gridDim.x / SEG * SEG, per_warp = per_cta / NW / SEG * SEG;
    const int cta = (blockIdx.x + shift) % gridDim.x;
    const char *src = base + (size_t)cta * per_cta + (size_t)warp * per_warp;
    for (size_t o = 0; o < per_warp; o += 8192) {
        const unsigned len = per_warp - o < 8192 ? (unsigned)(per_warp - o) : 8192u;
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src + o), "r"(len) : "memory");
    }
}
// paced: every warp issues one chunk, then sleeps `ns` nanoseconds
__global__ void pf_bulk_paced(const char *base, size_t bytes, unsigned chunk, unsigned ns) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane) return;
    const size_t per_cta = bytes / gridDim.x / SEG * SEG, per_warp = per_cta / NW / SEG * SEG;
    const char *src = base + (size_t)blockIdx.x * per_cta + (size_t)warp * per_warp;
    for (size_t o = 0; o < per_warp; o += chunk) {
        const unsigned len = per_warp - o < chunk ? (unsigned)(per_warp - o) : chunk;
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src + o), "r"(len) : "memory");
        if (ns) __nanosleep(ns);
    }
}
__global__ void pf_line(const char *base, size_t bytes, int shift) {
    const size_t per_cta = bytes / gridDim.x / SEG * SEG;
    const int cta = (blockIdx.x + shift) % gridDim.x;
    const char *src = base + (size_t)cta * per_cta;
    for (size_t o = (size_t)threadIdx.x * 128; o < per_cta; o += (size_t)blockDim.x * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(src + o));
}
__global__ void touch(const char *base, size_t bytes, int shift, unsigned *sink) {
    const size_t per_cta = bytes / gridDim.x / SEG * SEG;
    const int cta = (blockIdx.x + shift) % gridDim.x;
    const char *src = base + (size_t)cta * per_cta;
    unsigned a = 0;
    for (size_t o = (size_t)threadIdx.x * 32; o < per_cta; o += (size_t)blockDim.x * 32) a ^= __ldcg((const unsigned *)(src + o));
    if (a == 0x1234567u) *sink = a;
}
__global__ void spin(long long cycles) { const long long t0 = clock64(); while (clock64() - t0 < cycles) {} }
__global__ void flushk(const uint4 *p, size_t n16, uint4 *sink) {
    uint4 a = make_uint4(0, 0, 0, 0);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) { const uint4 v = __ldcg(p + i); a.x ^= v.x; a.y ^= v.y; }
    if ((a.x ^ a.y) == 0x12345678u) *sink = a;
}

int main() {
    const size_t big = (size_t)1 << 30;
    char *buf; unsigned long long *ts; unsigned *smids; uint4 *sink;
    CK(cudaMalloc(&buf, big)); CK(cudaMalloc(&ts, 148 * 16)); CK(cudaMalloc(&smids, 148 * 4)); CK(cudaMalloc(&sink, 64));
    CK(cudaMemset(buf, 1, big));
    CK(cudaFuncSetAttribute(consumer, cudaFuncAttributeMaxDynamicSharedMemorySize, NW * NS * SEG));
    auto flush = [&]() { flushk<<<148 * 4, 512>>>((const uint4 *)(buf + (512u << 20)), (400u << 20) / 16, sink); CK(cudaDeviceSynchronize()); };
    auto run_consumer = [&](size_t bytes) {
        consumer<<<148, 512, NW * NS * SEG>>>(buf, bytes, ts, smids);
        CK(cudaDeviceSynchronize());
        unsigned long long h[296]; CK(cudaMemcpy(h, ts, sizeof(h), cudaMemcpyDeviceToHost));
        unsigned long long t0 = ~0ull, t1 = 0;
        for (int i = 0; i < 148; i++) { if (h[2 * i] < t0) t0 = h[2 * i]; if (h[2 * i + 1] > t1) t1 = h[2 * i + 1]; }
        return (double)(t1 - t0) / 1e3;
    };
    printf("%-40s %6s %10s %10s\n", "variant", "MB", "cons_us", "GB/s");
    for (int mb : {28, 50}) {
        const size_t bytes = (size_t)mb << 20;
        for (unsigned chunk : {1024u, 4096u, 16384u}) {
            for (unsigned total_us : {0u, 5u, 10u, 20u, 40u}) {
                // per warp: per_warp/chunk issues; sleep so that the whole prefetch takes ~total_us
                const size_t per_warp = bytes / 148 / 16;
                const unsigned n_issue = (unsigned)(per_warp / chunk);
                const unsigned ns = total_us ? total_us * 1000u / (n_issue ? n_issue : 1) : 0;
                flush(); pf_bulk_paced<<<148, 512>>>(buf, bytes, chunk, ns); spin<<<1, 32>>>(60 * 1900); double us = run_consumer(bytes);
                char nm[64]; snprintf(nm, 64, "paced bulk pf chunk %u over ~%uus", chunk, total_us); printf("%-40s %6d %10.1f %10.0f\n", nm, mb, us, bytes / us / 1e3);
            }
        }
    }
    for (int mb : {28}) {
        const size_t bytes = (size_t)mb << 20;
        for (int rep = 0; rep < 2; rep++) {
            double us;
            flush(); us = run_consumer(bytes); printf("%-40s %6d %10.1f %10.0f\n", "cold", mb, us, bytes / us / 1e3);
            us = run_consumer(bytes); printf("%-40s %6d %10.1f %10.0f\n", "warm (consumer re-run)", mb, us, bytes / us / 1e3);
            for (int shift : {0, 74, 1}) {
                flush(); pf_bulk<<<148, 512>>>(buf, bytes, shift); spin<<<1, 32>>>(40 * 1900); us = run_consumer(bytes);
                char nm[64]; snprintf(nm, 64, "bulk prefetch, cta shift %d, gap 40us", shift); printf("%-40s %6d %10.1f %10.0f\n", nm, mb, us, bytes / us / 1e3);
            }
            flush(); pf_bulk<<<148, 512>>>(buf, bytes, 0); us = run_consumer(bytes); printf("%-40s %6d %10.1f %10.0f\n", "bulk prefetch, shift 0, no gap", mb, us, bytes / us / 1e3);
            for (int shift : {0, 74}) {
                flush(); pf_line<<<148, 512>>>(buf, bytes, shift); spin<<<1, 32>>>(40 * 1900); us = run_consumer(bytes);
                char nm[64]; snprintf(nm, 64, "line prefetch, cta shift %d, gap 40us", shift); printf("%-40s %6d %10.1f %10.0f\n", nm, mb, us, bytes / us / 1e3);
                flush(); touch<<<148, 512>>>(buf, bytes, shift, (unsigned *)sink); spin<<<1, 32>>>(40 * 1900); us = run_consumer(bytes);
                snprintf(nm, 64, "LDG touch, cta shift %d, gap 40us", shift); printf("%-40s %6d %10.1f %10.0f\n", nm, mb, us, bytes / us / 1e3);
            }
        }
    }
    unsigned hs[148]; CK(cudaMemcpy(hs, smids, sizeof(hs), cudaMemcpyDeviceToHost));
    printf("smid of CTA 0..15:"); for (int i = 0; i < 16; i++) printf(" %u", hs[i]); printf("\n");
    unsigned mx = 0; for (int i = 0; i < 148; i++) if (hs[i] > mx) mx = hs[i]; printf("max smid %u\n", mx);
    return 0;
}
