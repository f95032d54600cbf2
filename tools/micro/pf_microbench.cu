// Developer micro-benchmark: what does cp.async.bulk.prefetch.L2 (UBLKPF.L2) actually do on B200?
//   A: a kernel that ONLY issues bulk prefetches of X MB -- does the kernel's completion wait for the data? at what rate?
//   B: a streaming read of the same X MB right after (L2 hit rate => time), vs cold
//   C: the same with prefetch.global.L2 (one instruction per 128-byte line) instead of the bulk form
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pf_microbench pf_microbench.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__global__ void pf_bulk(const char *p, size_t bytes, unsigned chunk) {
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t nw = (size_t)gridDim.x * (blockDim.x >> 5), w = (size_t)blockIdx.x * (blockDim.x >> 5) + warp;
    if (lane) return;
    const size_t nch = bytes / chunk;
    for (size_t c = w; c < nch; c += nw) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p + c * chunk), "r"(chunk) : "memory");
}
__global__ void pf_line(const char *p, size_t bytes) {
    const size_t n = bytes / 128, stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + i * 128));
}
__global__ void stream_read(const uint4 *p, size_t n16, uint4 *sink) {
    uint4 a = make_uint4(0, 0, 0, 0);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
        const uint4 v = __ldcg(p + i);
        a.x ^= v.x; a.y ^= v.y; a.z ^= v.z; a.w ^= v.w;
    }
    if ((a.x ^ a.y ^ a.z ^ a.w) == 0x12345678u) *sink = a;
}
__global__ void spin(long long cycles) { const long long t0 = clock64(); while (clock64() - t0 < cycles) {} }

int main() {
    const size_t big = (size_t)1 << 30;
    char *buf; uint4 *sink;
    CK(cudaMalloc(&buf, big)); CK(cudaMalloc(&sink, 64));
    CK(cudaMemset(buf, 1, big));
    cudaEvent_t e0, e1, e2, e3; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreate(&e2)); CK(cudaEventCreate(&e3));
    auto flush = [&]() { stream_read<<<148 * 4, 512>>>((const uint4 *)(buf + (512u << 20)), (400u << 20) / 16, sink); CK(cudaDeviceSynchronize()); };
    float ms;
    printf("%-34s %8s %10s %10s %10s\n", "case", "MB", "issue_us", "read_us", "read_GB/s");
    for (int mb : {4, 16, 32, 48, 64, 96}) {
        const size_t bytes = (size_t)mb << 20;
        // cold read
        flush();
        CK(cudaEventRecord(e0)); stream_read<<<148 * 4, 512>>>((const uint4 *)buf, bytes / 16, sink); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("%-34s %8d %10s %10.1f %10.0f\n", "cold read", mb, "-", ms * 1e3, bytes / ms / 1e6);
        // warm read (just read)
        CK(cudaEventRecord(e0)); stream_read<<<148 * 4, 512>>>((const uint4 *)buf, bytes / 16, sink); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("%-34s %8d %10s %10.1f %10.0f\n", "re-read (L2 warm by reads)", mb, "-", ms * 1e3, bytes / ms / 1e6);
        for (unsigned chunk : {8192u, 65536u}) {
            for (int wait_us : {0, 30}) {
                flush();
                CK(cudaEventRecord(e0));
                pf_bulk<<<148, 512>>>(buf, bytes, chunk);
                CK(cudaEventRecord(e1));
                if (wait_us) spin<<<1, 32>>>((long long)wait_us * 1900);
                CK(cudaEventRecord(e2));
                stream_read<<<148 * 4, 512>>>((const uint4 *)buf, bytes / 16, sink);
                CK(cudaEventRecord(e3)); CK(cudaEventSynchronize(e3));
                float a, b; CK(cudaEventElapsedTime(&a, e0, e1)); CK(cudaEventElapsedTime(&b, e2, e3));
                char name[64]; snprintf(name, sizeof(name), "bulk pf chunk %uK, gap %dus", chunk / 1024, wait_us);
                printf("%-34s %8d %10.1f %10.1f %10.0f\n", name, mb, a * 1e3, b * 1e3, bytes / b / 1e6);
            }
        }
        for (int wait_us : {0, 30}) {
            flush();
            CK(cudaEventRecord(e0)); pf_line<<<148 * 2, 512>>>(buf, bytes); CK(cudaEventRecord(e1));
            if (wait_us) spin<<<1, 32>>>((long long)wait_us * 1900);
            CK(cudaEventRecord(e2)); stream_read<<<148 * 4, 512>>>((const uint4 *)buf, bytes / 16, sink); CK(cudaEventRecord(e3)); CK(cudaEventSynchronize(e3));
            float a, b; CK(cudaEventElapsedTime(&a, e0, e1)); CK(cudaEventElapsedTime(&b, e2, e3));
            char name[64]; snprintf(name, sizeof(name), "prefetch.global.L2 lines, gap %dus", wait_us);
            printf("%-34s %8d %10.1f %10.1f %10.0f\n", name, mb, a * 1e3, b * 1e3, bytes / b / 1e6);
        }
    }
    return 0;
}
