// b200_common.cuh -- shared device helpers for the sm_100a kernels.
//
// Block layouts follow the reference's on-disk == in-memory GGUF formats
// (/root/reference/ggml/src/ggml-common.h:144-335); we address bytes with explicit offsets
// instead of packed structs because several formats are only 2-byte aligned (q4_0 18 B,
// q8_0 34 B, q6_K 210 B) and the kernels read them out of shared memory with aligned-word
// loads + funnel shifts.
#pragma once

#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>

#include "../../include/b200_ggml.h"

namespace b200 {

constexpr int QK   = 32;   // QK4_0 == QK8_0 == QK8_1
constexpr int QK_K = 256;  // super-block

// bytes per block
constexpr int Q4_0_BYTES = 18;
constexpr int Q4_1_BYTES = 20;  // half d, half m, u8 qs[16]          (ggml-common.h:155-161)
constexpr int Q5_0_BYTES = 22;  // half d, u8 qh[4], u8 qs[16]         (:164-169)
constexpr int Q5_1_BYTES = 24;  // half d, half m, u8 qh[4], u8 qs[16] (:172-183)
constexpr int Q8_0_BYTES = 34;
constexpr int Q8_1_BYTES = 36;
constexpr int Q2_K_BYTES = 84;   // u8 scales[16], u8 qs[64], half d, half dmin   (ggml-common.h:257-268)
constexpr int Q3_K_BYTES = 110;  // u8 hmask[32], u8 qs[64], u8 scales[12], half d (:274-280)
constexpr int Q4_K_BYTES = 144;
constexpr int Q5_K_BYTES = 176;
constexpr int Q6_K_BYTES = 210;
constexpr int IQ4_NL_BYTES = 18;   // half d, u8 qs[16]: nibbles index 16 non-uniform int8 levels (ggml-common.h:406-410)
constexpr int IQ4_XS_BYTES = 136;  // half d, u16 scales_h, u8 scales_l[4], u8 qs[128] (:413-418)

__host__ __device__ constexpr int type_block_values(int t) {
    return (t == B200_TYPE_F32 || t == B200_TYPE_F16 || t == B200_TYPE_I32) ? 1
         : (t == B200_TYPE_Q4_0 || t == B200_TYPE_Q8_0 || t == B200_TYPE_Q8_1 || t == B200_TYPE_Q4_1 || t == B200_TYPE_Q5_0 || t == B200_TYPE_Q5_1 ||
            t == B200_TYPE_IQ4_NL) ? 32
         : (t == B200_TYPE_Q4_K || t == B200_TYPE_Q5_K || t == B200_TYPE_Q6_K || t == B200_TYPE_Q2_K || t == B200_TYPE_Q3_K || t == B200_TYPE_IQ4_XS) ? 256 : 0;
}
__host__ __device__ constexpr int type_block_bytes(int t) {
    return t == B200_TYPE_F32 ? 4 : t == B200_TYPE_F16 ? 2 : t == B200_TYPE_I32 ? 4
         : t == B200_TYPE_Q4_0 ? Q4_0_BYTES : t == B200_TYPE_Q8_0 ? Q8_0_BYTES : t == B200_TYPE_Q8_1 ? Q8_1_BYTES
         : t == B200_TYPE_Q4_1 ? Q4_1_BYTES : t == B200_TYPE_Q5_0 ? Q5_0_BYTES : t == B200_TYPE_Q5_1 ? Q5_1_BYTES
         : t == B200_TYPE_Q2_K ? Q2_K_BYTES : t == B200_TYPE_Q3_K ? Q3_K_BYTES
         : t == B200_TYPE_IQ4_NL ? IQ4_NL_BYTES : t == B200_TYPE_IQ4_XS ? IQ4_XS_BYTES
         : t == B200_TYPE_Q4_K ? Q4_K_BYTES : t == B200_TYPE_Q5_K ? Q5_K_BYTES : t == B200_TYPE_Q6_K ? Q6_K_BYTES : 0;
}
// bytes of 128 consecutive weights (one "slice": the unit one lane processes per step)
__host__ __device__ constexpr int type_slice_bytes(int t) { return type_block_bytes(t) * 128 / type_block_values(t); }

// ---------------------------------------------------------------- small device utilities
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ int dp4a(int a, int b, int c) { return __dp4a(a, b, c); }
__device__ __forceinline__ int dp4a_u(unsigned a, int b, int c) {  // unsigned a bytes x signed b bytes
    int r;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ float h2f(unsigned short h) { return __half2float(__ushort_as_half(h)); }
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

// word `i` (4 bytes) of a byte run that starts at shared address `base` which is only 2-byte aligned:
// two aligned LDS.32 + one funnel shift.  `base` must be even.
__device__ __forceinline__ unsigned lds_word_unaligned(const unsigned char *base, int i) {
    const uintptr_t a = (uintptr_t)base + 4 * i;
    const unsigned *w = (const unsigned *)(a & ~(uintptr_t)3);
    const unsigned sh = ((unsigned)a & 3u) * 8u;
    return __funnelshift_r(w[0], w[1], sh);  // sh == 0 -> w[0]
}

// consecutive 32-bit words from a shared-memory byte run that is only 2-byte aligned
template <int NW>
__device__ __forceinline__ void lds_words(const unsigned char *p, unsigned (&out)[NW]) {
    const unsigned *w = reinterpret_cast<const unsigned *>((uintptr_t)p & ~(uintptr_t)3);
    const unsigned sh = ((unsigned)(uintptr_t)p & 3u) * 8u;
    unsigned prev = w[0];
#pragma unroll
    for (int i = 0; i < NW; i++) {
        const unsigned nxt = w[i + 1];  // one word of over-read at most: stages carry 16 bytes of slack
        out[i] = __funnelshift_r(prev, nxt, sh);
        prev = nxt;
    }
}

// ---------------------------------------------------------------- q8_1 quantisation of one 32-value block
// held by 8 consecutive lanes (4 values each).  Arithmetic of quantize_row_q8_1_ref (ggml-quants.c:1160-1193):
//   d = amax/127, id = d ? 1/d : 0, q = roundf(x*id), y.d = fp16(d), y.s = fp16(sum(q)*d)
// Returns the lane's 4 quantised values packed little-endian.  d16/s16 are the fp16-rounded d and s converted back
// to float (what a consumer of the stored block sees); sum4 = sum of q over this lane's half (16 values).
__device__ __forceinline__ int quantize_q8_1_group8(float4 v, float &d16, float &s16, int &sum4) {
    float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    const int q0 = (int)roundf(__fmul_rn(v.x, id));
    const int q1 = (int)roundf(__fmul_rn(v.y, id));
    const int q2 = (int)roundf(__fmul_rn(v.z, id));
    const int q3 = (int)roundf(__fmul_rn(v.w, id));
    int s = q0 + q1 + q2 + q3;
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    sum4 = s;
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    d16 = __half2float(__float2half_rn(d));
    s16 = __half2float(__float2half_rn(__fmul_rn((float)s, d)));
    return (q0 & 0xff) | ((q1 & 0xff) << 8) | ((q2 & 0xff) << 16) | (int)((unsigned)q3 << 24);
}

// ---------------------------------------------------------------- mbarrier + bulk async copy (TMA engine, SASS UBLKCP)
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// global -> shared bulk copy; dst/src 16-byte aligned, bytes % 16 == 0; completes on `bar` (complete_tx)
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// order prior generic-proxy accesses to shared memory before subsequent async-proxy (bulk copy) writes
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// programmatic dependent launch: wait for the producer grid / allow the consumer grid to start
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ---------------------------------------------------------------- cp.async (LDGSTS)
__device__ __forceinline__ void cp_async16(void *dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8(void *dst, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(void *dst, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// ---------------------------------------------------------------- 6-bit scale/min unpack of q4_K / q5_K
// (get_scale_min_k4, ggml-quants.c:1899-1907) for sub-block j in [0,8) from the 12 scale bytes given as 3 words.
__device__ __forceinline__ void k4_scale_min(const unsigned sw[3], int j, int &sc, int &mn) {
    const unsigned char *s = (const unsigned char *)sw;
    if (j < 4) {
        sc = s[j] & 63;
        mn = s[j + 4] & 63;
    } else {
        sc = (s[j + 4] & 0x0F) | ((s[j - 4] >> 6) << 4);
        mn = (s[j + 4] >> 4) | ((s[j] >> 6) << 4);
    }
}

// ---------------------------------------------------------------- 6-bit scale unpack of q3_K
// The aux[] shuffle of dequantize_row_q3_K (ggml-quants.c:2345-2350): 12 scale bytes given as 3 words -> 16 scales, one per
// byte of aux[0..3], each still carrying the +32 offset of the stored form.
__device__ __forceinline__ void q3k_scales(const unsigned s0, const unsigned s1, const unsigned s2, unsigned (&aux)[4]) {
    const unsigned kmask1 = 0x03030303u, kmask2 = 0x0f0f0f0fu;
    aux[0] = (s0 & kmask2) | (((s2 >> 0) & kmask1) << 4);
    aux[1] = (s1 & kmask2) | (((s2 >> 2) & kmask1) << 4);
    aux[2] = ((s0 >> 4) & kmask2) | (((s2 >> 4) & kmask1) << 4);
    aux[3] = ((s1 >> 4) & kmask2) | (((s2 >> 6) & kmask1) << 4);
}

// ---------------------------------------------------------------- IQ4_NL / IQ4_XS codebook
// A nibble indexes 16 non-uniform int8 levels (kvalues_iq4nl, ggml-common.h): -127 -104 -83 -65 -49 -35 -22 -10 1 13 25 38 53 69 89
// 113, held here as four little-endian words.  iq4_levels() looks up the eight nibbles of one word with the byte permute unit
// (no memory table): PRMT selects among 8 table bytes with the low 3 bits of a selector nibble, once in the lower and once in
// the upper half of the table; bit 3 of each nibble then picks between the two results through a byte mask (PRMT's sign
// replication).  lo = levels of the low nibbles of bytes 0..3 (values j .. j+3 of a 32-value block), hi = those of the high
// nibbles (values j+16 .. j+19), as int8x4 ready for dp4a.
constexpr unsigned IQ4_T0 = 0xBFAD9881u, IQ4_T1 = 0xF6EADDCFu, IQ4_T2 = 0x26190D01u, IQ4_T3 = 0x71594535u;
__device__ __forceinline__ unsigned prmt(unsigned a, unsigned b, unsigned sel) {
    unsigned r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
}
__device__ __forceinline__ void iq4_levels(unsigned q, unsigned &lo, unsigned &hi) {
    const unsigned s = q & 0x77777777u, s2 = s >> 16;  // selector nibbles: [b0 lo, b0 hi, b1 lo, b1 hi], [b2 lo, b2 hi, b3 lo, b3 hi]
    const unsigned la = prmt(IQ4_T0, IQ4_T1, s), lb = prmt(IQ4_T0, IQ4_T1, s2);
    const unsigned ha = prmt(IQ4_T2, IQ4_T3, s), hb = prmt(IQ4_T2, IQ4_T3, s2);
    const unsigned l_lo = prmt(la, lb, 0x6420u), l_hi = prmt(la, lb, 0x7531u);  // table half 0..7
    const unsigned h_lo = prmt(ha, hb, 0x6420u), h_hi = prmt(ha, hb, 0x7531u);  // table half 8..15
    const unsigned m_lo = prmt(q << 4, 0u, 0xBA98u), m_hi = prmt(q, 0u, 0xBA98u);  // 0xFF where bit 3 of the nibble is set
    lo = l_lo ^ ((l_lo ^ h_lo) & m_lo);
    hi = l_hi ^ ((l_hi ^ h_hi) & m_hi);
}
// one level by index (0..15), for the scalar paths
__device__ __forceinline__ int iq4_level(unsigned n) {
    const unsigned x = (n & 8u) ? prmt(IQ4_T2, IQ4_T3, n & 7u) : prmt(IQ4_T0, IQ4_T1, n & 7u);
    return (int)(signed char)(x & 0xffu);
}

// last error text (host side, per thread)
void set_error(const char *fmt, ...);
int  cuda_fail(cudaError_t e, const char *what);

#define B200_CUDA_TRY(call)                                        \
    do {                                                           \
        cudaError_t e__ = (call);                                  \
        if (e__ != cudaSuccess) return ::b200::cuda_fail(e__, #call); \
    } while (0)

}  // namespace b200
