// mmq.cu -- (a7) prefill GEMM over GGUF block formats on the 5th-generation tensor cores (tcgen05 + TMEM).
//
//   dst[n][m] = sum_k W[m][k] * X[n][k],  W quantised (Q4_0, Q8_0, Q4_K, Q5_K, Q6_K), X f32, any N.
//
// Replaces ggml_cuda_op_mul_mat_q / mul_mat_q<type,...> (mma.sync m16n8k32 int8 tiles with ALU-side unpack and two CTA
// barriers per 128-K half step, /root/reference/ggml/src/ggml-cuda/mmq.cuh:2495-2905) and the cuBLAS dequantise-then-GEMM
// fallback that koboldcpp uses by default (ggml-cuda.cu:1186-1284: whole-tensor dequant to f16 in HBM, then cublasGemmEx).
//
// B200 design (the one dense contraction of the path: tensor-pipe bound once N is a few hundred):
//   * weights are dequantised ON THE FLY into bf16 operand tiles in shared memory (never to HBM) by 4 producer warps --
//     thread t owns weight row t of the 128-row tile and writes its 64-value K-step straight into the 128-byte-swizzled
//     K-major layout the tensor core reads; activations are converted f32 -> bf16 once per call by a pre-pass and staged
//     with 16-byte cp.async into the same swizzled layout;
//   * one elected thread issues tcgen05.mma (kind::f16, bf16 x bf16 -> f32, M=128, N<=256, K=16 x4 per stage); the f32
//     accumulator tile lives in TMEM (N columns x 128 lanes), so the producers' registers are free for unpacking;
//   * 4-stage shared-memory ring with mbarrier full/empty pairs; tcgen05.commit releases a stage when the MMAs that read
//     it retire -- no CTA-wide barrier in the K loop;
//   * epilogue: the same 4 warps read their TMEM lane quarter (tcgen05.ld 32x32b) and store f32 rows coalesced along m.
// Why bf16 and not kind::i8: per-32 (per-16 for Q6_K) group scales would force a TMEM read-out after every K=32 MMA,
// which is epilogue-bound; bf16 operands keep the whole K loop inside the tensor pipe.  Activations carry 8 mantissa bits
// (q8_1 carries 7 + sign per 32-block), weights are rounded once from their exact f32 dequantised value.
#include "b200_common.cuh"

#include <cuda_bf16.h>

#include <cstdlib>

namespace b200 {

constexpr int MQ_BM = 128, MQ_BK = 64, MQ_STAGES = 4;
constexpr int MQ_PRODUCERS = 128;              // threads; warps 0..3
constexpr int MQ_THREADS = MQ_PRODUCERS + 32;  // + the MMA warp

// ---------------------------------------------------------------- tcgen05 / TMEM wrappers (PTX ISA 8.6+, sm_100a)
__device__ __forceinline__ void tmem_alloc(unsigned *smem_dst, unsigned ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(unsigned taddr, unsigned ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(unsigned long long *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_mma_bf16(unsigned tmem_d, unsigned long long adesc, unsigned long long bdesc, unsigned idesc, unsigned accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 32 consecutive f32 accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld32(unsigned taddr, float (&v)[32]) {
    unsigned r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
          "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
          "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]),
          "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]);
}

// UMMA shared-memory descriptor, K-major operand tile with 128-byte swizzle (cute/arch/mma_sm100_desc.hpp SmemDescriptor):
// start address >> 4 in [0,14), LBO [16,30) (unused for swizzled K-major: 1), SBO [32,46) = 1024 B >> 4 (8 rows x 128 B),
// version [46,48) = 1 (sm_100), layout [61,64) = 2 (SWIZZLE_128B).
__device__ __forceinline__ unsigned long long umma_desc_sw128(unsigned smem_addr) {
    unsigned long long d = 0;
    d |= (unsigned long long)((smem_addr & 0x3FFFFu) >> 4);
    d |= (unsigned long long)1 << 16;
    d |= (unsigned long long)(1024 >> 4) << 32;
    d |= (unsigned long long)1 << 46;
    d |= (unsigned long long)2 << 61;
    return d;
}
// instruction descriptor (InstrDescriptor): D f32 [4,6)=1, A bf16 [7,10)=1, B bf16 [10,13)=1, both K-major, N>>3 [17,23), M>>4 [24,29)
__host__ __device__ constexpr unsigned umma_idesc_bf16(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((unsigned)(N >> 3) << 17) | ((unsigned)(M >> 4) << 24);
}

// ---------------------------------------------------------------- raw block loads from 2-byte aligned global addresses
// G = true: global memory through the read-only path; false: shared memory (raw blocks staged by the bulk-copy engine)
template <bool G> __device__ __forceinline__ unsigned ld32(const unsigned *q) { if constexpr (G) return __ldg(q); else return *q; }
template <bool G> __device__ __forceinline__ uint4 ld128(const uint4 *q) { if constexpr (G) return __ldg(q); else return *q; }
template <bool G, int NW>
__device__ __forceinline__ void ld_words(const unsigned char *p, unsigned (&w)[NW]) {
    const uintptr_t a = (uintptr_t)p;
    const unsigned *q = reinterpret_cast<const unsigned *>(a & ~(uintptr_t)3);
    const unsigned sh = (unsigned)(a & 3) * 8;
    if (sh == 0) {
#pragma unroll
        for (int i = 0; i < NW; i++) w[i] = ld32<G>(q + i);
    } else {
        unsigned prev = ld32<G>(q);
#pragma unroll
        for (int i = 0; i < NW; i++) {
            const unsigned nxt = ld32<G>(q + i + 1);
            w[i] = __funnelshift_r(prev, nxt, sh);
            prev = nxt;
        }
    }
}
__device__ __forceinline__ float f16lo(unsigned w) { return __half2float(__ushort_as_half((unsigned short)(w & 0xffff))); }
__device__ __forceinline__ float f16hi(unsigned w) { return __half2float(__ushort_as_half((unsigned short)(w >> 16))); }
__device__ __forceinline__ unsigned pack_bf16(float a, float b) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const unsigned *>(&h);
}

// ---------------------------------------------------------------- 64 consecutive weights of one row -> 32 packed bf16x2 words
// `row` points at the row's first block, kb is the 64-value step.  Arithmetic follows dequantize_row_* (ggml-quants.c:1523,
// 1617, 2556, 2764, 2978) in f32; the result is rounded once to bf16.
#define BYTE_OF(w, i) (((w)[(i) >> 2] >> (8 * ((i) & 3))) & 0xffu)

// bytes of the 256-value group a K-step of 64 belongs to (one K-quant super-block, or 8 q4_0 / q8_0 blocks)
__host__ __device__ constexpr int group256_bytes(int t) { return type_block_bytes(t) * 256 / type_block_values(t); }

template <int TYPE, bool G>
struct Dequant64;  // run(g, c, o): g -> first byte of the 256-value group, c = 64-value chunk 0..3 inside it

template <bool G>
struct Dequant64<B200_TYPE_Q4_0, G> {
    static __device__ __forceinline__ void run(const unsigned char *g, int c, unsigned (&o)[32]) {
        unsigned w[9];
        ld_words<G, 9>(g + c * 2 * Q4_0_BYTES, w);
#pragma unroll
        for (int b = 0; b < 2; b++) {
            const float d = __half2float(__ushort_as_half((unsigned short)(BYTE_OF(w, 18 * b) | (BYTE_OF(w, 18 * b + 1) << 8))));
#pragma unroll
            for (int j = 0; j < 16; j += 2) {
                const unsigned q0 = BYTE_OF(w, 18 * b + 2 + j), q1 = BYTE_OF(w, 18 * b + 3 + j);
                o[16 * b + j / 2]     = pack_bf16(d * (float)((int)(q0 & 15) - 8), d * (float)((int)(q1 & 15) - 8));
                o[16 * b + 8 + j / 2] = pack_bf16(d * (float)((int)(q0 >> 4) - 8), d * (float)((int)(q1 >> 4) - 8));
            }
        }
    }
};

// Q4_1 / Q5_0 / Q5_1 (dequantize_row_q4_1 / q5_0 / q5_1, ggml-quants.c:1543-1615): x d + m, or (x5 - 16) d; the fifth bit of value j
// is bit j of qh.  Byte offsets inside a block: d 0, [m 2], [qh 4 bytes], qs 16 bytes.
template <int TYPE>
struct LegacyLayout {
    static constexpr bool HAS_M = TYPE != B200_TYPE_Q5_0, HAS_H = TYPE != B200_TYPE_Q4_1;
    static constexpr int BB = type_block_bytes(TYPE), QH = 2 + (HAS_M ? 2 : 0), QS = QH + (HAS_H ? 4 : 0);
};
template <int TYPE, bool G>
struct Dequant64Legacy {
    static __device__ __forceinline__ void run(const unsigned char *g, int c, unsigned (&o)[32]) {
        using L = LegacyLayout<TYPE>;
        constexpr int NWL = (2 * L::BB + 3) / 4;
        unsigned w[NWL];
        ld_words<G, NWL>(g + c * 2 * L::BB, w);
#pragma unroll
        for (int b = 0; b < 2; b++) {
            const int base = L::BB * b;
            const float d = __half2float(__ushort_as_half((unsigned short)(BYTE_OF(w, base) | (BYTE_OF(w, base + 1) << 8))));
            float m = 0.f;
            if constexpr (L::HAS_M) m = __half2float(__ushort_as_half((unsigned short)(BYTE_OF(w, base + 2) | (BYTE_OF(w, base + 3) << 8))));
            unsigned qh = 0;
            if constexpr (L::HAS_H) qh = BYTE_OF(w, base + L::QH) | (BYTE_OF(w, base + L::QH + 1) << 8) | (BYTE_OF(w, base + L::QH + 2) << 16) | (BYTE_OF(w, base + L::QH + 3) << 24);
            if constexpr (TYPE == B200_TYPE_Q5_0) m = d * -16.0f;
#pragma unroll
            for (int j = 0; j < 16; j += 2) {
                const unsigned q0 = BYTE_OF(w, base + L::QS + j), q1 = BYTE_OF(w, base + L::QS + 1 + j);
                const unsigned l0 = (q0 & 15) | (((qh >> j) & 1u) << 4), l1 = (q1 & 15) | (((qh >> (j + 1)) & 1u) << 4);
                const unsigned h0 = (q0 >> 4) | (((qh >> (16 + j)) & 1u) << 4), h1 = (q1 >> 4) | (((qh >> (17 + j)) & 1u) << 4);
                o[16 * b + j / 2]     = pack_bf16(d * (float)l0 + m, d * (float)l1 + m);
                o[16 * b + 8 + j / 2] = pack_bf16(d * (float)h0 + m, d * (float)h1 + m);
            }
        }
    }
};
template <bool G> struct Dequant64<B200_TYPE_Q4_1, G> : Dequant64Legacy<B200_TYPE_Q4_1, G> {};
template <bool G> struct Dequant64<B200_TYPE_Q5_0, G> : Dequant64Legacy<B200_TYPE_Q5_0, G> {};
template <bool G> struct Dequant64<B200_TYPE_Q5_1, G> : Dequant64Legacy<B200_TYPE_Q5_1, G> {};

// IQ4_NL / IQ4_XS (dequantize_row_iq4_nl / iq4_xs, ggml-quants.c:3743-3785): dl * level[nibble], the levels looked up eight at a
// time with the byte permute unit (iq4_levels, b200_common.cuh).  q(k) must yield byte k of the sub-block's 16 nibble bytes;
// o16[0..7] receive values 0..15 (low nibbles), o16[8..15] values 16..31 (high nibbles).
template <int K> __device__ __forceinline__ float s8_f(unsigned x) { return (float)(int)(signed char)((x >> (8 * K)) & 0xffu); }
template <typename QFn>
__device__ __forceinline__ void iq4_unpack32(QFn q, float dl, unsigned *o16) {
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const unsigned w = q(4 * i) | (q(4 * i + 1) << 8) | (q(4 * i + 2) << 16) | (q(4 * i + 3) << 24);
        unsigned lo, hi;
        iq4_levels(w, lo, hi);
        o16[2 * i]         = pack_bf16(dl * s8_f<0>(lo), dl * s8_f<1>(lo));
        o16[2 * i + 1]     = pack_bf16(dl * s8_f<2>(lo), dl * s8_f<3>(lo));
        o16[8 + 2 * i]     = pack_bf16(dl * s8_f<0>(hi), dl * s8_f<1>(hi));
        o16[8 + 2 * i + 1] = pack_bf16(dl * s8_f<2>(hi), dl * s8_f<3>(hi));
    }
}
// 6-bit scale of IQ4_XS sub-block ib, minus 32: nibble ib of scales_l (as one word) | bits 2ib, 2ib+1 of scales_h << 4
__device__ __forceinline__ float iq4xs_scale(unsigned sl, unsigned sh, int ib) {
    return (float)((int)(((sl >> (4 * ib)) & 0xFu) | (((sh >> (2 * ib)) & 3u) << 4)) - 32);
}
template <bool G>
struct Dequant64<B200_TYPE_IQ4_NL, G> {
    static __device__ __forceinline__ void run(const unsigned char *g, int c, unsigned (&o)[32]) {
        unsigned w[9];
        ld_words<G, 9>(g + c * 2 * IQ4_NL_BYTES, w);
#pragma unroll
        for (int b = 0; b < 2; b++) {
            const int base = IQ4_NL_BYTES * b;
            const float d = __half2float(__ushort_as_half((unsigned short)(BYTE_OF(w, base) | (BYTE_OF(w, base + 1) << 8))));
            iq4_unpack32([&](int k) { return BYTE_OF(w, base + 2 + k); }, d, o + 16 * b);
        }
    }
};
template <bool G>
struct Dequant64<B200_TYPE_IQ4_XS, G> {
    static __device__ __forceinline__ void run(const unsigned char *blk, int c, unsigned (&o)[32]) {
        unsigned hdr[2], qw[8];
        ld_words<G, 2>(blk, hdr);
        ld_words<G, 8>(blk + 8 + 32 * c, qw);
        const float d = f16lo(hdr[0]);
#pragma unroll
        for (int b = 0; b < 2; b++)
            iq4_unpack32([&](int k) { return BYTE_OF(qw, 16 * b + k); }, d * iq4xs_scale(hdr[1], hdr[0] >> 16, 2 * c + b), o + 16 * b);
    }
};

template <bool G>
struct Dequant64<B200_TYPE_Q8_0, G> {
    static __device__ __forceinline__ void run(const unsigned char *g, int c, unsigned (&o)[32]) {
        unsigned w[17];
        ld_words<G, 17>(g + c * 2 * Q8_0_BYTES, w);
#pragma unroll
        for (int b = 0; b < 2; b++) {
            const float d = __half2float(__ushort_as_half((unsigned short)(BYTE_OF(w, 34 * b) | (BYTE_OF(w, 34 * b + 1) << 8))));
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
                const int q0 = (int)(signed char)BYTE_OF(w, 34 * b + 2 + j), q1 = (int)(signed char)BYTE_OF(w, 34 * b + 3 + j);
                o[16 * b + j / 2] = pack_bf16(d * (float)q0, d * (float)q1);
            }
        }
    }
};

template <int TYPE, bool G>
struct DequantK45 {  // Q4_K / Q5_K: half d, half dmin, u8 scales[12], (u8 qh[32],) u8 qs[128]
    static __device__ __forceinline__ void run(const unsigned char *blk, int c, unsigned (&o)[32]) {
        const uint4 hdr = ld128<G>(reinterpret_cast<const uint4 *>(blk));  // rows are 16-byte aligned (checked on the host)
        const float d = f16lo(hdr.x), dmin = f16hi(hdr.x);
        const unsigned sw[3] = {hdr.y, hdr.z, hdr.w};
        const uint4 *qp = reinterpret_cast<const uint4 *>(blk + (TYPE == B200_TYPE_Q4_K ? 16 : 48) + 32 * c);
        const uint4 qa = ld128<G>(qp), qb = ld128<G>(qp + 1);
        const unsigned q[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
        unsigned h[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if constexpr (TYPE == B200_TYPE_Q5_K) {
            const uint4 ha = ld128<G>(reinterpret_cast<const uint4 *>(blk + 16)), hb = ld128<G>(reinterpret_cast<const uint4 *>(blk + 32));
            h[0] = ha.x; h[1] = ha.y; h[2] = ha.z; h[3] = ha.w; h[4] = hb.x; h[5] = hb.y; h[6] = hb.z; h[7] = hb.w;
        }
#pragma unroll
        for (int hf = 0; hf < 2; hf++) {  // hf = 0: low nibbles (sub-block 2c), 1: high nibbles (sub-block 2c+1)
            const int sb = 2 * c + hf;
            int sc, mn;
            k4_scale_min(sw, sb, sc, mn);
            const float d1 = d * (float)sc, m1 = dmin * (float)mn;
#pragma unroll
            for (int l = 0; l < 32; l += 2) {
                unsigned q0 = (BYTE_OF(q, l) >> (4 * hf)) & 15, q1 = (BYTE_OF(q, l + 1) >> (4 * hf)) & 15;
                if constexpr (TYPE == B200_TYPE_Q5_K) {
                    q0 += ((BYTE_OF(h, l) >> sb) & 1) << 4;
                    q1 += ((BYTE_OF(h, l + 1) >> sb) & 1) << 4;
                }
                o[16 * hf + l / 2] = pack_bf16(d1 * (float)q0 - m1, d1 * (float)q1 - m1);
            }
        }
    }
};
template <bool G> struct Dequant64<B200_TYPE_Q4_K, G> : DequantK45<B200_TYPE_Q4_K, G> {};
template <bool G> struct Dequant64<B200_TYPE_Q5_K, G> : DequantK45<B200_TYPE_Q5_K, G> {};

// Q2_K / Q3_K (dequantize_row_q2_K / q3_K, ggml-quants.c:1980-2010, 2328-2373).  Chunk c (64 values) of a super-block is half
// n = c >> 1, bit fields jb = 2 (c & 1) and jb + 1 of the half's 32 quant bytes; value vv of the chunk: field jb + vv / 32, byte
// vv % 32, scale 8 n + 2 (jb + vv / 32) + (vv % 32) / 16.  W(i) must yield byte i of the block.
template <int TYPE, typename ByteFn>
__device__ __forceinline__ void dequant64_k23(ByteFn W, int c, unsigned (&o)[32]) {
    constexpr bool Q2 = TYPE == B200_TYPE_Q2_K;
    const int n = c >> 1, jb = 2 * (c & 1);
    unsigned aux[4] = {0, 0, 0, 0};
    float d, dmin = 0.f;
    if constexpr (Q2) {
        d = __half2float(__ushort_as_half((unsigned short)(W(80) | (W(81) << 8))));
        dmin = __half2float(__ushort_as_half((unsigned short)(W(82) | (W(83) << 8))));
    } else {
        d = __half2float(__ushort_as_half((unsigned short)(W(108) | (W(109) << 8))));
        unsigned sw[3];
#pragma unroll
        for (int k = 0; k < 3; k++) sw[k] = W(96 + 4 * k) | (W(97 + 4 * k) << 8) | (W(98 + 4 * k) << 16) | (W(99 + 4 * k) << 24);
        q3k_scales(sw[0], sw[1], sw[2], aux);
    }
#pragma unroll
    for (int gq = 0; gq < 4; gq++) {  // 16-value groups of the chunk
        const int jj = jb + (gq >> 1), is = 8 * n + 2 * jj + (gq & 1);
        float dl, ml = 0.f;
        if constexpr (Q2) {
            const unsigned sc = W(is);
            dl = d * (float)(sc & 0xF);
            ml = dmin * (float)(sc >> 4);
        } else {
            dl = d * (float)((int)((aux[is >> 2] >> (8 * (is & 3))) & 0xff) - 32);
        }
#pragma unroll
        for (int e = 0; e < 16; e += 2) {
            const int l = 16 * (gq & 1) + e;
            const unsigned b0 = W((Q2 ? 16 : 32) + 32 * n + l), b1 = W((Q2 ? 16 : 32) + 32 * n + l + 1);
            float x0 = (float)((b0 >> (2 * jj)) & 3), x1 = (float)((b1 >> (2 * jj)) & 3);
            if constexpr (!Q2) {
                x0 -= ((W(l) >> (4 * n + jj)) & 1) ? 0.f : 4.f;
                x1 -= ((W(l + 1) >> (4 * n + jj)) & 1) ? 0.f : 4.f;
            }
            o[8 * gq + e / 2] = pack_bf16(dl * x0 - ml, dl * x1 - ml);
        }
    }
}
template <int TYPE, bool G>
struct Dequant64K23 {
    static __device__ __forceinline__ void run(const unsigned char *blk, int c, unsigned (&o)[32]) {
        constexpr int NWB = (type_block_bytes(TYPE) + 3) / 4;
        unsigned w[NWB];
        ld_words<G, NWB>(blk, w);
        dequant64_k23<TYPE>([&](int i) { return BYTE_OF(w, i); }, c, o);
    }
};
template <bool G> struct Dequant64<B200_TYPE_Q2_K, G> : Dequant64K23<B200_TYPE_Q2_K, G> {};
template <bool G> struct Dequant64<B200_TYPE_Q3_K, G> : Dequant64K23<B200_TYPE_Q3_K, G> {};

template <bool G>
struct Dequant64<B200_TYPE_Q6_K, G> {  // u8 ql[128], u8 qh[64], i8 scales[16], half d   (2-byte aligned blocks)
    static __device__ __forceinline__ void run(const unsigned char *blk, int c, unsigned (&o)[32]) {
        const int half = c >> 1, hi = c & 1;
        unsigned ql[16], qh[8], sc[1], dw[1];
        ld_words<G, 16>(blk + 64 * half, ql);
        ld_words<G, 8>(blk + 128 + 32 * half, qh);
        ld_words<G, 1>(blk + 192 + 8 * half + 4 * hi, sc);
        ld_words<G, 1>(blk + 208, dw);
        const float d = f16lo(dw[0]);
#pragma unroll
        for (int g = 0; g < 4; g++) {  // 16-value groups: jj = 16g .. 16g+15, one int8 scale each
            const float ds = d * (float)(int)(signed char)BYTE_OF(sc, g);
            const int rlow = g >> 1, r = 2 * hi + rlow;
#pragma unroll
            for (int e = 0; e < 16; e += 2) {
                const int l = 16 * (g & 1) + e;
                const unsigned b0 = BYTE_OF(ql, 32 * rlow + l), b1 = BYTE_OF(ql, 32 * rlow + l + 1);
                const unsigned n0 = hi ? (b0 >> 4) : (b0 & 15), n1 = hi ? (b1 >> 4) : (b1 & 15);
                const unsigned h0 = (BYTE_OF(qh, l) >> (2 * r)) & 3, h1 = (BYTE_OF(qh, l + 1) >> (2 * r)) & 3;
                o[8 * g + e / 2] = pack_bf16(ds * (float)((int)(n0 | (h0 << 4)) - 32), ds * (float)((int)(n1 | (h1 << 4)) - 32));
            }
        }
    }
};

// ---------------------------------------------------------------- activation pre-pass: f32 [N][K] -> bf16 [N_pad][K] (pad rows zero)
__global__ void mmq_x_to_bf16_kernel(const float *__restrict__ x, int64_t x_stride, int64_t K, int64_t N, int64_t n_pad, __nv_bfloat16 *__restrict__ out) {
    const int64_t i4 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // float4 index within a row
    const int64_t n = blockIdx.y;
    if (i4 * 4 >= K) return;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (n < N) v = __ldg(reinterpret_cast<const float4 *>(x + n * x_stride) + i4);
    uint2 pk;
    pk.x = pack_bf16(v.x, v.y);
    pk.y = pack_bf16(v.z, v.w);
    *reinterpret_cast<uint2 *>(out + n * K + i4 * 4) = pk;
}

struct MmqParams {
    const unsigned char *W;
    size_t w_row_bytes;
    const __nv_bfloat16 *Xb;  // [n_pad][K]
    float *dst;
    int64_t dst_stride, K, M, N;
    // grouped mode (MUL_MAT_ID batches, b200_mul_mat_id_q): blockIdx.y indexes a device-built tile table {n_tiles, (expert, first
    // sorted row, end row) ...}; W advances by expert_bytes per expert; sorted row i is stored at dst[row_dst[i] + m]
    const int *tile_tab;
    const int *row_dst;
    size_t expert_bytes;
    // split-K (mmq_tc2_kernel, not grouped): blockIdx.z owns a contiguous share of the 256-value groups and ADDS its tile into a
    // zeroed dst (red.global.add.f32).  Only 1 or 2: the sum of two partials does not depend on their arrival order.
    int ksplit;
};

template <int TYPE, int BN>
__global__ void __launch_bounds__(MQ_THREADS, 1) mmq_tc_kernel(const __grid_constant__ MmqParams p) {
    constexpr int A_BYTES = MQ_BM * MQ_BK * 2, B_BYTES = BN * MQ_BK * 2, STAGE_BYTES = A_BYTES + B_BYTES;
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);  // SW128 atoms: 1024-byte aligned
    __shared__ __align__(8) unsigned long long full_bar[MQ_STAGES], empty_bar[MQ_STAGES], tmem_full_bar;
    __shared__ unsigned tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, tid = threadIdx.x;
    const int64_t m0 = (int64_t)blockIdx.x * MQ_BM, n0 = (int64_t)blockIdx.y * BN;
    const int nkb = (int)(p.K / MQ_BK);

    if (tid == 0) {
        for (int s = 0; s < MQ_STAGES; s++) { mbar_init(&full_bar[s], MQ_PRODUCERS); mbar_init(&empty_bar[s], 1); }
        mbar_init(&tmem_full_bar, 1);
        mbar_fence_init();
    }
    if (warp == 4) tmem_alloc(&tmem_base_s, BN);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const unsigned tmem_base = tmem_base_s;

    if (warp < 4) {
        // ===== producers: thread t owns weight row m0 + t and activation rows t, t + 128 of the tile
        const int64_t m = m0 + tid;
        const bool row_ok = m < p.M;
        const unsigned char *wrow = p.W + (size_t)(row_ok ? m : 0) * p.w_row_bytes;
        const unsigned a_off = (unsigned)((tid >> 3) * 1024 + (tid & 7) * 128);
        for (int kb = 0; kb < nkb; kb++) {
            const int s = kb % MQ_STAGES;
            const unsigned ph = (unsigned)(kb / MQ_STAGES) & 1u;
            mbar_wait(&empty_bar[s], ph ^ 1u);
            unsigned char *a_s = smem + (size_t)s * STAGE_BYTES, *b_s = a_s + A_BYTES;
#pragma unroll
            for (int r0 = 0; r0 < BN; r0 += MQ_PRODUCERS) {
                const int r = r0 + tid;
                if (r < BN) {
                    const unsigned char *src = reinterpret_cast<const unsigned char *>(p.Xb + (n0 + r) * p.K + (int64_t)kb * MQ_BK);
                    unsigned char *drow = b_s + (r >> 3) * 1024 + (r & 7) * 128;
#pragma unroll
                    for (int c = 0; c < 8; c++) cp_async16(drow + ((c ^ (r & 7)) << 4), src + 16 * c);
                }
            }
            cp_async_commit();
            unsigned o[32];
            if (row_ok) Dequant64<TYPE, true>::run(wrow + (size_t)(kb >> 2) * group256_bytes(TYPE), kb & 3, o);
            else {
#pragma unroll
                for (int i = 0; i < 32; i++) o[i] = 0u;
            }
#pragma unroll
            for (int c = 0; c < 8; c++)
                *reinterpret_cast<uint4 *>(a_s + a_off + ((c ^ (tid & 7)) << 4)) = make_uint4(o[4 * c], o[4 * c + 1], o[4 * c + 2], o[4 * c + 3]);
            cp_async_wait<0>();
            fence_proxy_async_smem();  // generic-proxy writes (st.shared, cp.async) -> visible to the tensor core's async proxy
            mbar_arrive(&full_bar[s]);
        }
        // ===== epilogue: TMEM lane quarter `warp` -> dst[n][m]
        mbar_wait(&tmem_full_bar, 0u);
        tc_fence_after();
        float *drow = p.dst + m;
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
            float v[32];
            tmem_ld32(tmem_base + ((unsigned)(warp * 32) << 16) + (unsigned)c0, v);
            if (row_ok) {
#pragma unroll
                for (int j = 0; j < 32; j++) {
                    const int64_t n = n0 + c0 + j;
                    if (n < p.N) drow[n * p.dst_stride] = v[j];
                }
            }
        }
        tc_fence_before();
    } else {
        // ===== MMA issuer
        constexpr unsigned idesc = umma_idesc_bf16(MQ_BM, BN);
        for (int kb = 0; kb < nkb; kb++) {
            const int s = kb % MQ_STAGES;
            const unsigned ph = (unsigned)(kb / MQ_STAGES) & 1u;
            mbar_wait(&full_bar[s], ph);
            tc_fence_after();
            if (lane == 0) {
                const unsigned a_addr = smem_u32(smem + (size_t)s * STAGE_BYTES), b_addr = a_addr + A_BYTES;
#pragma unroll
                for (int k = 0; k < MQ_BK / 16; k++) {
                    // 16 bf16 = 32 bytes further along K inside the swizzle atom: +2 in the (address >> 4) field
                    tc_mma_bf16(tmem_base, umma_desc_sw128(a_addr) + 2u * k, umma_desc_sw128(b_addr) + 2u * k, idesc, (kb | k) != 0 ? 1u : 0u);
                }
                tc_commit(&empty_bar[s]);
                if (kb == nkb - 1) tc_commit(&tmem_full_bar);
            }
            __syncwarp();
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        tmem_dealloc(tmem_base, BN);
    }
}

// ---------------------------------------------------------------- register-resident variant (software-pipelined producers)
// A thread keeps the raw bytes of one 256-value group of its row in registers (w[], block start at word 0) and unpacks chunk
// C (compile time) from them.  Small integers become floats without I2F: PRMT drops the byte into the mantissa of 2^23
// (0x4B0000bb = 8388608 + b), one exact FADD removes the bias.
__host__ __device__ constexpr int group256_words(int t) { return (group256_bytes(t) + 3) / 4; }
template <int K> __device__ __forceinline__ float byte_f(unsigned x) {  // float(byte K of x), exact
    return __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7650 + K)) - 8388608.0f;
}
__device__ __forceinline__ float f16_at(const unsigned *w, int byte_off) {  // fp16 at an even byte offset
    return (byte_off & 2) ? f16hi(w[byte_off >> 2]) : f16lo(w[byte_off >> 2]);
}

template <int TYPE, int C>
struct DequantR;

template <int C>
struct DequantR<B200_TYPE_Q4_0, C> {
    static __device__ __forceinline__ void run(const unsigned (&w)[36], unsigned (&o)[32]) {
#pragma unroll
        for (int b = 0; b < 2; b++) {
            constexpr int dummy = 0; (void)dummy;
            const int base = 36 * C + 18 * b;  // byte offset of the block
            const float d = f16_at(w, base), m8 = d * 8.0f;
#pragma unroll
            for (int j = 0; j < 16; j += 2) {
                const unsigned q0 = BYTE_OF(w, base + 2 + j), q1 = BYTE_OF(w, base + 3 + j);
                o[16 * b + j / 2]     = pack_bf16(d * byte_f<0>(q0 & 15) - m8, d * byte_f<0>(q1 & 15) - m8);
                o[16 * b + 8 + j / 2] = pack_bf16(d * byte_f<0>(q0 >> 4) - m8, d * byte_f<0>(q1 >> 4) - m8);
            }
        }
    }
};

template <int TYPE, int C>
struct DequantRLegacy {
    static __device__ __forceinline__ void run(const unsigned (&w)[group256_words(TYPE)], unsigned (&o)[32]) {
        using L = LegacyLayout<TYPE>;
#pragma unroll
        for (int b = 0; b < 2; b++) {
            const int base = 2 * L::BB * C + L::BB * b;  // byte offset of the block inside the 256-value group
            const float d = f16_at(w, base);
            float m = 0.f;
            if constexpr (L::HAS_M) m = f16_at(w, base + 2);
            if constexpr (TYPE == B200_TYPE_Q5_0) m = d * -16.0f;
            unsigned qh = 0;
            if constexpr (L::HAS_H) qh = BYTE_OF(w, base + L::QH) | (BYTE_OF(w, base + L::QH + 1) << 8) | (BYTE_OF(w, base + L::QH + 2) << 16) | (BYTE_OF(w, base + L::QH + 3) << 24);
#pragma unroll
            for (int j = 0; j < 16; j += 2) {
                const unsigned q0 = BYTE_OF(w, base + L::QS + j), q1 = BYTE_OF(w, base + L::QS + 1 + j);
                const unsigned l0 = (q0 & 15) | (((qh >> j) & 1u) << 4), l1 = (q1 & 15) | (((qh >> (j + 1)) & 1u) << 4);
                const unsigned h0 = (q0 >> 4) | (((qh >> (16 + j)) & 1u) << 4), h1 = (q1 >> 4) | (((qh >> (17 + j)) & 1u) << 4);
                o[16 * b + j / 2]     = pack_bf16(d * byte_f<0>(l0) + m, d * byte_f<0>(l1) + m);
                o[16 * b + 8 + j / 2] = pack_bf16(d * byte_f<0>(h0) + m, d * byte_f<0>(h1) + m);
            }
        }
    }
};
template <int C> struct DequantR<B200_TYPE_Q4_1, C> : DequantRLegacy<B200_TYPE_Q4_1, C> {};
template <int C> struct DequantR<B200_TYPE_Q5_0, C> : DequantRLegacy<B200_TYPE_Q5_0, C> {};
template <int C> struct DequantR<B200_TYPE_Q5_1, C> : DequantRLegacy<B200_TYPE_Q5_1, C> {};

template <int C>
struct DequantR<B200_TYPE_IQ4_NL, C> {
    static __device__ __forceinline__ void run(const unsigned (&w)[36], unsigned (&o)[32]) {
#pragma unroll
        for (int b = 0; b < 2; b++) {
            const int base = 2 * IQ4_NL_BYTES * C + IQ4_NL_BYTES * b;
            iq4_unpack32([&](int k) { return BYTE_OF(w, base + 2 + k); }, f16_at(w, base), o + 16 * b);
        }
    }
};
template <int C>
struct DequantR<B200_TYPE_IQ4_XS, C> {
    static __device__ __forceinline__ void run(const unsigned (&w)[34], unsigned (&o)[32]) {
        const float d = f16lo(w[0]);
#pragma unroll
        for (int b = 0; b < 2; b++)
            iq4_unpack32([&](int k) { return BYTE_OF(w, 8 + 32 * C + 16 * b + k); }, d * iq4xs_scale(w[1], w[0] >> 16, 2 * C + b), o + 16 * b);
    }
};

template <int C>
struct DequantR<B200_TYPE_Q8_0, C> {
    static __device__ __forceinline__ void run(const unsigned (&w)[68], unsigned (&o)[32]) {
#pragma unroll
        for (int b = 0; b < 2; b++) {
            const int base = 68 * C + 34 * b;
            const float d = f16_at(w, base), m128 = d * 128.0f;
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
                // int8 -> biased uint8 (x ^ 0x80 = x + 128), unbias after the multiply
                const unsigned q0 = BYTE_OF(w, base + 2 + j) ^ 0x80u, q1 = BYTE_OF(w, base + 3 + j) ^ 0x80u;
                o[16 * b + j / 2] = pack_bf16(d * byte_f<0>(q0) - m128, d * byte_f<0>(q1) - m128);
            }
        }
    }
};

template <int TYPE, int C>
struct DequantRK45 {
    static constexpr int NW = TYPE == B200_TYPE_Q4_K ? 36 : 44;
    static __device__ __forceinline__ void run(const unsigned (&w)[NW], unsigned (&o)[32]) {
        constexpr int QS = (TYPE == B200_TYPE_Q4_K ? 4 : 12) + 8 * C;  // first word of this chunk's 32 qs bytes
        const float d = f16lo(w[0]), dmin = f16hi(w[0]);
        const unsigned sw[3] = {w[1], w[2], w[3]};
#pragma unroll
        for (int hf = 0; hf < 2; hf++) {
            constexpr int dummy = 0; (void)dummy;
            const int sb = 2 * C + hf;
            int sc, mn;
            k4_scale_min(sw, sb, sc, mn);
            const float d1 = d * (float)sc, m1 = dmin * (float)mn;
#pragma unroll
            for (int i = 0; i < 8; i++) {  // word i holds bytes l = 4i .. 4i+3
                unsigned x = (w[QS + i] >> (4 * hf)) & 0x0F0F0F0Fu;
                if constexpr (TYPE == B200_TYPE_Q5_K) x |= ((w[4 + i] >> sb) & 0x01010101u) << 4;
                o[16 * hf + 2 * i]     = pack_bf16(d1 * byte_f<0>(x) - m1, d1 * byte_f<1>(x) - m1);
                o[16 * hf + 2 * i + 1] = pack_bf16(d1 * byte_f<2>(x) - m1, d1 * byte_f<3>(x) - m1);
            }
        }
    }
};
template <int C> struct DequantR<B200_TYPE_Q4_K, C> : DequantRK45<B200_TYPE_Q4_K, C> {};
template <int C> struct DequantR<B200_TYPE_Q5_K, C> : DequantRK45<B200_TYPE_Q5_K, C> {};

template <int TYPE, int C>
struct DequantRK23 {
    static __device__ __forceinline__ void run(const unsigned (&w)[group256_words(TYPE)], unsigned (&o)[32]) {
        dequant64_k23<TYPE>([&](int i) { return BYTE_OF(w, i); }, C, o);
    }
};
template <int C> struct DequantR<B200_TYPE_Q2_K, C> : DequantRK23<B200_TYPE_Q2_K, C> {};
template <int C> struct DequantR<B200_TYPE_Q3_K, C> : DequantRK23<B200_TYPE_Q3_K, C> {};

template <int C>
struct DequantR<B200_TYPE_Q6_K, C> {
    static __device__ __forceinline__ void run(const unsigned (&w)[53], unsigned (&o)[32]) {
        constexpr int half = C >> 1, hi = C & 1;
        const float d = f16lo(w[52]);
        const unsigned scw = w[48 + 2 * half + hi];  // 4 int8 scales of this chunk
#pragma unroll
        for (int g = 0; g < 4; g++) {
            const int rlow = g >> 1, r = 2 * hi + rlow;
            const float ds = d * (float)(int)(signed char)((scw >> (8 * g)) & 0xff), m32 = ds * 32.0f;
#pragma unroll
            for (int i = 0; i < 4; i++) {  // 4 values per word: l = 16*(g&1) + 4i ..
                const int lw = 4 * (g & 1) + i;  // word index inside the 32-byte row
                const unsigned lo = (w[16 * half + 8 * rlow + lw] >> (4 * hi)) & 0x0F0F0F0Fu;
                const unsigned h2 = ((w[32 + 8 * half + lw] >> (2 * r)) & 0x03030303u) << 4;
                const unsigned x = lo | h2;
                o[8 * g + 2 * i]     = pack_bf16(ds * byte_f<0>(x) - m32, ds * byte_f<1>(x) - m32);
                o[8 * g + 2 * i + 1] = pack_bf16(ds * byte_f<2>(x) - m32, ds * byte_f<3>(x) - m32);
            }
        }
    }
};

// ================================================================ v2: asynchronous operand supply
// Same tensor-core pipeline, but neither operand waits on a global load inside the producers:
//   * B (bf16 activations) arrives by TMA (cp.async.bulk.tensor.2d, 128-byte swizzle applied by the copy engine) issued by
//     a dedicated warp as soon as a stage is free;
//   * the raw quantised blocks of the NEXT 256-value group of each of the 128 rows are staged in shared memory by 1-D bulk
//     copies (one per row, issued by the row's own thread, aligned-down windows like the decode GEMV) two groups ahead, so the
//     producers unpack from shared memory.  A thread only ever reads its own row's slot: re-arming a slot needs no CTA barrier.
constexpr int MQ2_STAGES = 4;
constexpr int MQ2_THREADS = MQ_PRODUCERS + 64;  // + MMA warp + TMA warp
// per-row slot stride: >= group bytes + 14 bytes of head misalignment (2-byte aligned formats), rounded to 16, and an ODD
// number of 16-byte units so that the 8 threads of a quarter-warp hit distinct bank groups with 16-byte accesses
__host__ __device__ constexpr int raw_stride(int t) {
    return t == B200_TYPE_Q4_K ? 144 : t == B200_TYPE_Q5_K ? 176 : t == B200_TYPE_Q6_K ? 240 : t == B200_TYPE_Q8_0 ? 304 : 176;
}

__device__ __forceinline__ void tma_load_2d(void *dst, const void *tmap, int c0, int c1, unsigned long long *bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(dst)),
                 "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar))
                 : "memory");
}

struct alignas(64) TmapBytes { unsigned char b[128]; };  // CUtensorMap without pulling <cuda.h> into device code

template <int TYPE, int BN, bool GROUPED = false>
__global__ void __launch_bounds__(MQ2_THREADS, 1) mmq_tc2_kernel(const __grid_constant__ TmapBytes tmB, const __grid_constant__ MmqParams p) {
    constexpr int A_BYTES = MQ_BM * MQ_BK * 2, B_BYTES = BN * MQ_BK * 2, STAGE_BYTES = A_BYTES + B_BYTES;
    constexpr int G256 = group256_bytes(TYPE);
    int64_t n0 = (int64_t)blockIdx.y * BN, n_end = p.N;
    const unsigned char *Wbase = p.W;
    if constexpr (GROUPED) {  // uniform over the CTA: tiles past the table's end have nothing to do
        if ((int)blockIdx.y >= __ldg(p.tile_tab)) return;
        const int *te = p.tile_tab + 1 + 3 * blockIdx.y;
        Wbase += (size_t)__ldg(te) * p.expert_bytes;
        n0 = __ldg(te + 1);
        n_end = __ldg(te + 2);
    }
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) unsigned long long full_bar[MQ2_STAGES], empty_bar[MQ2_STAGES], tmem_full_bar;
    __shared__ unsigned tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, tid = threadIdx.x;
    const int64_t m0 = (int64_t)blockIdx.x * MQ_BM;
    const int nsb_all = (int)(p.K / 256);
    // split-K: this CTA's 256-value groups [sb0, sb0 + nsb)
    const int ksplit = GROUPED ? 1 : p.ksplit;
    const int sb0 = (int)(((int64_t)blockIdx.z * nsb_all) / ksplit);
    const int nsb = (int)(((int64_t)(blockIdx.z + 1) * nsb_all) / ksplit) - sb0, nkb = nsb * 4;

    if (tid == 0) {
        for (int s = 0; s < MQ2_STAGES; s++) { mbar_init(&full_bar[s], MQ_PRODUCERS + 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&tmem_full_bar, 1);
        mbar_fence_init();
    }
    if (warp == 4) tmem_alloc(&tmem_base_s, BN);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const unsigned tmem_base = tmem_base_s;

    if (warp < 4) {
        const int64_t m = m0 + tid;
        const bool row_ok = m < p.M;
        const unsigned char *wrow = Wbase + (size_t)(row_ok ? m : 0) * p.w_row_bytes;
        const unsigned a_off = (unsigned)((tid >> 3) * 1024 + (tid & 7) * 128);
        // software pipeline: the raw bytes of group sb+1 are in flight (registers) while group sb is unpacked
        constexpr int NW = group256_words(TYPE);
        unsigned cur[NW], nxt[NW];
        auto fetch = [&](unsigned (&w)[NW], int sb) {
            if (row_ok) {
                if constexpr (TYPE == B200_TYPE_Q4_K || TYPE == B200_TYPE_Q5_K) {  // 16-byte aligned rows: LDG.128
                    const uint4 *q4 = reinterpret_cast<const uint4 *>(wrow + (size_t)sb * G256);
#pragma unroll
                    for (int i = 0; i < NW / 4; i++) {
                        const uint4 v = __ldg(q4 + i);
                        w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
                    }
                } else {
                    // 2-byte aligned rows: 16-byte loads from the aligned-down address (a quarter of the LSU sector traffic of
                    // 4-byte loads), then a word select (offset 0..3 words) and a funnel shift (0 or 2 bytes) per word
                    const uintptr_t a = (uintptr_t)(wrow + (size_t)sb * G256);
                    const uint4 *q4 = reinterpret_cast<const uint4 *>(a & ~(uintptr_t)15);
                    const unsigned off = (unsigned)(a & 15) >> 2, sh = (unsigned)(a & 3) * 8;
                    constexpr int NV = (NW + 1 + 3 + 3) / 4;
                    unsigned t[NV * 4];
#pragma unroll
                    for (int i = 0; i < NV; i++) {
                        const uint4 v = __ldg(q4 + i);
                        t[4 * i] = v.x; t[4 * i + 1] = v.y; t[4 * i + 2] = v.z; t[4 * i + 3] = v.w;
                    }
                    unsigned prev = (off & 2) ? ((off & 1) ? t[3] : t[2]) : ((off & 1) ? t[1] : t[0]);
#pragma unroll
                    for (int i = 0; i < NW; i++) {
                        const unsigned nx = (off & 2) ? ((off & 1) ? t[i + 4] : t[i + 3]) : ((off & 1) ? t[i + 2] : t[i + 1]);
                        w[i] = __funnelshift_r(prev, nx, sh);
                        prev = nx;
                    }
                }
            } else {
#pragma unroll
                for (int i = 0; i < NW; i++) w[i] = 0u;
            }
        };
        auto publish = [&](int kb, const unsigned (&o)[32]) {
            const int s = kb % MQ2_STAGES;
            mbar_wait(&empty_bar[s], ((unsigned)(kb / MQ2_STAGES) & 1u) ^ 1u);
            unsigned char *a_s = smem + (size_t)s * STAGE_BYTES;
#pragma unroll
            for (int q = 0; q < 8; q++)
                *reinterpret_cast<uint4 *>(a_s + a_off + ((q ^ (tid & 7)) << 4)) = make_uint4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
            fence_proxy_async_smem();
            mbar_arrive(&full_bar[s]);
        };
        fetch(cur, sb0);
#pragma unroll 1
        for (int sb = 0; sb < nsb; sb++) {
            if (sb + 1 < nsb) fetch(nxt, sb0 + sb + 1);
            unsigned o[32];
            DequantR<TYPE, 0>::run(cur, o); publish(sb * 4 + 0, o);
            DequantR<TYPE, 1>::run(cur, o); publish(sb * 4 + 1, o);
            DequantR<TYPE, 2>::run(cur, o); publish(sb * 4 + 2, o);
            DequantR<TYPE, 3>::run(cur, o); publish(sb * 4 + 3, o);
#pragma unroll
            for (int i = 0; i < NW; i++) cur[i] = nxt[i];
        }
        mbar_wait(&tmem_full_bar, 0u);
        tc_fence_after();
        float *drow = p.dst + m;
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
            float v[32];
            tmem_ld32(tmem_base + ((unsigned)(warp * 32) << 16) + (unsigned)c0, v);
            if (row_ok) {
#pragma unroll
                for (int j = 0; j < 32; j++) {
                    const int64_t n = n0 + c0 + j;
                    if constexpr (GROUPED) {
                        if (n < n_end) drow[(int64_t)__ldg(p.row_dst + n)] = v[j];  // scatter: sorted row -> (token, slot)
                    } else {
                        if (n < n_end) {
                            if (ksplit > 1) atomicAdd(drow + n * p.dst_stride, v[j]);  // (result unused: RED.ADD.F32)
                            else drow[n * p.dst_stride] = v[j];
                        }
                    }
                }
            }
        }
        tc_fence_before();
    } else if (warp == 4) {
        constexpr unsigned idesc = umma_idesc_bf16(MQ_BM, BN);
        for (int kb = 0; kb < nkb; kb++) {
            const int s = kb % MQ2_STAGES;
            mbar_wait(&full_bar[s], (unsigned)(kb / MQ2_STAGES) & 1u);
            tc_fence_after();
            if (lane == 0) {
                const unsigned a_addr = smem_u32(smem + (size_t)s * STAGE_BYTES), b_addr = a_addr + A_BYTES;
#pragma unroll
                for (int k = 0; k < MQ_BK / 16; k++)
                    tc_mma_bf16(tmem_base, umma_desc_sw128(a_addr) + 2u * k, umma_desc_sw128(b_addr) + 2u * k, idesc, (kb | k) != 0 ? 1u : 0u);
                tc_commit(&empty_bar[s]);
                if (kb == nkb - 1) tc_commit(&tmem_full_bar);
            }
            __syncwarp();
        }
        tc_fence_before();
    } else {
        // ===== TMA warp: activation tiles [BN rows x 64 bf16] into the swizzled B slot of each free stage
        if (lane == 0) {
            for (int kb = 0; kb < nkb; kb++) {
                const int s = kb % MQ2_STAGES;
                mbar_wait(&empty_bar[s], ((unsigned)(kb / MQ2_STAGES) & 1u) ^ 1u);
                mbar_expect_tx(&full_bar[s], (unsigned)B_BYTES);
                tma_load_2d(smem + (size_t)s * STAGE_BYTES + A_BYTES, &tmB, (sb0 * 4 + kb) * MQ_BK, (int)n0, &full_bar[s]);
            }
        }
    }
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        tmem_dealloc(tmem_base, BN);
    }
}

// ---------------------------------------------------------------- host side
static int64_t mmq_bn(int64_t N) { return N <= 32 ? 32 : N <= 64 ? 64 : N <= 128 ? 128 : 256; }
static int64_t mmq_n_pad(int64_t N) { const int64_t bn = mmq_bn(N); return (N + bn - 1) / bn * bn; }

template <int TYPE, int BN>
static int mmq_launch(const MmqParams &p, cudaStream_t st) {
    constexpr size_t smem = (size_t)MQ_STAGES * (MQ_BM * MQ_BK * 2 + BN * MQ_BK * 2) + 1024;
    static bool attr_set_dev[16] = {};  // cudaFuncSetAttribute is per device
    int cur_dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&cur_dev));
    bool &attr_set = attr_set_dev[cur_dev & 15];
    if (!attr_set) {
        B200_CUDA_TRY(cudaFuncSetAttribute(mmq_tc_kernel<TYPE, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    dim3 grid((unsigned)((p.M + MQ_BM - 1) / MQ_BM), (unsigned)((p.N + BN - 1) / BN));
    mmq_tc_kernel<TYPE, BN><<<grid, MQ_THREADS, smem, st>>>(p);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time dependency on a particular libcuda)
typedef int (*EncodeTiledFn)(void *, int, unsigned, void *, const unsigned long long *, const unsigned long long *, const unsigned *, const unsigned *, int, int, int, int);
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *sym = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess) fn = (EncodeTiledFn)sym;
    }
    return fn;
}

template <int TYPE, int BN, bool GROUPED = false>
static int mmq2_launch(const MmqParams &p, int64_t n_pad, cudaStream_t st, unsigned grid_y = 0) {
    constexpr size_t smem = (size_t)MQ2_STAGES * (MQ_BM * MQ_BK * 2 + BN * MQ_BK * 2) + 1024;
    static bool attr_set_dev[16] = {};  // cudaFuncSetAttribute is per device
    int cur_dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&cur_dev));
    bool &attr_set = attr_set_dev[cur_dev & 15];
    if (!attr_set) {
        B200_CUDA_TRY(cudaFuncSetAttribute(mmq_tc2_kernel<TYPE, BN, GROUPED>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) { set_error("b200_mul_mat_q: cuTensorMapEncodeTiled not available"); return B200_ERR_CUDA; }
    TmapBytes tm;
    const unsigned long long gdim[2] = {(unsigned long long)p.K, (unsigned long long)n_pad};
    const unsigned long long gstr[1] = {(unsigned long long)p.K * 2};
    const unsigned box[2] = {(unsigned)MQ_BK, (unsigned)BN};
    const unsigned estr[2] = {1, 1};
    // CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 = 9, INTERLEAVE_NONE = 0, SWIZZLE_128B = 3, L2_PROMOTION_L2_256B = 3, FLOAT_OOB_FILL_NONE = 0
    const int rc = enc(&tm, 9, 2, (void *)p.Xb, gdim, gstr, box, estr, 0, 3, 3, 0);
    if (rc != 0) { set_error("b200_mul_mat_q: cuTensorMapEncodeTiled failed (%d)", rc); return B200_ERR_CUDA; }
    dim3 grid((unsigned)((p.M + MQ_BM - 1) / MQ_BM), GROUPED ? grid_y : (unsigned)((p.N + BN - 1) / BN));
    MmqParams q = p;
    q.ksplit = 1;
    if constexpr (!GROUPED) {
        // fewer tiles than half the SMs (M = 4096 x N = 512: 64 CTAs, every one bound by its own 128 dequantising threads): two
        // CTAs share a tile's K range and add into a zeroed dst.  Two partials only: the sum is independent of the arrival order.
        static const bool no_split = getenv("B200_MMQ_NO_SPLITK") != nullptr;
        if (!no_split && (int64_t)grid.x * grid.y * 2 <= 148 && p.K >= 2048) {
            q.ksplit = 2;
            grid.z = 2;
            if (p.dst_stride == p.M) B200_CUDA_TRY(cudaMemsetAsync(p.dst, 0, (size_t)p.M * p.N * sizeof(float), st));
            else B200_CUDA_TRY(cudaMemset2DAsync(p.dst, (size_t)p.dst_stride * sizeof(float), 0, (size_t)p.M * sizeof(float), (size_t)p.N, st));
        }
    }
    mmq_tc2_kernel<TYPE, BN, GROUPED><<<grid, MQ2_THREADS, smem, st>>>(tm, q);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

template <int TYPE>
static int mmq_dispatch_bn(const MmqParams &p, cudaStream_t st) {
    if (p.K % 256 == 0 && !getenv("B200_MMQ_V1")) {
        const int64_t n_pad = mmq_n_pad(p.N);
        switch (mmq_bn(p.N)) {
            case 32: return mmq2_launch<TYPE, 32>(p, n_pad, st);
            case 64: return mmq2_launch<TYPE, 64>(p, n_pad, st);
            case 128: return mmq2_launch<TYPE, 128>(p, n_pad, st);
            default: return mmq2_launch<TYPE, 256>(p, n_pad, st);
        }
    }
    switch (mmq_bn(p.N)) {
        case 32: return mmq_launch<TYPE, 32>(p, st);
        case 64: return mmq_launch<TYPE, 64>(p, st);
        case 128: return mmq_launch<TYPE, 128>(p, st);
        default: return mmq_launch<TYPE, 256>(p, st);
    }
}


// ================================================================ (a9) MUL_MAT_ID for batches: group by expert on the device
// Replaces the batch branch of ggml_cuda_mul_mat_id (/root/reference/ggml/src/ggml-cuda.cu:2066-2139: ids copied to the host +
// stream sync, then per expert k_copy_src1_to_contiguous -> mul_mat -> k_copy_dst_from_contiguous, :1954-2001).  Here nothing
// leaves the device: one small kernel counts the (token, slot) pairs per expert, prefix-sums them, writes the sorted row lists and
// the tile table; a gather pass converts the selected activation rows to bf16 in sorted order; ONE grouped tensor-core launch
// walks the tile table (every tile belongs to one expert) and scatters its f32 results straight to dst[(token, slot)].
constexpr int MID_MAX_EXPERTS = 256;
struct MidTables {      // int offsets into the table area of the workspace
    int *tile_tab;      // 1 + 3 * max_tiles
    int *row_src;       // n_rows: float offset of the activation row of sorted row i
    int *row_dst;       // n_rows: float offset of dst row of sorted row i
};
__global__ void __launch_bounds__(1024, 1) mid_group_kernel(const int32_t *__restrict__ ids, int64_t ids_stride, int n_used, int T, int n_expert, int n_b_rows,
                                                            int64_t b_stride_row, int64_t b_stride_tok, int64_t d_stride_row, int64_t d_stride_tok, int BN,
                                                            MidTables t) {
    __shared__ int count[MID_MAX_EXPERTS], off[MID_MAX_EXPERTS + 1], cursor[MID_MAX_EXPERTS];
    for (int e = threadIdx.x; e < n_expert; e += blockDim.x) { count[e] = 0; cursor[e] = 0; }
    __syncthreads();
    const int n = n_used * T;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int e = ids[(int64_t)(i / n_used) * ids_stride + i % n_used];
        if (e >= 0 && e < n_expert) atomicAdd(&count[e], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int acc = 0, nt = 0;
        for (int e = 0; e < n_expert; e++) {
            off[e] = acc;
            for (int r0 = acc; r0 < acc + count[e]; r0 += BN) {
                t.tile_tab[1 + 3 * nt] = e;
                t.tile_tab[2 + 3 * nt] = r0;
                t.tile_tab[3 + 3 * nt] = acc + count[e];
                nt++;
            }
            acc += count[e];
        }
        off[n_expert] = acc;
        t.tile_tab[0] = nt;
    }
    __syncthreads();
    // rows of invalid expert ids are left out (their dst rows stay untouched; the reference asserts on them); the slots behind
    // the last valid row are given a harmless source so the gather pass needs no count
    for (int i = off[n_expert] + threadIdx.x; i < n; i += blockDim.x) { t.row_src[i] = 0; t.row_dst[i] = 0; }
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int tok = i / n_used, s = i % n_used;
        const int e = ids[(int64_t)tok * ids_stride + s];
        if (e < 0 || e >= n_expert) continue;
        const int pos = off[e] + atomicAdd(&cursor[e], 1);
        t.row_src[pos] = (int)(tok * b_stride_tok + (s % n_b_rows) * b_stride_row);
        t.row_dst[pos] = (int)(tok * d_stride_tok + s * d_stride_row);
    }
}
// gather + convert: sorted row i <- bf16(b[row_src[i]]), K values
__global__ void mid_gather_bf16_kernel(const float *__restrict__ b, const int *__restrict__ row_src, int64_t K, __nv_bfloat16 *__restrict__ out) {
    const int64_t i4 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n = blockIdx.y;
    if (i4 * 4 >= K) return;
    const float4 v = __ldg(reinterpret_cast<const float4 *>(b + (int64_t)__ldg(row_src + n)) + i4);
    uint2 pk;
    pk.x = pack_bf16(v.x, v.y);
    pk.y = pack_bf16(v.z, v.w);
    *reinterpret_cast<uint2 *>(out + n * K + i4 * 4) = pk;
}

static int64_t mid_bn(int64_t n_rows, int n_expert) { return mmq_bn((n_rows + n_expert - 1) / n_expert); }
static int64_t mid_max_tiles(int64_t n_rows, int n_expert) { return (n_rows + mid_bn(n_rows, n_expert) - 1) / mid_bn(n_rows, n_expert) + n_expert; }
static size_t mid_xb_bytes(int64_t n_rows, int64_t K) { return ((size_t)n_rows * (size_t)K * 2 + 255) / 256 * 256; }

template <int TYPE>
static int mid_dispatch_bn(const MmqParams &p, int64_t bn, int64_t n_rows, unsigned grid_y, cudaStream_t st) {
    switch (bn) {
        case 32: return mmq2_launch<TYPE, 32, true>(p, n_rows, st, grid_y);
        case 64: return mmq2_launch<TYPE, 64, true>(p, n_rows, st, grid_y);
        case 128: return mmq2_launch<TYPE, 128, true>(p, n_rows, st, grid_y);
        default: return mmq2_launch<TYPE, 256, true>(p, n_rows, st, grid_y);
    }
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_mul_mat_q_workspace_bytes(int type, int64_t K, int64_t M, int64_t N) {
    (void)type; (void)M;
    if (K <= 0 || N <= 0) return 0;
    return (size_t)mmq_n_pad(N) * (size_t)K * 2 + 256;
}

extern "C" int b200_mul_mat_q(int type, const void *W, size_t w_row_bytes, const float *X, int64_t x_stride, float *dst, int64_t dst_stride,
                              int64_t K, int64_t M, int64_t N, void *workspace, size_t workspace_bytes, b200_stream_t stream) {
    if (!W || !dst) { set_error("b200_mul_mat_q: NULL argument"); return B200_ERR_INVALID; }
    const int bv = type_block_values(type);
    if (!(type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0 || type == B200_TYPE_Q4_K || type == B200_TYPE_Q5_K || type == B200_TYPE_Q6_K ||
          type == B200_TYPE_Q4_1 || type == B200_TYPE_Q5_0 || type == B200_TYPE_Q5_1 || type == B200_TYPE_Q2_K || type == B200_TYPE_Q3_K ||
          type == B200_TYPE_IQ4_NL || type == B200_TYPE_IQ4_XS)) {
        set_error("b200_mul_mat_q: unsupported weight type %d", type);
        return B200_ERR_UNSUPPORTED;
    }
    if (K <= 0 || K % bv || M < 0 || N < 0) { set_error("b200_mul_mat_q: bad shape K=%lld M=%lld N=%lld", (long long)K, (long long)M, (long long)N); return B200_ERR_INVALID; }
    if (K % MQ_BK) { set_error("b200_mul_mat_q: K=%lld must be a multiple of %d", (long long)K, MQ_BK); return B200_ERR_UNSUPPORTED; }
    if (M == 0 || N == 0) return B200_OK;
    const size_t row_bytes = (size_t)(K / bv) * type_block_bytes(type);
    if (w_row_bytes < row_bytes || (w_row_bytes & 1) || ((uintptr_t)W & 1)) { set_error("b200_mul_mat_q: weight stride/alignment"); return B200_ERR_INVALID; }
    if ((type == B200_TYPE_Q4_K || type == B200_TYPE_Q5_K) && ((w_row_bytes & 15) || ((uintptr_t)W & 15))) {
        set_error("b200_mul_mat_q: q4_K/q5_K rows must be 16-byte aligned");
        return B200_ERR_UNSUPPORTED;
    }
    if (X && (((uintptr_t)X & 15) || (x_stride & 3))) { set_error("b200_mul_mat_q: X must be 16-byte aligned with a row stride multiple of 4"); return B200_ERR_UNSUPPORTED; }
    const size_t need = b200_mul_mat_q_workspace_bytes(type, K, M, N);
    if (!workspace || workspace_bytes < need || ((uintptr_t)workspace & 15)) {
        set_error("b200_mul_mat_q: workspace of %zu bytes (16-byte aligned) required, %zu given", need, workspace_bytes);
        return B200_ERR_INVALID;
    }
    const cudaStream_t st = (cudaStream_t)stream;
    const int64_t n_pad = mmq_n_pad(N);
    if (n_pad > 65535) { set_error("b200_mul_mat_q: N too large"); return B200_ERR_UNSUPPORTED; }
    __nv_bfloat16 *xb = (__nv_bfloat16 *)workspace;
    if (X) {  // (X == NULL: the workspace holds the bf16 activations of the previous call with this K and N)
        dim3 g1((unsigned)((K / 4 + 255) / 256), (unsigned)n_pad);
        mmq_x_to_bf16_kernel<<<g1, 256, 0, st>>>(X, x_stride, K, N, n_pad, xb);
        B200_CUDA_TRY(cudaGetLastError());
    }
    MmqParams p;
    p.W = (const unsigned char *)W; p.w_row_bytes = w_row_bytes; p.Xb = xb; p.dst = dst; p.dst_stride = dst_stride; p.K = K; p.M = M; p.N = N;
    p.tile_tab = nullptr; p.row_dst = nullptr; p.expert_bytes = 0;
    switch (type) {
        case B200_TYPE_Q4_0: return mmq_dispatch_bn<B200_TYPE_Q4_0>(p, st);
        case B200_TYPE_Q8_0: return mmq_dispatch_bn<B200_TYPE_Q8_0>(p, st);
        case B200_TYPE_Q4_K: return mmq_dispatch_bn<B200_TYPE_Q4_K>(p, st);
        case B200_TYPE_Q5_K: return mmq_dispatch_bn<B200_TYPE_Q5_K>(p, st);
        case B200_TYPE_Q6_K: return mmq_dispatch_bn<B200_TYPE_Q6_K>(p, st);
        case B200_TYPE_Q4_1: return mmq_dispatch_bn<B200_TYPE_Q4_1>(p, st);
        case B200_TYPE_Q5_0: return mmq_dispatch_bn<B200_TYPE_Q5_0>(p, st);
        case B200_TYPE_Q5_1: return mmq_dispatch_bn<B200_TYPE_Q5_1>(p, st);
        case B200_TYPE_Q2_K: return mmq_dispatch_bn<B200_TYPE_Q2_K>(p, st);
        case B200_TYPE_Q3_K: return mmq_dispatch_bn<B200_TYPE_Q3_K>(p, st);
        case B200_TYPE_IQ4_NL: return mmq_dispatch_bn<B200_TYPE_IQ4_NL>(p, st);
        case B200_TYPE_IQ4_XS: return mmq_dispatch_bn<B200_TYPE_IQ4_XS>(p, st);
    }
    return B200_ERR_UNSUPPORTED;
}

extern "C" size_t b200_mul_mat_id_q_workspace_bytes(int type, int64_t K, int64_t M, int64_t n_used, int64_t T, int32_t n_expert) {
    (void)type; (void)M;
    const int64_t n_rows = n_used * T;
    if (K <= 0 || n_rows <= 0 || n_expert <= 0) return 0;
    return mid_xb_bytes(n_rows, K) + sizeof(int) * (size_t)(1 + 3 * mid_max_tiles(n_rows, n_expert) + 2 * n_rows) + 256;
}

extern "C" int b200_mul_mat_id_q(int type, const void *as, size_t w_row_bytes, size_t expert_bytes, int32_t n_expert, const int32_t *ids,
                                 int64_t ids_stride, const float *b, int64_t b_stride_row, int64_t b_stride_tok, float *dst, int64_t dst_stride_row,
                                 int64_t dst_stride_tok, int64_t K, int64_t M, int64_t n_used, int64_t n_b_rows, int64_t T, void *workspace,
                                 size_t workspace_bytes, b200_stream_t stream) {
    if (!as || !ids || !b || !dst) { set_error("b200_mul_mat_id_q: NULL argument"); return B200_ERR_INVALID; }
    if (!(type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0 || type == B200_TYPE_Q4_K || type == B200_TYPE_Q5_K || type == B200_TYPE_Q6_K ||
          type == B200_TYPE_Q4_1 || type == B200_TYPE_Q5_0 || type == B200_TYPE_Q5_1 || type == B200_TYPE_Q2_K || type == B200_TYPE_Q3_K ||
          type == B200_TYPE_IQ4_NL || type == B200_TYPE_IQ4_XS)) {
        set_error("b200_mul_mat_id_q: unsupported weight type %d", type);
        return B200_ERR_UNSUPPORTED;
    }
    if (K <= 0 || K % 256 || M < 0 || T < 0 || n_used < 1 || n_b_rows < 1 || n_expert < 1 || n_expert > MID_MAX_EXPERTS) {
        set_error("b200_mul_mat_id_q: bad shape (K must be a multiple of 256, at most %d experts)", MID_MAX_EXPERTS);
        return K % 256 ? B200_ERR_UNSUPPORTED : B200_ERR_INVALID;
    }
    if (M == 0 || T == 0) return B200_OK;
    const int bv = type_block_values(type);
    const size_t row_bytes = (size_t)(K / bv) * type_block_bytes(type);
    if (w_row_bytes < row_bytes || ((w_row_bytes | expert_bytes | (uintptr_t)as) & 1)) { set_error("b200_mul_mat_id_q: weight stride/alignment"); return B200_ERR_INVALID; }
    if ((type == B200_TYPE_Q4_K || type == B200_TYPE_Q5_K) && ((w_row_bytes | expert_bytes | (uintptr_t)as) & 15)) {
        set_error("b200_mul_mat_id_q: q4_K/q5_K rows must be 16-byte aligned");
        return B200_ERR_UNSUPPORTED;
    }
    if (((uintptr_t)b & 15) || ((b_stride_row | b_stride_tok) & 3)) { set_error("b200_mul_mat_id_q: b rows must be 16-byte aligned"); return B200_ERR_UNSUPPORTED; }
    const int64_t n_rows = n_used * T;
    const int64_t span_b = (T - 1) * b_stride_tok + (n_b_rows - 1) * b_stride_row, span_d = (T - 1) * dst_stride_tok + (n_used - 1) * dst_stride_row;
    if (n_rows > 65535 || span_b > 0x7fffffff || span_d > 0x7fffffff) { set_error("b200_mul_mat_id_q: batch too large"); return B200_ERR_UNSUPPORTED; }
    const size_t need = b200_mul_mat_id_q_workspace_bytes(type, K, M, n_used, T, n_expert);
    if (!workspace || workspace_bytes < need || ((uintptr_t)workspace & 255)) {
        set_error("b200_mul_mat_id_q: workspace of %zu bytes (256-byte aligned) required, %zu given", need, workspace_bytes);
        return B200_ERR_INVALID;
    }
    const cudaStream_t st = (cudaStream_t)stream;
    const int64_t bn = mid_bn(n_rows, n_expert), max_tiles = mid_max_tiles(n_rows, n_expert);
    __nv_bfloat16 *xb = (__nv_bfloat16 *)workspace;
    MidTables t;
    t.tile_tab = (int *)((unsigned char *)workspace + mid_xb_bytes(n_rows, K));
    t.row_src = t.tile_tab + 1 + 3 * max_tiles;
    t.row_dst = t.row_src + n_rows;
    mid_group_kernel<<<1, 1024, 0, st>>>(ids, ids_stride, (int)n_used, (int)T, (int)n_expert, (int)n_b_rows, b_stride_row, b_stride_tok, dst_stride_row,
                                         dst_stride_tok, (int)bn, t);
    B200_CUDA_TRY(cudaGetLastError());
    dim3 g1((unsigned)((K / 4 + 255) / 256), (unsigned)n_rows);
    mid_gather_bf16_kernel<<<g1, 256, 0, st>>>(b, t.row_src, K, xb);
    B200_CUDA_TRY(cudaGetLastError());
    MmqParams p;
    p.W = (const unsigned char *)as; p.w_row_bytes = w_row_bytes; p.Xb = xb; p.dst = dst; p.dst_stride = 0; p.K = K; p.M = M; p.N = n_rows;
    p.tile_tab = t.tile_tab; p.row_dst = t.row_dst; p.expert_bytes = expert_bytes;
    switch (type) {
        case B200_TYPE_Q4_0: return mid_dispatch_bn<B200_TYPE_Q4_0>(p, bn, n_rows, (unsigned)max_tiles, st);
        case B200_TYPE_Q8_0: return mid_dispatch_bn<B200_TYPE_Q8_0>(p, bn, n_rows, (unsigned)max_tiles, st);
        case B200_TYPE_Q4_K: return mid_dispatch_bn<B200_TYPE_Q4_K>(p, bn, n_rows, (unsigned)max_tiles, st);
        case B200_TYPE_Q5_K: return mid_dispatch_bn<B200_TYPE_Q5_K>(p, bn, n_rows, (unsigned)max_tiles, st);
        case B200_TYPE_Q6_K: return mid_dispatch_bn<B200_TYPE_Q6_K>(p, bn, n_rows, (unsigned)max_tiles, st);
        case B200_TYPE_Q4_1: return mid_dispatch_bn<B200_TYPE_Q4_1>(p, bn, n_rows, (unsigned)max_tiles, st);
        case B200_TYPE_Q5_0: return mid_dispatch_bn<B200_TYPE_Q5_0>(p, bn, n_rows, (unsigned)max_tiles, st);
        case B200_TYPE_Q5_1: return mid_dispatch_bn<B200_TYPE_Q5_1>(p, bn, n_rows, (unsigned)max_tiles, st);
        case B200_TYPE_Q2_K: return mid_dispatch_bn<B200_TYPE_Q2_K>(p, bn, n_rows, (unsigned)max_tiles, st);
        case B200_TYPE_Q3_K: return mid_dispatch_bn<B200_TYPE_Q3_K>(p, bn, n_rows, (unsigned)max_tiles, st);
        case B200_TYPE_IQ4_NL: return mid_dispatch_bn<B200_TYPE_IQ4_NL>(p, bn, n_rows, (unsigned)max_tiles, st);
        case B200_TYPE_IQ4_XS: return mid_dispatch_bn<B200_TYPE_IQ4_XS>(p, bn, n_rows, (unsigned)max_tiles, st);
    }
    return B200_ERR_UNSUPPORTED;
}
