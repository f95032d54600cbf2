// quantize.cu -- (a2) bit-exact block dequantisation and (a3) f32 -> q8_1 activation quantisation.
//
// Reference semantics (paths under /root/reference/ggml/src/):
//   dequantize_row_{q4_0,q8_0,q4_K,q5_K,q6_K}   ggml-quants.c:1523, 1617, 2556, 2764, 2978
//   quantize_row_q8_1_ref                       ggml-quants.c:1160-1193
// The reference C is compiled without FMA contraction (-std=c11), so every multiply / subtract is its own
// IEEE rounding; the kernels use __fmul_rn/__fsub_rn/__fdiv_rn to pin exactly that.
#include "b200_common.cuh"

namespace b200 {

__global__ void __launch_bounds__(256) quantize_q8_1_kernel(const float *__restrict__ x, int64_t x_stride,
                                                            unsigned char *__restrict__ y, int64_t K) {
    const int64_t n   = blockIdx.y;
    const int64_t i4  = (int64_t)blockIdx.x * 256 + threadIdx.x;  // float4 index within the row
    const int64_t nb  = K / QK;
    const bool active = i4 * 4 < K;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (active) v = *reinterpret_cast<const float4 *>(x + n * x_stride + i4 * 4);
    float d, s;
    int sum4;
    const int packed = quantize_q8_1_group8(v, d, s, sum4);
    if (!active) return;
    const int64_t blk = i4 / 8;
    const int part    = (int)(i4 & 7);
    unsigned char *yb = y + (n * nb + blk) * Q8_1_BYTES;
    *reinterpret_cast<int *>(yb + 4 + 4 * part) = packed;  // 36-byte blocks: every field is 4-byte aligned
    if (part == 0) {
        *reinterpret_cast<__half2 *>(yb) = __halves2half2(__float2half_rn(d), __float2half_rn(s));
    }
}

// ------------------------------------------------------------------------------------------------
// dequantisation: one thread per 4 consecutive output values (so q4/q5/q6 sub-block scales are uniform per thread)
template <typename T>
__device__ __forceinline__ void store4(T *dst, int64_t i, float a, float b, float c, float d);
template <>
__device__ __forceinline__ void store4<float>(float *dst, int64_t i, float a, float b, float c, float d) {
    *reinterpret_cast<float4 *>(dst + i) = make_float4(a, b, c, d);
}
template <>
__device__ __forceinline__ void store4<__half>(__half *dst, int64_t i, float a, float b, float c, float d) {
    __half2 lo = __halves2half2(__float2half_rn(a), __float2half_rn(b));
    __half2 hi = __halves2half2(__float2half_rn(c), __float2half_rn(d));
    *reinterpret_cast<__half2 *>(dst + i)     = lo;
    *reinterpret_cast<__half2 *>(dst + i + 2) = hi;
}

__device__ __forceinline__ float ldh(const unsigned char *p) { return h2f(*reinterpret_cast<const unsigned short *>(p)); }

template <int TYPE, typename T>
__global__ void __launch_bounds__(256) dequantize_kernel(const unsigned char *__restrict__ src, T *__restrict__ dst, int64_t n) {
    const int64_t i = ((int64_t)blockIdx.x * 256 + threadIdx.x) * 4;  // first of 4 values
    if (i >= n) return;
    float o[4];
    if constexpr (TYPE == B200_TYPE_Q4_0) {
        const unsigned char *b = src + (i / QK) * Q4_0_BYTES;
        const float d = ldh(b);
        const int j = (int)(i % QK);  // values j..j+3; low nibbles hold 0..15, high nibbles 16..31 (ggml-quants.c:1533-1538)
#pragma unroll
        for (int l = 0; l < 4; l++) {
            const int jj = j + l;
            const int q = jj < 16 ? (b[2 + jj] & 0x0F) : (b[2 + jj - 16] >> 4);
            o[l] = __fmul_rn((float)(q - 8), d);
        }
    } else if constexpr (TYPE == B200_TYPE_Q4_1 || TYPE == B200_TYPE_Q5_0 || TYPE == B200_TYPE_Q5_1) {
        // dequantize_row_q4_1 / q5_0 / q5_1 (ggml-quants.c:1543-1615): x*d + m (a product, then a sum); the fifth bit of value
        // j is bit j of qh
        constexpr bool HAS_M = TYPE != B200_TYPE_Q5_0, HAS_H = TYPE != B200_TYPE_Q4_1;
        constexpr int BB = type_block_bytes(TYPE), QS = 2 + (HAS_M ? 2 : 0) + (HAS_H ? 4 : 0);
        const unsigned char *b = src + (i / QK) * BB;
        const float d = ldh(b), m = HAS_M ? ldh(b + 2) : 0.f;
        unsigned qh = 0;
        if constexpr (HAS_H) qh = (unsigned)b[QS - 4] | ((unsigned)b[QS - 3] << 8) | ((unsigned)b[QS - 2] << 16) | ((unsigned)b[QS - 1] << 24);
        const int j = (int)(i % QK);
#pragma unroll
        for (int l = 0; l < 4; l++) {
            const int jj = j + l;
            int q = jj < 16 ? (b[QS + jj] & 0x0F) : (b[QS + jj - 16] >> 4);
            if constexpr (HAS_H) q |= (int)((qh >> jj) & 1u) << 4;
            if constexpr (TYPE == B200_TYPE_Q5_0) o[l] = __fmul_rn((float)(q - 16), d);
            else o[l] = __fadd_rn(__fmul_rn((float)q, d), m);
        }
    } else if constexpr (TYPE == B200_TYPE_IQ4_NL) {
        // dequantize_row_iq4_nl (ggml-quants.c:3743-3760): d * level[nibble]; low nibbles are values 0..15, high nibbles 16..31
        const unsigned char *b = src + (i / QK) * IQ4_NL_BYTES;
        const float d = ldh(b);
        const int j = (int)(i % QK);
#pragma unroll
        for (int l = 0; l < 4; l++) {
            const int jj = j + l;
            const unsigned q = jj < 16 ? (b[2 + jj] & 0x0Fu) : (unsigned)(b[2 + jj - 16] >> 4);
            o[l] = __fmul_rn(d, (float)iq4_level(q));
        }
    } else if constexpr (TYPE == B200_TYPE_IQ4_XS) {
        // dequantize_row_iq4_xs (ggml-quants.c:3762-3785): (d * (ls - 32)) * level[nibble], ls = 6-bit scale of the 32-value sub-block
        const unsigned char *b = src + (i / QK_K) * IQ4_XS_BYTES;
        const int v = (int)(i % QK_K), ib = v / 32, j = v % 32;
        const unsigned sh = (unsigned)b[2] | ((unsigned)b[3] << 8);
        const int ls = (int)(((b[4 + ib / 2] >> (4 * (ib % 2))) & 0xFu) | (((sh >> (2 * ib)) & 3u) << 4)) - 32;
        const float dl = __fmul_rn(ldh(b), (float)ls);
        const unsigned char *qs = b + 8 + 16 * ib;
#pragma unroll
        for (int l = 0; l < 4; l++) {
            const int jj = j + l;
            const unsigned q = jj < 16 ? (qs[jj] & 0x0Fu) : (unsigned)(qs[jj - 16] >> 4);
            o[l] = __fmul_rn(dl, (float)iq4_level(q));
        }
    } else if constexpr (TYPE == B200_TYPE_Q8_0) {
        const unsigned char *b = src + (i / QK) * Q8_0_BYTES;
        const float d = ldh(b);
        const int j = (int)(i % QK);
#pragma unroll
        for (int l = 0; l < 4; l++) o[l] = __fmul_rn((float)(signed char)b[2 + j + l], d);
    } else if constexpr (TYPE == B200_TYPE_Q4_K || TYPE == B200_TYPE_Q5_K) {
        constexpr int BB = TYPE == B200_TYPE_Q4_K ? Q4_K_BYTES : Q5_K_BYTES;
        const unsigned char *b = src + (i / QK_K) * BB;
        const int j  = (int)(i % QK_K);
        const int sb = j / 32;        // sub-block 0..7
        const int l0 = j % 32;
        const int c  = sb / 2;        // 64-value chunk: qs bytes [32c, 32c+32), low nibble = even sub-block
        unsigned sw[3];
        sw[0] = *reinterpret_cast<const unsigned *>(b + 4);
        sw[1] = *reinterpret_cast<const unsigned *>(b + 8);
        sw[2] = *reinterpret_cast<const unsigned *>(b + 12);
        int sc, mn;
        k4_scale_min(sw, sb, sc, mn);
        const float d1 = __fmul_rn(ldh(b), (float)sc);
        const float m1 = __fmul_rn(ldh(b + 2), (float)mn);
        const unsigned char *qs = b + (TYPE == B200_TYPE_Q4_K ? 16 : 48) + 32 * c + l0;
#pragma unroll
        for (int l = 0; l < 4; l++) {
            int q = (sb & 1) ? (qs[l] >> 4) : (qs[l] & 0x0F);
            if constexpr (TYPE == B200_TYPE_Q5_K) q += ((b[16 + l0 + l] >> sb) & 1) ? 16 : 0;  // qh bit sb (ggml-quants.c:2783-2786)
            o[l] = __fsub_rn(__fmul_rn(d1, (float)q), m1);
        }
    } else if constexpr (TYPE == B200_TYPE_Q2_K || TYPE == B200_TYPE_Q3_K) {
        // dequantize_row_q2_K / q3_K (ggml-quants.c:1980-2010, 2328-2373): value v sits in qs[32 (v / 128) + v % 32] at bits
        // 2 ((v % 128) / 32); its scale is number v / 16
        constexpr bool Q2 = TYPE == B200_TYPE_Q2_K;
        const unsigned char *b = src + (i / QK_K) * (Q2 ? Q2_K_BYTES : Q3_K_BYTES);
        const int v = (int)(i % QK_K), is = v / 16, sh = 2 * ((v % 128) / 32);
        const unsigned char *qs = b + (Q2 ? 16 : 32) + 32 * (v / 128) + v % 32;
        if constexpr (Q2) {
            const float dl = __fmul_rn(ldh(b + 80), (float)(b[is] & 0xF)), ml = __fmul_rn(ldh(b + 82), (float)(b[is] >> 4));
#pragma unroll
            for (int l = 0; l < 4; l++) o[l] = __fsub_rn(__fmul_rn(dl, (float)((qs[l] >> sh) & 3)), ml);
        } else {
            unsigned sw[3], aux[4];
#pragma unroll
            for (int k = 0; k < 3; k++) sw[k] = (unsigned)b[96 + 4 * k] | ((unsigned)b[97 + 4 * k] << 8) | ((unsigned)b[98 + 4 * k] << 16) | ((unsigned)b[99 + 4 * k] << 24);
            q3k_scales(sw[0], sw[1], sw[2], aux);
            const int sc = (int)((aux[is >> 2] >> (8 * (is & 3))) & 0xff) - 32;
            const float dl = __fmul_rn(ldh(b + 108), (float)sc);
            const unsigned char *hm = b + v % 32;
#pragma unroll
            for (int l = 0; l < 4; l++) o[l] = __fmul_rn(dl, (float)((int)((qs[l] >> sh) & 3) - (((hm[l] >> (v / 32)) & 1) ? 0 : 4)));
        }
    } else if constexpr (TYPE == B200_TYPE_Q6_K) {
        const unsigned char *b = src + (i / QK_K) * Q6_K_BYTES;
        const int j    = (int)(i % QK_K);
        const int half = j / 128;       // which 128-value half
        const int r    = (j % 128) / 32;  // 0: ql lo, 1: ql+32 lo, 2: ql hi, 3: ql+32 hi (ggml-quants.c:2990-2999)
        const int l0   = j % 32;
        const float d  = ldh(b + 208);
        const int is   = 8 * half + 2 * r + l0 / 16;
        const float ds = __fmul_rn(d, (float)(signed char)b[192 + is]);
        const unsigned char *ql = b + 64 * half + 32 * (r & 1) + l0;
        const unsigned char *qh = b + 128 + 32 * half + l0;
#pragma unroll
        for (int l = 0; l < 4; l++) {
            const int lo = (r & 2) ? (ql[l] >> 4) : (ql[l] & 0x0F);
            const int hi = (qh[l] >> (2 * r)) & 3;
            o[l] = __fmul_rn(ds, (float)((signed char)(lo | (hi << 4)) - 32));
        }
    }
    store4<T>(dst, i, o[0], o[1], o[2], o[3]);
}

template <typename T>
static int launch_dequant(int type, const void *src, T *dst, int64_t n, cudaStream_t st) {
    const unsigned grid = (unsigned)((n / 4 + 255) / 256);
    const unsigned char *s = (const unsigned char *)src;
    switch (type) {
        case B200_TYPE_Q4_0: dequantize_kernel<B200_TYPE_Q4_0, T><<<grid, 256, 0, st>>>(s, dst, n); break;
        case B200_TYPE_Q8_0: dequantize_kernel<B200_TYPE_Q8_0, T><<<grid, 256, 0, st>>>(s, dst, n); break;
        case B200_TYPE_Q4_1: dequantize_kernel<B200_TYPE_Q4_1, T><<<grid, 256, 0, st>>>(s, dst, n); break;
        case B200_TYPE_Q5_0: dequantize_kernel<B200_TYPE_Q5_0, T><<<grid, 256, 0, st>>>(s, dst, n); break;
        case B200_TYPE_Q5_1: dequantize_kernel<B200_TYPE_Q5_1, T><<<grid, 256, 0, st>>>(s, dst, n); break;
        case B200_TYPE_Q2_K: dequantize_kernel<B200_TYPE_Q2_K, T><<<grid, 256, 0, st>>>(s, dst, n); break;
        case B200_TYPE_Q3_K: dequantize_kernel<B200_TYPE_Q3_K, T><<<grid, 256, 0, st>>>(s, dst, n); break;
        case B200_TYPE_IQ4_NL: dequantize_kernel<B200_TYPE_IQ4_NL, T><<<grid, 256, 0, st>>>(s, dst, n); break;
        case B200_TYPE_IQ4_XS: dequantize_kernel<B200_TYPE_IQ4_XS, T><<<grid, 256, 0, st>>>(s, dst, n); break;
        case B200_TYPE_Q4_K: dequantize_kernel<B200_TYPE_Q4_K, T><<<grid, 256, 0, st>>>(s, dst, n); break;
        case B200_TYPE_Q5_K: dequantize_kernel<B200_TYPE_Q5_K, T><<<grid, 256, 0, st>>>(s, dst, n); break;
        case B200_TYPE_Q6_K: dequantize_kernel<B200_TYPE_Q6_K, T><<<grid, 256, 0, st>>>(s, dst, n); break;
        default: set_error("b200_dequantize: unsupported type %d", type); return B200_ERR_UNSUPPORTED;
    }
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_dequantize(int type, const void *src, void *dst, int dst_type, int64_t n, b200_stream_t stream) {
    if (!src || !dst || n < 0) { set_error("b200_dequantize: bad arguments"); return B200_ERR_INVALID; }
    const int bv = type_block_values(type);
    if (bv <= 1) { set_error("b200_dequantize: unsupported type %d", type); return B200_ERR_UNSUPPORTED; }
    if (n % bv) { set_error("b200_dequantize: n=%lld not a multiple of the block size %d", (long long)n, bv); return B200_ERR_INVALID; }
    if (n == 0) return B200_OK;
    if (dst_type == B200_TYPE_F32) return launch_dequant<float>(type, src, (float *)dst, n, (cudaStream_t)stream);
    if (dst_type == B200_TYPE_F16) return launch_dequant<__half>(type, src, (__half *)dst, n, (cudaStream_t)stream);
    set_error("b200_dequantize: dst_type must be F32 or F16");
    return B200_ERR_UNSUPPORTED;
}

extern "C" int b200_quantize_q8_1(const float *x, int64_t x_stride, void *y, int64_t K, int64_t N, b200_stream_t stream) {
    if (!x || !y || K < 0 || N < 0) { set_error("b200_quantize_q8_1: bad arguments"); return B200_ERR_INVALID; }
    if (K % QK) { set_error("b200_quantize_q8_1: K=%lld not a multiple of 32", (long long)K); return B200_ERR_INVALID; }
    if (((uintptr_t)x & 15) || (x_stride & 3)) { set_error("b200_quantize_q8_1: x must be 16-byte aligned with a row stride multiple of 4"); return B200_ERR_UNSUPPORTED; }
    if (K == 0 || N == 0) return B200_OK;
    if (N > 65535) { set_error("b200_quantize_q8_1: N too large"); return B200_ERR_UNSUPPORTED; }
    dim3 grid((unsigned)((K / 4 + 255) / 256), (unsigned)N);
    quantize_q8_1_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, x_stride, (unsigned char *)y, K);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}
