// ops.cu -- the hot-path-adjacent ops of a llama / mixtral graph, so a whole decode graph runs on
// the device with no CPU fallback.  All of them are HBM/L2-bound elementwise or row-reduction kernels; the design rule
// is vector access + one pass.  Semantics cite /root/reference/ggml/src/ggml.c; see include/b200_ggml_ops.h.
#include "b200_common.cuh"
#include "rope_kv.cuh"
#include "../../include/b200_ggml_ops.h"

namespace b200 {

struct TD {  // device-side tensor descriptor
    char   *data;
    int64_t ne[4];
    int64_t nb[4];
};
static TD td(const b200_tensor *t) {
    TD d;
    d.data = (char *)t->data;
    for (int i = 0; i < 4; i++) { d.ne[i] = t->ne[i]; d.nb[i] = (int64_t)t->nb[i]; }
    return d;
}
static int64_t nelements(const b200_tensor *t) { return t->ne[0] * t->ne[1] * t->ne[2] * t->ne[3]; }
static int64_t nrows(const b200_tensor *t) { return t->ne[1] * t->ne[2] * t->ne[3]; }
static bool contiguous(const b200_tensor *t) {
    const int bs = type_block_values(t->type), by = type_block_bytes(t->type);
    if (!bs) return false;
    return t->nb[0] == (size_t)by && t->nb[1] == t->nb[0] * (size_t)(t->ne[0] / bs) && t->nb[2] == t->nb[1] * (size_t)t->ne[1] &&
           t->nb[3] == t->nb[2] * (size_t)t->ne[2];
}
static bool same_shape(const b200_tensor *a, const b200_tensor *b) {
    return a->ne[0] == b->ne[0] && a->ne[1] == b->ne[1] && a->ne[2] == b->ne[2] && a->ne[3] == b->ne[3];
}
#define REQUIRE(cond, code, ...)                  \
    do {                                          \
        if (!(cond)) {                            \
            set_error(__VA_ARGS__);               \
            return code;                          \
        }                                         \
    } while (0)

// Launch with the programmatic-dependent-launch attribute when the backend asked for it: the kernel may be scheduled while
// its producer drains (it begins with griddepcontrol.wait) and lets its consumer's prologue start at once.
extern bool g_mmvq_pdl;
#define LAUNCH_PDL(kernel, grid, block, st, ...)                                       \
    do {                                                                               \
        cudaLaunchConfig_t cfg__ = {};                                                 \
        cfg__.gridDim = dim3(grid);                                                    \
        cfg__.blockDim = dim3(block);                                                  \
        cfg__.stream = (st);                                                           \
        cudaLaunchAttribute at__[1];                                                   \
        at__[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;               \
        at__[0].val.programmaticStreamSerializationAllowed = 1;                        \
        cfg__.attrs = at__;                                                            \
        cfg__.numAttrs = g_mmvq_pdl ? 1 : 0;                                           \
        B200_CUDA_TRY(cudaLaunchKernelEx(&cfg__, kernel, __VA_ARGS__));                \
    } while (0)
__device__ __forceinline__ void pdl_enter() { pdl_wait(); pdl_trigger(); }

__device__ __forceinline__ float block_sum(float v, float *sh) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
    if (nw == 1) return v;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    v = l < nw ? sh[l] : 0.f;
    return warp_sum(v);
}
__device__ __forceinline__ float block_max(float v, float *sh) {
    v = warp_max(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
    if (nw == 1) return v;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    v = l < nw ? sh[l] : -INFINITY;
    return warp_max(v);
}

// ------------------------------------------------------------------------------------------------ RMS_NORM (+MUL)
__global__ void rms_norm_kernel(TD x, const float *__restrict__ w, TD dst, float eps, float *__restrict__ norm_out) {
    __shared__ float sh[32];
    const int64_t r = blockIdx.x;
    const int64_t i1 = r % x.ne[1], i2 = (r / x.ne[1]) % x.ne[2], i3 = r / (x.ne[1] * x.ne[2]);
    const float *xr = (const float *)(x.data + i1 * x.nb[1] + i2 * x.nb[2] + i3 * x.nb[3]);
    float *yr = (float *)(dst.data + i1 * dst.nb[1] + i2 * dst.nb[2] + i3 * dst.nb[3]);
    const int n = (int)x.ne[0];
    float *nr = norm_out ? norm_out + r * n : nullptr;  // contiguous rows
    float s = 0.f;
    const bool vec = (n & 3) == 0 && (((uintptr_t)xr | (uintptr_t)yr | (uintptr_t)nr) & 15) == 0;
    if (vec) {
        for (int i = threadIdx.x; i < n / 4; i += blockDim.x) {
            const float4 v = reinterpret_cast<const float4 *>(xr)[i];
            s += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
        }
    } else {
        for (int i = threadIdx.x; i < n; i += blockDim.x) s += xr[i] * xr[i];
    }
    s = block_sum(s, sh);
    const float scale = rsqrtf(s / (float)n + eps);
    if (vec) {
        for (int i = threadIdx.x; i < n / 4; i += blockDim.x) {
            float4 v = reinterpret_cast<const float4 *>(xr)[i];
            v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
            if (nr) reinterpret_cast<float4 *>(nr)[i] = v;
            if (w) {
                const float4 ww = reinterpret_cast<const float4 *>(w)[i];
                v.x *= ww.x; v.y *= ww.y; v.z *= ww.z; v.w *= ww.w;
            }
            reinterpret_cast<float4 *>(yr)[i] = v;
        }
    } else {
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            const float v = xr[i] * scale;
            if (nr) nr[i] = v;
            yr[i] = w ? v * w[i] : v;
        }
    }
}

// fused SILU(gate) * up; optionally also stores SILU(gate) (the intermediate ggml node)
__global__ void silu_mul_kernel(const float *__restrict__ g, const float *__restrict__ u, float *__restrict__ act, float *__restrict__ dst, int64_t n) {
    pdl_enter();
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x = g[i];
    const float a = x / (1.0f + expf(-x));
    if (act) act[i] = a;
    dst[i] = a * u[i];
}

// ------------------------------------------------------------------------------------------------ binary with broadcast
template <int OP>
__global__ void binary_kernel(TD a, TD b, TD dst, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t i0 = i % dst.ne[0], i1 = (i / dst.ne[0]) % dst.ne[1], i2 = (i / (dst.ne[0] * dst.ne[1])) % dst.ne[2],
                  i3 = i / (dst.ne[0] * dst.ne[1] * dst.ne[2]);
    const float x = *(const float *)(a.data + i0 * a.nb[0] + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
    const float y = *(const float *)(b.data + (i0 % b.ne[0]) * b.nb[0] + (i1 % b.ne[1]) * b.nb[1] + (i2 % b.ne[2]) * b.nb[2] +
                                     (i3 % b.ne[3]) * b.nb[3]);
    float r;
    if (OP == B200_BIN_ADD) r = x + y;
    else if (OP == B200_BIN_SUB) r = x - y;
    else if (OP == B200_BIN_MUL) r = x * y;
    else r = x / y;
    *(float *)(dst.data + i0 * dst.nb[0] + i1 * dst.nb[1] + i2 * dst.nb[2] + i3 * dst.nb[3]) = r;
}
// fast path: everything contiguous, same shape (or b a single row), ne0 % 4 == 0
template <int OP>
__global__ void binary_vec_kernel(const float4 *__restrict__ a, const float4 *__restrict__ b, float4 *__restrict__ dst, int64_t n4,
                                  int64_t b_n4) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const float4 x = a[i], y = b[i % b_n4];
    float4 r;
    if (OP == B200_BIN_ADD) r = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
    else if (OP == B200_BIN_SUB) r = make_float4(x.x - y.x, x.y - y.y, x.z - y.z, x.w - y.w);
    else if (OP == B200_BIN_MUL) r = make_float4(x.x * y.x, x.y * y.y, x.z * y.z, x.w * y.w);
    else r = make_float4(x.x / y.x, x.y / y.y, x.z / y.z, x.w / y.w);
    dst[i] = r;
}

// ------------------------------------------------------------------------------------------------ unary
__device__ __forceinline__ float unary_apply(int op, float x, float p0, float p1) {
    switch (op) {
        case B200_UN_SILU: return x / (1.0f + expf(-x));                                                  // ggml.c:2574
        case B200_UN_GELU: return 0.5f * x * (1.0f + tanhf(0.79788456080286535587989211986876f * x * (1.0f + 0.044715f * x * x)));
        case B200_UN_GELU_QUICK: return x * (1.0f / (1.0f + expf(-1.702f * x)));
        case B200_UN_RELU: return fmaxf(x, 0.f);
        case B200_UN_TANH: return tanhf(x);
        case B200_UN_SIGMOID: return 1.0f / (1.0f + expf(-x));
        case B200_UN_SQR: return x * x;
        case B200_UN_SCALE: return x * p0;
        case B200_UN_CLAMP: return x < p0 ? p0 : (x > p1 ? p1 : x);
    }
    return x;
}
__global__ void unary_kernel(int op, const float *__restrict__ x, float *__restrict__ dst, int64_t n, float p0, float p1) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = unary_apply(op, x[i], p0, p1);
}

// ------------------------------------------------------------------------------------------------ ROPE
// T = float, or __half for the in-place K-shift over an f16 cache (build_k_shift, src/llama.cpp:10144-10192; reference rope.cu f16 path)
template <typename T> struct RopeIO;
template <> struct RopeIO<float> {
    static __device__ __forceinline__ float2 ld2(const char *p) { return *(const float2 *)p; }
    static __device__ __forceinline__ void st2(char *p, float2 v) { *(float2 *)p = v; }
    static __device__ __forceinline__ float ld(const char *p) { return *(const float *)p; }
    static __device__ __forceinline__ void st(char *p, float v) { *(float *)p = v; }
};
template <> struct RopeIO<__half> {
    static __device__ __forceinline__ float2 ld2(const char *p) { return __half22float2(*(const __half2 *)p); }
    static __device__ __forceinline__ void st2(char *p, float2 v) { *(__half2 *)p = __floats2half2_rn(v.x, v.y); }
    static __device__ __forceinline__ float ld(const char *p) { return __half2float(*(const __half *)p); }
    static __device__ __forceinline__ void st(char *p, float v) { *(__half *)p = __float2half_rn(v); }
};
template <typename T>
__global__ void rope_kernel(TD x, const int32_t *__restrict__ pos, const float *__restrict__ ff, TD dst, RopeDev rp) {
    // one thread per (pair, head, token)
    constexpr int ES = (int)sizeof(T);
    using IO = RopeIO<T>;
    const int64_t half = x.ne[0] / 2;
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t total = half * x.ne[1] * x.ne[2] * x.ne[3];
    if (idx >= total) return;
    const int64_t ip = idx % half, i1 = (idx / half) % x.ne[1], i2 = (idx / (half * x.ne[1])) % x.ne[2], i3 = idx / (half * x.ne[1] * x.ne[2]);
    const char *xs = x.data + i1 * x.nb[1] + i2 * x.nb[2] + i3 * x.nb[3];
    char *ds = dst.data + i1 * dst.nb[1] + i2 * dst.nb[2] + i3 * dst.nb[3];
    const int64_t i0 = 2 * ip;
    if (i0 >= rp.n_dims) {  // pass-through tail (ggml.c:14386-14392)
        IO::st2(ds + i0 * ES, IO::ld2(xs + i0 * ES));
        return;
    }
    // theta = pos * theta_scale^(i0/2) / freq_factor; YaRN mix (rope_yarn, ggml.c:14223-14240)
    const float theta_extrap = (float)pos[i2] * powf(rp.theta_scale, (float)ip) / (ff ? ff[ip] : 1.0f);
    float theta = rp.freq_scale * theta_extrap;
    float mscale = rp.attn_factor;
    if (rp.ext_factor != 0.0f) {
        const float y = ((float)ip - rp.corr0) / fmaxf(0.001f, rp.corr1 - rp.corr0);
        const float ramp_mix = (1.0f - fminf(1.0f, fmaxf(0.0f, y))) * rp.ext_factor;
        theta = theta * (1.0f - ramp_mix) + theta_extrap * ramp_mix;
        mscale *= 1.0f + 0.1f * logf(1.0f / rp.freq_scale);
    }
    float sn, cs;
    sincosf(theta, &sn, &cs);
    sn *= mscale;
    cs *= mscale;
    if (!rp.neox) {
        const float2 v = IO::ld2(xs + i0 * ES);
        IO::st2(ds + i0 * ES, make_float2(v.x * cs - v.y * sn, v.x * sn + v.y * cs));
    } else {
        const float x0 = IO::ld(xs + ip * ES), x1 = IO::ld(xs + (ip + rp.n_dims / 2) * ES);
        IO::st(ds + ip * ES, x0 * cs - x1 * sn);
        IO::st(ds + (ip + rp.n_dims / 2) * ES, x0 * sn + x1 * cs);
    }
}

__global__ void rope_kv_store_kernel(RopeKvDev a, const int32_t *__restrict__ pos, const float *__restrict__ ff, RopeDev rp) {
    pdl_enter();
    rope_kv_store_item(a, pos, ff, rp, (int64_t)blockIdx.x * blockDim.x + threadIdx.x);
}

// ------------------------------------------------------------------------------------------------ CPY
template <typename S, typename D>
__device__ __forceinline__ D cvt(S v);
template <> __device__ __forceinline__ float  cvt<float, float>(float v) { return v; }
template <> __device__ __forceinline__ __half cvt<float, __half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ float  cvt<__half, float>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ __half cvt<__half, __half>(__half v) { return v; }
template <> __device__ __forceinline__ int    cvt<int, int>(int v) { return v; }

// `ind` (optional): device slot holding the destination base pointer, read at run time -- lets a captured CUDA graph
// follow a KV-cache write position that moves every token without being re-captured or patched.
template <typename S, typename D>
__global__ void cpy_kernel(TD src, TD dst, int64_t n, char *const *ind) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (ind) dst.data = *ind;
    // same linear element order, possibly different shapes (ggml_cpy only requires equal element counts)
    const int64_t s0 = i % src.ne[0], s1 = (i / src.ne[0]) % src.ne[1], s2 = (i / (src.ne[0] * src.ne[1])) % src.ne[2],
                  s3 = i / (src.ne[0] * src.ne[1] * src.ne[2]);
    const int64_t d0 = i % dst.ne[0], d1 = (i / dst.ne[0]) % dst.ne[1], d2 = (i / (dst.ne[0] * dst.ne[1])) % dst.ne[2],
                  d3 = i / (dst.ne[0] * dst.ne[1] * dst.ne[2]);
    const S v = *(const S *)(src.data + s0 * src.nb[0] + s1 * src.nb[1] + s2 * src.nb[2] + s3 * src.nb[3]);
    *(D *)(dst.data + d0 * dst.nb[0] + d1 * dst.nb[1] + d2 * dst.nb[2] + d3 * dst.nb[3]) = cvt<S, D>(v);
}

// f32 -> q8_0 (KV-cache store): 8 lanes per 32-value block; quantize_row_q8_0_ref arithmetic (ggml-quants.c:849-872)
__global__ void cpy_f32_q8_0_kernel(TD src, TD dst, int64_t nblocks, char *const *ind) {
    if (ind) dst.data = *ind;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t blk = t >> 3;
    const int part = (int)(t & 7);
    const bool active = blk < nblocks;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    unsigned char *out = nullptr;
    if (active) {
        const int64_t e = blk * 32;  // linear element index of the block start
        const int64_t s0 = e % src.ne[0], s1 = (e / src.ne[0]) % src.ne[1], s2 = (e / (src.ne[0] * src.ne[1])) % src.ne[2],
                      s3 = e / (src.ne[0] * src.ne[1] * src.ne[2]);
        v = *(const float4 *)(src.data + (s0 + 4 * part) * 4 + s1 * src.nb[1] + s2 * src.nb[2] + s3 * src.nb[3]);
        const int64_t bpr = dst.ne[0] / 32;  // blocks per dst row
        const int64_t d0 = blk % bpr, d1 = (blk / bpr) % dst.ne[1], d2 = (blk / (bpr * dst.ne[1])) % dst.ne[2], d3 = blk / (bpr * dst.ne[1] * dst.ne[2]);
        out = (unsigned char *)(dst.data + d0 * Q8_0_BYTES + d1 * dst.nb[1] + d2 * dst.nb[2] + d3 * dst.nb[3]);
    }
    float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if (!active) return;
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    const int q0 = (int)roundf(__fmul_rn(v.x, id)) & 0xff, q1 = (int)roundf(__fmul_rn(v.y, id)) & 0xff;
    const int q2 = (int)roundf(__fmul_rn(v.z, id)) & 0xff, q3 = (int)roundf(__fmul_rn(v.w, id)) & 0xff;
    unsigned short *o16 = reinterpret_cast<unsigned short *>(out + 2 + 4 * part);  // 34-byte blocks: 2-byte aligned
    o16[0] = (unsigned short)(q0 | (q1 << 8));
    o16[1] = (unsigned short)(q2 | (q3 << 8));
    if (part == 0) *reinterpret_cast<__half *>(out) = __float2half_rn(d);
}

// f32 -> q4_0 (KV-cache store with a q4_0 cache): 8 lanes per 32-value block; quantize_row_q4_0_ref arithmetic (ggml-quants.c:669-704):
// d = (the value of largest magnitude, first one on ties) / -8, q = min(15, trunc(x / d + 8.5)); byte j packs values j and j + 16
__global__ void cpy_f32_q4_0_kernel(TD src, TD dst, int64_t nblocks, char *const *ind) {
    if (ind) dst.data = *ind;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t blk = t >> 3;
    const int part = (int)(t & 7);
    const bool active = blk < nblocks;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    unsigned char *out = nullptr;
    if (active) {
        const int64_t e = blk * 32;
        const int64_t s0 = e % src.ne[0], s1 = (e / src.ne[0]) % src.ne[1], s2 = (e / (src.ne[0] * src.ne[1])) % src.ne[2],
                      s3 = e / (src.ne[0] * src.ne[1] * src.ne[2]);
        v = *(const float4 *)(src.data + (s0 + 4 * part) * 4 + s1 * src.nb[1] + s2 * src.nb[2] + s3 * src.nb[3]);
        const int64_t bpr = dst.ne[0] / 32;
        const int64_t d0 = blk % bpr, d1 = (blk / bpr) % dst.ne[1], d2 = (blk / (bpr * dst.ne[1])) % dst.ne[2], d3 = blk / (bpr * dst.ne[1] * dst.ne[2]);
        out = (unsigned char *)(dst.data + d0 * Q4_0_BYTES + d1 * dst.nb[1] + d2 * dst.nb[2] + d3 * dst.nb[3]);
    }
    // (|x|, x, index) of the first value of largest magnitude: in the lane, then over the block's 8 lanes
    const float xs[4] = {v.x, v.y, v.z, v.w};
    float amax = 0.0f, vmax = 0.0f;
    int imax = 32;
#pragma unroll
    for (int l = 0; l < 4; l++) {
        if (amax < fabsf(xs[l])) { amax = fabsf(xs[l]); vmax = xs[l]; imax = 4 * part + l; }
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
        const float a2 = __shfl_xor_sync(0xffffffffu, amax, o), v2 = __shfl_xor_sync(0xffffffffu, vmax, o);
        const int i2 = __shfl_xor_sync(0xffffffffu, imax, o);
        if (a2 > amax || (a2 == amax && i2 < imax)) { amax = a2; vmax = v2; imax = i2; }
    }
    const float d  = __fdiv_rn(vmax, -8.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    unsigned q = 0;
#pragma unroll
    for (int l = 0; l < 4; l++) {
        const int qi = min(15, __float2int_rz(__fadd_rn(__fmul_rn(xs[l], id), 8.5f)) & 0xff);
        q |= (unsigned)qi << (8 * l);
    }
    const unsigned qhi = __shfl_down_sync(0xffffffffu, q, 4);  // lanes 0..3 of the block take the values 16.. of lanes 4..7
    if (!active || part >= 4) return;
    const unsigned w = q | (qhi << 4);
    unsigned short *o16 = reinterpret_cast<unsigned short *>(out + 2 + 4 * part);  // 18-byte blocks: 2-byte aligned
    o16[0] = (unsigned short)(w & 0xffff);
    o16[1] = (unsigned short)(w >> 16);
    if (part == 0) *reinterpret_cast<__half *>(out) = __float2half_rn(d);
}

// q8_0 -> f32 (ggml_cast of a quantised K cache before the K-shift RoPE, src/llama.cpp:10165-10168; reference cpy.cu q8_0->f32):
// one thread per element of the linear element order; dequantize_row_q8_0 arithmetic (d * q, ggml-quants.c:1617-1632)
__global__ void cpy_q8_0_f32_kernel(TD src, TD dst, int64_t n, char *const *ind) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (ind) dst.data = *ind;
    const int64_t s0 = i % src.ne[0], s1 = (i / src.ne[0]) % src.ne[1], s2 = (i / (src.ne[0] * src.ne[1])) % src.ne[2],
                  s3 = i / (src.ne[0] * src.ne[1] * src.ne[2]);
    const int64_t d0 = i % dst.ne[0], d1 = (i / dst.ne[0]) % dst.ne[1], d2 = (i / (dst.ne[0] * dst.ne[1])) % dst.ne[2],
                  d3 = i / (dst.ne[0] * dst.ne[1] * dst.ne[2]);
    const unsigned char *b = (const unsigned char *)(src.data + (s0 >> 5) * Q8_0_BYTES + s1 * src.nb[1] + s2 * src.nb[2] + s3 * src.nb[3]);
    const float d = h2f(*reinterpret_cast<const unsigned short *>(b));
    *(float *)(dst.data + d0 * dst.nb[0] + d1 * dst.nb[1] + d2 * dst.nb[2] + d3 * dst.nb[3]) = __fmul_rn((float)(signed char)b[2 + (s0 & 31)], d);
}

// ------------------------------------------------------------------------------------------------ GET_ROWS
__device__ __forceinline__ float dq_value(int type, const unsigned char *row, int64_t i) {
    // scalar dequantisation of element i of a packed row (same arithmetic as quantize.cu's bit-exact kernel)
    switch (type) {
        case B200_TYPE_F32: return reinterpret_cast<const float *>(row)[i];
        case B200_TYPE_F16: return __half2float(reinterpret_cast<const __half *>(row)[i]);
        case B200_TYPE_Q4_0: {
            const unsigned char *b = row + (i / 32) * Q4_0_BYTES;
            const int j = (int)(i % 32);
            const int q = j < 16 ? (b[2 + j] & 0x0F) : (b[2 + j - 16] >> 4);
            return __fmul_rn((float)(q - 8), h2f(*reinterpret_cast<const unsigned short *>(b)));
        }
        case B200_TYPE_Q8_0: {
            const unsigned char *b = row + (i / 32) * Q8_0_BYTES;
            return __fmul_rn((float)(signed char)b[2 + i % 32], h2f(*reinterpret_cast<const unsigned short *>(b)));
        }
        case B200_TYPE_Q2_K: {  // dequantize_row_q2_K (ggml-quants.c:1980-2010)
            const unsigned char *b = row + (i / 256) * Q2_K_BYTES;
            const int v = (int)(i % 256), is = v / 16;
            const int q = (b[16 + 32 * (v / 128) + v % 32] >> (2 * ((v % 128) / 32))) & 3;
            const float dl = __fmul_rn(h2f(*reinterpret_cast<const unsigned short *>(b + 80)), (float)(b[is] & 0xF));
            const float ml = __fmul_rn(h2f(*reinterpret_cast<const unsigned short *>(b + 82)), (float)(b[is] >> 4));
            return __fsub_rn(__fmul_rn(dl, (float)q), ml);
        }
        case B200_TYPE_Q3_K: {  // dequantize_row_q3_K (:2328-2373)
            const unsigned char *b = row + (i / 256) * Q3_K_BYTES;
            const int v = (int)(i % 256), is = v / 16;
            unsigned sw[3], aux[4];
            for (int k = 0; k < 3; k++) sw[k] = (unsigned)b[96 + 4 * k] | ((unsigned)b[97 + 4 * k] << 8) | ((unsigned)b[98 + 4 * k] << 16) | ((unsigned)b[99 + 4 * k] << 24);
            q3k_scales(sw[0], sw[1], sw[2], aux);
            const int sc = (int)((aux[is >> 2] >> (8 * (is & 3))) & 0xff) - 32;
            const int q = (int)((b[32 + 32 * (v / 128) + v % 32] >> (2 * ((v % 128) / 32))) & 3) - (((b[v % 32] >> (v / 32)) & 1) ? 0 : 4);
            return __fmul_rn(__fmul_rn(h2f(*reinterpret_cast<const unsigned short *>(b + 108)), (float)sc), (float)q);
        }
        case B200_TYPE_IQ4_NL: {  // dequantize_row_iq4_nl (ggml-quants.c:3743-3760)
            const unsigned char *b = row + (i / 32) * IQ4_NL_BYTES;
            const int j = (int)(i % 32);
            const unsigned q = j < 16 ? (b[2 + j] & 0x0Fu) : (unsigned)(b[2 + j - 16] >> 4);
            return __fmul_rn(h2f(*reinterpret_cast<const unsigned short *>(b)), (float)iq4_level(q));
        }
        case B200_TYPE_IQ4_XS: {  // dequantize_row_iq4_xs (:3762-3785)
            const unsigned char *b = row + (i / 256) * IQ4_XS_BYTES;
            const int v = (int)(i % 256), ib = v / 32, j = v % 32;
            const unsigned sh = (unsigned)b[2] | ((unsigned)b[3] << 8);
            const int ls = (int)(((b[4 + ib / 2] >> (4 * (ib % 2))) & 0xFu) | (((sh >> (2 * ib)) & 3u) << 4)) - 32;
            const unsigned q = j < 16 ? (b[8 + 16 * ib + j] & 0x0Fu) : (unsigned)(b[8 + 16 * ib + j - 16] >> 4);
            return __fmul_rn(__fmul_rn(h2f(*reinterpret_cast<const unsigned short *>(b)), (float)ls), (float)iq4_level(q));
        }
        case B200_TYPE_Q4_1:
        case B200_TYPE_Q5_0:
        case B200_TYPE_Q5_1: {  // dequantize_row_q4_1 / q5_0 / q5_1 (ggml-quants.c:1543-1615)
            const bool has_m = type != B200_TYPE_Q5_0, has_h = type != B200_TYPE_Q4_1;
            const int qs = 2 + (has_m ? 2 : 0) + (has_h ? 4 : 0);
            const unsigned char *b = row + (i / 32) * type_block_bytes(type);
            const int j = (int)(i % 32);
            int q = j < 16 ? (b[qs + j] & 0x0F) : (b[qs + j - 16] >> 4);
            if (has_h) q |= ((b[qs - 4 + (j >> 3)] >> (j & 7)) & 1) << 4;
            const float d = h2f(*reinterpret_cast<const unsigned short *>(b));
            if (!has_m) return __fmul_rn((float)(q - 16), d);
            return __fadd_rn(__fmul_rn((float)q, d), h2f(*reinterpret_cast<const unsigned short *>(b + 2)));
        }
        case B200_TYPE_Q4_K:
        case B200_TYPE_Q5_K: {
            const int bb = type == B200_TYPE_Q4_K ? Q4_K_BYTES : Q5_K_BYTES;
            const unsigned char *b = row + (i / 256) * bb;
            const int j = (int)(i % 256), sb = j / 32, l = j % 32, c = sb / 2;
            unsigned sw[3];
            sw[0] = *reinterpret_cast<const unsigned *>(b + 4);
            sw[1] = *reinterpret_cast<const unsigned *>(b + 8);
            sw[2] = *reinterpret_cast<const unsigned *>(b + 12);
            int sc, mn;
            k4_scale_min(sw, sb, sc, mn);
            const unsigned char qb = b[(type == B200_TYPE_Q4_K ? 16 : 48) + 32 * c + l];
            int q = (sb & 1) ? (qb >> 4) : (qb & 0x0F);
            if (type == B200_TYPE_Q5_K) q += ((b[16 + l] >> sb) & 1) ? 16 : 0;
            const float d1 = __fmul_rn(h2f(*reinterpret_cast<const unsigned short *>(b)), (float)sc);
            const float m1 = __fmul_rn(h2f(*reinterpret_cast<const unsigned short *>(b + 2)), (float)mn);
            return __fsub_rn(__fmul_rn(d1, (float)q), m1);
        }
        case B200_TYPE_Q6_K: {
            const unsigned char *b = row + (i / 256) * Q6_K_BYTES;
            const int j = (int)(i % 256), half = j / 128, r = (j % 128) / 32, l = j % 32;
            const unsigned char qlb = b[64 * half + 32 * (r & 1) + l], qhb = b[128 + 32 * half + l];
            const int lo = (r & 2) ? (qlb >> 4) : (qlb & 0x0F), hi = (qhb >> (2 * r)) & 3;
            const float ds = __fmul_rn(h2f(*reinterpret_cast<const unsigned short *>(b + 208)), (float)(signed char)b[192 + 8 * half + 2 * r + l / 16]);
            return __fmul_rn(ds, (float)((lo | (hi << 4)) - 32));
        }
    }
    return 0.f;
}
__global__ void get_rows_kernel(TD src, int type, TD ids, TD dst) {
    // grid: (row index i10, i11, i12); threads over columns
    const int64_t i10 = blockIdx.x, i11 = blockIdx.y, i12 = blockIdx.z;
    const int32_t r = *(const int32_t *)(ids.data + i10 * ids.nb[0] + i11 * ids.nb[1] + i12 * ids.nb[2]);
    float *out = (float *)(dst.data + i10 * dst.nb[1] + i11 * dst.nb[2] + i12 * dst.nb[3]);
    const bool ok = r >= 0 && r < src.ne[1];
    const unsigned char *row = (const unsigned char *)(src.data + (int64_t)(ok ? r : 0) * src.nb[1] + i11 * src.nb[2] + i12 * src.nb[3]);
    for (int64_t c = threadIdx.x; c < dst.ne[0]; c += blockDim.x) out[c] = ok ? dq_value(type, row, c) : __int_as_float(0x7fc00000);
}

// ------------------------------------------------------------------------------------------------ SOFT_MAX
__global__ void soft_max_kernel(TD x, TD mask, int mask_f16, TD dst, float scale, float max_bias, float m0, float m1, int n_head_log2) {
    extern __shared__ float vals[];
    __shared__ float sh[32];
    const int64_t row = blockIdx.x;  // over ne1*ne2*ne3
    const int64_t i1 = row % x.ne[1], i2 = (row / x.ne[1]) % x.ne[2], i3 = row / (x.ne[1] * x.ne[2]);
    const float *xr = (const float *)(x.data + i1 * x.nb[1] + i2 * x.nb[2] + i3 * x.nb[3]);
    float *yr = (float *)(dst.data + i1 * dst.nb[1] + i2 * dst.nb[2] + i3 * dst.nb[3]);
    const int n = (int)x.ne[0];
    const unsigned h = (unsigned)i2;  // head (ggml.c:13958: (i1/ne01)%ne02 over the flattened row index)
    const float slope = max_bias > 0.0f ? (h < (unsigned)n_head_log2 ? powf(m0, (float)(h + 1)) : powf(m1, (float)(2 * (h - n_head_log2) + 1))) : 1.0f;
    const char *mrow = mask.data ? mask.data + i1 * mask.nb[1] : nullptr;
    float mx = -INFINITY;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        float v = xr[i] * scale;
        if (mrow) v += slope * (mask_f16 ? __half2float(((const __half *)mrow)[i]) : ((const float *)mrow)[i]);
        vals[i] = v;
        mx = fmaxf(mx, v);
    }
    mx = block_max(mx, sh);
    float s = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float e = expf(vals[i] - mx);
        vals[i] = e;
        s += e;
    }
    s = block_sum(s, sh);
    const float inv = 1.0f / s;
    for (int i = threadIdx.x; i < n; i += blockDim.x) yr[i] = vals[i] * inv;
}

// ------------------------------------------------------------------------------------------------ MoE router (fused)
// soft_max -> argsort(desc)[:n_used] -> get_rows -> sum_rows -> div of llm_build_moe_ffn (src/llama.cpp:9416-9470) for one token
// per warp: the five launches of the node chain become one.  Every intermediate the graph defines is still written (probs, the
// first n_used entries of the argsort row, the gathered weights, their sum), so readers of those tensors see what they expect.
constexpr int ROUTER_MAX_EXPERT = 256, ROUTER_WARPS = 4;
__global__ void __launch_bounds__(ROUTER_WARPS * 32) moe_router_kernel(const float *__restrict__ logits, int64_t ld_logits, float *__restrict__ probs,
                                                                       int64_t ld_probs, int32_t *__restrict__ ids, int64_t ld_ids, float *__restrict__ w_gather,
                                                                       float *__restrict__ wsum, float *__restrict__ w_out, int n_expert, int n_used,
                                                                       int64_t n_tok) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int64_t tok = (int64_t)blockIdx.x * ROUTER_WARPS + (threadIdx.x >> 5);
    if (tok >= n_tok) return;
    constexpr int U = ROUTER_MAX_EXPERT / 32;
    float v[U];
    float mx = -INFINITY;
#pragma unroll
    for (int u = 0; u < U; u++) {
        const int e = lane + 32 * u;
        v[u] = e < n_expert ? logits[tok * ld_logits + e] : -INFINITY;
        mx = fmaxf(mx, v[u]);
    }
    mx = warp_max(mx);
    float s = 0.f;
#pragma unroll
    for (int u = 0; u < U; u++) {
        const int e = lane + 32 * u;
        const float ev = e < n_expert ? expf(v[u] - mx) : 0.f;
        v[u] = ev;
        s += ev;
    }
    s = warp_sum(s);
    const float inv = 1.0f / s;
#pragma unroll
    for (int u = 0; u < U; u++) {
        const int e = lane + 32 * u;
        const float pr = v[u] * inv;
        if (e < n_expert) probs[tok * ld_probs + e] = pr;
        v[u] = e < n_expert ? pr : -INFINITY;
    }
    float mine = 0.f, tot = 0.f;  // lane k keeps the k-th selected weight
    for (int k = 0; k < n_used; k++) {
        float best = -INFINITY;
        int bi = 0x7fffffff;
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (v[u] > best) { best = v[u]; bi = lane + 32 * u; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (bi == lane + 32 * u) v[u] = -INFINITY;
        }
        if (lane == k) mine = best;
        if (lane == 0) {
            ids[tok * ld_ids + k] = bi;
            w_gather[tok * n_used + k] = best;
        }
        tot += best;
    }
    if (lane == 0) wsum[tok] = tot;
    if (lane < n_used) w_out[tok * n_used + lane] = mine / tot;
}

// ------------------------------------------------------------------------------------------------ ARGSORT (bitonic, one CTA per row)
__global__ void argsort_kernel(TD x, TD dst, int order, int npad) {
    extern __shared__ int idx[];
    const int64_t row = blockIdx.x;
    const float *xr = (const float *)(x.data + row * x.nb[1]);
    int *out = (int *)(dst.data + row * dst.nb[1]);
    const int n = (int)x.ne[0];
    for (int i = threadIdx.x; i < npad; i += blockDim.x) idx[i] = i;
    __syncthreads();
    for (int k = 2; k <= npad; k *= 2) {
        for (int j = k / 2; j > 0; j /= 2) {
            for (int i = threadIdx.x; i < npad; i += blockDim.x) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const int a = idx[i], b = idx[ixj];
                    // padding indices (>= n) sort last; ties break on the index so the result is deterministic
                    bool a_before_b;
                    if (a >= n) a_before_b = false;
                    else if (b >= n) a_before_b = true;
                    else {
                        const float va = xr[a], vb = xr[b];
                        a_before_b = order == 0 ? (va < vb || (va == vb && a < b)) : (va > vb || (va == vb && a < b));
                    }
                    const bool up = (i & k) == 0;
                    if (up ? !a_before_b : a_before_b) { idx[i] = b; idx[ixj] = a; }
                }
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < n; i += blockDim.x) out[i] = idx[i];
}

// ------------------------------------------------------------------------------------------------ SUM_ROWS
__global__ void sum_rows_kernel(TD x, TD dst) {
    __shared__ float sh[32];
    const int64_t row = blockIdx.x;
    const int64_t i1 = row % x.ne[1], i2 = (row / x.ne[1]) % x.ne[2], i3 = row / (x.ne[1] * x.ne[2]);
    const float *xr = (const float *)(x.data + i1 * x.nb[1] + i2 * x.nb[2] + i3 * x.nb[3]);
    float s = 0.f;
    for (int i = threadIdx.x; i < x.ne[0]; i += blockDim.x) s += xr[i];
    s = block_sum(s, sh);
    if (threadIdx.x == 0) *(float *)(dst.data + i1 * dst.nb[1] + i2 * dst.nb[2] + i3 * dst.nb[3]) = s;
}

// ------------------------------------------------------------------------------------------------ MUL_MAT, float weights
template <typename W>
__global__ void mul_mat_float_kernel(TD w, TD x, TD dst) {
    // one warp per output element (m, n, i2, i3); ggml broadcast: w dims 2,3 divide x dims 2,3
    const int64_t gw = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int64_t M = dst.ne[0], N = dst.ne[1];
    const int64_t total = M * N * dst.ne[2] * dst.ne[3];
    if (gw >= total) return;
    const int64_t m = gw % M, n = (gw / M) % N, i2 = (gw / (M * N)) % dst.ne[2], i3 = gw / (M * N * dst.ne[2]);
    const int64_t w2 = i2 / (x.ne[2] / w.ne[2]), w3 = i3 / (x.ne[3] / w.ne[3]);
    const W *wr = (const W *)(w.data + m * w.nb[1] + w2 * w.nb[2] + w3 * w.nb[3]);
    const float *xr = (const float *)(x.data + n * x.nb[1] + i2 * x.nb[2] + i3 * x.nb[3]);
    float s = 0.f;
    for (int64_t k = lane; k < w.ne[0]; k += 32) s += cvt<W, float>(wr[k]) * xr[k];
    s = warp_sum(s);
    if (lane == 0) *(float *)(dst.data + m * dst.nb[0] + n * dst.nb[1] + i2 * dst.nb[2] + i3 * dst.nb[3]) = s;
}

static unsigned grid1d(int64_t n, int block) { return (unsigned)((n + block - 1) / block); }

}  // namespace b200

using namespace b200;

extern "C" int b200_op_rms_norm(const b200_tensor *x, const float *w, const b200_tensor *dst, float eps, b200_stream_t stream) {
    REQUIRE(x && dst && x->data && dst->data, B200_ERR_INVALID, "b200_op_rms_norm: NULL argument");
    REQUIRE(x->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && x->nb[0] == 4 && dst->nb[0] == 4 && same_shape(x, dst), B200_ERR_UNSUPPORTED,
            "b200_op_rms_norm: f32 rows only");
    const int64_t rows = nrows(x);
    if (rows == 0 || x->ne[0] == 0) return B200_OK;
    const int block = x->ne[0] >= 4096 ? 512 : (x->ne[0] >= 1024 ? 256 : 128);
    rms_norm_kernel<<<(unsigned)rows, block, 0, (cudaStream_t)stream>>>(td(x), w, td(dst), eps, nullptr);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_op_rms_norm_mul(const b200_tensor *x, const float *w, float *norm_out, const b200_tensor *dst, float eps, b200_stream_t stream) {
    REQUIRE(x && dst && w && x->data && dst->data, B200_ERR_INVALID, "b200_op_rms_norm_mul: NULL argument");
    REQUIRE(x->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && x->nb[0] == 4 && dst->nb[0] == 4 && same_shape(x, dst), B200_ERR_UNSUPPORTED,
            "b200_op_rms_norm_mul: f32 rows only");
    const int64_t rows = nrows(x);
    if (rows == 0 || x->ne[0] == 0) return B200_OK;
    const int block = x->ne[0] >= 4096 ? 512 : (x->ne[0] >= 1024 ? 256 : 128);
    rms_norm_kernel<<<(unsigned)rows, block, 0, (cudaStream_t)stream>>>(td(x), w, td(dst), eps, norm_out);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_op_silu_mul(const float *gate, const float *up, float *act_out, float *dst, int64_t n, b200_stream_t stream) {
    REQUIRE(gate && up && dst && n >= 0, B200_ERR_INVALID, "b200_op_silu_mul: NULL argument");
    if (n == 0) return B200_OK;
    LAUNCH_PDL(silu_mul_kernel, grid1d(n, 256), 256, (cudaStream_t)stream, gate, up, act_out, dst, n);
    return B200_OK;
}

extern "C" int b200_op_binary(int op, const b200_tensor *a, const b200_tensor *b, const b200_tensor *dst, b200_stream_t stream) {
    REQUIRE(a && b && dst && a->data && b->data && dst->data, B200_ERR_INVALID, "b200_op_binary: NULL argument");
    REQUIRE(a->type == B200_TYPE_F32 && b->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32, B200_ERR_UNSUPPORTED, "b200_op_binary: f32 only");
    REQUIRE(same_shape(a, dst), B200_ERR_INVALID, "b200_op_binary: src0 and dst shapes differ");
    for (int i = 0; i < 4; i++) REQUIRE(b->ne[i] > 0 && a->ne[i] % b->ne[i] == 0, B200_ERR_INVALID, "b200_op_binary: src1 cannot be broadcast over src0");
    const int64_t n = nelements(dst);
    if (n == 0) return B200_OK;
    const cudaStream_t st = (cudaStream_t)stream;
    const int64_t nb_el = nelements(b);
    const bool fast = contiguous(a) && contiguous(b) && contiguous(dst) && (n & 3) == 0 && (nb_el & 3) == 0 && (a->ne[0] & 3) == 0 &&
                      ((((uintptr_t)a->data | (uintptr_t)b->data | (uintptr_t)dst->data) & 15) == 0) &&
                      (nb_el == n || (nb_el == b->ne[0] && b->ne[0] == a->ne[0]));
    if (fast) {
        const unsigned g = grid1d(n / 4, 256);
        switch (op) {
            case B200_BIN_ADD: binary_vec_kernel<B200_BIN_ADD><<<g, 256, 0, st>>>((const float4 *)a->data, (const float4 *)b->data, (float4 *)dst->data, n / 4, nb_el / 4); break;
            case B200_BIN_SUB: binary_vec_kernel<B200_BIN_SUB><<<g, 256, 0, st>>>((const float4 *)a->data, (const float4 *)b->data, (float4 *)dst->data, n / 4, nb_el / 4); break;
            case B200_BIN_MUL: binary_vec_kernel<B200_BIN_MUL><<<g, 256, 0, st>>>((const float4 *)a->data, (const float4 *)b->data, (float4 *)dst->data, n / 4, nb_el / 4); break;
            case B200_BIN_DIV: binary_vec_kernel<B200_BIN_DIV><<<g, 256, 0, st>>>((const float4 *)a->data, (const float4 *)b->data, (float4 *)dst->data, n / 4, nb_el / 4); break;
            default: REQUIRE(false, B200_ERR_INVALID, "b200_op_binary: bad op %d", op);
        }
    } else {
        const unsigned g = grid1d(n, 256);
        switch (op) {
            case B200_BIN_ADD: binary_kernel<B200_BIN_ADD><<<g, 256, 0, st>>>(td(a), td(b), td(dst), n); break;
            case B200_BIN_SUB: binary_kernel<B200_BIN_SUB><<<g, 256, 0, st>>>(td(a), td(b), td(dst), n); break;
            case B200_BIN_MUL: binary_kernel<B200_BIN_MUL><<<g, 256, 0, st>>>(td(a), td(b), td(dst), n); break;
            case B200_BIN_DIV: binary_kernel<B200_BIN_DIV><<<g, 256, 0, st>>>(td(a), td(b), td(dst), n); break;
            default: REQUIRE(false, B200_ERR_INVALID, "b200_op_binary: bad op %d", op);
        }
    }
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_op_unary(int op, const b200_tensor *x, const b200_tensor *dst, float p0, float p1, b200_stream_t stream) {
    REQUIRE(x && dst && x->data && dst->data, B200_ERR_INVALID, "b200_op_unary: NULL argument");
    REQUIRE(x->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && contiguous(x) && contiguous(dst) && nelements(x) == nelements(dst),
            B200_ERR_UNSUPPORTED, "b200_op_unary: contiguous f32 only");
    REQUIRE(op >= 0 && op <= B200_UN_CLAMP, B200_ERR_INVALID, "b200_op_unary: bad op %d", op);
    const int64_t n = nelements(x);
    if (n == 0) return B200_OK;
    unary_kernel<<<grid1d(n, 256), 256, 0, (cudaStream_t)stream>>>(op, (const float *)x->data, (float *)dst->data, n, p0, p1);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_op_rope(const b200_tensor *x, const int32_t *pos, const float *freq_factors, const b200_tensor *dst,
                            const b200_rope_params *rp, b200_stream_t stream) {
    REQUIRE(x && dst && pos && rp && x->data && dst->data, B200_ERR_INVALID, "b200_op_rope: NULL argument");
    const bool f16 = x->type == B200_TYPE_F16;
    REQUIRE((x->type == B200_TYPE_F32 || f16) && dst->type == x->type && x->nb[0] == (f16 ? 2u : 4u) && dst->nb[0] == x->nb[0] && same_shape(x, dst),
            B200_ERR_UNSUPPORTED, "b200_op_rope: f32 or f16, same type in and out");
    REQUIRE(!f16 || ((((uintptr_t)x->data | (uintptr_t)dst->data | x->nb[1] | x->nb[2] | x->nb[3] | dst->nb[1] | dst->nb[2] | dst->nb[3]) & 3) == 0), B200_ERR_UNSUPPORTED,
            "b200_op_rope: f16 rows must be 4-byte aligned");
    REQUIRE(rp->n_dims > 0 && rp->n_dims % 2 == 0 && rp->n_dims <= x->ne[0] && x->ne[0] % 2 == 0, B200_ERR_INVALID, "b200_op_rope: bad n_dims");
    REQUIRE((rp->mode & ~2) == 0, B200_ERR_UNSUPPORTED, "b200_op_rope: mode %d not supported", rp->mode);
    RopeDev d;
    d.n_dims = rp->n_dims;
    d.neox = (rp->mode & 2) != 0;
    d.freq_scale = rp->freq_scale;
    d.ext_factor = rp->ext_factor;
    d.attn_factor = rp->attn_factor;
    d.theta_scale = powf(rp->freq_base, -2.0f / rp->n_dims);
    // ggml_rope_yarn_corr_dims, ggml.c:14262-14270
    auto corr_dim = [&](float n_rot) { return rp->n_dims * logf(rp->n_ctx_orig / (n_rot * 2 * 3.14159265358979323846f)) / (2 * logf(rp->freq_base)); };
    const float start = floorf(corr_dim(rp->beta_fast)), end = ceilf(corr_dim(rp->beta_slow));
    d.corr0 = start > 0 ? start : 0;
    d.corr1 = end < rp->n_dims - 1 ? end : (float)(rp->n_dims - 1);
    const int64_t total = x->ne[0] / 2 * x->ne[1] * x->ne[2] * x->ne[3];
    if (total == 0) return B200_OK;
    if (f16) rope_kernel<__half><<<grid1d(total, 256), 256, 0, (cudaStream_t)stream>>>(td(x), pos, freq_factors, td(dst), d);
    else rope_kernel<float><<<grid1d(total, 256), 256, 0, (cudaStream_t)stream>>>(td(x), pos, freq_factors, td(dst), d);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

static RopeDev rope_dev(const b200_rope_params *rp) {
    RopeDev d;
    d.n_dims = rp->n_dims;
    d.neox = (rp->mode & 2) != 0;
    d.freq_scale = rp->freq_scale;
    d.ext_factor = rp->ext_factor;
    d.attn_factor = rp->attn_factor;
    d.theta_scale = powf(rp->freq_base, -2.0f / rp->n_dims);
    auto corr_dim = [&](float n_rot) { return rp->n_dims * logf(rp->n_ctx_orig / (n_rot * 2 * 3.14159265358979323846f)) / (2 * logf(rp->freq_base)); };
    const float start = floorf(corr_dim(rp->beta_fast)), end = ceilf(corr_dim(rp->beta_slow));
    d.corr0 = start > 0 ? start : 0;
    d.corr1 = end < rp->n_dims - 1 ? end : (float)(rp->n_dims - 1);
    return d;
}

extern "C" int b200_op_rope_kv_store(const b200_tensor *q, const b200_tensor *q_dst, const b200_tensor *k, const b200_tensor *k_dst,
                                     const b200_tensor *v, const int32_t *pos, const float *freq_factors, const b200_rope_params *rp,
                                     int kv_type, void *k_cache, void *v_cache, void *const *k_slot, void *const *v_slot, b200_stream_t stream) {
    REQUIRE(q && q_dst && k && k_dst && v && pos && rp, B200_ERR_INVALID, "b200_op_rope_kv_store: NULL argument");
    REQUIRE((k_cache || k_slot) && (v_cache || v_slot), B200_ERR_INVALID, "b200_op_rope_kv_store: no cache destination");
    REQUIRE(kv_type == B200_TYPE_F16 || kv_type == B200_TYPE_Q8_0, B200_ERR_UNSUPPORTED, "b200_op_rope_kv_store: cache type %d", kv_type);
    for (const b200_tensor *t : {q, q_dst, k, k_dst, v})
        REQUIRE(t->type == B200_TYPE_F32 && contiguous(t) && t->data && ((uintptr_t)t->data & 15) == 0, B200_ERR_UNSUPPORTED, "b200_op_rope_kv_store: contiguous aligned f32 only");
    REQUIRE(same_shape(q, q_dst) && same_shape(k, k_dst) && q->ne[0] == k->ne[0] && q->ne[2] == k->ne[2] && q->ne[3] == 1 && k->ne[3] == 1, B200_ERR_INVALID,
            "b200_op_rope_kv_store: shapes");
    REQUIRE(nelements(v) == nelements(k), B200_ERR_INVALID, "b200_op_rope_kv_store: V and K sizes differ");
    REQUIRE((rp->mode & ~2) == 0 && (rp->mode & 2) == 0, B200_ERR_UNSUPPORTED, "b200_op_rope_kv_store: NORM rope ordering only");
    REQUIRE(q->ne[0] % 32 == 0 && rp->n_dims > 0 && rp->n_dims % 2 == 0 && rp->n_dims <= q->ne[0], B200_ERR_UNSUPPORTED, "b200_op_rope_kv_store: head size");
    RopeKvDev a;
    a.q = (const float *)q->data; a.k = (const float *)k->data; a.v = (const float *)v->data;
    a.q_out = (float *)q_dst->data; a.k_out = (float *)k_dst->data;
    a.k_cache = (char *)k_cache; a.v_cache = (char *)v_cache;
    a.k_slot = (char *const *)k_slot; a.v_slot = (char *const *)v_slot;
    a.D = (int)q->ne[0]; a.H = (int)q->ne[1]; a.HK = (int)k->ne[1]; a.T = (int)q->ne[2]; a.kv_type = kv_type;
    const int64_t nq = nelements(q) / 4, nk = nelements(k) / 4;
    if (nq + nk == 0) return B200_OK;
    const int64_t total = (nq + 31) / 32 * 32 + 2 * ((nk + 31) / 32 * 32);
    const RopeDev rd = rope_dev(rp);
    LAUNCH_PDL(rope_kv_store_kernel, grid1d(total, 256), 256, (cudaStream_t)stream, a, pos, freq_factors, rd);
    return B200_OK;
}

extern "C" int b200_op_cpy(const b200_tensor *src, const b200_tensor *dst, b200_stream_t stream) {
    return b200_op_cpy_indirect(src, dst, nullptr, stream);
}

extern "C" int b200_op_cpy_indirect(const b200_tensor *src, const b200_tensor *dst, void *const *dst_slot, b200_stream_t stream) {
    REQUIRE(src && dst && src->data && (dst->data || dst_slot), B200_ERR_INVALID, "b200_op_cpy: NULL argument");
    char *const *ind = (char *const *)dst_slot;
    const int64_t n = nelements(src);
    REQUIRE(n == nelements(dst), B200_ERR_INVALID, "b200_op_cpy: element counts differ");
    if (n == 0) return B200_OK;
    const cudaStream_t st = (cudaStream_t)stream;
    const int s = src->type, d = dst->type;
    if (s == d && contiguous(src) && contiguous(dst) && !ind) {
        const size_t bytes = (size_t)(n / type_block_values(s)) * type_block_bytes(s);
        B200_CUDA_TRY(cudaMemcpyAsync(dst->data, src->data, bytes, cudaMemcpyDeviceToDevice, st));
        return B200_OK;
    }
    const unsigned g = grid1d(n, 256);
    if (s == B200_TYPE_F32 && d == B200_TYPE_F32) cpy_kernel<float, float><<<g, 256, 0, st>>>(td(src), td(dst), n, ind);
    else if (s == B200_TYPE_F32 && d == B200_TYPE_F16) cpy_kernel<float, __half><<<g, 256, 0, st>>>(td(src), td(dst), n, ind);
    else if (s == B200_TYPE_F16 && d == B200_TYPE_F32) cpy_kernel<__half, float><<<g, 256, 0, st>>>(td(src), td(dst), n, ind);
    else if (s == B200_TYPE_F16 && d == B200_TYPE_F16) cpy_kernel<__half, __half><<<g, 256, 0, st>>>(td(src), td(dst), n, ind);
    else if (s == B200_TYPE_I32 && d == B200_TYPE_I32) cpy_kernel<int, int><<<g, 256, 0, st>>>(td(src), td(dst), n, ind);
    else if (s == B200_TYPE_F32 && d == B200_TYPE_Q8_0) {
        REQUIRE(src->nb[0] == 4 && src->ne[0] % 32 == 0 && dst->ne[0] % 32 == 0 && dst->nb[0] == Q8_0_BYTES && (((uintptr_t)src->data | src->nb[1] | src->nb[2] | src->nb[3]) & 15) == 0,
                B200_ERR_UNSUPPORTED, "b200_op_cpy: f32->q8_0 needs contiguous 16-byte aligned source rows of a multiple of 32 values");
        const int64_t nblk = n / 32;
        cpy_f32_q8_0_kernel<<<grid1d(nblk * 8, 256), 256, 0, st>>>(td(src), td(dst), nblk, ind);
    } else if (s == B200_TYPE_F32 && d == B200_TYPE_Q4_0) {
        REQUIRE(src->nb[0] == 4 && src->ne[0] % 32 == 0 && dst->ne[0] % 32 == 0 && dst->nb[0] == Q4_0_BYTES && (((uintptr_t)src->data | src->nb[1] | src->nb[2] | src->nb[3]) & 15) == 0,
                B200_ERR_UNSUPPORTED, "b200_op_cpy: f32->q4_0 needs contiguous 16-byte aligned source rows of a multiple of 32 values");
        const int64_t nblk = n / 32;
        cpy_f32_q4_0_kernel<<<grid1d(nblk * 8, 256), 256, 0, st>>>(td(src), td(dst), nblk, ind);
    } else if (s == B200_TYPE_Q8_0 && d == B200_TYPE_F32) {
        REQUIRE(src->nb[0] == Q8_0_BYTES && src->ne[0] % 32 == 0 && dst->nb[0] == 4 && ((src->nb[1] | src->nb[2] | src->nb[3] | (uintptr_t)src->data) & 1) == 0,
                B200_ERR_UNSUPPORTED, "b200_op_cpy: q8_0->f32 needs whole 2-byte aligned source blocks");
        cpy_q8_0_f32_kernel<<<g, 256, 0, st>>>(td(src), td(dst), n, ind);
    } else {
        REQUIRE(false, B200_ERR_UNSUPPORTED, "b200_op_cpy: type pair (%d -> %d) not supported", s, d);
    }
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_op_get_rows(const b200_tensor *src, const b200_tensor *ids, const b200_tensor *dst, b200_stream_t stream) {
    REQUIRE(src && ids && dst && src->data && ids->data && dst->data, B200_ERR_INVALID, "b200_op_get_rows: NULL argument");
    REQUIRE(ids->type == B200_TYPE_I32 && dst->type == B200_TYPE_F32 && dst->nb[0] == 4, B200_ERR_UNSUPPORTED, "b200_op_get_rows: i32 ids, f32 dst");
    const int t = src->type;
    REQUIRE(t == B200_TYPE_F32 || t == B200_TYPE_F16 || t == B200_TYPE_Q4_0 || t == B200_TYPE_Q8_0 || t == B200_TYPE_Q4_K || t == B200_TYPE_Q5_K || t == B200_TYPE_Q6_K ||
            t == B200_TYPE_Q4_1 || t == B200_TYPE_Q5_0 || t == B200_TYPE_Q5_1 || t == B200_TYPE_Q2_K || t == B200_TYPE_Q3_K ||
            t == B200_TYPE_IQ4_NL || t == B200_TYPE_IQ4_XS,
            B200_ERR_UNSUPPORTED, "b200_op_get_rows: source type %d not supported", t);
    REQUIRE(dst->ne[0] == src->ne[0] && dst->ne[1] == ids->ne[0], B200_ERR_INVALID, "b200_op_get_rows: shape mismatch");
    if (nelements(dst) == 0) return B200_OK;
    REQUIRE(ids->ne[1] <= 65535 && ids->ne[2] <= 65535, B200_ERR_UNSUPPORTED, "b200_op_get_rows: too many index rows");
    dim3 grid((unsigned)ids->ne[0], (unsigned)ids->ne[1], (unsigned)ids->ne[2]);
    get_rows_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(td(src), t, td(ids), td(dst));
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_op_soft_max(const b200_tensor *x, const b200_tensor *mask, const b200_tensor *dst, float scale, float max_bias,
                                b200_stream_t stream) {
    REQUIRE(x && dst && x->data && dst->data, B200_ERR_INVALID, "b200_op_soft_max: NULL argument");
    REQUIRE(x->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && x->nb[0] == 4 && dst->nb[0] == 4 && same_shape(x, dst), B200_ERR_UNSUPPORTED,
            "b200_op_soft_max: f32 only");
    REQUIRE(x->ne[0] * 4 <= 200 * 1024, B200_ERR_UNSUPPORTED, "b200_op_soft_max: row of %lld values too long", (long long)x->ne[0]);
    TD m = {};
    int mf16 = 0;
    if (mask && mask->data) {
        REQUIRE(mask->type == B200_TYPE_F16 || mask->type == B200_TYPE_F32, B200_ERR_UNSUPPORTED, "b200_op_soft_max: mask must be f16 or f32");
        REQUIRE(mask->ne[0] == x->ne[0] && mask->ne[1] >= x->ne[1], B200_ERR_INVALID, "b200_op_soft_max: mask shape");
        m = td(mask);
        mf16 = mask->type == B200_TYPE_F16;
    }
    const int64_t rows = nrows(x);
    if (rows == 0 || x->ne[0] == 0) return B200_OK;
    int n_head_log2 = 1;
    while (n_head_log2 * 2 <= x->ne[2]) n_head_log2 *= 2;
    const float m0 = powf(2.0f, -max_bias / n_head_log2), m1 = powf(2.0f, -(max_bias / 2.0f) / n_head_log2);
    const size_t smem = (size_t)x->ne[0] * 4;
    if (smem > 48 * 1024) B200_CUDA_TRY(cudaFuncSetAttribute(soft_max_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int block = x->ne[0] >= 1024 ? 256 : 128;
    soft_max_kernel<<<(unsigned)rows, block, smem, (cudaStream_t)stream>>>(td(x), m, mf16, td(dst), scale, max_bias, m0, m1, n_head_log2);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_op_moe_router(const float *logits, int64_t ld_logits, float *probs, int64_t ld_probs, int32_t *ids, int64_t ld_ids, float *w_gather,
                                  float *wsum, float *w_out, int32_t n_expert, int32_t n_used, int64_t n_tok, b200_stream_t stream) {
    REQUIRE(logits && probs && ids && w_gather && wsum && w_out, B200_ERR_INVALID, "b200_op_moe_router: NULL argument");
    REQUIRE(n_expert >= 1 && n_expert <= ROUTER_MAX_EXPERT && n_used >= 1 && n_used <= 32 && n_used <= n_expert && n_tok >= 0, B200_ERR_UNSUPPORTED,
            "b200_op_moe_router: up to %d experts, up to 32 used", ROUTER_MAX_EXPERT);
    if (n_tok == 0) return B200_OK;
    LAUNCH_PDL(moe_router_kernel, (unsigned)((n_tok + ROUTER_WARPS - 1) / ROUTER_WARPS), ROUTER_WARPS * 32, (cudaStream_t)stream, logits, ld_logits, probs, ld_probs, ids,
               ld_ids, w_gather, wsum, w_out, (int)n_expert, (int)n_used, n_tok);
    return B200_OK;
}

extern "C" int b200_op_argsort(const b200_tensor *x, const b200_tensor *dst, int order, b200_stream_t stream) {
    REQUIRE(x && dst && x->data && dst->data, B200_ERR_INVALID, "b200_op_argsort: NULL argument");
    REQUIRE(x->type == B200_TYPE_F32 && dst->type == B200_TYPE_I32 && contiguous(x) && contiguous(dst) && same_shape(x, dst), B200_ERR_UNSUPPORTED,
            "b200_op_argsort: contiguous f32 -> i32");
    REQUIRE(x->ne[0] <= 1024, B200_ERR_UNSUPPORTED, "b200_op_argsort: rows longer than 1024");
    const int64_t rows = nrows(x);
    if (rows == 0 || x->ne[0] == 0) return B200_OK;
    int npad = 1;
    while (npad < x->ne[0]) npad *= 2;
    b200_tensor xf = *x, df = *dst;  // flatten rows
    xf.ne[1] = rows; df.ne[1] = rows;
    argsort_kernel<<<(unsigned)rows, npad < 32 ? 32 : (npad > 256 ? 256 : npad), npad * sizeof(int), (cudaStream_t)stream>>>(td(&xf), td(&df), order, npad);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_op_sum_rows(const b200_tensor *x, const b200_tensor *dst, b200_stream_t stream) {
    REQUIRE(x && dst && x->data && dst->data, B200_ERR_INVALID, "b200_op_sum_rows: NULL argument");
    REQUIRE(x->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && x->nb[0] == 4 && dst->ne[0] == 1 && dst->ne[1] == x->ne[1] && dst->ne[2] == x->ne[2] && dst->ne[3] == x->ne[3],
            B200_ERR_UNSUPPORTED, "b200_op_sum_rows: f32 rows");
    const int64_t rows = nrows(x);
    if (rows == 0) return B200_OK;
    sum_rows_kernel<<<(unsigned)rows, x->ne[0] >= 1024 ? 256 : 32, 0, (cudaStream_t)stream>>>(td(x), td(dst));
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_op_mul_mat_float(const b200_tensor *w, const b200_tensor *x, const b200_tensor *dst, b200_stream_t stream) {
    REQUIRE(w && x && dst && w->data && x->data && dst->data, B200_ERR_INVALID, "b200_op_mul_mat_float: NULL argument");
    REQUIRE((w->type == B200_TYPE_F32 || w->type == B200_TYPE_F16) && x->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32, B200_ERR_UNSUPPORTED,
            "b200_op_mul_mat_float: f32/f16 weights, f32 activations");
    REQUIRE(w->nb[0] == (size_t)type_block_bytes(w->type) && x->nb[0] == 4, B200_ERR_UNSUPPORTED, "b200_op_mul_mat_float: rows must be contiguous");
    REQUIRE(w->ne[0] == x->ne[0] && dst->ne[0] == w->ne[1] && dst->ne[1] == x->ne[1] && dst->ne[2] == x->ne[2] && dst->ne[3] == x->ne[3] &&
            w->ne[2] > 0 && w->ne[3] > 0 && x->ne[2] % w->ne[2] == 0 && x->ne[3] % w->ne[3] == 0, B200_ERR_INVALID, "b200_op_mul_mat_float: shape mismatch");
    const int64_t total = nelements(dst);
    if (total == 0) return B200_OK;
    const unsigned g = grid1d(total * 32, 256);
    if (w->type == B200_TYPE_F32) mul_mat_float_kernel<float><<<g, 256, 0, (cudaStream_t)stream>>>(td(w), td(x), td(dst));
    else mul_mat_float_kernel<__half><<<g, 256, 0, (cudaStream_t)stream>>>(td(w), td(x), td(dst));
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}
