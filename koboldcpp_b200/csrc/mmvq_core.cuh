// mmvq_core.cuh -- shared device code of the decode GEMV kernels (mmvq.cu: one launch per matmul; chain.cu: the persistent
// decode chain): activation quantisation into the lane-major shared-memory layout, and the per-type 128-weight slice dots.
#pragma once
#include "b200_common.cuh"

namespace b200 {

constexpr int MMVQ_NW        = 16;                 // warps per CTA
constexpr int MMVQ_THREADS   = MMVQ_NW * 32;
constexpr int MMVQ_MAX_STAGE = 8;
constexpr int MMVQ_BAR_BYTES = 128 * ((MMVQ_NW * MMVQ_MAX_STAGE * 8 + 127) / 128);

struct MmvqParams {
    b200_gemv_job  jobs[B200_GEMV_MAX_JOBS];
    int            n_jobs;
    int            n_stage;
    int            dbg;
    int            prefill;       // ring stages filled before griddepcontrol.wait
    int            n_it;          // ceil(K / 4096): steps per row
    int            n_expert;
    int            rows_q, rows_rem;   // rows per CTA = rows_q (+1 for the first rows_rem CTAs)
    int            tab_rows;           // MULTI: capacity of the per-CTA row table
    int            grp, part_rows;     // mmvq3 (K > 4096): warps per K-slice, rows of the per-slice partial-sum table
    unsigned       t2_mask;            // mmvq3 mixed-type launch: bit j = job j is of the second weight type
    unsigned       row_bytes2;         //   ... whose rows are row_bytes2 long and contiguous
    int64_t        K;
    int64_t        total_rows;
    size_t         w_row_bytes;   // row stride
    unsigned       row_bytes;     // payload bytes of one row
    int64_t        x_stride, dst_stride;
    const int32_t *ids;
    size_t         expert_bytes;
    // fused prologue / epilogue (b200_mul_mat_vec_q_fused): X is RMS-normalised and scaled by norm_w while it is quantised;
    // residual[j] (same layout as jobs[j].dst) is added to the result of job j
    const float   *norm_w;
    float          norm_eps;
    int            silu_pair;     // jobs[0] = gate, jobs[1] = up: rows_q / rows_rem count PAIRS; dst = silu(gate . x) * (up . x)
    const float   *residual[B200_GEMV_MAX_JOBS];
    // the folded nodes' own values (b200_gemv_fuse: never elide a graph node's output); NULL = not wanted
    float         *raw_dst[B200_GEMV_MAX_JOBS];
    float         *norm_out, *norm_mul_out;
    float         *gate_out, *up_out, *silu_out;
};

template <int TYPE>
struct Geo {
    static constexpr int SLICE_BYTES = type_slice_bytes(TYPE);          // 128 weights
    static constexpr int SEG_BYTES   = SLICE_BYTES * 32;               // 4096 weights: one warp step
    static constexpr int STAGE_BYTES = ((SEG_BYTES + 32 + 15) / 16) * 16;  // head misalignment (<16) + word over-read slack
};

// per-column activation storage for n_it steps: qs [it][8][32] int4, d [it][32] float4, s [it][32] float4, g [it][32] int4
__host__ __device__ constexpr int act_col_bytes(int n_it) { return n_it * (4096 + 3 * 512); }

struct ActView {
    const int4   *q;
    const float4 *d, *s;
    const int4   *g;
};
__device__ __forceinline__ ActView act_view(unsigned char *base, int n_it, int col) {
    unsigned char *p = base + (size_t)col * act_col_bytes(n_it);
    ActView v;
    v.q = (const int4 *)p;
    v.d = (const float4 *)(p + (size_t)n_it * 4096);
    v.s = (const float4 *)(p + (size_t)n_it * (4096 + 512));
    v.g = (const int4 *)(p + (size_t)n_it * (4096 + 1024));
    return v;
}

// rsqrt(mean(x^2) + eps) of one activation row, computed by the whole CTA (RMS_NORM, ggml.c:12060-12110)
// BAR == 0: all threads of the CTA take part (__syncthreads); BAR > 0: the first MMVQ_THREADS threads only (named barrier BAR)
template <int BAR, int NT>
__device__ __forceinline__ void cta_sync() {
    if constexpr (BAR == 0) __syncthreads();
    else asm volatile("bar.sync %0, %1;" ::"n"(BAR), "n"(NT) : "memory");
}
template <int BAR = 0, int NT = MMVQ_THREADS>
__device__ __forceinline__ float rms_scale(const float *__restrict__ xs, int64_t K, float norm_eps, float *red) {
    const int n4 = (int)(K / 4);
    float ss = 0.f;
    for (int i4 = threadIdx.x; i4 < n4; i4 += NT) {
        const float4 v = __ldcg(reinterpret_cast<const float4 *>(xs) + i4);
        ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    ss = warp_sum(ss);
    cta_sync<BAR, NT>();  // red[] may still be read by a previous call
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    cta_sync<BAR, NT>();
    float t = (threadIdx.x & 31) < NT / 32 ? red[threadIdx.x & 31] : 0.f;
    t = warp_sum(t);
    return rsqrtf(t / (float)K + norm_eps);
}

// Quantise one row x[0..K) (f32, global) into the column store `p` (q8_1 arithmetic of quantize_row_q8_1_ref).  With norm_w the
// value quantised is (x * nscale) * w, the order of the RMS_NORM -> MUL node pair.  All threads of the CTA call this.
template <int NT = MMVQ_THREADS>
__device__ __forceinline__ void quantize_activation_row(unsigned char *p, const float *__restrict__ x, int64_t K, int n_it,
                                                        const float *__restrict__ norm_w, float nscale,
                                                        float *__restrict__ norm_out = nullptr, float *__restrict__ norm_mul_out = nullptr) {
    const int n4 = (int)(K / 4);
    int   *aq = (int *)p;
    float *ad = (float *)(p + (size_t)n_it * 4096);
    float *as = (float *)(p + (size_t)n_it * (4096 + 512));
    int   *ag = (int *)(p + (size_t)n_it * (4096 + 1024));
    // 4 independent 16-byte loads per thread are issued before any of them is consumed (one L2 round trip per
    // 8192 activations instead of one per 2048)
    for (int base = 0; base < n4; base += 4 * NT) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int i4 = base + u * NT + threadIdx.x;
            v[u] = i4 < n4 ? __ldcg(reinterpret_cast<const float4 *>(x) + i4) : make_float4(0.f, 0.f, 0.f, 0.f);
            if (norm_w != nullptr && i4 < n4) {
                const float4 ww = __ldg(reinterpret_cast<const float4 *>(norm_w) + i4);
                v[u].x = v[u].x * nscale; v[u].y = v[u].y * nscale; v[u].z = v[u].z * nscale; v[u].w = v[u].w * nscale;
                if (norm_out) reinterpret_cast<float4 *>(norm_out)[i4] = v[u];
                v[u].x = v[u].x * ww.x; v[u].y = v[u].y * ww.y; v[u].z = v[u].z * ww.z; v[u].w = v[u].w * ww.w;
                if (norm_mul_out) reinterpret_cast<float4 *>(norm_mul_out)[i4] = v[u];
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int i4 = base + u * NT + threadIdx.x;
            if (base + u * NT >= n4) break;  // warp-uniform: the whole batch is past the end
            const bool active = i4 < n4;
            float d, s;
            int sum4;
            const int packed = quantize_q8_1_group8(v[u], d, s, sum4);
            const int hi4 = __shfl_xor_sync(0xffffffffu, sum4, 4);
            if (active) {
                const int blk = i4 >> 3, part = i4 & 7;
                const int slice = blk >> 2, b = blk & 3;
                const int it = slice >> 5, lane = slice & 31;
                const int j = b * 2 + (part >> 2);
                aq[(((it * 8 + j) * 32 + lane) << 2) + (part & 3)] = packed;
                if (part == 0) {
                    const int o = ((it * 32 + lane) << 2) + b;
                    ad[o] = d;
                    as[o] = s;
                    ag[o] = (sum4 & 0xffff) | (hi4 << 16);  // 16-wide integer sums (q6_K offset term)
                }
            }
        }
    }
}

// Quantise X[n][0..K) (f32, global) into the CTA's shared activation store.  All threads of the CTA call this.
template <int NCOLS>
__device__ void load_activations(unsigned char *act_base, const float *__restrict__ X, int64_t x_stride, int64_t K, int n_it,
                                 const float *__restrict__ norm_w, float norm_eps, float *red, float *norm_out = nullptr,
                                 float *norm_mul_out = nullptr) {
    for (int n = 0; n < NCOLS; n++) {
        const float *x = X + (int64_t)n * x_stride;
        const float nscale = norm_w != nullptr ? rms_scale(x, K, norm_eps, red) : 1.0f;
        quantize_activation_row(act_base + (size_t)n * act_col_bytes(n_it), x, K, n_it, norm_w, nscale,
                                norm_out ? norm_out + (int64_t)n * K : nullptr, norm_mul_out ? norm_mul_out + (int64_t)n * K : nullptr);
    }
}

// Register-resident, two-phase variant for one column (decode): act_issue() puts every load of the row (and of the norm weights)
// in flight and returns; act_finish() computes the RMS statistics from the registers, quantises and stores.  The GEMV issues the
// activation loads BEFORE it fills (most of) its weight ring: an SM ingests ~64 B/clk, so a load issued behind 190 KB of bulk
// copies waits ~1.5 us for its turn (measured with tools/trace_mmvq.py), while issued in front of them it returns in one L2 round
// trip and the quantisation overlaps the arrival of the weights.
// A thread owns 8 consecutive values (two float4), four lanes one q8_1 block: the per-block work (two IEEE divisions, the
// reductions, the fp16 roundings) is paid once per 8 values -- the quantisation is replicated in every CTA and sits on the critical
// path of every launch (1.4 us of a 3.9 us 4096 x 4096 launch when a thread owned 4 values, profiles/r02_gemv_trace_timeline.log).
constexpr int ACT_RR     = 8;                               // float4 per thread
constexpr int ACT_RR_MAX = ACT_RR * 4 * MMVQ_THREADS;       // 16384 values
template <int RR>
struct ActRegsT {
    float4 v[RR];
    float4 w[RR];
};
using ActRegs = ActRegsT<ACT_RR>;
// RR float4 per thread (even): RR * 4 * MMVQ_THREADS >= K (the decode kernel instantiates RR = 2 for K <= 4096)
template <int RR>
__device__ __forceinline__ void act_issue(ActRegsT<RR> &r, const float *__restrict__ x, const float *__restrict__ norm_w, int64_t K) {
    static_assert(RR % 2 == 0, "a thread owns pairs of float4");
    constexpr int NT = MMVQ_THREADS;
    const int n4 = (int)(K / 4);  // K is a multiple of 32
#pragma unroll
    for (int u = 0; u < RR / 2; u++) {
        const int i4 = 2 * (u * NT + threadIdx.x);
        if (2 * u * NT < n4) {  // warp-uniform: skip whole batches past the end
            r.v[2 * u]     = i4 < n4 ? __ldcg(reinterpret_cast<const float4 *>(x) + i4) : make_float4(0.f, 0.f, 0.f, 0.f);
            r.v[2 * u + 1] = i4 < n4 ? __ldcg(reinterpret_cast<const float4 *>(x) + i4 + 1) : make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
            r.v[2 * u] = r.v[2 * u + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    if (norm_w != nullptr) {
#pragma unroll
        for (int u = 0; u < RR / 2; u++) {
            const int i4 = 2 * (u * NT + threadIdx.x);
            if (2 * u * NT < n4) {
                r.w[2 * u]     = i4 < n4 ? __ldg(reinterpret_cast<const float4 *>(norm_w) + i4) : make_float4(0.f, 0.f, 0.f, 0.f);
                r.w[2 * u + 1] = i4 < n4 ? __ldg(reinterpret_cast<const float4 *>(norm_w) + i4 + 1) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
    }
}
// q8_1 quantisation of one 32-value block held by 4 consecutive lanes (8 values each): same arithmetic as quantize_q8_1_group8
// (quantize_row_q8_1_ref, ggml-quants.c:1160-1193).  lo / hi: the lane's 8 quantised values; sum16: sum of q over this lane's
// half of the block (16 values).
__device__ __forceinline__ void quantize_q8_1_group4(const float4 a, const float4 b, int &lo, int &hi, float &d16, float &s16, int &sum16) {
    float amax = fmaxf(fmaxf(fmaxf(fabsf(a.x), fabsf(a.y)), fmaxf(fabsf(a.z), fabsf(a.w))),
                       fmaxf(fmaxf(fabsf(b.x), fabsf(b.y)), fmaxf(fabsf(b.z), fabsf(b.w))));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    const int q0 = (int)roundf(__fmul_rn(a.x, id)), q1 = (int)roundf(__fmul_rn(a.y, id));
    const int q2 = (int)roundf(__fmul_rn(a.z, id)), q3 = (int)roundf(__fmul_rn(a.w, id));
    const int q4 = (int)roundf(__fmul_rn(b.x, id)), q5 = (int)roundf(__fmul_rn(b.y, id));
    const int q6 = (int)roundf(__fmul_rn(b.z, id)), q7 = (int)roundf(__fmul_rn(b.w, id));
    int s = (q0 + q1 + q2 + q3) + (q4 + q5 + q6 + q7);
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    sum16 = s;
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    d16 = __half2float(__float2half_rn(d));
    s16 = __half2float(__float2half_rn(__fmul_rn((float)s, d)));
    lo = (q0 & 0xff) | ((q1 & 0xff) << 8) | ((q2 & 0xff) << 16) | (int)((unsigned)q3 << 24);
    hi = (q4 & 0xff) | ((q5 & 0xff) << 8) | ((q6 & 0xff) << 16) | (int)((unsigned)q7 << 24);
}
// All threads of the CTA call this (it contains CTA barriers when norm is set).
template <int RR>
// it0: the registers hold the values [4096 it0, 4096 it0 + K) of a longer row of n_it steps (rows above ACT_RR_MAX values are
// quantised in chunks; not with norm, whose statistic needs the whole row first).
__device__ __forceinline__ void act_finish(ActRegsT<RR> &r, unsigned char *p, int64_t K, int n_it, bool norm, float norm_eps, float *red,
                                           float *__restrict__ norm_out = nullptr, float *__restrict__ norm_mul_out = nullptr, int it0 = 0) {
    constexpr int NT = MMVQ_THREADS;
    const int n4 = (int)(K / 4);
    int   *aq = (int *)p + (size_t)it0 * 1024;
    float *ad = (float *)(p + (size_t)n_it * 4096) + (size_t)it0 * 128;
    float *as = (float *)(p + (size_t)n_it * (4096 + 512)) + (size_t)it0 * 128;
    int   *ag = (int *)(p + (size_t)n_it * (4096 + 1024)) + (size_t)it0 * 128;
    if (norm) {
        float ss = 0.f;
#pragma unroll
        for (int u = 0; u < RR; u++) ss += r.v[u].x * r.v[u].x + r.v[u].y * r.v[u].y + r.v[u].z * r.v[u].z + r.v[u].w * r.v[u].w;
        ss = warp_sum(ss);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
        __syncthreads();
        float t = (threadIdx.x & 31) < NT / 32 ? red[threadIdx.x & 31] : 0.f;
        t = warp_sum(t);
        const float nscale = rsqrtf(t / (float)K + norm_eps);
#pragma unroll
        for (int u = 0; u < RR; u++) {
            const int i4 = 2 * ((u >> 1) * NT + threadIdx.x) + (u & 1);
            if ((u & ~1) * NT >= n4) break;  // warp-uniform
            r.v[u].x = r.v[u].x * nscale; r.v[u].y = r.v[u].y * nscale; r.v[u].z = r.v[u].z * nscale; r.v[u].w = r.v[u].w * nscale;
            if (norm_out && i4 < n4) reinterpret_cast<float4 *>(norm_out)[i4] = r.v[u];
            r.v[u].x = r.v[u].x * r.w[u].x; r.v[u].y = r.v[u].y * r.w[u].y; r.v[u].z = r.v[u].z * r.w[u].z; r.v[u].w = r.v[u].w * r.w[u].w;
            if (norm_mul_out && i4 < n4) reinterpret_cast<float4 *>(norm_mul_out)[i4] = r.v[u];
        }
    }
#pragma unroll
    for (int u = 0; u < RR / 2; u++) {
        const int i8 = u * NT + threadIdx.x;  // 8-value group
        if (2 * u * NT >= n4) break;          // warp-uniform: the whole batch is past the end
        const bool active = 2 * i8 < n4;
        float d, s;
        int lo, hi, sum16;
        quantize_q8_1_group4(r.v[2 * u], r.v[2 * u + 1], lo, hi, d, s, sum16);
        const int hi16 = __shfl_xor_sync(0xffffffffu, sum16, 2);
        if (active) {
            const int blk = i8 >> 2, part = i8 & 3;   // q8_1 block, quarter of the block
            const int slice = blk >> 2, b = blk & 3;  // 128-value slice (one lane's share of a step), block inside it
            const int it = slice >> 5, lane = slice & 31;
            const int j = b * 2 + (part >> 1);        // int4 inside the slice
            *reinterpret_cast<int2 *>(&aq[(((it * 8 + j) * 32 + lane) << 2) + 2 * (part & 1)]) = make_int2(lo, hi);
            if (part == 0) {
                const int o = ((it * 32 + lane) << 2) + b;
                ad[o] = d;
                as[o] = s;
                ag[o] = (sum16 & 0xffff) | (hi16 << 16);  // 16-wide integer sums (q6_K offset term)
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ per-type slice dot
// Each returns, for every column, the contribution of this lane's 128-weight slice.
template <int TYPE, int NCOLS>
struct SliceDot;

// ---- Q4_K: half a super-block per lane.  block: half d, half dmin, u8 scales[12], u8 qs[128]  (ggml-common.h:286-298)
template <int NCOLS>
struct SliceDot<B200_TYPE_Q4_K, NCOLS> {
    static __device__ __forceinline__ void run(const unsigned char *seg, int lane, int it, unsigned char *act_base, int n_it,
                                               int /*nblk*/, float acc[NCOLS]) {
        const unsigned char *b = seg + (lane >> 1) * Q4_K_BYTES;
        const int h = lane & 1;
        const uint4 hdr = *reinterpret_cast<const uint4 *>(b);
        const float d = h2f((unsigned short)(hdr.x & 0xffff)), dmin = h2f((unsigned short)(hdr.x >> 16));
        unsigned scw, mnw;  // 4 scales / 4 mins of sub-blocks 4h..4h+3, one per byte (get_scale_min_k4)
        if (h == 0) {
            scw = hdr.y & 0x3f3f3f3fu;
            mnw = hdr.z & 0x3f3f3f3fu;
        } else {
            scw = (hdr.w & 0x0f0f0f0fu) | ((hdr.y >> 2) & 0x30303030u);
            mnw = ((hdr.w >> 4) & 0x0f0f0f0fu) | ((hdr.z >> 2) & 0x30303030u);
        }
        const uint4 *qp = reinterpret_cast<const uint4 *>(b + 16 + 64 * h);
        unsigned v[16];
        *reinterpret_cast<uint4 *>(v + 0)  = qp[0];
        *reinterpret_cast<uint4 *>(v + 4)  = qp[1];
        *reinterpret_cast<uint4 *>(v + 8)  = qp[2];
        *reinterpret_cast<uint4 *>(v + 12) = qp[3];
#pragma unroll
        for (int n = 0; n < NCOLS; n++) {
            const ActView a = act_view(act_base, n_it, n);
            int sumi[4] = {0, 0, 0, 0};
#pragma unroll
            for (int cc = 0; cc < 2; cc++) {       // 64-value chunk: low nibbles -> sub-block 2cc, high -> 2cc+1
#pragma unroll
                for (int hq = 0; hq < 2; hq++) {   // which 16 bytes of the chunk
                    const int4 ulo = a.q[((it * 8 + (2 * cc) * 2 + hq) * 32) + lane];
                    const int4 uhi = a.q[((it * 8 + (2 * cc + 1) * 2 + hq) * 32) + lane];
                    const unsigned *w = v + 8 * cc + 4 * hq;
                    sumi[2 * cc]     = dp4a((int)(w[0] & 0x0f0f0f0fu), ulo.x, sumi[2 * cc]);
                    sumi[2 * cc]     = dp4a((int)(w[1] & 0x0f0f0f0fu), ulo.y, sumi[2 * cc]);
                    sumi[2 * cc]     = dp4a((int)(w[2] & 0x0f0f0f0fu), ulo.z, sumi[2 * cc]);
                    sumi[2 * cc]     = dp4a((int)(w[3] & 0x0f0f0f0fu), ulo.w, sumi[2 * cc]);
                    // high nibbles stay in place (value * 16, unsigned byte): exact >> 4 after the accumulation
                    sumi[2 * cc + 1] = dp4a_u(w[0] & 0xf0f0f0f0u, uhi.x, sumi[2 * cc + 1]);
                    sumi[2 * cc + 1] = dp4a_u(w[1] & 0xf0f0f0f0u, uhi.y, sumi[2 * cc + 1]);
                    sumi[2 * cc + 1] = dp4a_u(w[2] & 0xf0f0f0f0u, uhi.z, sumi[2 * cc + 1]);
                    sumi[2 * cc + 1] = dp4a_u(w[3] & 0xf0f0f0f0u, uhi.w, sumi[2 * cc + 1]);
                }
            }
            sumi[1] >>= 4;
            sumi[3] >>= 4;
            const float4 d8 = a.d[it * 32 + lane], s8 = a.s[it * 32 + lane];
            const float sd = d8.x * (float)(sumi[0] * (int)(scw & 0xff)) + d8.y * (float)(sumi[1] * (int)((scw >> 8) & 0xff)) +
                             d8.z * (float)(sumi[2] * (int)((scw >> 16) & 0xff)) + d8.w * (float)(sumi[3] * (int)(scw >> 24));
            const float sm = s8.x * (float)(mnw & 0xff) + s8.y * (float)((mnw >> 8) & 0xff) + s8.z * (float)((mnw >> 16) & 0xff) +
                             s8.w * (float)(mnw >> 24);
            acc[n] += d * sd - dmin * sm;
        }
    }
};

// ---- Q5_K: half d, half dmin, u8 scales[12], u8 qh[32], u8 qs[128]  (ggml-common.h:303-316)
template <int NCOLS>
struct SliceDot<B200_TYPE_Q5_K, NCOLS> {
    static __device__ __forceinline__ void run(const unsigned char *seg, int lane, int it, unsigned char *act_base, int n_it,
                                               int /*nblk*/, float acc[NCOLS]) {
        const unsigned char *b = seg + (lane >> 1) * Q5_K_BYTES;
        const int h = lane & 1;
        const uint4 hdr = *reinterpret_cast<const uint4 *>(b);
        const float d = h2f((unsigned short)(hdr.x & 0xffff)), dmin = h2f((unsigned short)(hdr.x >> 16));
        unsigned scw, mnw;
        if (h == 0) {
            scw = hdr.y & 0x3f3f3f3fu;
            mnw = hdr.z & 0x3f3f3f3fu;
        } else {
            scw = (hdr.w & 0x0f0f0f0fu) | ((hdr.y >> 2) & 0x30303030u);
            mnw = ((hdr.w >> 4) & 0x0f0f0f0fu) | ((hdr.z >> 2) & 0x30303030u);
        }
        unsigned qh[8], v[16];
        *reinterpret_cast<uint4 *>(qh + 0) = *reinterpret_cast<const uint4 *>(b + 16);
        *reinterpret_cast<uint4 *>(qh + 4) = *reinterpret_cast<const uint4 *>(b + 32);
        const uint4 *qp = reinterpret_cast<const uint4 *>(b + 48 + 64 * h);
        *reinterpret_cast<uint4 *>(v + 0)  = qp[0];
        *reinterpret_cast<uint4 *>(v + 4)  = qp[1];
        *reinterpret_cast<uint4 *>(v + 8)  = qp[2];
        *reinterpret_cast<uint4 *>(v + 12) = qp[3];
        // 5-bit values: sub-block sb = 4h + 2cc + {0: low nibble, 1: high nibble}; fifth bit = bit sb of qh[l] (ggml-quants.c:2783-2786)
        unsigned q5[4][8];
#pragma unroll
        for (int cc = 0; cc < 2; cc++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const unsigned w = v[8 * cc + i], hb = qh[i] >> (4 * h + 2 * cc);
                q5[2 * cc][i]     = (w & 0x0f0f0f0fu) | ((hb << 4) & 0x10101010u);
                q5[2 * cc + 1][i] = ((w >> 4) & 0x0f0f0f0fu) | ((hb << 3) & 0x10101010u);
            }
        }
#pragma unroll
        for (int n = 0; n < NCOLS; n++) {
            const ActView a = act_view(act_base, n_it, n);
            int sumi[4] = {0, 0, 0, 0};
#pragma unroll
            for (int sb = 0; sb < 4; sb++) {
#pragma unroll
                for (int hq = 0; hq < 2; hq++) {
                    const int4 u = a.q[((it * 8 + sb * 2 + hq) * 32) + lane];
                    sumi[sb] = dp4a((int)q5[sb][4 * hq + 0], u.x, sumi[sb]);
                    sumi[sb] = dp4a((int)q5[sb][4 * hq + 1], u.y, sumi[sb]);
                    sumi[sb] = dp4a((int)q5[sb][4 * hq + 2], u.z, sumi[sb]);
                    sumi[sb] = dp4a((int)q5[sb][4 * hq + 3], u.w, sumi[sb]);
                }
            }
            const float4 d8 = a.d[it * 32 + lane], s8 = a.s[it * 32 + lane];
            const float sd = d8.x * (float)(sumi[0] * (int)(scw & 0xff)) + d8.y * (float)(sumi[1] * (int)((scw >> 8) & 0xff)) +
                             d8.z * (float)(sumi[2] * (int)((scw >> 16) & 0xff)) + d8.w * (float)(sumi[3] * (int)(scw >> 24));
            const float sm = s8.x * (float)(mnw & 0xff) + s8.y * (float)((mnw >> 8) & 0xff) + s8.z * (float)((mnw >> 16) & 0xff) +
                             s8.w * (float)(mnw >> 24);
            acc[n] += d * sd - dmin * sm;
        }
    }
};

// ---- Q6_K: u8 ql[128], u8 qh[64], i8 scales[16], half d  (ggml-common.h:321-327); 210 bytes => 2-byte aligned
template <int NCOLS>
struct SliceDot<B200_TYPE_Q6_K, NCOLS> {
    static __device__ __forceinline__ void run(const unsigned char *seg, int lane, int it, unsigned char *act_base, int n_it,
                                               int /*nblk*/, float acc[NCOLS]) {
        const unsigned char *b = seg + (lane >> 1) * Q6_K_BYTES;
        const int h = lane & 1;
        unsigned ql[16], qh[8], scw[2];
        lds_words<16>(b + 64 * h, ql);
        lds_words<8>(b + 128 + 32 * h, qh);
        lds_words<2>(b + 192 + 8 * h, scw);
        const float d = h2f(*reinterpret_cast<const unsigned short *>(b + 208));
        // q[r][i]: 4 values of row-group r (= q8_1 block r of this slice), word i (ggml-quants.c:2990-2999)
        unsigned q[4][8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const unsigned l0 = ql[i], l1 = ql[8 + i], hb = qh[i];
            q[0][i] = (l0 & 0x0f0f0f0fu) | ((hb << 4) & 0x30303030u);
            q[1][i] = (l1 & 0x0f0f0f0fu) | ((hb << 2) & 0x30303030u);
            q[2][i] = ((l0 >> 4) & 0x0f0f0f0fu) | (hb & 0x30303030u);
            q[3][i] = ((l1 >> 4) & 0x0f0f0f0fu) | ((hb >> 2) & 0x30303030u);
        }
        const signed char *sc = reinterpret_cast<const signed char *>(scw);
#pragma unroll
        for (int n = 0; n < NCOLS; n++) {
            const ActView a = act_view(act_base, n_it, n);
            const float4 d8 = a.d[it * 32 + lane];
            const int4 g = a.g[it * 32 + lane];
            const float d8v[4] = {d8.x, d8.y, d8.z, d8.w};
            const int gv[4] = {g.x, g.y, g.z, g.w};
            float sd = 0.f;
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int4 u0 = a.q[((it * 8 + r * 2 + 0) * 32) + lane];
                const int4 u1 = a.q[((it * 8 + r * 2 + 1) * 32) + lane];
                int s0 = dp4a((int)q[r][0], u0.x, 0);
                s0 = dp4a((int)q[r][1], u0.y, s0);
                s0 = dp4a((int)q[r][2], u0.z, s0);
                s0 = dp4a((int)q[r][3], u0.w, s0);
                int s1 = dp4a((int)q[r][4], u1.x, 0);
                s1 = dp4a((int)q[r][5], u1.y, s1);
                s1 = dp4a((int)q[r][6], u1.z, s1);
                s1 = dp4a((int)q[r][7], u1.w, s1);
                // stored values are q+32: subtract 32 * (sum of the 16 activations of the group)
                s0 -= 32 * (int)(short)(gv[r] & 0xffff);
                s1 -= 32 * (gv[r] >> 16);
                sd += d8v[r] * (float)(s0 * (int)sc[2 * r] + s1 * (int)sc[2 * r + 1]);
            }
            acc[n] += d * sd;
        }
    }
};

// ---- Q8_0: half d, i8 qs[32] (ggml-common.h:186-191); 4 blocks (136 bytes) per lane
template <int NCOLS>
struct SliceDot<B200_TYPE_Q8_0, NCOLS> {
    static __device__ __forceinline__ void run(const unsigned char *seg, int lane, int it, unsigned char *act_base, int n_it,
                                               int nblk, float acc[NCOLS]) {
        const unsigned char *b = seg + lane * (4 * Q8_0_BYTES);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (k < nblk) {
                const unsigned char *bk = b + k * Q8_0_BYTES;
                const float d = h2f(*reinterpret_cast<const unsigned short *>(bk));
                unsigned w[8];
                lds_words<8>(bk + 2, w);
#pragma unroll
                for (int n = 0; n < NCOLS; n++) {
                    const ActView a = act_view(act_base, n_it, n);
                    const int4 u0 = a.q[((it * 8 + k * 2 + 0) * 32) + lane];
                    const int4 u1 = a.q[((it * 8 + k * 2 + 1) * 32) + lane];
                    int s = dp4a((int)w[0], u0.x, 0);
                    s = dp4a((int)w[1], u0.y, s);
                    s = dp4a((int)w[2], u0.z, s);
                    s = dp4a((int)w[3], u0.w, s);
                    s = dp4a((int)w[4], u1.x, s);
                    s = dp4a((int)w[5], u1.y, s);
                    s = dp4a((int)w[6], u1.z, s);
                    s = dp4a((int)w[7], u1.w, s);
                    const float d8 = reinterpret_cast<const float *>(a.d)[((it * 32 + lane) << 2) + k];
                    acc[n] += d * d8 * (float)s;
                }
            }
        }
    }
};

// ---- Q4_0: half d, u8 qs[16] (ggml-common.h:144-149); values 0..15 in low nibbles, 16..31 in high nibbles; 4 blocks (72 B) per lane
template <int NCOLS>
struct SliceDot<B200_TYPE_Q4_0, NCOLS> {
    static __device__ __forceinline__ void run(const unsigned char *seg, int lane, int it, unsigned char *act_base, int n_it,
                                               int nblk, float acc[NCOLS]) {
        const unsigned char *b = seg + lane * (4 * Q4_0_BYTES);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (k < nblk) {
                const unsigned char *bk = b + k * Q4_0_BYTES;
                const float d = h2f(*reinterpret_cast<const unsigned short *>(bk));
                unsigned w[4];
                lds_words<4>(bk + 2, w);
#pragma unroll
                for (int n = 0; n < NCOLS; n++) {
                    const ActView a = act_view(act_base, n_it, n);
                    const int4 u0 = a.q[((it * 8 + k * 2 + 0) * 32) + lane];
                    const int4 u1 = a.q[((it * 8 + k * 2 + 1) * 32) + lane];
                    int s = dp4a((int)(w[0] & 0x0f0f0f0fu), u0.x, 0);
                    s = dp4a((int)(w[1] & 0x0f0f0f0fu), u0.y, s);
                    s = dp4a((int)(w[2] & 0x0f0f0f0fu), u0.z, s);
                    s = dp4a((int)(w[3] & 0x0f0f0f0fu), u0.w, s);
                    int sh = dp4a_u(w[0] & 0xf0f0f0f0u, u1.x, 0);
                    sh = dp4a_u(w[1] & 0xf0f0f0f0u, u1.y, sh);
                    sh = dp4a_u(w[2] & 0xf0f0f0f0u, u1.z, sh);
                    sh = dp4a_u(w[3] & 0xf0f0f0f0u, u1.w, sh);
                    s += sh >> 4;
                    const int o = ((it * 32 + lane) << 2) + k;
                    const float d8 = reinterpret_cast<const float *>(a.d)[o], s8 = reinterpret_cast<const float *>(a.s)[o];
                    acc[n] += d * (d8 * (float)s - 8.0f * s8);  // sum (v-8)*u*d8 = d8*sum(v*u) - 8*s8  (vecdotq.cuh:23-42)
                }
            }
        }
    }
};

// ---- Q4_1 / Q5_0 / Q5_1 (ggml-common.h:155-183): 32-value blocks, 4 blocks per lane.  One block against one q8_1 block
// (u0 = its values 0..15, u1 = 16..31), the arithmetic of vec_dot_q4_1_q8_1 / q5_0_q8_1 / q5_1_q8_1 (vecdotq.cuh:44-149):
//   Q4_1: d d8 sum(v u) + m s8      Q5_0: d (d8 sum(v5 u) - 16 s8)      Q5_1: d d8 sum(v5 u) + m s8
// The fifth bit of value j is bit j of qh: four bits are spread to bit 4 of four bytes with one multiply.
template <int TYPE>
__device__ __forceinline__ float legacy_block_dot(const unsigned char *bk, const int4 u0, const int4 u1, float d8, float s8) {
    // IQ4_NL (ggml-common.h:406-410) has the layout of q4_0 -- d, 16 bytes of nibbles -- and no offset term: the nibbles go through
    // the codebook (iq4_levels) and the block is d d8 sum(level u)   (vec_dot_iq4_nl_q8_1, vecdotq.cuh:1086-1105)
    constexpr bool NL = TYPE == B200_TYPE_IQ4_NL;
    constexpr bool HAS_M = TYPE != B200_TYPE_Q5_0 && !NL, HAS_H = TYPE != B200_TYPE_Q4_1 && !NL;
    constexpr int NW = 4 + (HAS_M ? 1 : 0) + (HAS_H ? 1 : 0) + 1;  // d (+ m) (+ qh) + 16 bytes of nibbles, as 32-bit words
    unsigned w[NW];
    lds_words<NW>(bk, w);  // (reads up to 2 bytes past the block: the stage carries slack)
    const float d = h2f((unsigned short)(w[0] & 0xffff));
    // the block starts with d (2 bytes); everything behind it is 2 bytes off the word grid unless m fills the first word
    unsigned q[4], qh = 0;
    float m = 0.f;
    if constexpr (HAS_M) {
        m = h2f((unsigned short)(w[0] >> 16));
        if constexpr (HAS_H) qh = w[1];
#pragma unroll
        for (int i = 0; i < 4; i++) q[i] = w[(HAS_H ? 2 : 1) + i];
    } else if constexpr (NL) {  // d, then qs straddles the words
#pragma unroll
        for (int i = 0; i < 4; i++) q[i] = __funnelshift_r(w[i], w[1 + i], 16);
    } else {  // Q5_0: d, then qh and qs straddle the words
        qh = __funnelshift_r(w[0], w[1], 16);
#pragma unroll
        for (int i = 0; i < 4; i++) q[i] = __funnelshift_r(w[1 + i], w[2 + i], 16);
    }
    const int ua[4] = {u0.x, u0.y, u0.z, u0.w}, ub[4] = {u1.x, u1.y, u1.z, u1.w};
    int s = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        unsigned lo, hi;
        if constexpr (NL) {
            iq4_levels(q[i], lo, hi);
        } else {
            lo = q[i] & 0x0f0f0f0fu;
            hi = (q[i] >> 4) & 0x0f0f0f0fu;
        }
        if constexpr (HAS_H) {
            lo |= ((((qh >> (4 * i)) & 0xFu) * 0x00204081u) & 0x01010101u) << 4;
            hi |= ((((qh >> (16 + 4 * i)) & 0xFu) * 0x00204081u) & 0x01010101u) << 4;
        }
        s = dp4a((int)lo, ua[i], s);
        s = dp4a((int)hi, ub[i], s);
    }
    if constexpr (NL) return d * d8 * (float)s;
    else if constexpr (TYPE == B200_TYPE_Q5_0) return d * (d8 * (float)s - 16.0f * s8);
    else return d * d8 * (float)s + m * s8;
}
template <int TYPE, int NCOLS>
struct LegacySliceDot {
    static __device__ __forceinline__ void run(const unsigned char *seg, int lane, int it, unsigned char *act_base, int n_it,
                                               int nblk, float acc[NCOLS]) {
        constexpr int BB = type_block_bytes(TYPE);
        const unsigned char *b = seg + lane * (4 * BB);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (k < nblk) {
#pragma unroll
                for (int n = 0; n < NCOLS; n++) {
                    const ActView a = act_view(act_base, n_it, n);
                    const int4 u0 = a.q[((it * 8 + k * 2 + 0) * 32) + lane];
                    const int4 u1 = a.q[((it * 8 + k * 2 + 1) * 32) + lane];
                    const int o = ((it * 32 + lane) << 2) + k;
                    acc[n] += legacy_block_dot<TYPE>(b + k * BB, u0, u1, reinterpret_cast<const float *>(a.d)[o], reinterpret_cast<const float *>(a.s)[o]);
                }
            }
        }
    }
};
template <int NCOLS> struct SliceDot<B200_TYPE_Q4_1, NCOLS> : LegacySliceDot<B200_TYPE_Q4_1, NCOLS> {};
template <int NCOLS> struct SliceDot<B200_TYPE_Q5_0, NCOLS> : LegacySliceDot<B200_TYPE_Q5_0, NCOLS> {};
template <int NCOLS> struct SliceDot<B200_TYPE_Q5_1, NCOLS> : LegacySliceDot<B200_TYPE_Q5_1, NCOLS> {};
template <int NCOLS> struct SliceDot<B200_TYPE_IQ4_NL, NCOLS> : LegacySliceDot<B200_TYPE_IQ4_NL, NCOLS> {};

// ---- Q2_K / Q3_K (ggml-common.h:257-280): half a super-block (128 values) per lane.  Value v of the half sits in byte v % 32 of the
// half's 32 quant bytes at bits 2 (v / 32); its scale is number v / 16 of the half's eight.  The 2-bit fields (and the q3_K high
// bits out of hmask) are contracted IN PLACE with dp4a -- a field at bit position s only scales the sum by 2^s, removed by an exact
// shift afterwards -- against the q8_1 activations; per 16-wide group (vec_dot_q2_K_q8_1 / q3_K_q8_1, vecdotq.cuh:230-355):
//   Q2_K: d sc d8 sum(v u) - dmin m d8 sum(u)          Q3_K: d (sc - 32) d8 (sum(v2 u) + 4 sum(h u) - 4 sum(u))
template <int TYPE>
__device__ __forceinline__ float k23_half_dot(const unsigned char *b, int n, const int4 (&u)[8], const float4 d8, const int4 g) {
    constexpr bool Q2 = TYPE == B200_TYPE_Q2_K;
    const float d8v[4] = {d8.x, d8.y, d8.z, d8.w};
    if constexpr (TYPE == B200_TYPE_IQ4_XS) {
        // IQ4_XS (ggml-common.h:413-418): half d, u16 scales_h, u8 scales_l[4], then 8 sub-blocks of 16 nibble bytes.  Sub-block ib has
        // the 6-bit scale ls = nibble ib of scales_l | bits 2ib..2ib+1 of scales_h << 4; the half n holds sub-blocks 4n .. 4n+3:
        //   d sum_j (ls_j - 32) d8_j sum(level u)      (vec_dot_iq4_xs_q8_1, vecdotq.cuh:1110-1133)
        unsigned hw[2], qw[16];
        lds_words<2>(b, hw);
        lds_words<16>(b + 8 + 64 * n, qw);
        const float d = h2f((unsigned short)(hw[0] & 0xffff));
        const unsigned sl = hw[1] >> (16 * n), sh = hw[0] >> (16 + 8 * n);
        float sd = 0.f;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int ua[4] = {u[2 * j].x, u[2 * j].y, u[2 * j].z, u[2 * j].w}, ub[4] = {u[2 * j + 1].x, u[2 * j + 1].y, u[2 * j + 1].z, u[2 * j + 1].w};
            int sum = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                unsigned lo, hi;
                iq4_levels(qw[4 * j + i], lo, hi);
                sum = dp4a((int)lo, ua[i], sum);
                sum = dp4a((int)hi, ub[i], sum);
            }
            const int ls = (int)(((sl >> (4 * j)) & 0xFu) | (((sh >> (2 * j)) & 3u) << 4)) - 32;
            sd += d8v[j] * (float)(sum * ls);
        }
        return d * sd;
    }
    const int gv[4] = {g.x, g.y, g.z, g.w};
    unsigned qw[8];
    lds_words<8>(b + (Q2 ? 16 : 32) + 32 * n, qw);
    if constexpr (Q2) {
        unsigned scw[2], dm[1];
        lds_words<2>(b + 8 * n, scw);
        lds_words<1>(b + 80, dm);
        const float d = h2f((unsigned short)(dm[0] & 0xffff)), dmin = h2f((unsigned short)(dm[0] >> 16));
        float sd = 0.f, sm = 0.f;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const unsigned mq = 0x03030303u << (2 * j);
            const int ua[4] = {u[2 * j].x, u[2 * j].y, u[2 * j].z, u[2 * j].w}, ub[4] = {u[2 * j + 1].x, u[2 * j + 1].y, u[2 * j + 1].z, u[2 * j + 1].w};
            int d0 = 0, d1 = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                d0 = dp4a_u(qw[i] & mq, ua[i], d0);
                d1 = dp4a_u(qw[4 + i] & mq, ub[i], d1);
            }
            d0 >>= 2 * j;
            d1 >>= 2 * j;
            const unsigned s0 = (scw[j >> 1] >> (16 * (j & 1))) & 0xff, s1 = (scw[j >> 1] >> (16 * (j & 1) + 8)) & 0xff;
            const int g0 = (int)(short)(gv[j] & 0xffff), g1 = gv[j] >> 16;
            sd += d8v[j] * (float)((int)(s0 & 0xF) * d0 + (int)(s1 & 0xF) * d1);
            sm += d8v[j] * (float)((int)(s0 >> 4) * g0 + (int)(s1 >> 4) * g1);
        }
        return d * sd - dmin * sm;
    } else {
        unsigned hm[8], tail[4], aux[4];
        lds_words<8>(b, hm);
        lds_words<4>(b + 96, tail);  // 12 scale bytes, then d
        q3k_scales(tail[0], tail[1], tail[2], aux);
        const float d = h2f((unsigned short)(tail[3] & 0xffff));
        float sd = 0.f;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const unsigned mq = 0x03030303u << (2 * j), mh = 0x01010101u << (4 * n + j);
            const int ua[4] = {u[2 * j].x, u[2 * j].y, u[2 * j].z, u[2 * j].w}, ub[4] = {u[2 * j + 1].x, u[2 * j + 1].y, u[2 * j + 1].z, u[2 * j + 1].w};
            int q0 = 0, q1 = 0, h0 = 0, h1 = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                q0 = dp4a_u(qw[i] & mq, ua[i], q0);
                h0 = dp4a_u(hm[i] & mh, ua[i], h0);
                q1 = dp4a_u(qw[4 + i] & mq, ub[i], q1);
                h1 = dp4a_u(hm[4 + i] & mh, ub[i], h1);
            }
            const int g0 = (int)(short)(gv[j] & 0xffff), g1 = gv[j] >> 16;
            const int v0 = (q0 >> (2 * j)) + 4 * (h0 >> (4 * n + j)) - 4 * g0;
            const int v1 = (q1 >> (2 * j)) + 4 * (h1 >> (4 * n + j)) - 4 * g1;
            const unsigned aw = n ? (j < 2 ? aux[2] : aux[3]) : (j < 2 ? aux[0] : aux[1]);
            const int s0 = (int)((aw >> (16 * (j & 1))) & 0xff) - 32, s1 = (int)((aw >> (16 * (j & 1) + 8)) & 0xff) - 32;
            sd += d8v[j] * (float)(s0 * v0 + s1 * v1);
        }
        return d * sd;
    }
}
template <int TYPE, int NCOLS>
struct K23SliceDot {
    static __device__ __forceinline__ void run(const unsigned char *seg, int lane, int it, unsigned char *act_base, int n_it,
                                               int /*nblk*/, float acc[NCOLS]) {
        const unsigned char *b = seg + (lane >> 1) * type_block_bytes(TYPE);
#pragma unroll
        for (int n = 0; n < NCOLS; n++) {
            const ActView a = act_view(act_base, n_it, n);
            int4 u[8];
#pragma unroll
            for (int j = 0; j < 8; j++) u[j] = a.q[((it * 8 + j) * 32) + lane];
            acc[n] += k23_half_dot<TYPE>(b, lane & 1, u, a.d[it * 32 + lane], a.g[it * 32 + lane]);
        }
    }
};
template <int NCOLS> struct SliceDot<B200_TYPE_Q2_K, NCOLS> : K23SliceDot<B200_TYPE_Q2_K, NCOLS> {};
template <int NCOLS> struct SliceDot<B200_TYPE_Q3_K, NCOLS> : K23SliceDot<B200_TYPE_Q3_K, NCOLS> {};
template <int NCOLS> struct SliceDot<B200_TYPE_IQ4_XS, NCOLS> : K23SliceDot<B200_TYPE_IQ4_XS, NCOLS> {};

// ------------------------------------------------------------------------------------------------ row dots on preloaded activations
// One column (decode), several weight rows per step: the lane's 128 activations of step `it` are loaded into registers ONCE
// (LaneActs) and contracted with the same 128-weight slice of every row of the step.  Same arithmetic, in the same order, as
// SliceDot<TYPE, 1>: results are bit-identical.
struct LaneActs {
    int4   u[8];   // 8 x 16 int8
    float4 d8;     // q8_1 d of the 4 blocks
    float4 s8;     // q8_1 s of the 4 blocks      (types with an offset/min term)
    int4   g;      // 16-wide integer sums          (q6_K)
};
template <int TYPE>
__device__ __forceinline__ void lane_acts_load(LaneActs &A, unsigned char *act_base, int n_it, int it, int lane) {
    const ActView a = act_view(act_base, n_it, 0);
#pragma unroll
    for (int j = 0; j < 8; j++) A.u[j] = a.q[((it * 8 + j) * 32) + lane];
    A.d8 = a.d[it * 32 + lane];
    if constexpr (TYPE != B200_TYPE_Q6_K && TYPE != B200_TYPE_Q8_0) A.s8 = a.s[it * 32 + lane];
    if constexpr (TYPE == B200_TYPE_Q6_K || TYPE == B200_TYPE_Q2_K || TYPE == B200_TYPE_Q3_K) A.g = a.g[it * 32 + lane];
}

template <int TYPE>
struct RowDot;

template <>
struct RowDot<B200_TYPE_Q4_K> {
    static __device__ __forceinline__ void run(const unsigned char *seg, int lane, const LaneActs &A, int /*nblk*/, float &acc) {
        const unsigned char *b = seg + (lane >> 1) * Q4_K_BYTES;
        const int h = lane & 1;
        const uint4 hdr = *reinterpret_cast<const uint4 *>(b);
        const float d = h2f((unsigned short)(hdr.x & 0xffff)), dmin = h2f((unsigned short)(hdr.x >> 16));
        unsigned scw, mnw;
        if (h == 0) {
            scw = hdr.y & 0x3f3f3f3fu;
            mnw = hdr.z & 0x3f3f3f3fu;
        } else {
            scw = (hdr.w & 0x0f0f0f0fu) | ((hdr.y >> 2) & 0x30303030u);
            mnw = ((hdr.w >> 4) & 0x0f0f0f0fu) | ((hdr.z >> 2) & 0x30303030u);
        }
        const uint4 *qp = reinterpret_cast<const uint4 *>(b + 16 + 64 * h);
        unsigned v[16];
        *reinterpret_cast<uint4 *>(v + 0)  = qp[0];
        *reinterpret_cast<uint4 *>(v + 4)  = qp[1];
        *reinterpret_cast<uint4 *>(v + 8)  = qp[2];
        *reinterpret_cast<uint4 *>(v + 12) = qp[3];
        int sumi[4] = {0, 0, 0, 0};
#pragma unroll
        for (int cc = 0; cc < 2; cc++) {
#pragma unroll
            for (int hq = 0; hq < 2; hq++) {
                const int4 ulo = A.u[(2 * cc) * 2 + hq];
                const int4 uhi = A.u[(2 * cc + 1) * 2 + hq];
                const unsigned *w = v + 8 * cc + 4 * hq;
                sumi[2 * cc]     = dp4a((int)(w[0] & 0x0f0f0f0fu), ulo.x, sumi[2 * cc]);
                sumi[2 * cc]     = dp4a((int)(w[1] & 0x0f0f0f0fu), ulo.y, sumi[2 * cc]);
                sumi[2 * cc]     = dp4a((int)(w[2] & 0x0f0f0f0fu), ulo.z, sumi[2 * cc]);
                sumi[2 * cc]     = dp4a((int)(w[3] & 0x0f0f0f0fu), ulo.w, sumi[2 * cc]);
                sumi[2 * cc + 1] = dp4a_u(w[0] & 0xf0f0f0f0u, uhi.x, sumi[2 * cc + 1]);
                sumi[2 * cc + 1] = dp4a_u(w[1] & 0xf0f0f0f0u, uhi.y, sumi[2 * cc + 1]);
                sumi[2 * cc + 1] = dp4a_u(w[2] & 0xf0f0f0f0u, uhi.z, sumi[2 * cc + 1]);
                sumi[2 * cc + 1] = dp4a_u(w[3] & 0xf0f0f0f0u, uhi.w, sumi[2 * cc + 1]);
            }
        }
        sumi[1] >>= 4;
        sumi[3] >>= 4;
        const float4 d8 = A.d8, s8 = A.s8;
        const float sd = d8.x * (float)(sumi[0] * (int)(scw & 0xff)) + d8.y * (float)(sumi[1] * (int)((scw >> 8) & 0xff)) +
                         d8.z * (float)(sumi[2] * (int)((scw >> 16) & 0xff)) + d8.w * (float)(sumi[3] * (int)(scw >> 24));
        const float sm = s8.x * (float)(mnw & 0xff) + s8.y * (float)((mnw >> 8) & 0xff) + s8.z * (float)((mnw >> 16) & 0xff) +
                         s8.w * (float)(mnw >> 24);
        acc += d * sd - dmin * sm;
    }
};

template <>
struct RowDot<B200_TYPE_Q5_K> {
    static __device__ __forceinline__ void run(const unsigned char *seg, int lane, const LaneActs &A, int /*nblk*/, float &acc) {
        const unsigned char *b = seg + (lane >> 1) * Q5_K_BYTES;
        const int h = lane & 1;
        const uint4 hdr = *reinterpret_cast<const uint4 *>(b);
        const float d = h2f((unsigned short)(hdr.x & 0xffff)), dmin = h2f((unsigned short)(hdr.x >> 16));
        unsigned scw, mnw;
        if (h == 0) {
            scw = hdr.y & 0x3f3f3f3fu;
            mnw = hdr.z & 0x3f3f3f3fu;
        } else {
            scw = (hdr.w & 0x0f0f0f0fu) | ((hdr.y >> 2) & 0x30303030u);
            mnw = ((hdr.w >> 4) & 0x0f0f0f0fu) | ((hdr.z >> 2) & 0x30303030u);
        }
        unsigned qh[8], v[16];
        *reinterpret_cast<uint4 *>(qh + 0) = *reinterpret_cast<const uint4 *>(b + 16);
        *reinterpret_cast<uint4 *>(qh + 4) = *reinterpret_cast<const uint4 *>(b + 32);
        const uint4 *qp = reinterpret_cast<const uint4 *>(b + 48 + 64 * h);
        *reinterpret_cast<uint4 *>(v + 0)  = qp[0];
        *reinterpret_cast<uint4 *>(v + 4)  = qp[1];
        *reinterpret_cast<uint4 *>(v + 8)  = qp[2];
        *reinterpret_cast<uint4 *>(v + 12) = qp[3];
        int sumi[4] = {0, 0, 0, 0};
#pragma unroll
        for (int cc = 0; cc < 2; cc++) {
#pragma unroll
            for (int hq = 0; hq < 2; hq++) {
                const int4 ulo = A.u[(2 * cc) * 2 + hq];
                const int4 uhi = A.u[(2 * cc + 1) * 2 + hq];
                const int ul[4] = {ulo.x, ulo.y, ulo.z, ulo.w}, uh[4] = {uhi.x, uhi.y, uhi.z, uhi.w};
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const unsigned w = v[8 * cc + 4 * hq + i], hb = qh[4 * hq + i] >> (4 * h + 2 * cc);
                    sumi[2 * cc]     = dp4a((int)((w & 0x0f0f0f0fu) | ((hb << 4) & 0x10101010u)), ul[i], sumi[2 * cc]);
                    sumi[2 * cc + 1] = dp4a((int)(((w >> 4) & 0x0f0f0f0fu) | ((hb << 3) & 0x10101010u)), uh[i], sumi[2 * cc + 1]);
                }
            }
        }
        const float4 d8 = A.d8, s8 = A.s8;
        const float sd = d8.x * (float)(sumi[0] * (int)(scw & 0xff)) + d8.y * (float)(sumi[1] * (int)((scw >> 8) & 0xff)) +
                         d8.z * (float)(sumi[2] * (int)((scw >> 16) & 0xff)) + d8.w * (float)(sumi[3] * (int)(scw >> 24));
        const float sm = s8.x * (float)(mnw & 0xff) + s8.y * (float)((mnw >> 8) & 0xff) + s8.z * (float)((mnw >> 16) & 0xff) +
                         s8.w * (float)(mnw >> 24);
        acc += d * sd - dmin * sm;
    }
};

template <>
struct RowDot<B200_TYPE_Q6_K> {
    static __device__ __forceinline__ void run(const unsigned char *seg, int lane, const LaneActs &A, int /*nblk*/, float &acc) {
        const unsigned char *b = seg + (lane >> 1) * Q6_K_BYTES;
        const int h = lane & 1;
        unsigned ql[16], qh[8], scw[2];
        lds_words<16>(b + 64 * h, ql);
        lds_words<8>(b + 128 + 32 * h, qh);
        lds_words<2>(b + 192 + 8 * h, scw);
        const float d = h2f(*reinterpret_cast<const unsigned short *>(b + 208));
        const signed char *sc = reinterpret_cast<const signed char *>(scw);
        const float d8v[4] = {A.d8.x, A.d8.y, A.d8.z, A.d8.w};
        const int gv[4] = {A.g.x, A.g.y, A.g.z, A.g.w};
        float sd = 0.f;
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const int4 u0 = A.u[r * 2 + 0], u1 = A.u[r * 2 + 1];
            const int ua[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
            int s0 = 0, s1 = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const unsigned l = r & 1 ? ql[8 + i] : ql[i], hb = qh[i];
                unsigned q;  // row-group r (= q8_1 block r of this slice), word i (ggml-quants.c:2990-2999)
                if (r == 0)      q = (l & 0x0f0f0f0fu) | ((hb << 4) & 0x30303030u);
                else if (r == 1) q = (l & 0x0f0f0f0fu) | ((hb << 2) & 0x30303030u);
                else if (r == 2) q = ((l >> 4) & 0x0f0f0f0fu) | (hb & 0x30303030u);
                else             q = ((l >> 4) & 0x0f0f0f0fu) | ((hb >> 2) & 0x30303030u);
                if (i < 4) s0 = dp4a((int)q, ua[i], s0);
                else       s1 = dp4a((int)q, ua[i], s1);
            }
            s0 -= 32 * (int)(short)(gv[r] & 0xffff);
            s1 -= 32 * (gv[r] >> 16);
            sd += d8v[r] * (float)(s0 * (int)sc[2 * r] + s1 * (int)sc[2 * r + 1]);
        }
        acc += d * sd;
    }
};

template <>
struct RowDot<B200_TYPE_Q8_0> {
    static __device__ __forceinline__ void run(const unsigned char *seg, int lane, const LaneActs &A, int nblk, float &acc) {
        const unsigned char *b = seg + lane * (4 * Q8_0_BYTES);
        const float d8v[4] = {A.d8.x, A.d8.y, A.d8.z, A.d8.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (k < nblk) {
                const unsigned char *bk = b + k * Q8_0_BYTES;
                const float d = h2f(*reinterpret_cast<const unsigned short *>(bk));
                unsigned w[8];
                lds_words<8>(bk + 2, w);
                const int4 u0 = A.u[k * 2 + 0], u1 = A.u[k * 2 + 1];
                int s = dp4a((int)w[0], u0.x, 0);
                s = dp4a((int)w[1], u0.y, s);
                s = dp4a((int)w[2], u0.z, s);
                s = dp4a((int)w[3], u0.w, s);
                s = dp4a((int)w[4], u1.x, s);
                s = dp4a((int)w[5], u1.y, s);
                s = dp4a((int)w[6], u1.z, s);
                s = dp4a((int)w[7], u1.w, s);
                acc += d * d8v[k] * (float)s;
            }
        }
    }
};

template <>
struct RowDot<B200_TYPE_Q4_0> {
    static __device__ __forceinline__ void run(const unsigned char *seg, int lane, const LaneActs &A, int nblk, float &acc) {
        const unsigned char *b = seg + lane * (4 * Q4_0_BYTES);
        const float d8v[4] = {A.d8.x, A.d8.y, A.d8.z, A.d8.w}, s8v[4] = {A.s8.x, A.s8.y, A.s8.z, A.s8.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (k < nblk) {
                const unsigned char *bk = b + k * Q4_0_BYTES;
                const float d = h2f(*reinterpret_cast<const unsigned short *>(bk));
                unsigned w[4];
                lds_words<4>(bk + 2, w);
                const int4 u0 = A.u[k * 2 + 0], u1 = A.u[k * 2 + 1];
                int s = dp4a((int)(w[0] & 0x0f0f0f0fu), u0.x, 0);
                s = dp4a((int)(w[1] & 0x0f0f0f0fu), u0.y, s);
                s = dp4a((int)(w[2] & 0x0f0f0f0fu), u0.z, s);
                s = dp4a((int)(w[3] & 0x0f0f0f0fu), u0.w, s);
                int sh = dp4a_u(w[0] & 0xf0f0f0f0u, u1.x, 0);
                sh = dp4a_u(w[1] & 0xf0f0f0f0u, u1.y, sh);
                sh = dp4a_u(w[2] & 0xf0f0f0f0u, u1.z, sh);
                sh = dp4a_u(w[3] & 0xf0f0f0f0u, u1.w, sh);
                s += sh >> 4;
                acc += d * (d8v[k] * (float)s - 8.0f * s8v[k]);
            }
        }
    }
};


template <int TYPE>
struct LegacyRowDot {
    static __device__ __forceinline__ void run(const unsigned char *seg, int lane, const LaneActs &A, int nblk, float &acc) {
        constexpr int BB = type_block_bytes(TYPE);
        const unsigned char *b = seg + lane * (4 * BB);
        const float d8v[4] = {A.d8.x, A.d8.y, A.d8.z, A.d8.w}, s8v[4] = {A.s8.x, A.s8.y, A.s8.z, A.s8.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (k < nblk) acc += legacy_block_dot<TYPE>(b + k * BB, A.u[k * 2 + 0], A.u[k * 2 + 1], d8v[k], s8v[k]);
        }
    }
};
template <> struct RowDot<B200_TYPE_Q4_1> : LegacyRowDot<B200_TYPE_Q4_1> {};
template <> struct RowDot<B200_TYPE_Q5_0> : LegacyRowDot<B200_TYPE_Q5_0> {};
template <> struct RowDot<B200_TYPE_Q5_1> : LegacyRowDot<B200_TYPE_Q5_1> {};
template <> struct RowDot<B200_TYPE_IQ4_NL> : LegacyRowDot<B200_TYPE_IQ4_NL> {};


template <int TYPE>
struct K23RowDot {
    static __device__ __forceinline__ void run(const unsigned char *seg, int lane, const LaneActs &A, int /*nblk*/, float &acc) {
        acc += k23_half_dot<TYPE>(seg + (lane >> 1) * type_block_bytes(TYPE), lane & 1, A.u, A.d8, A.g);
    }
};
template <> struct RowDot<B200_TYPE_Q2_K> : K23RowDot<B200_TYPE_Q2_K> {};
template <> struct RowDot<B200_TYPE_Q3_K> : K23RowDot<B200_TYPE_Q3_K> {};
template <> struct RowDot<B200_TYPE_IQ4_XS> : K23RowDot<B200_TYPE_IQ4_XS> {};

}  // namespace b200
