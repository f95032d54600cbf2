// mmvq.cu -- (a6, a9) decode GEMV over GGUF block formats: dst[n][m] = W[m][:] . X[n][:], N <= 8,
// plus the expert-indexed (MUL_MAT_ID) and multi-matrix (fused Q/K/V, gate/up) forms.
//
// Replaces the reference's quantize_q8_1 + mul_mat_vec_q<type,ncols_y> pair
// (/root/reference/ggml/src/ggml-cuda/quantize.cu:4-38, mmvq.cu:50-202, vecdotq.cuh:23-787) and, for MUL_MAT_ID,
// the host round trip in ggml_cuda_mul_mat_id (ggml-cuda.cu:2003-2139).
//
// B200 design (HBM-bound: the job is to keep >= ~40 KB of weight bytes in flight per SM and never touch a byte twice):
//   * one persistent CTA per SM; CTA c owns a contiguous, byte-balanced range of the (concatenated) weight rows;
//   * every warp streams whole rows through its OWN ring of shared-memory stages filled by the TMA engine with
//     1-D bulk copies (cp.async.bulk ... mbarrier::complete_tx, SASS UBLKCP): the raw 144/176/210/34/18-byte blocks
//     land in shared memory untouched, so no lane ever issues a misaligned or sub-32-bit global load, and no
//     registers or LSU slots are spent on in-flight data.  Stage reuse is warp-private: no CTA barrier in the loop;
//   * the f32 activation row(s) are quantised to q8_1 by the CTA itself while the first weight stages are in flight
//     (same arithmetic as quantize_row_q8_1_ref) and kept in shared memory in a lane-major layout so every
//     activation read is a conflict-free LDS.128;
//   * a lane handles one 128-weight slice per step: K-quant 6-bit scales/mins are unpacked in registers, nibbles are
//     split with two LOP3s per word and contracted with dp4a against the int8 activations; the q4_K/q5_K "min"
//     term and the q4_0 "-8" offset use the q8_1 block sums, the q6_K "-32" offset uses 16-wide integer sums;
//   * weight streaming starts BEFORE griddepcontrol.wait, so under programmatic dependent launch the ramp-up of
//     this kernel overlaps the tail of its producer.
#include "b200_common.cuh"
#include "rope_kv.cuh"

#include <cstdio>

namespace b200 {

constexpr int MMVQ_NW        = 16;                 // warps per CTA
constexpr int MMVQ_THREADS   = MMVQ_NW * 32;
constexpr int MMVQ_MAX_STAGE = 8;
constexpr int MMVQ_BAR_BYTES = 128 * ((MMVQ_NW * MMVQ_MAX_STAGE * 8 + 127) / 128);

struct MmvqParams {
    b200_gemv_job  jobs[B200_GEMV_MAX_JOBS];
    int            n_jobs;
    int            n_stage;
    int            n_it;          // ceil(K / 4096): steps per row
    int            n_expert;
    int            rows_q, rows_rem;   // rows per CTA = rows_q (+1 for the first rows_rem CTAs)
    int            tab_rows;           // MULTI: capacity of the per-CTA row table
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
    const float   *residual[B200_GEMV_MAX_JOBS];
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

// Quantise X[n][0..K) (f32, global) into the CTA's shared activation store.  All threads of the CTA call this.
template <int NCOLS>
__device__ void load_activations(unsigned char *act_base, const float *__restrict__ X, int64_t x_stride, int64_t K, int n_it,
                                 const float *__restrict__ norm_w, float norm_eps, float *red) {
    const int n4 = (int)(K / 4);
    for (int n = 0; n < NCOLS; n++) {
        float nscale = 1.0f;
        if (norm_w != nullptr) {  // RMS_NORM folded into the quantiser: scale = rsqrt(mean(x^2) + eps)  (ggml.c:12060-12110)
            const float *xs = X + (int64_t)n * x_stride;
            float ss = 0.f;
            for (int i4 = threadIdx.x; i4 < n4; i4 += MMVQ_THREADS) {
                const float4 v = __ldcg(reinterpret_cast<const float4 *>(xs) + i4);
                ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
            }
            ss = warp_sum(ss);
            __syncthreads();  // red[] may still be read by the previous column
            if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
            __syncthreads();
            float t = (threadIdx.x & 31) < MMVQ_NW ? red[threadIdx.x & 31] : 0.f;
            t = warp_sum(t);
            nscale = rsqrtf(t / (float)K + norm_eps);
        }
        unsigned char *p = act_base + (size_t)n * act_col_bytes(n_it);
        int   *aq = (int *)p;
        float *ad = (float *)(p + (size_t)n_it * 4096);
        float *as = (float *)(p + (size_t)n_it * (4096 + 512));
        int   *ag = (int *)(p + (size_t)n_it * (4096 + 1024));
        const float *x = X + (int64_t)n * x_stride;
        // 4 independent 16-byte loads per thread are issued before any of them is consumed (one L2 round trip per
        // 8192 activations instead of one per 2048)
        for (int base = 0; base < n4; base += 4 * MMVQ_THREADS) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int i4 = base + u * MMVQ_THREADS + threadIdx.x;
                v[u] = i4 < n4 ? __ldcg(reinterpret_cast<const float4 *>(x) + i4) : make_float4(0.f, 0.f, 0.f, 0.f);
                if (norm_w != nullptr && i4 < n4) {  // (x * scale) * w, the order of the RMS_NORM -> MUL node pair
                    const float4 ww = __ldg(reinterpret_cast<const float4 *>(norm_w) + i4);
                    v[u].x = v[u].x * nscale * ww.x; v[u].y = v[u].y * nscale * ww.y;
                    v[u].z = v[u].z * nscale * ww.z; v[u].w = v[u].w * nscale * ww.w;
                }
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int i4 = base + u * MMVQ_THREADS + threadIdx.x;
                if (base + u * MMVQ_THREADS >= n4) break;  // warp-uniform: the whole batch is past the end
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

// ------------------------------------------------------------------------------------------------ the kernel
// CTA c owns the contiguous row range [r0, r1) of the concatenated jobs; warp w owns local rows w, w+NW, ...
// A row is consumed in n_it steps of 4096 weights; step u of a warp lives in ring stage u % n_stage.
//
// MULTI (several jobs, or expert ids): a per-CTA row table in shared memory (source pointer, destination pointer,
// job) replaces index arithmetic, so the streaming loop is identical and prefetch runs across job boundaries.
constexpr int MMVQ_MAX_TAB = 1024;  // rows per CTA in MULTI mode (17 bytes each)

#ifdef B200_MMVQ_TRACE
__device__ unsigned long long g_mmvq_trace[148 * 8];
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define TRACE(i) do { if (threadIdx.x == 0 && blockIdx.x < 148) g_mmvq_trace[blockIdx.x * 8 + (i)] = gtime(); } while (0)
#else
#define TRACE(i) do {} while (0)
#endif

template <int TYPE, int NCOLS, bool MULTI>
__global__ void __launch_bounds__(MMVQ_THREADS, 1) mmvq_kernel(const __grid_constant__ MmvqParams p) {
    using G = Geo<TYPE>;
    TRACE(0);
    constexpr bool ALIGNED16 = (TYPE == B200_TYPE_Q4_K || TYPE == B200_TYPE_Q5_K);  // host guarantees 16-byte rows
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ float red[MMVQ_NW];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_stage = p.n_stage, n_it = p.n_it;
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(smem);  // [NW][MAX_STAGE]
    unsigned char *act_base  = smem + MMVQ_BAR_BYTES;
    unsigned char *ring      = act_base + ((size_t)NCOLS * act_col_bytes(n_it) + 127) / 128 * 128;
    unsigned char *my_ring   = ring + (size_t)warp * n_stage * G::STAGE_BYTES;
    unsigned long long *my_bars = bars + warp * MMVQ_MAX_STAGE;
    // MULTI row table lives behind the ring
    unsigned char *tab_base = ring + (size_t)MMVQ_NW * n_stage * G::STAGE_BYTES;
    const unsigned char **rsrc = reinterpret_cast<const unsigned char **>(tab_base);
    float **rdst = reinterpret_cast<float **>(tab_base + 8 * (size_t)p.tab_rows);
    signed char *rjob = reinterpret_cast<signed char *>(tab_base + 16 * (size_t)p.tab_rows);

    // contiguous row range, balanced to one row: the first rows_rem CTAs take rows_q + 1 rows
    const int bid = blockIdx.x;
    const int r0 = bid * p.rows_q + (bid < p.rows_rem ? bid : p.rows_rem);
    const int nrows = p.rows_q + (bid < p.rows_rem ? 1 : 0);
    const int nmy = warp < nrows ? (nrows - warp + MMVQ_NW - 1) / MMVQ_NW : 0;  // rows of this warp

    if (lane == 0) {
        for (int s = 0; s < n_stage; s++) mbar_init(&my_bars[s], 1);
        mbar_fence_init();
    }
    int jfirst = 0, jlast = 0;
    if constexpr (MULTI) {
        if (p.ids != nullptr) pdl_wait();  // expert ids are written by the producer kernel
        // row table: every thread resolves a few rows (<= 16 jobs: linear scan)
        for (int lr = threadIdx.x; lr < nrows; lr += MMVQ_THREADS) {
            int64_t row = r0 + lr;
            int j = 0;
            while (row >= p.jobs[j].M) { row -= p.jobs[j].M; j++; }
            const unsigned char *w = (const unsigned char *)p.jobs[j].W;
            if (p.ids != nullptr && p.jobs[j].id_index >= 0) {
                const int e = p.ids[p.jobs[j].id_index];
                w = (e >= 0 && e < p.n_expert) ? w + (size_t)e * p.expert_bytes : nullptr;
            }
            rsrc[lr] = w ? w + (size_t)row * p.w_row_bytes : nullptr;
            rdst[lr] = p.jobs[j].dst + row;
            rjob[lr] = (signed char)j;
        }
        __syncthreads();
        if (nrows > 0) { jfirst = rjob[0]; jlast = rjob[nrows - 1]; }
    }
    pdl_trigger();  // dependents may start their own (independent) prologue as our CTAs retire

    // ---- producer side of the warp-private ring (lane 0 issues)
    const unsigned char *w0 = (const unsigned char *)p.jobs[0].W;
    const unsigned tail_bytes = p.row_bytes - (unsigned)(n_it - 1) * G::SEG_BYTES;
    int ik = 0, iit = 0, istage = 0;  // next unit to issue: row counter, step, stage
    auto issue = [&]() {
        if (lane == 0) {
            const int lr = warp + ik * MMVQ_NW;
            const unsigned char *src;
            if constexpr (MULTI) src = rsrc[lr];
            else src = w0 + (size_t)(r0 + lr) * p.w_row_bytes;
            const unsigned bytes = iit == n_it - 1 ? tail_bytes : (unsigned)G::SEG_BYTES;
            if (!MULTI || src != nullptr) {
                src += (size_t)iit * G::SEG_BYTES;
                // aligned-down source, length rounded up to 16: the slack lands in the stage's head/tail padding
                const unsigned mis = ALIGNED16 ? 0u : (unsigned)((uintptr_t)src & 15);
                const unsigned len = (mis + bytes + 15u) & ~15u;
                mbar_expect_tx(&my_bars[istage], len);
                bulk_g2s(my_ring + (size_t)istage * G::STAGE_BYTES, src - mis, len, &my_bars[istage]);
            } else {
                // invalid expert id: nothing to copy; complete the phase so the consumer does not hang
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&my_bars[istage])) : "memory");
            }
        }
        istage = istage + 1 == n_stage ? 0 : istage + 1;
        if (++iit == n_it) { iit = 0; ik++; }
    };
    TRACE(1);
    for (int s = 0; s < n_stage && ik < nmy; s++) issue();
    TRACE(2);

    // ---- everything above is independent of the producer kernel's output
    if (!MULTI || p.ids == nullptr) pdl_wait();
    TRACE(3);

    // lanes past the end of the row in the last step (and, for q4_0/q8_0, a partial last slice)
    int last_nblk = (int)(p.K / 32) - 4 * ((n_it - 1) * 32 + lane);
    last_nblk = last_nblk < 0 ? 0 : (last_nblk > 4 ? 4 : last_nblk);

    int cstage = 0;
    unsigned cphase = 0;
    const float *loaded_x = nullptr;
    float acc[NCOLS];
#pragma unroll
    for (int n = 0; n < NCOLS; n++) acc[n] = 0.f;

    int k = 0;
    for (int job = jfirst; job <= jlast; job++) {
        const float *xj = p.jobs[job].X;
        if (loaded_x != xj && nrows > 0) {
            if (loaded_x != nullptr) __syncthreads();  // everyone is done with the previous activations
            load_activations<NCOLS>(act_base, xj, p.x_stride, p.K, n_it, p.norm_w, p.norm_eps, red);
            __syncthreads();
            loaded_x = xj;
            TRACE(4);
        }
        for (; k < nmy; k++) {
            const int lr = warp + k * MMVQ_NW;
            uintptr_t src;
            float *dst;
            const float *res;
            if constexpr (MULTI) {
                if (rjob[lr] != job) break;
                src = (uintptr_t)rsrc[lr];
                dst = rdst[lr];
                res = p.residual[job] ? p.residual[job] + (dst - p.jobs[job].dst) : nullptr;
            } else {
                src = (uintptr_t)w0 + (size_t)(r0 + lr) * p.w_row_bytes;
                dst = p.jobs[0].dst + (r0 + lr);
                res = p.residual[0] ? p.residual[0] + (r0 + lr) : nullptr;
            }
            for (int it = 0; it < n_it; it++) {
                mbar_wait(&my_bars[cstage], cphase);
                if (k == 0 && it == 0) TRACE(5);
                const int nblk = it == n_it - 1 ? last_nblk : 4;
                if ((!MULTI || src != 0) && nblk > 0) {
                    const unsigned mis = ALIGNED16 ? 0u : (unsigned)((src + (size_t)it * G::SEG_BYTES) & 15);
                    SliceDot<TYPE, NCOLS>::run(my_ring + (size_t)cstage * G::STAGE_BYTES + mis, lane, it, act_base, n_it, nblk, acc);
                }
                __syncwarp();
                if (lane == 0) fence_proxy_async_smem();  // our generic reads of the stage precede the next bulk write to it
                cstage = cstage + 1 == n_stage ? 0 : cstage + 1;
                cphase ^= (cstage == 0);
                if (ik < nmy) issue();
            }
#pragma unroll
            for (int n = 0; n < NCOLS; n++) {
                float v = warp_sum(acc[n]);
                if (MULTI && src == 0) v = __int_as_float(0x7fc00000);
                if (lane == 0) dst[(int64_t)n * p.dst_stride] = res ? v + res[(int64_t)n * p.dst_stride] : v;
                acc[n] = 0.f;
            }
            if (k == 0) TRACE(6);
        }
    }
    __syncthreads();
    TRACE(7);
}

// ------------------------------------------------------------------------------------------------ host side
static int g_sm_count = 0;
static size_t g_smem_optin = 0;

static int device_props() {
    if (g_sm_count) return B200_OK;
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    int sm = 0, smem = 0;
    B200_CUDA_TRY(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
    B200_CUDA_TRY(cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    g_sm_count = sm;
    g_smem_optin = (size_t)smem;
    return B200_OK;
}

template <int TYPE, int NCOLS, bool MULTI>
static int launch_mmvq(MmvqParams &p, cudaStream_t st, bool pdl) {
    using G = Geo<TYPE>;
    int rc = device_props();
    if (rc) return rc;
    const size_t bar_bytes = MMVQ_BAR_BYTES;
    const size_t act_bytes = ((size_t)NCOLS * act_col_bytes(p.n_it) + 127) / 128 * 128;
    const int64_t grid = p.total_rows < g_sm_count ? p.total_rows : g_sm_count;
    const int64_t rows_per_cta = (p.total_rows + grid - 1) / grid;
    if (MULTI && rows_per_cta > MMVQ_MAX_TAB) {
        set_error("b200_mul_mat_vec_q_multi: %lld rows exceed the per-CTA row table", (long long)p.total_rows);
        return B200_ERR_UNSUPPORTED;
    }
    p.tab_rows = MULTI ? (int)((rows_per_cta + 63) / 64 * 64) : 0;
    const size_t tab_bytes = MULTI ? (size_t)17 * p.tab_rows + 64 : 0;
    // (measured: 8 warps + half-SM rings so that consecutive launches overlap under PDL lost 15 %: a 93 KB ring cannot cover
    // the HBM latency-bandwidth product of one SM on the large matrices)
    const size_t min_need = bar_bytes + act_bytes + 2 * (size_t)MMVQ_NW * G::STAGE_BYTES + tab_bytes + 256;
    const size_t cap = g_smem_optin;
    if (min_need > cap) {
        set_error("b200_mul_mat_vec_q: K=%lld x N=%d activations do not fit in shared memory", (long long)p.K, NCOLS);
        return B200_ERR_UNSUPPORTED;
    }
    const size_t budget = cap - 256 - tab_bytes;
    size_t ns = (budget - bar_bytes - act_bytes) / ((size_t)MMVQ_NW * G::STAGE_BYTES);
    // no point in more stages than a warp has steps: rows per warp ~ total_rows / (SMs * NW)
    const int64_t steps = ((p.total_rows + g_sm_count - 1) / g_sm_count + MMVQ_NW - 1) / MMVQ_NW * p.n_it;
    if ((int64_t)ns > steps) ns = (size_t)(steps < 2 ? 2 : steps);
    if (ns > (size_t)MMVQ_MAX_STAGE) ns = MMVQ_MAX_STAGE;
    p.n_stage = (int)ns;
    const size_t smem = bar_bytes + act_bytes + ns * MMVQ_NW * G::STAGE_BYTES + tab_bytes;
    static bool attr_set_dev[16] = {};  // cudaFuncSetAttribute is per device
    int cur_dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&cur_dev));
    bool &attr_set = attr_set_dev[cur_dev & 15];
    if (!attr_set) {
        B200_CUDA_TRY(cudaFuncSetAttribute(mmvq_kernel<TYPE, NCOLS, MULTI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)(g_smem_optin - 256)));
        attr_set = true;
    }
    p.rows_q = (int)(p.total_rows / grid);
    p.rows_rem = (int)(p.total_rows % grid);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(MMVQ_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, mmvq_kernel<TYPE, NCOLS, MULTI>, p));
    return B200_OK;
}

template <int TYPE>
static int dispatch_cols(MmvqParams &p, int N, cudaStream_t st, bool pdl) {
    const bool multi = p.n_jobs > 1 || p.ids != nullptr;
    if (multi) {  // fused / expert launches are decode-only: 1..4 columns
        switch (N) {
            case 1: return launch_mmvq<TYPE, 1, true>(p, st, pdl);
            case 2: return launch_mmvq<TYPE, 2, true>(p, st, pdl);
            case 3: return launch_mmvq<TYPE, 3, true>(p, st, pdl);
            case 4: return launch_mmvq<TYPE, 4, true>(p, st, pdl);
        }
        set_error("b200_mul_mat_vec_q_multi: N=%d out of range 1..4", N);
        return B200_ERR_UNSUPPORTED;
    }
    switch (N) {
        case 1: return launch_mmvq<TYPE, 1, false>(p, st, pdl);
        case 2: return launch_mmvq<TYPE, 2, false>(p, st, pdl);
        case 3: return launch_mmvq<TYPE, 3, false>(p, st, pdl);
        case 4: return launch_mmvq<TYPE, 4, false>(p, st, pdl);
        case 5: return launch_mmvq<TYPE, 5, false>(p, st, pdl);
        case 6: return launch_mmvq<TYPE, 6, false>(p, st, pdl);
        case 7: return launch_mmvq<TYPE, 7, false>(p, st, pdl);
        case 8: return launch_mmvq<TYPE, 8, false>(p, st, pdl);
    }
    set_error("b200_mul_mat_vec_q: N=%d out of range 1..8", N);
    return B200_ERR_INVALID;
}

bool g_mmvq_pdl = false;  // set by the backend when it launches chains of dependent kernels

int mmvq_launch(int type, const b200_gemv_job *jobs, int n_jobs, size_t w_row_bytes, int64_t x_stride, int64_t dst_stride,
                int64_t K, int64_t N, const int32_t *ids, size_t expert_bytes, int32_t n_expert, cudaStream_t st,
                const b200_gemv_fuse *fuse = nullptr) {
    if (!jobs || n_jobs < 1 || n_jobs > B200_GEMV_MAX_JOBS) { set_error("b200_mul_mat_vec_q: n_jobs=%d out of range", n_jobs); return B200_ERR_INVALID; }
    const int bv = type_block_values(type);
    if (!(type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0 || type == B200_TYPE_Q4_K || type == B200_TYPE_Q5_K || type == B200_TYPE_Q6_K)) {
        set_error("b200_mul_mat_vec_q: unsupported weight type %d", type);
        return B200_ERR_UNSUPPORTED;
    }
    if (K <= 0 || K % bv) { set_error("b200_mul_mat_vec_q: K=%lld must be a positive multiple of %d", (long long)K, bv); return B200_ERR_INVALID; }
    if (N < 1 || N > 8) { set_error("b200_mul_mat_vec_q: N=%lld out of range 1..8", (long long)N); return B200_ERR_INVALID; }
    const size_t row_bytes = (size_t)(K / bv) * type_block_bytes(type);
    if (w_row_bytes < row_bytes) { set_error("b200_mul_mat_vec_q: row stride %zu < row size %zu", w_row_bytes, row_bytes); return B200_ERR_INVALID; }
    MmvqParams p = {};
    p.n_jobs = n_jobs;
    p.K = K;
    p.n_it = (int)((K + 4095) / 4096);
    p.w_row_bytes = w_row_bytes;
    p.row_bytes = (unsigned)row_bytes;
    p.x_stride = x_stride;
    p.dst_stride = dst_stride;
    p.ids = ids;
    p.expert_bytes = expert_bytes;
    p.n_expert = n_expert;
    if (fuse) {
        p.norm_w = fuse->norm_weight;
        p.norm_eps = fuse->norm_eps;
        if (fuse->norm_weight && ((uintptr_t)fuse->norm_weight & 15)) { set_error("b200_mul_mat_vec_q_fused: norm weight must be 16-byte aligned"); return B200_ERR_UNSUPPORTED; }
        for (int j = 0; j < n_jobs; j++) p.residual[j] = fuse->residual[j];
    }
    int64_t total = 0;
    const bool need16 = (type == B200_TYPE_Q4_K || type == B200_TYPE_Q5_K);
    for (int j = 0; j < n_jobs; j++) {
        p.jobs[j] = jobs[j];
        if (!jobs[j].W || !jobs[j].X || !jobs[j].dst || jobs[j].M < 0) { set_error("b200_mul_mat_vec_q: job %d has a NULL pointer", j); return B200_ERR_INVALID; }
        if (((uintptr_t)jobs[j].W & (need16 ? 15 : 1)) || (need16 && ((w_row_bytes & 15) || (expert_bytes & 15)))) {
            set_error("b200_mul_mat_vec_q: weight pointer/stride alignment (need %d bytes)", need16 ? 16 : 2);
            return B200_ERR_UNSUPPORTED;
        }
        if (((uintptr_t)jobs[j].X & 15) || (x_stride & 3)) { set_error("b200_mul_mat_vec_q: X must be 16-byte aligned"); return B200_ERR_UNSUPPORTED; }
        total += jobs[j].M;
    }
    if ((w_row_bytes & 1) || (expert_bytes & 1)) { set_error("b200_mul_mat_vec_q: odd strides"); return B200_ERR_UNSUPPORTED; }
    p.total_rows = total;
    if (total == 0) return B200_OK;
    if (total > 0x7fffffff) { set_error("b200_mul_mat_vec_q: too many rows"); return B200_ERR_UNSUPPORTED; }
    const bool pdl = g_mmvq_pdl;
    switch (type) {
        case B200_TYPE_Q4_0: return dispatch_cols<B200_TYPE_Q4_0>(p, (int)N, st, pdl);
        case B200_TYPE_Q8_0: return dispatch_cols<B200_TYPE_Q8_0>(p, (int)N, st, pdl);
        case B200_TYPE_Q4_K: return dispatch_cols<B200_TYPE_Q4_K>(p, (int)N, st, pdl);
        case B200_TYPE_Q5_K: return dispatch_cols<B200_TYPE_Q5_K>(p, (int)N, st, pdl);
        case B200_TYPE_Q6_K: return dispatch_cols<B200_TYPE_Q6_K>(p, (int)N, st, pdl);
    }
    return B200_ERR_UNSUPPORTED;
}


// ================================================================================================ decode chain
// One persistent CTA per SM runs a whole chain of dependent decode phases (b200_ggml.h: b200_decode_chain).  The streaming
// loop is the one above; what changes is that the producer cursor of every warp walks across phase boundaries, so the TMA
// ring is already full of the NEXT phase's weights while the grid barrier and the activation reload of that phase happen.
struct ChainProgram {
    b200_chain_phase ph[B200_CHAIN_MAX_PHASES];
    int n_phases, n_stage, stage_bytes, act_bytes;
    unsigned *barrier;
};

__device__ __forceinline__ int chain_seg_bytes(int type) {
    return type == B200_TYPE_Q4_K ? Geo<B200_TYPE_Q4_K>::SEG_BYTES : type == B200_TYPE_Q5_K ? Geo<B200_TYPE_Q5_K>::SEG_BYTES
         : type == B200_TYPE_Q6_K ? Geo<B200_TYPE_Q6_K>::SEG_BYTES : type == B200_TYPE_Q8_0 ? Geo<B200_TYPE_Q8_0>::SEG_BYTES : Geo<B200_TYPE_Q4_0>::SEG_BYTES;
}
struct ChainUnit {
    const unsigned char *src;
    float *dst;
    const float *res;
    int type;
};
__device__ __forceinline__ void chain_unit(const b200_chain_phase &ph, int row, ChainUnit &u) {
    int j = 0, m = row;
    if (ph.silu_pair) { j = row & 1; m = row >> 1; }
    else { while (j < ph.n_jobs - 1 && m >= ph.u.jobs[j].M) { m -= ph.u.jobs[j].M; j++; } }
    const b200_chain_job &jb = ph.u.jobs[j];
    u.src = (const unsigned char *)jb.W + (size_t)m * jb.w_row_bytes;
    u.dst = jb.dst + m;
    u.res = jb.residual ? jb.residual + m : nullptr;
    u.type = jb.type;
}
// rows (units) of CTA `bid` in a GEMV phase, and the units of warp `warp` among them
struct ChainGeom { int r0, nrows, nmy, n_it, pair; };
__device__ __forceinline__ ChainGeom chain_geom(const b200_chain_phase &ph, int bid, int nblk, int warp) {
    ChainGeom g;
    g.pair = ph.silu_pair != 0;
    g.n_it = (ph.K + 4095) / 4096;
    int total = 0;
    if (g.pair) total = ph.u.jobs[0].M;  // pairs
    else for (int j = 0; j < ph.n_jobs; j++) total += ph.u.jobs[j].M;
    const int q = total / nblk, rem = total % nblk;
    const int start = bid * q + (bid < rem ? bid : rem), cnt = q + (bid < rem ? 1 : 0);
    const int mine = warp < cnt ? (cnt - warp + MMVQ_NW - 1) / MMVQ_NW : 0;
    if (g.pair) { g.r0 = 2 * start; g.nrows = 2 * cnt; g.nmy = 2 * mine; }
    else { g.r0 = start; g.nrows = cnt; g.nmy = mine; }
    return g;
}
__device__ __forceinline__ int chain_lr(const ChainGeom &g, int warp, int k) {
    return g.pair ? 2 * (warp + (k >> 1) * MMVQ_NW) + (k & 1) : warp + k * MMVQ_NW;
}
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void grid_barrier(unsigned *counter, unsigned target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1u);
        while (ld_acquire_u32(counter) < target) { }
        __threadfence();
    }
    __syncthreads();
}

__global__ void __launch_bounds__(MMVQ_THREADS, 1) chain_kernel(const __grid_constant__ ChainProgram prog) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ float red[MMVQ_NW];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, bid = blockIdx.x, nblk = gridDim.x;
    const int n_stage = prog.n_stage, stage_bytes = prog.stage_bytes;
    unsigned long long *my_bars = reinterpret_cast<unsigned long long *>(smem) + warp * MMVQ_MAX_STAGE;
    unsigned char *act_base = smem + MMVQ_BAR_BYTES;
    unsigned char *my_ring = act_base + prog.act_bytes + (size_t)warp * n_stage * stage_bytes;
    if (lane == 0) {
        for (int s = 0; s < n_stage; s++) mbar_init(&my_bars[s], 1);
        mbar_fence_init();
    }
    __syncwarp();

    // ---- producer cursor (lane 0 issues; every lane tracks it)
    int ip = 0, ik = 0, iit = 0, istage = 0;
    ChainGeom ig;
    ig.nmy = 0; ig.n_it = 1; ig.r0 = 0; ig.nrows = 0; ig.pair = 0;
    bool ig_valid = false;
    auto issue = [&]() {
        for (;;) {
            if (ip >= prog.n_phases) return;
            if (prog.ph[ip].kind != B200_CHAIN_GEMV) { ip++; ig_valid = false; continue; }
            if (!ig_valid) { ig = chain_geom(prog.ph[ip], bid, nblk, warp); ig_valid = true; ik = 0; iit = 0; }
            if (ik >= ig.nmy) { ip++; ig_valid = false; continue; }
            break;
        }
        if (lane == 0) {
            const b200_chain_phase &ph = prog.ph[ip];
            ChainUnit u;
            chain_unit(ph, ig.r0 + chain_lr(ig, warp, ik), u);
            const int seg = chain_seg_bytes(u.type);
            const unsigned row_bytes = (unsigned)(ph.K / type_block_values(u.type) * type_block_bytes(u.type));
            const unsigned bytes = iit == ig.n_it - 1 ? row_bytes - (unsigned)(ig.n_it - 1) * seg : (unsigned)seg;
            const unsigned char *src = u.src + (size_t)iit * seg;
            const unsigned mis = (unsigned)((uintptr_t)src & 15);
            const unsigned len = (mis + bytes + 15u) & ~15u;
            mbar_expect_tx(&my_bars[istage], len);
            bulk_g2s(my_ring + (size_t)istage * stage_bytes, src - mis, len, &my_bars[istage]);
        }
        istage = istage + 1 == n_stage ? 0 : istage + 1;
        if (++iit == ig.n_it) { iit = 0; ik++; }
    };
    for (int s = 0; s < n_stage; s++) issue();

    int cstage = 0;
    unsigned cphase = 0, n_bar = 0;
    for (int p = 0; p < prog.n_phases; p++) {
        const b200_chain_phase &ph = prog.ph[p];
        if (p > 0) grid_barrier(prog.barrier, ++n_bar * (unsigned)nblk);  // results of phase p-1 are visible
        if (ph.kind == B200_CHAIN_ROPE_KV) {
            RopeKvDev a;
            a.q = ph.u.rope.q; a.k = ph.u.rope.k; a.v = ph.u.rope.v; a.q_out = ph.u.rope.q_out; a.k_out = ph.u.rope.k_out;
            a.k_cache = (char *)ph.u.rope.k_cache; a.v_cache = (char *)ph.u.rope.v_cache;
            a.k_slot = (char *const *)ph.u.rope.k_slot; a.v_slot = (char *const *)ph.u.rope.v_slot;
            a.D = ph.u.rope.D; a.H = ph.u.rope.H; a.HK = ph.u.rope.HK; a.T = 1; a.kv_type = ph.u.rope.kv_type;
            RopeDev rd;  // the host entry stored the derived constants in place of the raw parameters
            rd.n_dims = ph.u.rope.n_dims; rd.neox = 0; rd.freq_scale = ph.u.rope.freq_scale; rd.ext_factor = ph.u.rope.ext_factor;
            rd.attn_factor = ph.u.rope.attn_factor; rd.theta_scale = ph.u.rope.freq_base; rd.corr0 = ph.u.rope.beta_fast; rd.corr1 = ph.u.rope.beta_slow;
            const int64_t items = rope_kv_items(a.D, a.H, a.HK, 1);
            for (int64_t t4 = (int64_t)bid * MMVQ_THREADS + threadIdx.x; t4 < items; t4 += (int64_t)nblk * MMVQ_THREADS)
                rope_kv_store_item(a, ph.u.rope.pos, ph.u.rope.freq_factors, rd, t4);
            continue;
        }
        const ChainGeom g = chain_geom(ph, bid, nblk, warp);
        const int n_it = g.n_it;
        if (g.nrows > 0) {
            __syncthreads();  // every warp is done with the previous phase's activations
            load_activations<1>(act_base, ph.X, 0, ph.K, n_it, ph.norm_weight, ph.norm_eps, red);
            __syncthreads();
        }
        int last_nblk = ph.K / 32 - 4 * ((n_it - 1) * 32 + lane);
        last_nblk = last_nblk < 0 ? 0 : (last_nblk > 4 ? 4 : last_nblk);
        float gate_val = 0.f;
        for (int k = 0; k < g.nmy; k++) {
            ChainUnit u;
            chain_unit(ph, g.r0 + chain_lr(g, warp, k), u);
            const int seg = chain_seg_bytes(u.type);
            float acc[1] = {0.f};
            for (int it = 0; it < n_it; it++) {
                mbar_wait(&my_bars[cstage], cphase);
                const int nb = it == n_it - 1 ? last_nblk : 4;
                if (nb > 0) {
                    const unsigned mis = (unsigned)(((uintptr_t)u.src + (size_t)it * seg) & 15);
                    const unsigned char *sp = my_ring + (size_t)cstage * stage_bytes + mis;
                    switch (u.type) {
                        case B200_TYPE_Q4_K: SliceDot<B200_TYPE_Q4_K, 1>::run(sp, lane, it, act_base, n_it, nb, acc); break;
                        case B200_TYPE_Q5_K: SliceDot<B200_TYPE_Q5_K, 1>::run(sp, lane, it, act_base, n_it, nb, acc); break;
                        case B200_TYPE_Q6_K: SliceDot<B200_TYPE_Q6_K, 1>::run(sp, lane, it, act_base, n_it, nb, acc); break;
                        case B200_TYPE_Q8_0: SliceDot<B200_TYPE_Q8_0, 1>::run(sp, lane, it, act_base, n_it, nb, acc); break;
                        default: SliceDot<B200_TYPE_Q4_0, 1>::run(sp, lane, it, act_base, n_it, nb, acc); break;
                    }
                }
                __syncwarp();
                if (lane == 0) fence_proxy_async_smem();
                cstage = cstage + 1 == n_stage ? 0 : cstage + 1;
                cphase ^= (cstage == 0);
                issue();
            }
            const float v = warp_sum(acc[0]);
            if (g.pair) {
                if ((k & 1) == 0) gate_val = v;
                else if (lane == 0) ph.u.jobs[0].dst[u.dst - ph.u.jobs[1].dst] = (gate_val / (1.0f + expf(-gate_val))) * v;  // the pair writes jobs[0].dst
            } else if (lane == 0) {
                *u.dst = u.res ? v + __ldcg(u.res) : v;
            }
        }
    }
    // leave the barrier word zero for the next launch: the last CTA to pass resets it
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned old = atomicAdd(prog.barrier, 1u);
        if (old == (n_bar + 1) * (unsigned)nblk - 1) atomicExch(prog.barrier, 0u);
    }
}

int chain_launch(const b200_chain_phase *phases, int n_phases, unsigned *barrier, cudaStream_t st) {
    if (!phases || n_phases < 1 || n_phases > B200_CHAIN_MAX_PHASES || !barrier) { set_error("b200_decode_chain: bad arguments"); return B200_ERR_INVALID; }
    int rc = device_props();
    if (rc) return rc;
    ChainProgram prog;
    memset(&prog, 0, sizeof(prog));
    prog.n_phases = n_phases;
    prog.barrier = barrier;
    int max_nit = 1, stage_bytes = 0, n_gemv = 0;
    int64_t min_steps = 1 << 30;
    for (int p = 0; p < n_phases; p++) {
        b200_chain_phase ph = phases[p];
        if (ph.kind == B200_CHAIN_GEMV) {
            if (ph.n_jobs < 1 || ph.n_jobs > B200_CHAIN_MAX_JOBS || ph.K <= 0 || !ph.X || ((uintptr_t)ph.X & 15)) { set_error("b200_decode_chain: phase %d: bad GEMV header", p); return B200_ERR_INVALID; }
            if (ph.norm_weight && ((uintptr_t)ph.norm_weight & 15)) { set_error("b200_decode_chain: phase %d: norm weight alignment", p); return B200_ERR_UNSUPPORTED; }
            if (ph.silu_pair && (ph.n_jobs != 2 || ph.u.jobs[0].M != ph.u.jobs[1].M)) { set_error("b200_decode_chain: phase %d: silu_pair needs two jobs of equal M", p); return B200_ERR_INVALID; }
            int64_t rows = 0;
            for (int j = 0; j < ph.n_jobs; j++) {
                const b200_chain_job &jb = ph.u.jobs[j];
                const int t = jb.type, bv = type_block_values(t);
                if (!(t == B200_TYPE_Q4_0 || t == B200_TYPE_Q8_0 || t == B200_TYPE_Q4_K || t == B200_TYPE_Q5_K || t == B200_TYPE_Q6_K)) { set_error("b200_decode_chain: weight type %d", t); return B200_ERR_UNSUPPORTED; }
                if (ph.K % bv || !jb.W || !jb.dst || jb.M <= 0 || (jb.w_row_bytes & 1) || ((uintptr_t)jb.W & 1)) { set_error("b200_decode_chain: phase %d job %d", p, j); return B200_ERR_INVALID; }
                if (jb.w_row_bytes < (size_t)(ph.K / bv) * type_block_bytes(t)) { set_error("b200_decode_chain: row stride"); return B200_ERR_INVALID; }
                const int sb = t == B200_TYPE_Q4_K ? Geo<B200_TYPE_Q4_K>::STAGE_BYTES : t == B200_TYPE_Q5_K ? Geo<B200_TYPE_Q5_K>::STAGE_BYTES
                             : t == B200_TYPE_Q6_K ? Geo<B200_TYPE_Q6_K>::STAGE_BYTES : t == B200_TYPE_Q8_0 ? Geo<B200_TYPE_Q8_0>::STAGE_BYTES : Geo<B200_TYPE_Q4_0>::STAGE_BYTES;
                if (sb > stage_bytes) stage_bytes = sb;
                rows += jb.M;
            }
            const int nit = (ph.K + 4095) / 4096;
            if (nit > max_nit) max_nit = nit;
            n_gemv++;
            (void)rows; (void)min_steps;
        } else if (ph.kind == B200_CHAIN_ROPE_KV) {
            auto &r = ph.u.rope;
            if (!r.q || !r.k || !r.v || !r.q_out || !r.k_out || !r.pos || (!(r.k_cache || r.k_slot)) || (!(r.v_cache || r.v_slot))) { set_error("b200_decode_chain: phase %d: rope pointers", p); return B200_ERR_INVALID; }
            if ((r.mode & 2) || r.D % 32 || (r.kv_type != B200_TYPE_F16 && r.kv_type != B200_TYPE_Q8_0)) { set_error("b200_decode_chain: phase %d: rope mode / head size / cache type", p); return B200_ERR_UNSUPPORTED; }
            // derived constants (ggml_rope_yarn_corr_dims, ggml.c:14262-14270) stored in place of the raw parameters
            const float theta_scale = powf(r.freq_base, -2.0f / r.n_dims);
            auto corr_dim = [&](float n_rot) { return r.n_dims * logf(r.n_ctx_orig / (n_rot * 2 * 3.14159265358979323846f)) / (2 * logf(r.freq_base)); };
            const float start = floorf(corr_dim(r.beta_fast)), end = ceilf(corr_dim(r.beta_slow));
            r.freq_base = theta_scale;
            r.beta_fast = start > 0 ? start : 0;
            r.beta_slow = end < r.n_dims - 1 ? end : (float)(r.n_dims - 1);
        } else { set_error("b200_decode_chain: phase %d: unknown kind %d", p, ph.kind); return B200_ERR_INVALID; }
        prog.ph[p] = ph;
    }
    if (n_gemv == 0) { set_error("b200_decode_chain: no GEMV phase"); return B200_ERR_INVALID; }
    prog.stage_bytes = stage_bytes;
    prog.act_bytes = (int)(((size_t)act_col_bytes(max_nit) + 127) / 128 * 128);
    const size_t budget = g_smem_optin - 256;
    const size_t fixed = MMVQ_BAR_BYTES + (size_t)prog.act_bytes;
    if (fixed + 2 * (size_t)MMVQ_NW * stage_bytes > budget) { set_error("b200_decode_chain: activations do not fit in shared memory"); return B200_ERR_UNSUPPORTED; }
    size_t ns = (budget - fixed) / ((size_t)MMVQ_NW * stage_bytes);
    if (ns > (size_t)MMVQ_MAX_STAGE) ns = MMVQ_MAX_STAGE;
    prog.n_stage = (int)ns;
    const size_t smem = fixed + ns * MMVQ_NW * stage_bytes;
    static bool attr_set_dev[16] = {};
    int cur_dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&cur_dev));
    if (!attr_set_dev[cur_dev & 15]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(g_smem_optin - 256)));
        attr_set_dev[cur_dev & 15] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)g_sm_count);
    cfg.blockDim = dim3(MMVQ_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;  // all CTAs co-resident: the grid barrier cannot deadlock
    at[0].val.cooperative = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, chain_kernel, prog));
    return B200_OK;
}

}  // namespace b200

using namespace b200;

#ifdef B200_MMVQ_TRACE
extern "C" __attribute__((visibility("default"))) int b200_mmvq_trace(unsigned long long *out) {
    B200_CUDA_TRY(cudaMemcpyFromSymbol(out, g_mmvq_trace, sizeof(unsigned long long) * 148 * 8));
    return B200_OK;
}
#endif

extern "C" int b200_mul_mat_vec_q(int type, const void *W, size_t w_row_bytes, const float *X, int64_t x_stride, float *dst,
                                  int64_t dst_stride, int64_t K, int64_t M, int64_t N, b200_stream_t stream) {
    b200_gemv_job job = {W, X, dst, M, -1, 0};
    return mmvq_launch(type, &job, 1, w_row_bytes, x_stride, dst_stride, K, N, nullptr, 0, 0, (cudaStream_t)stream);
}

extern "C" int b200_mul_mat_vec_q_multi(int type, const b200_gemv_job *jobs, int n_jobs, size_t w_row_bytes, int64_t x_stride,
                                        int64_t dst_stride, int64_t K, int64_t N, const int32_t *ids, size_t expert_bytes,
                                        int32_t n_expert, b200_stream_t stream) {
    return mmvq_launch(type, jobs, n_jobs, w_row_bytes, x_stride, dst_stride, K, N, ids, expert_bytes, n_expert, (cudaStream_t)stream);
}

extern "C" int b200_decode_chain(const b200_chain_phase *phases, int n_phases, unsigned *barrier, b200_stream_t stream) {
    return chain_launch(phases, n_phases, barrier, (cudaStream_t)stream);
}

extern "C" int b200_mul_mat_vec_q_fused(int type, const b200_gemv_job *jobs, int n_jobs, size_t w_row_bytes, int64_t x_stride,
                                        int64_t dst_stride, int64_t K, int64_t N, const b200_gemv_fuse *fuse, b200_stream_t stream) {
    return mmvq_launch(type, jobs, n_jobs, w_row_bytes, x_stride, dst_stride, K, N, nullptr, 0, 0, (cudaStream_t)stream, fuse);
}

extern "C" int b200_mul_mat_id_vec_q(int type, const void *as, size_t w_row_bytes, size_t expert_bytes, int32_t n_expert,
                                     const int32_t *ids, int64_t ids_stride, const float *b, int64_t b_stride_row,
                                     int64_t b_stride_tok, float *dst, int64_t dst_stride_row, int64_t dst_stride_tok, int64_t K,
                                     int64_t M, int64_t n_used, int64_t n_b_rows, int64_t T, b200_stream_t stream) {
    if (!as || !ids || !b || !dst) { set_error("b200_mul_mat_id_vec_q: NULL argument"); return B200_ERR_INVALID; }
    if (T < 0 || n_used < 1 || n_b_rows < 1) { set_error("b200_mul_mat_id_vec_q: bad shape"); return B200_ERR_INVALID; }
    if (T * n_used > B200_GEMV_MAX_JOBS) {
        set_error("b200_mul_mat_id_vec_q: T*n_used=%lld exceeds %d (decode regime only)", (long long)(T * n_used), B200_GEMV_MAX_JOBS);
        return B200_ERR_UNSUPPORTED;
    }
    if (ids_stride * (T - 1) + n_used > 0x7fffffff) { set_error("b200_mul_mat_id_vec_q: ids too large"); return B200_ERR_UNSUPPORTED; }
    b200_gemv_job jobs[B200_GEMV_MAX_JOBS];
    int nj = 0;
    for (int64_t t = 0; t < T; t++) {
        for (int64_t e = 0; e < n_used; e++) {
            jobs[nj].W = as;
            jobs[nj].X = b + t * b_stride_tok + (e % n_b_rows) * b_stride_row;
            jobs[nj].dst = dst + t * dst_stride_tok + e * dst_stride_row;
            jobs[nj].M = M;
            jobs[nj].id_index = (int32_t)(t * ids_stride + e);
            jobs[nj]._pad = 0;
            nj++;
        }
    }
    if (nj == 0) return B200_OK;
    return mmvq_launch(type, jobs, nj, w_row_bytes, 0, 0, K, 1, ids, expert_bytes, n_expert, (cudaStream_t)stream);
}
