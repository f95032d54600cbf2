// fattn.cu -- (a10-a12) FLASH_ATTN_EXT over an f16 or q8_0 KV cache: split-K flash-decode with online softmax.
//
// Replaces flash_attn_vec_ext_f16/f32 + flash_attn_combine_results and (for now also) the wmma prefill kernel of
// the reference (/root/reference/ggml/src/ggml-cuda/fattn-vec-f16.cuh:8-299, fattn-common.cuh:519-708,
// fattn-wmma-f16.cuh:13-506).  Semantics follow ggml_compute_forward_flash_attn_ext_f16 (ggml/src/ggml.c:15667-15877).
//
// B200 design (HBM-bound at decode: every K/V byte must be read exactly once and >= 40 KB kept in flight per SM):
//   * grid = (split, kv-head, column tile).  One CTA serves ALL q-heads of a kv-head (GQA) and up to 8 query
//     columns at once, so K/V rows are fetched once per kv-head instead of once per q-head;
//   * the KV range of a CTA is walked in 16-row tiles, round-robin over its 4 warps; each warp owns a private
//     3-deep ring of (K tile, V tile) buffers filled with cp.async (16-byte LDGSTS for f16 rows, 8-byte for the
//     136-byte q8_0 rows): no CTA-wide barrier inside the KV loop, two tiles always in flight per warp;
//   * lane (r, j) holds 8 head dims (chunk j) of row-class r; per tile the 16 partial K.Q sums are reduced with a
//     transposing butterfly (8 shuffles instead of 64); q8_0 K uses int8 Q (quantised like the CPU reference,
//     quantize_row_q8_0) and dp4a; softmax state and the V accumulator stay in fp32 registers;
//   * the number of splits is chosen so that #CTAs is a near-multiple of 2 x SM count; the last CTA to finish a
//     (kv-head, column tile) merges the splits itself (ticket counter), so there is no combine launch.
#include <cstdlib>

#include "b200_common.cuh"

namespace b200 {

constexpr int FA_NW    = 4;   // warps per CTA
constexpr int FA_NBUF  = 3;   // tiles in the per-warp ring
constexpr int FA_TP    = 16;  // KV rows per tile
constexpr int FA_MAXC  = 8;   // query columns per CTA
constexpr int FA_MAX_SPLITS = 32;
constexpr int FA_RS_PAD = 4;   // floats behind the D accumulators of a partial result: m, l, 2 pad

struct FaParams {
    const unsigned char *q, *k, *v, *mask;
    float *dst;
    float *part;       // [pair][split][col][D + FA_RS_PAD]
    int   *tickets;    // [pair]
    int    D, N, n_head, n_head_kv, n_kv, gqa;
    int    n_cols;     // N * gqa query columns per kv-head
    int    n_splits, tiles_per_split;
    int    cols_per_cta;  // fattn_mma_kernel: query columns per CTA (<= 16, runtime)
    int    trace_slot;    // developer timeline (B200_FA_TRACE builds)
    int    dyn_tiles;     // fattn_mma_kernel, decode inside a cluster: KV tiles handed out dynamically to the warps of the cluster
    int    cluster_merge; // fattn_mma_kernel: the n_splits CTAs of a (kv-head, column tile) are one cluster and merge through DSMEM
    size_t q_nb1, q_nb2, k_nb1, k_nb2, v_nb1, v_nb2, mask_nb1;
    float  scale, max_bias, softcap, m0, m1;
    int    n_head_log2;
};

template <int TYPE, int D>
struct KvGeo {
    static constexpr int ROW_BYTES  = TYPE == B200_TYPE_F16 ? D * 2 : D / 32 * Q8_0_BYTES;
    static constexpr int ROW_STRIDE = TYPE == B200_TYPE_F16 ? ROW_BYTES : (ROW_BYTES + 15) / 16 * 16 + 8;  // q8_0: 8-byte aligned rows, staggered banks
    static constexpr int TILE_BYTES = (FA_TP * ROW_STRIDE + 16 + 127) / 128 * 128;
};

// copy a tile of rows [row0, row0+16) of one head into shared memory; rows >= n_kv are zero-filled
template <int TYPE, int D>
__device__ __forceinline__ void load_tile(unsigned char *dst, const unsigned char *base, size_t nb1, int row0, int n_kv, int lane) {
    using G = KvGeo<TYPE, D>;
    if constexpr (TYPE == B200_TYPE_F16) {
        constexpr int CPR = G::ROW_BYTES / 16;  // 16-byte chunks per row
#pragma unroll
        for (int i = lane; i < FA_TP * CPR; i += 32) {
            const int r = i / CPR, c = i % CPR;
            unsigned char *d = dst + r * G::ROW_STRIDE + c * 16;
            if (row0 + r < n_kv) cp_async16(d, base + (size_t)(row0 + r) * nb1 + c * 16);
            else *reinterpret_cast<uint4 *>(d) = make_uint4(0, 0, 0, 0);
        }
    } else if constexpr (G::ROW_BYTES % 8 == 0) {
        constexpr int CPR = G::ROW_BYTES / 8;  // 8-byte chunks per row (D=128: 136 = 17 * 8)
        for (int i = lane; i < FA_TP * CPR; i += 32) {
            const int r = i / CPR, c = i % CPR;
            unsigned char *d = dst + r * G::ROW_STRIDE + c * 8;
            if (row0 + r < n_kv) cp_async8(d, base + (size_t)(row0 + r) * nb1 + c * 8);
            else *reinterpret_cast<uint2 *>(d) = make_uint2(0, 0);
        }
    } else {
        constexpr int CPR = G::ROW_BYTES / 4;  // 4-byte chunks per row (D=64: 68 = 17 * 4)
        for (int i = lane; i < FA_TP * CPR; i += 32) {
            const int r = i / CPR, c = i % CPR;
            unsigned char *d = dst + r * G::ROW_STRIDE + c * 4;
            if (row0 + r < n_kv) cp_async4(d, base + (size_t)(row0 + r) * nb1 + c * 4);
            else *reinterpret_cast<unsigned *>(d) = 0u;
        }
    }
}

__device__ __forceinline__ void unpack8(const uint4 &u, float f[8]) {
    const __half2 *h = reinterpret_cast<const __half2 *>(&u);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float2 t = __half22float2(h[i]);
        f[2 * i] = t.x;
        f[2 * i + 1] = t.y;
    }
}

// transposing butterfly: every lane of an LPR-lane group enters with NV = LPR/2 partial sums (one per tile row of
// its row class) and leaves with the complete sum of row (j >> 1), duplicated in lanes j and j^1.
template <int LPR>
__device__ __forceinline__ float butterfly(float (&v)[LPR / 2], int lane) {
#pragma unroll
    for (int o = LPR / 2, n = LPR / 2; n >= 2; o >>= 1, n >>= 1) {
        const bool up = (lane & o) != 0;
#pragma unroll
        for (int i = 0; i < n / 2; i++) {
            const float send = up ? v[i] : v[i + n / 2];
            const float keep = up ? v[i + n / 2] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
    return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 1);
}

template <int TYPE, int D, int C>
__global__ void __launch_bounds__(FA_NW * 32, 2) fattn_kernel(const __grid_constant__ FaParams p) {
    using G = KvGeo<TYPE, D>;
    constexpr int LPR = D / 8;        // lanes per row (each lane: 8 dims)
    constexpr int RPW = 32 / LPR;     // rows per warp wave
    constexpr int NV  = FA_TP / RPW;  // rows per lane per tile == LPR / 2
    static_assert(NV == LPR / 2, "tile geometry");

    extern __shared__ __align__(128) unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int j = lane % LPR, rsel = lane / LPR;
    const int split = blockIdx.x, hk = blockIdx.y, ct = blockIdx.z;
    const int col0 = ct * C;
    const int ncol = min(C, p.n_cols - col0);

    unsigned char *my_tiles = smem + (size_t)warp * FA_NBUF * 2 * G::TILE_BYTES;

    const int n_tiles = (p.n_kv + FA_TP - 1) / FA_TP;
    const int t_begin = split * p.tiles_per_split;
    const int t_end   = min(n_tiles, t_begin + p.tiles_per_split);
    const unsigned char *kbase = p.k + (size_t)hk * p.k_nb2;
    const unsigned char *vbase = p.v + (size_t)hk * p.v_nb2;

    // K/V do not depend on the producer kernel in a decode graph only if the cache rows were written earlier;
    // the current token's row is written by the immediately preceding kernels, so wait first.
    pdl_wait();
    pdl_trigger();

    // ---- prologue: start the first tiles
    int t_issue = t_begin + warp;
    int ibuf = 0;
    auto issue = [&]() {
        if (t_issue < t_end) {
            unsigned char *kb = my_tiles + (size_t)ibuf * 2 * G::TILE_BYTES;
            load_tile<TYPE, D>(kb, kbase, p.k_nb1, t_issue * FA_TP, p.n_kv, lane);
            load_tile<TYPE, D>(kb + G::TILE_BYTES, vbase, p.v_nb1, t_issue * FA_TP, p.n_kv, lane);
        }
        cp_async_commit();  // always commit (possibly empty) so the group accounting stays uniform
        t_issue += FA_NW;
        ibuf = ibuf + 1 == FA_NBUF ? 0 : ibuf + 1;
    };
#pragma unroll
    for (int s = 0; s < FA_NBUF - 1; s++) issue();

    // ---- Q for this lane: 8 dims x C columns (f16 path: f32 pre-scaled; q8_0 path: int8 + per-32 scale)
    float qf[C][8];
    int   qi[C][2];
    float qd[C];
    float slope[C];
    int   col_n[C];
#pragma unroll
    for (int c = 0; c < C; c++) {
        const int col = col0 + (c < ncol ? c : 0);
        const int n = col / p.gqa, g = col % p.gqa;
        const int h = hk * p.gqa + g;
        col_n[c] = n;
        slope[c] = 1.0f;
        if (p.max_bias > 0.0f) {  // ALiBi, ggml.c:15743-15766
            slope[c] = h < p.n_head_log2 ? powf(p.m0, (float)(h + 1)) : powf(p.m1, (float)(2 * (h - p.n_head_log2) + 1));
        }
        const float *qrow = reinterpret_cast<const float *>(p.q + (size_t)n * p.q_nb1 + (size_t)h * p.q_nb2) + 8 * j;
        const float4 a = *reinterpret_cast<const float4 *>(qrow), b = *reinterpret_cast<const float4 *>(qrow + 4);
        const float x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        if constexpr (TYPE == B200_TYPE_F16) {
            // the CPU reference rounds Q to f16 before the dot (ggml.c:15788, vec_dot_type of an f16 K is f16)
#pragma unroll
            for (int i = 0; i < 8; i++) qf[c][i] = __half2float(__float2half_rn(x[i]));
        } else {
            // quantize_row_q8_0_ref on the 32-dim block this lane belongs to (4 lanes per block)
            float amax = 0.f;
#pragma unroll
            for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(x[i]));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
            const float d  = __fdiv_rn(amax, 127.0f);
            const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
            int q[8];
#pragma unroll
            for (int i = 0; i < 8; i++) q[i] = (int)roundf(__fmul_rn(x[i], id)) & 0xff;
            qi[c][0] = q[0] | (q[1] << 8) | (q[2] << 16) | (q[3] << 24);
            qi[c][1] = q[4] | (q[5] << 8) | (q[6] << 16) | (q[7] << 24);
            qd[c] = __half2float(__float2half_rn(d));
        }
    }
    float scale = p.scale;
    if (p.softcap != 0.0f) scale /= p.softcap;

    float m[C], l[C], acc[C][8];
#pragma unroll
    for (int c = 0; c < C; c++) {
        m[c] = -INFINITY;
        l[c] = 0.f;
#pragma unroll
        for (int i = 0; i < 8; i++) acc[c][i] = 0.f;
    }

    // ---- main loop over this warp's tiles
    int cbuf = 0;
    for (int t = t_begin + warp; t < t_end; t += FA_NW) {
        issue();
        cp_async_wait<FA_NBUF - 1>();
        __syncwarp();
        const unsigned char *kt = my_tiles + (size_t)cbuf * 2 * G::TILE_BYTES;
        const unsigned char *vt = kt + G::TILE_BYTES;

        // scores: partial sums for this lane's NV rows
        float s[C][NV];
#pragma unroll
        for (int i = 0; i < NV; i++) {
            const int r = i * RPW + rsel;
            if constexpr (TYPE == B200_TYPE_F16) {
                float kf[8];
                unpack8(*reinterpret_cast<const uint4 *>(kt + r * G::ROW_STRIDE + 16 * j), kf);
#pragma unroll
                for (int c = 0; c < C; c++) {
                    float a = 0.f;
#pragma unroll
                    for (int e = 0; e < 8; e++) a = fmaf(kf[e], qf[c][e], a);
                    s[c][i] = a;
                }
            } else {
                const unsigned char *blk = kt + r * G::ROW_STRIDE + (j >> 2) * Q8_0_BYTES;
                const float dk = h2f(*reinterpret_cast<const unsigned short *>(blk));
                const unsigned *w2 = reinterpret_cast<const unsigned *>(blk + 2 + 8 * (j & 3) - ((j >> 2) & 1 ? 0 : 2));
                // blk + 2 + 8*(j&3) is 4-byte aligned for odd blocks, 2 mod 4 for even blocks (rows are 8-byte aligned)
                unsigned w0, w1;
                if ((j >> 2) & 1) { w0 = w2[0]; w1 = w2[1]; }
                else { w0 = __funnelshift_r(w2[0], w2[1], 16); w1 = __funnelshift_r(w2[1], w2[2], 16); }
#pragma unroll
                for (int c = 0; c < C; c++) {
                    const int si = dp4a((int)w1, qi[c][1], dp4a((int)w0, qi[c][0], 0));
                    s[c][i] = (float)si * (dk * qd[c]);
                }
            }
        }
        const int my_row = t * FA_TP + (j >> 1) * RPW + rsel;  // the row whose full score this lane ends up with
        float pr[C];
#pragma unroll
        for (int c = 0; c < C; c++) {
            float sc = butterfly<LPR>(s[c], lane) * scale;
            if (p.softcap != 0.0f) sc = p.softcap * tanhf(sc);
            float mv = 0.f;
            if (p.mask != nullptr && my_row < p.n_kv) {
                mv = slope[c] * __half2float(*reinterpret_cast<const __half *>(p.mask + (size_t)col_n[c] * p.mask_nb1 + 2 * (size_t)my_row));
            }
            sc = my_row < p.n_kv ? sc + mv : -INFINITY;
            const float tmax = warp_max(sc);
            const float mn = fmaxf(m[c], tmax);
            const float alpha = mn == -INFINITY ? 1.0f : __expf(m[c] - mn);
            const float pv = mn == -INFINITY ? 0.0f : __expf(sc - mn);
            const float psum = warp_sum((j & 1) ? 0.f : pv);  // every row score is held by two lanes
            l[c] = l[c] * alpha + psum;
            m[c] = mn;
            pr[c] = pv;
#pragma unroll
            for (int e = 0; e < 8; e++) acc[c][e] *= alpha;
        }
        // P.V: lane accumulates its 8 dims over its NV rows
#pragma unroll
        for (int i = 0; i < NV; i++) {
            const int r = i * RPW + rsel;
            float vf[8];
            if constexpr (TYPE == B200_TYPE_F16) {
                unpack8(*reinterpret_cast<const uint4 *>(vt + r * G::ROW_STRIDE + 16 * j), vf);
            } else {
                const unsigned char *blk = vt + r * G::ROW_STRIDE + (j >> 2) * Q8_0_BYTES;
                const float dv = h2f(*reinterpret_cast<const unsigned short *>(blk));
                const unsigned *w2 = reinterpret_cast<const unsigned *>(blk + 2 + 8 * (j & 3) - ((j >> 2) & 1 ? 0 : 2));
                unsigned w0, w1;
                if ((j >> 2) & 1) { w0 = w2[0]; w1 = w2[1]; }
                else { w0 = __funnelshift_r(w2[0], w2[1], 16); w1 = __funnelshift_r(w2[1], w2[2], 16); }
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    vf[e]     = dv * (float)(signed char)((w0 >> (8 * e)) & 0xff);
                    vf[4 + e] = dv * (float)(signed char)((w1 >> (8 * e)) & 0xff);
                }
            }
#pragma unroll
            for (int c = 0; c < C; c++) {
                const float pw = __shfl_sync(0xffffffffu, pr[c], rsel * LPR + 2 * i);
#pragma unroll
                for (int e = 0; e < 8; e++) acc[c][e] = fmaf(pw, vf[e], acc[c][e]);
            }
        }
        __syncwarp();  // all lanes are done with this buffer before it is refilled two iterations from now
        cbuf = cbuf + 1 == FA_NBUF ? 0 : cbuf + 1;
    }
    cp_async_wait<0>();

    // ---- merge the row classes inside the warp (lanes with equal j hold disjoint rows)
#pragma unroll
    for (int c = 0; c < C; c++) {
#pragma unroll
        for (int o = LPR; o < 32; o <<= 1) {
#pragma unroll
            for (int e = 0; e < 8; e++) acc[c][e] += __shfl_xor_sync(0xffffffffu, acc[c][e], o);
        }
    }
    // ---- merge the warps of the CTA through shared memory (reuse the tile ring)
    __syncthreads();
    float *red = reinterpret_cast<float *>(smem);  // [warp][c][RS]
    constexpr int RS = D + FA_RS_PAD;  // m, l behind the D values; padded so that rows stay 16-byte aligned
    if (rsel == 0) {
#pragma unroll
        for (int c = 0; c < C; c++) {
            float *o = red + ((size_t)warp * C + c) * RS;
#pragma unroll
            for (int e = 0; e < 8; e++) o[8 * j + e] = acc[c][e];
            if (j == 0) { o[D] = m[c]; o[D + 1] = l[c]; }
        }
    }
    __syncthreads();
    const int pair = (ct * p.n_head_kv + hk);
    float *part = p.part + ((size_t)pair * p.n_splits + split) * C * RS;
    for (int idx = threadIdx.x; idx < C * D; idx += FA_NW * 32) {
        const int c = idx / D, d = idx % D;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < FA_NW; w++) M = fmaxf(M, red[((size_t)w * C + c) * RS + D]);
        float a = 0.f, L = 0.f;
#pragma unroll
        for (int w = 0; w < FA_NW; w++) {
            const float *o = red + ((size_t)w * C + c) * RS;
            const float f = o[D] == -INFINITY ? 0.f : __expf(o[D] - M);
            a += f * o[d];
            L += f * o[D + 1];
        }
        if (p.n_splits == 1) {
            if (c < ncol) {
                const int col = col0 + c, n = col / p.gqa, h = hk * p.gqa + col % p.gqa;
                p.dst[((size_t)n * p.n_head + h) * D + d] = a / L;
            }
        } else {
            part[(size_t)c * RS + d] = a;
            if (d == 0) { part[(size_t)c * RS + D] = M; part[(size_t)c * RS + D + 1] = L; }
        }
    }
    if (p.n_splits == 1) return;

    // ---- split-K merge by the last CTA of this (kv-head, column tile)
    __shared__ int is_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const int tk = atomicAdd(&p.tickets[pair], 1);
        is_last = tk == p.n_splits - 1;
        if (is_last) p.tickets[pair] = 0;  // leave the counter ready for the next launch
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    // The merging CTA is alone on the critical path: keep its loads independent and wide.  Phase 1 fetches the (m, l) pairs of
    // all splits in one round trip and turns them into normalised weights in shared memory; phase 2 streams the partial
    // accumulators with 16 independent 16-byte loads in flight per thread.
    const float *all = p.part + (size_t)pair * p.n_splits * C * RS;
    float *wts = reinterpret_cast<float *>(smem);            // [C][FA_MAX_SPLITS]: exp(m_s - M) / L
    float *lss = wts + C * FA_MAX_SPLITS;                     // [C][FA_MAX_SPLITS]: l_s
    const int ns = p.n_splits;
    for (int idx = threadIdx.x; idx < ncol * ns; idx += FA_NW * 32) {
        const int c = idx / ns, s2 = idx - c * ns;
        const float2 ml = __ldcg(reinterpret_cast<const float2 *>(all + ((size_t)s2 * C + c) * RS + D));
        wts[c * FA_MAX_SPLITS + s2] = ml.x;
        lss[c * FA_MAX_SPLITS + s2] = ml.y;
    }
    __syncthreads();
    if (threadIdx.x < ncol) {
        const int c = threadIdx.x;
        float M = -INFINITY;
        for (int s2 = 0; s2 < ns; s2++) M = fmaxf(M, wts[c * FA_MAX_SPLITS + s2]);
        float L = 0.f;
        for (int s2 = 0; s2 < ns; s2++) {
            const float ms = wts[c * FA_MAX_SPLITS + s2];
            const float f = ms == -INFINITY ? 0.f : __expf(ms - M);
            wts[c * FA_MAX_SPLITS + s2] = f;
            L += f * lss[c * FA_MAX_SPLITS + s2];
        }
        lss[c * FA_MAX_SPLITS] = L;
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < ncol * (D / 4); idx += FA_NW * 32) {
        const int c = idx / (D / 4), d4 = idx - c * (D / 4);
        const float4 *src = reinterpret_cast<const float4 *>(all + (size_t)c * RS) + d4;
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 16
        for (int s2 = 0; s2 < ns; s2++) {
            const float4 v = __ldcg(src + (size_t)s2 * C * RS / 4);
            const float f = wts[c * FA_MAX_SPLITS + s2];
            a.x += f * v.x; a.y += f * v.y; a.z += f * v.z; a.w += f * v.w;
        }
        const float L = lss[c * FA_MAX_SPLITS];
        const int col = col0 + c, n = col / p.gqa, h = hk * p.gqa + col % p.gqa;
        *reinterpret_cast<float4 *>(p.dst + ((size_t)n * p.n_head + h) * D + 4 * d4) = make_float4(a.x / L, a.y / L, a.z / L, a.w / L);
    }
}

#ifdef B200_FA_TRACE
// developer timeline: [launch slot][CTA][stamp] (globaltimer ns), read back by tools/trace_fattn.py
constexpr int FA_TRACE_SLOTS = 64, FA_TRACE_CTAS = 320;
__device__ unsigned long long g_fa_trace[FA_TRACE_SLOTS * FA_TRACE_CTAS * 8];
__device__ __forceinline__ unsigned long long fa_gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define FTRACE(i) do { if (threadIdx.x == 0) { const int cta__ = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z); \
    if (cta__ < FA_TRACE_CTAS) g_fa_trace[((size_t)p.trace_slot * FA_TRACE_CTAS + cta__) * 8 + (i)] = fa_gtime(); } } while (0)
static int g_fa_trace_slot = 0;
#define FTRACE_SLOT(p) do { (p).trace_slot = g_fa_trace_slot++ % FA_TRACE_SLOTS; } while (0)
#else
#define FTRACE(i) do {} while (0)
#define FTRACE_SLOT(p) do {} while (0)
#endif
}  // namespace b200
#include "fattn_mma.cuh"
namespace b200 {

// ------------------------------------------------------------------------------------------------ host side
struct FaPlan {
    bool prefill;  // tensor-core kernel, prompt form: 64 columns per CTA, tiles shared by its warps, no splits
    bool cluster;  // tensor-core kernel: splits merged through distributed shared memory
    int C, n_ct, n_splits, tiles_per_split;
    size_t part_bytes, ticket_bytes;
};

static int fa_sm_count() {
    static int sm = 0;
    if (!sm) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
            cudaGetLastError();
            sm = 148;
        }
    }
    return sm;
}

static bool fa_use_mma(const b200_fattn_args *a) {
    static const bool off = getenv("B200_FA_NO_MMA") != nullptr;  // developer switch: the scalar kernel
    if (a->D != 64 && a->D != 128) return true;  // head sizes 80 / 96 / 112 / 256 exist on the tensor-core kernels only
    if (a->type_k == B200_TYPE_Q4_0) return true;  // a q4_0 cache too (tiles dequantised to f16 in shared memory)
    if (off) return false;
    // q8_0 cache: the tile is dequantised to f16 in shared memory first (~2 instructions per value); that pays as soon as several
    // query columns share it (GQA group, prompt batch): 32/8 heads @ 32k 67 -> 54 us, but one column per kv-head 28 -> 31 us
    if (a->type_k == B200_TYPE_Q8_0) return a->N * (a->n_head / a->n_head_kv) >= 2;
    // a tiny cache with one or two kv-heads (a rank's share under 8-way tensor parallelism: 8 q-heads on 1 kv-head, 2 MB) is
    // pure latency: the scalar kernel spreads the q-heads over 8x more CTAs and finishes in 10.6 us against 12.6 us here
    if (a->N == 1 && a->n_head / a->n_head_kv <= 8 && (int64_t)a->n_head_kv * a->n_kv * a->D * 4 < (6 << 20)) return false;
    return a->type_k == B200_TYPE_F16;
}

static FaPlan fa_plan(const b200_fattn_args *a) {
    FaPlan pl;
    const int64_t n_cols = a->N * (a->n_head / a->n_head_kv);
    pl.prefill = false;
    static const bool no_prefill = getenv("B200_FA_NO_PREFILL") != nullptr;  // developer switch
    if (fa_use_mma(a) && a->type_k == B200_TYPE_F16 && n_cols > 2 * FM_MAXC && !no_prefill) {
        // prompt batch: 64 columns per CTA, every warp 16 of them over the whole KV range
        pl.prefill = true;
        pl.cluster = false;
        pl.C = FP_COLS;
        pl.n_ct = (int)((n_cols + FP_COLS - 1) / FP_COLS);
        pl.n_splits = 1;
        pl.tiles_per_split = (int)((a->n_kv + FM_TP - 1) / FM_TP);
        pl.ticket_bytes = 256;
        pl.part_bytes = 0;
        return pl;
    }
    if (fa_use_mma(a)) {
        // tensor-core kernel: a CTA costs the same for 1..16 columns, so a column tile is as wide as it can be
        pl.C = (int)(n_cols < FM_MAXC ? n_cols : FM_MAXC);
        pl.n_ct = (int)((n_cols + pl.C - 1) / pl.C);
        const int n_tiles = (int)((a->n_kv + FM_TP - 1) / FM_TP);
        const int64_t pairs = (int64_t)pl.n_ct * a->n_head_kv;
        const int slots = 2 * fa_sm_count();
        int best = 1;
        int64_t best_cost = -1;
        for (int s = 1; s <= FA_MAX_SPLITS; s++) {
            const int tps = (n_tiles + s - 1) / s;
            if (s > 1 && tps < 2 * FM_NW) break;
            if ((int64_t)(s - 1) * tps >= n_tiles) continue;  // would leave an empty split
            const int64_t waves = (pairs * s + slots - 1) / slots;
            const int64_t cost = waves * tps;
            if (best_cost < 0 || cost * 100 < best_cost * 97) { best = s; best_cost = cost; }
        }
        // 2, 4, 8 or 16 splits merge on chip (one thread-block cluster per (kv-head, column tile); other sizes pack badly into
        // the GPCs: 9-CTA clusters ran 32 heads @ 4k in 32.8 us instead of 18.6); more splits go through global memory, which costs
        // ~6 us more: take the best cluster size unless the unrestricted split count is much shorter
        static const bool no_cluster = getenv("B200_FA_NO_CLUSTER") != nullptr;
        if (best > 1 && !no_cluster) {
            int bestc = 0;
            int64_t costc = -1;
            for (int s = 2; s <= FM_MAX_CLUSTER; s *= 2) {
                const int tps = (n_tiles + s - 1) / s;
                if (tps < 2 * FM_NW) break;
                if ((int64_t)(s - 1) * tps >= n_tiles) continue;
                if (s == 16 && pairs * s > slots) continue;  // 16-CTA clusters only when all of them are resident at once
                const int64_t waves = (pairs * s + slots - 1) / slots;
                const int64_t cost = waves * tps;
                if (costc < 0 || cost * 100 < costc * 97) { bestc = s; costc = cost; }
            }
            // HBM-bound, a CTA streams a (K, V) tile in ~0.37 us (8 KB at 1/296 of 6.5 TB/s): 16 tiles ~ the 6 us saved
            if (bestc > 0 && costc - best_cost < 16) best = bestc;
        }
        pl.n_splits = best;
        pl.cluster = best > 1 && best <= FM_MAX_CLUSTER && (best & (best - 1)) == 0 && !no_cluster;
        pl.tiles_per_split = (n_tiles + best - 1) / best;
        pl.ticket_bytes = ((size_t)pairs * sizeof(int) + 255) / 256 * 256;
        // (the global-memory partials are sized for cluster launches too: if the device cannot place a cluster the launch falls back)
        pl.part_bytes = best > 1 ? (size_t)pairs * best * pl.C * (a->D + FA_RS_PAD) * sizeof(float) : 0;
        return pl;
    }
    pl.cluster = false;
    pl.C = n_cols >= 8 ? 8 : n_cols >= 4 ? 4 : n_cols >= 2 ? 2 : 1;
    if (n_cols == 3) pl.C = 4;
    if (n_cols > 4 && n_cols < 8) pl.C = 8;
    // Few kv-heads and a short cache (GQA decode, or 1-2 kv-heads per rank under tensor parallelism): even FA_MAX_SPLITS splits
    // leave SMs idle and every CTA carries all the q-heads of the group.  Such a launch is latency-bound, not HBM-bound, so spread
    // the q-heads over more CTAs (smaller C, more column tiles); the extra K/V reads hit L2.  Measured on B200 @ n_kv = 4096:
    // 64/8 heads 26.3 -> 21.4 us (C 8 -> 4), 32/4 22.8 -> 16.4 (C 2), 8/1 22.6 -> 10.6 (C 1); at 32k rows it loses, hence the
    // tiles-per-warp bound.
    {
        const int n_tiles0 = (int)((a->n_kv + FA_TP - 1) / FA_TP);
        const int slots0 = 2 * fa_sm_count();
        while (pl.C > 1 && n_tiles0 <= 4 * FA_MAX_SPLITS * FA_NW &&
               ((n_cols + pl.C - 1) / pl.C) * a->n_head_kv * FA_MAX_SPLITS < slots0) pl.C /= 2;
    }
    if (const char *e = getenv("B200_FA_C")) {  // developer override (tools/bench_kernels.py)
        const int c = atoi(e);
        if ((c == 1 || c == 2 || c == 4 || c == 8) && c < pl.C) pl.C = c;
    }
    pl.n_ct = (int)((n_cols + pl.C - 1) / pl.C);
    const int n_tiles = (int)((a->n_kv + FA_TP - 1) / FA_TP);
    const int64_t pairs = (int64_t)pl.n_ct * a->n_head_kv;
    const int slots = 2 * fa_sm_count();
    // choose the split count that minimises (waves x tiles per split), i.e. the length of the critical path;
    // every warp should still see >= 2 tiles; ties go to fewer splits (less merge traffic)
    int best = 1;
    int64_t best_cost = -1;
    for (int s = 1; s <= FA_MAX_SPLITS; s++) {
        const int tps = (n_tiles + s - 1) / s;
        if (s > 1 && tps < 2 * FA_NW) break;
        if ((int64_t)(s - 1) * tps >= n_tiles) continue;  // would leave an empty split
        const int64_t ctas = pairs * s;
        const int64_t waves = (ctas + slots - 1) / slots;
        const int64_t cost = waves * tps;
        if (best_cost < 0 || cost * 100 < best_cost * 97) { best = s; best_cost = cost; }
    }
    pl.n_splits = best;
    pl.tiles_per_split = (n_tiles + best - 1) / best;
    pl.ticket_bytes = ((size_t)pairs * sizeof(int) + 255) / 256 * 256;
    pl.part_bytes = best > 1 ? (size_t)pairs * best * pl.C * (a->D + FA_RS_PAD) * sizeof(float) : 0;
    return pl;
}

template <int TYPE, int D, int C>
static int fa_launch(const FaParams &p, const FaPlan &pl, cudaStream_t st, bool pdl) {
    using G = KvGeo<TYPE, D>;
    size_t smem = (size_t)FA_NW * FA_NBUF * 2 * G::TILE_BYTES;
    const size_t red = (size_t)FA_NW * C * (D + FA_RS_PAD) * sizeof(float);
    if (red > smem) smem = red;
    static bool attr_set_dev[16] = {};  // cudaFuncSetAttribute is per device
    int cur_dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&cur_dev));
    bool &attr_set = attr_set_dev[cur_dev & 15];
    if (!attr_set) {
        B200_CUDA_TRY(cudaFuncSetAttribute(fattn_kernel<TYPE, D, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)pl.n_splits, (unsigned)p.n_head_kv, (unsigned)pl.n_ct);
    cfg.blockDim = dim3(FA_NW * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, fattn_kernel<TYPE, D, C>, p));
    return B200_OK;
}

template <int TYPE, int D>
static int fa_launch_mma(const FaParams &p, const FaPlan &pl, cudaStream_t st, bool pdl) {
    using G = MmaGeo<TYPE, D>;
    size_t smem = (size_t)FM_NW * G::WARP_BYTES;
    const size_t merge = ((size_t)(FM_NW + 1) * FM_MAXC * (D + FA_RS_PAD) + 2 * FM_MAXC * FM_MAX_CLUSTER + 2 * FM_MAXC * FA_MAX_SPLITS) * sizeof(float);
    if (merge > smem) smem = merge;
    static bool attr_set_dev[16] = {};  // cudaFuncSetAttribute is per device
    int cur_dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&cur_dev));
    bool &attr_set = attr_set_dev[cur_dev & 15];
    if (!attr_set) {
        B200_CUDA_TRY(cudaFuncSetAttribute(fattn_mma_kernel<TYPE, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        B200_CUDA_TRY(cudaFuncSetAttribute(fattn_mma_kernel<TYPE, D>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)pl.n_splits, (unsigned)p.n_head_kv, (unsigned)pl.n_ct);
    cfg.blockDim = dim3(FM_NW * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    int na = 0;
    if (pdl) {
        at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    if (pl.cluster) {
        at[na].id = cudaLaunchAttributeClusterDimension;
        at[na].val.clusterDim.x = (unsigned)pl.n_splits;
        at[na].val.clusterDim.y = 1;
        at[na].val.clusterDim.z = 1;
        na++;
    }
    cfg.attrs = at;
    cfg.numAttrs = na;
    cudaError_t e = cudaLaunchKernelEx(&cfg, fattn_mma_kernel<TYPE, D>, p);
    if (e != cudaSuccess && pl.cluster) {
        // the device could not place the cluster (size, shared memory, a partitioned GPU): same splits, merged through global memory
        cudaGetLastError();
        FaParams q = p;
        q.cluster_merge = 0;
        q.dyn_tiles = 0;
        cfg.numAttrs = pdl ? 1 : 0;
        e = cudaLaunchKernelEx(&cfg, fattn_mma_kernel<TYPE, D>, q);
    }
    if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(fattn_mma_kernel)");
    return B200_OK;
}

template <int D>
static int fa_launch_prefill(const FaParams &p, const FaPlan &pl, cudaStream_t st, bool pdl) {
    using G = MmaGeo<B200_TYPE_F16, D>;
    const size_t smem = (size_t)FP_NBUF * 2 * G::TILE_BYTES;
    static bool attr_set_dev[16] = {};  // cudaFuncSetAttribute is per device
    int cur_dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&cur_dev));
    bool &attr_set = attr_set_dev[cur_dev & 15];
    if (!attr_set) {
        B200_CUDA_TRY(cudaFuncSetAttribute(fattn_mma_prefill_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)pl.n_ct, (unsigned)p.n_head_kv, 1);
    cfg.blockDim = dim3(FM_NW * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, fattn_mma_prefill_kernel<D>, p));
    return B200_OK;
}

template <int TYPE, int D>
static int fa_dispatch_c(const FaParams &p, const FaPlan &pl, cudaStream_t st, bool pdl) {
    switch (pl.C) {
        case 1: return fa_launch<TYPE, D, 1>(p, pl, st, pdl);
        case 2: return fa_launch<TYPE, D, 2>(p, pl, st, pdl);
        case 4: return fa_launch<TYPE, D, 4>(p, pl, st, pdl);
        case 8: return fa_launch<TYPE, D, 8>(p, pl, st, pdl);
    }
    return B200_ERR_INVALID;
}

extern bool g_mmvq_pdl;

static int fa_check(const b200_fattn_args *a) {
    if (!a || !a->q || !a->k || !a->v || !a->dst) { set_error("b200_flash_attn_ext: NULL argument"); return B200_ERR_INVALID; }
    const bool d_main = a->D == 64 || a->D == 128, d_more = a->D == 80 || a->D == 96 || a->D == 112 || a->D == 256;
    if (!d_main && !(d_more && a->type_k == B200_TYPE_F16)) {
        set_error("b200_flash_attn_ext: head size %lld not supported (64, 128; 80, 96, 112, 256 with an f16 cache)", (long long)a->D);
        return B200_ERR_UNSUPPORTED;
    }
    if (a->type_k != a->type_v || (a->type_k != B200_TYPE_F16 && a->type_k != B200_TYPE_Q8_0 && a->type_k != B200_TYPE_Q4_0)) {
        set_error("b200_flash_attn_ext: K/V types (%d, %d) not supported (f16/f16, q8_0/q8_0, q4_0/q4_0)", a->type_k, a->type_v);
        return B200_ERR_UNSUPPORTED;
    }
    if (a->n_head_kv <= 0 || a->n_head % a->n_head_kv || a->N <= 0 || a->n_kv <= 0) { set_error("b200_flash_attn_ext: bad shape"); return B200_ERR_INVALID; }
    const int align = a->type_k == B200_TYPE_F16 ? 16 : (a->D % 128 == 0 ? 8 : 4);
    if (((uintptr_t)a->k | (uintptr_t)a->v | a->k_nb1 | a->k_nb2 | a->v_nb1 | a->v_nb2) & (align - 1)) {
        set_error("b200_flash_attn_ext: K/V pointers and strides must be %d-byte aligned", align);
        return B200_ERR_UNSUPPORTED;
    }
    if (((uintptr_t)a->q | a->q_nb1 | a->q_nb2) & 15) { set_error("b200_flash_attn_ext: Q must be 16-byte aligned"); return B200_ERR_UNSUPPORTED; }
    if (a->n_head_kv > 65535 || a->N * (a->n_head / a->n_head_kv) > 8 * 65535) { set_error("b200_flash_attn_ext: grid too large"); return B200_ERR_UNSUPPORTED; }
    return B200_OK;
}

// bytes of split arrival counters at the front of the workspace (fattn_static.cu clears exactly these before a launch)
size_t fa_ticket_bytes(const b200_fattn_args *a) { return fa_check(a) == B200_OK ? fa_plan(a).ticket_bytes : 0; }

}  // namespace b200

using namespace b200;

#ifdef B200_FA_TRACE
extern "C" __attribute__((visibility("default"))) int b200_fattn_trace(unsigned long long *out) {
    B200_CUDA_TRY(cudaMemcpyFromSymbol(out, g_fa_trace, sizeof(unsigned long long) * FA_TRACE_SLOTS * FA_TRACE_CTAS * 8));
    g_fa_trace_slot = 0;
    return B200_OK;
}
#endif

extern "C" size_t b200_flash_attn_workspace_bytes(const b200_fattn_args *a) {
    if (fa_check(a) != B200_OK) return 0;
    const FaPlan pl = fa_plan(a);
    return pl.ticket_bytes + pl.part_bytes;
}

// The first ticket_bytes of the workspace are split-K arrival counters: they must be zero on entry and are left
// zero on exit (the backend zeroes its workspace once, at allocation).
extern "C" int b200_flash_attn_ext(const b200_fattn_args *a, void *workspace, size_t workspace_bytes, b200_stream_t stream) {
    int rc = fa_check(a);
    if (rc) return rc;
    const FaPlan pl = fa_plan(a);
    if (pl.n_splits > 1 && ((((uintptr_t)workspace) | (uintptr_t)a->dst) & 15)) {
        set_error("b200_flash_attn_ext: dst and workspace must be 16-byte aligned");
        return B200_ERR_UNSUPPORTED;
    }
    if (pl.n_splits > 1 && (!workspace || workspace_bytes < pl.ticket_bytes + pl.part_bytes)) {
        set_error("b200_flash_attn_ext: workspace too small (%zu < %zu)", workspace_bytes, pl.ticket_bytes + pl.part_bytes);
        return B200_ERR_INVALID;
    }
    FaParams p = {};
    p.q = (const unsigned char *)a->q; p.k = (const unsigned char *)a->k; p.v = (const unsigned char *)a->v;
    p.mask = (const unsigned char *)a->mask;
    p.dst = a->dst;
    p.tickets = (int *)workspace;
    p.part = workspace ? (float *)((unsigned char *)workspace + pl.ticket_bytes) : nullptr;
    p.D = (int)a->D; p.N = (int)a->N; p.n_head = (int)a->n_head; p.n_head_kv = (int)a->n_head_kv; p.n_kv = (int)a->n_kv;
    p.gqa = (int)(a->n_head / a->n_head_kv);
    p.n_cols = p.N * p.gqa;
    p.n_splits = pl.n_splits; p.tiles_per_split = pl.tiles_per_split;
    p.cols_per_cta = pl.C;
    p.cluster_merge = pl.cluster ? 1 : 0;
    // dynamic tile hand-out inside the cluster: measured SLOWER (7B @ 4k: 18.7 -> 20.1 us; the CTAs' finish times close up from
    // 4.3 to 2.6 us but 32 warps fetching from one shared-memory counter of CTA 0 every iteration delay everybody), so it is
    // opt-in (B200_FA_DYN=1) until the hand-out is chunked
    static const bool use_dyn = getenv("B200_FA_DYN") != nullptr;
    p.dyn_tiles = pl.cluster && a->N == 1 && use_dyn ? 1 : 0;
    FTRACE_SLOT(p);
    p.q_nb1 = a->q_nb1; p.q_nb2 = a->q_nb2; p.k_nb1 = a->k_nb1; p.k_nb2 = a->k_nb2; p.v_nb1 = a->v_nb1; p.v_nb2 = a->v_nb2;
    p.mask_nb1 = a->mask_nb1;
    p.scale = a->scale; p.max_bias = a->max_bias; p.softcap = a->logit_softcap;
    p.n_head_log2 = 1;
    while (p.n_head_log2 * 2 <= p.n_head) p.n_head_log2 *= 2;
    p.m0 = powf(2.0f, -(a->max_bias) / p.n_head_log2);
    p.m1 = powf(2.0f, -(a->max_bias / 2.0f) / p.n_head_log2);
    const cudaStream_t st = (cudaStream_t)stream;
    const bool pdl = g_mmvq_pdl;
    if (pl.prefill) {
        if (((uintptr_t)a->q | a->q_nb1 | a->q_nb2) & 7) { set_error("b200_flash_attn_ext: Q must be 8-byte aligned"); return B200_ERR_UNSUPPORTED; }
        switch (a->D) {
            case 64: return fa_launch_prefill<64>(p, pl, st, pdl);
            case 80: return fa_launch_prefill<80>(p, pl, st, pdl);
            case 96: return fa_launch_prefill<96>(p, pl, st, pdl);
            case 112: return fa_launch_prefill<112>(p, pl, st, pdl);
            case 256: return fa_launch_prefill<256>(p, pl, st, pdl);
            default: return fa_launch_prefill<128>(p, pl, st, pdl);
        }
    }
    if (fa_use_mma(a)) {
        if (((uintptr_t)a->q | a->q_nb1 | a->q_nb2) & 7) { set_error("b200_flash_attn_ext: Q must be 8-byte aligned"); return B200_ERR_UNSUPPORTED; }
        if (a->type_k == B200_TYPE_Q8_0) return a->D == 128 ? fa_launch_mma<B200_TYPE_Q8_0, 128>(p, pl, st, pdl) : fa_launch_mma<B200_TYPE_Q8_0, 64>(p, pl, st, pdl);
        if (a->type_k == B200_TYPE_Q4_0) return a->D == 128 ? fa_launch_mma<B200_TYPE_Q4_0, 128>(p, pl, st, pdl) : fa_launch_mma<B200_TYPE_Q4_0, 64>(p, pl, st, pdl);
        switch (a->D) {
            case 64: return fa_launch_mma<B200_TYPE_F16, 64>(p, pl, st, pdl);
            case 80: return fa_launch_mma<B200_TYPE_F16, 80>(p, pl, st, pdl);
            case 96: return fa_launch_mma<B200_TYPE_F16, 96>(p, pl, st, pdl);
            case 112: return fa_launch_mma<B200_TYPE_F16, 112>(p, pl, st, pdl);
            case 256: return fa_launch_mma<B200_TYPE_F16, 256>(p, pl, st, pdl);
            default: return fa_launch_mma<B200_TYPE_F16, 128>(p, pl, st, pdl);
        }
    }
    if (a->type_k == B200_TYPE_F16) {
        return a->D == 128 ? fa_dispatch_c<B200_TYPE_F16, 128>(p, pl, st, pdl) : fa_dispatch_c<B200_TYPE_F16, 64>(p, pl, st, pdl);
    }
    return a->D == 128 ? fa_dispatch_c<B200_TYPE_Q8_0, 128>(p, pl, st, pdl) : fa_dispatch_c<B200_TYPE_Q8_0, 64>(p, pl, st, pdl);
}
