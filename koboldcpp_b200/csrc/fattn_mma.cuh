// fattn_mma.cuh -- FLASH_ATTN_EXT over an f16 KV cache with the K.Q and P.V contractions on the tensor cores.
//
// The scalar flash-decode kernel (fattn_kernel) spent 782 warp instructions per (16 keys x K+V) tile -- 128 FFMA plus 128
// f16->f32 conversions for one query column, ~70 shared loads, a 27-shuffle butterfly -- and ran out of issue slots at 3.0 TB/s with
// two warps per scheduler (ncu: profiles/r02_ncu_fattn_f16_scalar.txt).  Here a warp keeps up to 16 query columns (all q-heads of a
// GQA group x the tokens of a small batch) as the M side of mma.sync.m16n8k16 tiles:
//   S (16 q x 16 keys) = Q (16 x D, f16 A fragments resident in registers) . K^T   -- K rows come out of shared memory with
//                                                                                     ldmatrix (B fragments, keys = n)
//   O (16 q x D)      += P (the S accumulators, repacked as A fragments) . V        -- V through ldmatrix.trans
// ~190 instructions per tile whatever the number of columns <= 16, so a GQA group or a q8_0... (f16 here) long-context decode is
// no longer issue-bound; the KV stream (per-warp cp.async rings, split-K over CTAs, ticket merge) is the one of fattn.cu.
// Semantics: ggml_compute_forward_flash_attn_ext_f16 (ggml/src/ggml.c:15667-15877): Q rounded to f16, scores scaled in f32,
// optional softcap / mask / ALiBi slope, online softmax, f32 accumulation of P.V with P rounded to f16 (the reference rounds
// nothing there; the f16 rounding of P is within the tolerance stated in the tests).
#pragma once
#include "b200_common.cuh"

namespace b200 {

constexpr int FM_NW   = 4;    // warps per CTA
constexpr int FM_NBUF = 3;    // tiles in the per-warp ring
constexpr int FM_TP   = 16;   // KV rows per tile
constexpr int FM_MAXC = 16;   // query columns per CTA
constexpr int FM_MAX_CLUSTER = 8;  // splits merged on chip (portable cluster size)

template <int D>
struct MmaGeo {
    static constexpr int ROW_BYTES  = D * 2;
    static constexpr int ROW_STRIDE = ROW_BYTES + 16;  // ldmatrix: 8 rows x 16 bytes must hit 8 different bank groups
    static constexpr int TILE_BYTES = FM_TP * ROW_STRIDE;
};

__device__ __forceinline__ void ldsm_x4(unsigned addr, unsigned &r0, unsigned &r1, unsigned &r2, unsigned &r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(unsigned addr, unsigned &r0, unsigned &r1, unsigned &r2, unsigned &r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_f16(float (&c)[4], const unsigned (&a)[4], unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ unsigned pack_h2(float lo, float hi) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const unsigned *>(&h);
}

// thread-block cluster: the splits of one (kv-head, column tile) exchange their partial results through distributed shared memory
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float ld_cluster_f32(const float *local, unsigned rank) {
    unsigned ra;
    float v;
    // (not volatile, no memory clobber: the loads of a merge are independent and must overlap -- ordering comes from the
    // cluster barriers around them)
    asm("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(local)), "r"(rank));
    asm("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(ra));
    return v;
}

// copy rows [row0, row0+16) of one head (f16, D values) into a padded tile; rows >= n_kv are zero-filled
template <int D>
__device__ __forceinline__ void mma_load_tile(unsigned char *dst, const unsigned char *base, size_t nb1, int row0, int n_kv, int lane) {
    using G = MmaGeo<D>;
    constexpr int CPR = G::ROW_BYTES / 16;       // 16-byte chunks per row
    constexpr int RPP = 32 / CPR;                // rows per pass of the warp
    const int c = lane % CPR, r0 = lane / CPR;
    const unsigned char *src = base + (size_t)(row0 + r0) * nb1 + c * 16;
    unsigned char *d = dst + r0 * G::ROW_STRIDE + c * 16;
#pragma unroll
    for (int i = 0; i < FM_TP / RPP; i++) {
        if (row0 + r0 + i * RPP < n_kv) cp_async16(d, src);
        else *reinterpret_cast<uint4 *>(d) = make_uint4(0, 0, 0, 0);
        src += (size_t)RPP * nb1;
        d += RPP * G::ROW_STRIDE;
    }
}

template <int D>
__global__ void __launch_bounds__(FM_NW * 32, 2) fattn_mma_kernel(const __grid_constant__ FaParams p) {
    using G = MmaGeo<D>;
    constexpr int KS = D / 16;  // k-steps of the K.Q contraction == 16-dim blocks of the P.V output
    constexpr int RS = D + FA_RS_PAD;

    extern __shared__ __align__(128) unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t4 = lane & 3;  // fragment coordinates: row group / thread in group
    const int split = blockIdx.x, hk = blockIdx.y, ct = blockIdx.z;
    const int C = p.cols_per_cta;
    const int col0 = ct * C;
    const int ncol = min(C, p.n_cols - col0);

    unsigned char *my_tiles = smem + (size_t)warp * FM_NBUF * 2 * G::TILE_BYTES;
    const int n_tiles = (p.n_kv + FM_TP - 1) / FM_TP;
    const int t_begin = split * p.tiles_per_split;
    const int t_end   = min(n_tiles, t_begin + p.tiles_per_split);
    const unsigned char *kbase = p.k + (size_t)hk * p.k_nb2;
    const unsigned char *vbase = p.v + (size_t)hk * p.v_nb2;

    FTRACE(0);
    // the current token's K/V rows are written by the immediately preceding kernels
    pdl_wait();
    pdl_trigger();
    FTRACE(1);

    // ---- Q: this lane holds columns g and g + 8 (rows of the M side); columns >= ncol are zero.  The loads go out BEFORE the
    // first K/V tiles: behind 128 KB of cp.async per SM they came back 3.5 us later (profiles/r02_fattn_trace_timeline.log)
    float2 qraw[2][KS][2];
    float slope[2];
    const unsigned char *mrow[2];
#pragma unroll
    for (int hh = 0; hh < 2; hh++) {
        const int c = g + 8 * hh;
        const bool ok = c < ncol;
        const int col = col0 + (ok ? c : 0);
        const int n = col / p.gqa, h = hk * p.gqa + col % p.gqa;
        slope[hh] = 1.0f;
        if (p.max_bias > 0.0f)  // ALiBi, ggml.c:15743-15766
            slope[hh] = h < p.n_head_log2 ? powf(p.m0, (float)(h + 1)) : powf(p.m1, (float)(2 * (h - p.n_head_log2) + 1));
        mrow[hh] = p.mask != nullptr ? p.mask + (size_t)n * p.mask_nb1 : nullptr;
        const float *qrow = reinterpret_cast<const float *>(p.q + (size_t)n * p.q_nb1 + (size_t)h * p.q_nb2);
#pragma unroll
        for (int s = 0; s < KS; s++) {
            qraw[hh][s][0] = qraw[hh][s][1] = make_float2(0.f, 0.f);
            if (ok) {
                qraw[hh][s][0] = __ldcg(reinterpret_cast<const float2 *>(qrow + 16 * s + 2 * t4));
                qraw[hh][s][1] = __ldcg(reinterpret_cast<const float2 *>(qrow + 16 * s + 8 + 2 * t4));
            }
        }
    }

    int t_issue = t_begin + warp;
    int ibuf = 0;
    auto issue = [&]() {
        if (t_issue < t_end) {
            unsigned char *kb = my_tiles + (size_t)ibuf * 2 * G::TILE_BYTES;
            mma_load_tile<D>(kb, kbase, p.k_nb1, t_issue * FM_TP, p.n_kv, lane);
            mma_load_tile<D>(kb + G::TILE_BYTES, vbase, p.v_nb1, t_issue * FM_TP, p.n_kv, lane);
        }
        cp_async_commit();  // always commit (possibly empty) so the group accounting stays uniform
        t_issue += FM_NW;
        ibuf = ibuf + 1 == FM_NBUF ? 0 : ibuf + 1;
    };
#pragma unroll
    for (int s = 0; s < FM_NBUF - 1; s++) issue();

    // Q as f16 A fragments (the reference rounds Q to f16, ggml.c:15788)
    unsigned qa[KS][4];
#pragma unroll
    for (int hh = 0; hh < 2; hh++) {
#pragma unroll
        for (int s = 0; s < KS; s++) {
            qa[s][hh]     = pack_h2(qraw[hh][s][0].x, qraw[hh][s][0].y);  // a0 / a1: row g (+8), k = 2 t4, 2 t4 + 1
            qa[s][2 + hh] = pack_h2(qraw[hh][s][1].x, qraw[hh][s][1].y);  // a2 / a3: row g (+8), k = 8 + 2 t4, ...
        }
    }
    FTRACE(2);
    const bool mask_al4 = (((uintptr_t)p.mask | p.mask_nb1) & 3) == 0;
    float scale = p.scale;
    if (p.softcap != 0.0f) scale /= p.softcap;
    constexpr float LOG2E = 1.4426950408889634f;

    float m[2] = {-INFINITY, -INFINITY}, l[2] = {0.f, 0.f};  // l: this lane's share (its 4 keys of every tile)
    float o[2 * KS][4];
#pragma unroll
    for (int i = 0; i < 2 * KS; i++) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;

    // ldmatrix lane addresses inside a tile: matrix mi = lane >> 3, row r = lane & 7
    const int mi = lane >> 3, mr = lane & 7;
    const unsigned k_off = (unsigned)(((mi >> 1) * 8 + mr) * G::ROW_STRIDE + (mi & 1) * 16);  // K: keys (mi>>1)*8.., dims (mi&1)*8..
    const unsigned v_off = (unsigned)(((mi & 1) * 8 + mr) * G::ROW_STRIDE + (mi >> 1) * 16);  // V: keys (mi&1)*8.., dims (mi>>1)*8..

    int cbuf = 0;
    for (int t = t_begin + warp; t < t_end; t += FM_NW) {
        issue();
        cp_async_wait<FM_NBUF - 1>();
        __syncwarp();
        if (t == t_begin) FTRACE(3);
        const unsigned kt = smem_u32(my_tiles + (size_t)cbuf * 2 * G::TILE_BYTES);
        const unsigned vt = kt + G::TILE_BYTES;

        float s[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
        for (int ks = 0; ks < KS; ks++) {
            unsigned b0, b1, b2, b3;
            ldsm_x4(kt + k_off + ks * 32, b0, b1, b2, b3);
            mma_f16(s[0], qa[ks], b0, b1);  // keys 0-7
            mma_f16(s[1], qa[ks], b2, b3);  // keys 8-15
        }
        // s[nb][2 hh + e]: column g + 8 hh, key 16 t + 8 nb + 2 t4 + e
        const int key0 = t * FM_TP + 2 * t4;
        float tmax[2];
#pragma unroll
        for (int hh = 0; hh < 2; hh++) {
#pragma unroll
            for (int nb = 0; nb < 2; nb++) {
                const int key = key0 + 8 * nb;
                float s0 = s[nb][2 * hh] * scale, s1 = s[nb][2 * hh + 1] * scale;
                if (p.softcap != 0.0f) { s0 = p.softcap * tanhf(s0); s1 = p.softcap * tanhf(s1); }
                if (mrow[hh] != nullptr && key < p.n_kv) {
                    float2 mf;
                    if (mask_al4) {  // (an odd n_kv reads one half past the row: it is masked out below; buffers carry a tail)
                        mf = __half22float2(*reinterpret_cast<const __half2 *>(mrow[hh] + 2 * (size_t)key));
                    } else {
                        mf.x = __half2float(*reinterpret_cast<const __half *>(mrow[hh] + 2 * (size_t)key));
                        mf.y = key + 1 < p.n_kv ? __half2float(*reinterpret_cast<const __half *>(mrow[hh] + 2 * (size_t)key + 2)) : 0.f;
                    }
                    s0 += slope[hh] * mf.x;
                    s1 += slope[hh] * mf.y;
                }
                s[nb][2 * hh]     = key < p.n_kv ? s0 : -INFINITY;
                s[nb][2 * hh + 1] = key + 1 < p.n_kv ? s1 : -INFINITY;
            }
            float mx = fmaxf(fmaxf(s[0][2 * hh], s[0][2 * hh + 1]), fmaxf(s[1][2 * hh], s[1][2 * hh + 1]));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
            tmax[hh] = mx;
        }
        unsigned pa[4];
        float alpha[2];
#pragma unroll
        for (int hh = 0; hh < 2; hh++) {
            const float mn = fmaxf(m[hh], tmax[hh]);
            const float mb = mn == -INFINITY ? 0.f : mn * LOG2E;
            alpha[hh] = mn == -INFINITY ? 1.0f : exp2f(m[hh] * LOG2E - mb);
            const float p0 = exp2f(s[0][2 * hh] * LOG2E - mb), p1 = exp2f(s[0][2 * hh + 1] * LOG2E - mb);
            const float p2 = exp2f(s[1][2 * hh] * LOG2E - mb), p3 = exp2f(s[1][2 * hh + 1] * LOG2E - mb);
            const float q0 = mn == -INFINITY ? 0.f : p0, q1 = mn == -INFINITY ? 0.f : p1;
            const float q2 = mn == -INFINITY ? 0.f : p2, q3 = mn == -INFINITY ? 0.f : p3;
            l[hh] = l[hh] * alpha[hh] + ((q0 + q1) + (q2 + q3));
            m[hh] = mn;
            pa[hh]     = pack_h2(q0, q1);  // a0 / a1: row g (+8), keys 2 t4, 2 t4 + 1
            pa[2 + hh] = pack_h2(q2, q3);  // a2 / a3: keys 8 + 2 t4, ...
        }
        if (__any_sync(0xffffffffu, alpha[0] != 1.0f || alpha[1] != 1.0f)) {
#pragma unroll
            for (int i = 0; i < 2 * KS; i++) {
                o[i][0] *= alpha[0]; o[i][1] *= alpha[0];
                o[i][2] *= alpha[1]; o[i][3] *= alpha[1];
            }
        }
#pragma unroll
        for (int j = 0; j < KS; j++) {
            unsigned b0, b1, b2, b3;
            ldsm_x4_t(vt + v_off + j * 32, b0, b1, b2, b3);
            mma_f16(o[2 * j], pa, b0, b1);      // dims 16 j .. 16 j + 7
            mma_f16(o[2 * j + 1], pa, b2, b3);  // dims 16 j + 8 .. 16 j + 15
        }
        __syncwarp();  // all lanes are done with this buffer before it is refilled two iterations from now
        cbuf = cbuf + 1 == FM_NBUF ? 0 : cbuf + 1;
    }
    cp_async_wait<0>();
    FTRACE(4);

    // every lane of a quad holds a share of l
#pragma unroll
    for (int hh = 0; hh < 2; hh++) {
        l[hh] += __shfl_xor_sync(0xffffffffu, l[hh], 1);
        l[hh] += __shfl_xor_sync(0xffffffffu, l[hh], 2);
    }
    // ---- merge the warps of the CTA through shared memory (reuse the tile ring): red[warp][c][RS]
    __syncthreads();
    float *red = reinterpret_cast<float *>(smem);
#pragma unroll
    for (int hh = 0; hh < 2; hh++) {
        const int c = g + 8 * hh;
        if (c < ncol) {
            float *orow = red + ((size_t)warp * C + c) * RS;
#pragma unroll
            for (int i = 0; i < 2 * KS; i++) *reinterpret_cast<float2 *>(orow + 8 * i + 2 * t4) = make_float2(o[i][2 * hh], o[i][2 * hh + 1]);
            if (t4 == 0) { orow[D] = m[hh]; orow[D + 1] = l[hh]; }
        }
    }
    __syncthreads();
    const int pair = (ct * p.n_head_kv + hk);
    float *part = p.part + ((size_t)pair * p.n_splits + split) * C * RS;
    float *cpart = red + (size_t)FM_NW * C * RS;  // cluster merge: this CTA's merged (a[D], M, L) per column, read by its peers
    for (int idx = threadIdx.x; idx < ncol * D; idx += FM_NW * 32) {
        const int c = idx / D, d = idx % D;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < FM_NW; w++) M = fmaxf(M, red[((size_t)w * C + c) * RS + D]);
        float a = 0.f, L = 0.f;
#pragma unroll
        for (int w = 0; w < FM_NW; w++) {
            const float *orow = red + ((size_t)w * C + c) * RS;
            const float f = orow[D] == -INFINITY ? 0.f : __expf(orow[D] - M);
            a += f * orow[d];
            L += f * orow[D + 1];
        }
        if (p.n_splits == 1) {
            const int col = col0 + c, n = col / p.gqa, h = hk * p.gqa + col % p.gqa;
            p.dst[((size_t)n * p.n_head + h) * D + d] = a / L;
        } else if (p.cluster_merge) {
            cpart[(size_t)c * RS + d] = a;
            if (d == 0) { cpart[(size_t)c * RS + D] = M; cpart[(size_t)c * RS + D + 1] = L; }
        } else {
            part[(size_t)c * RS + d] = a;
            if (d == 0) { part[(size_t)c * RS + D] = M; part[(size_t)c * RS + D + 1] = L; }
        }
    }
    if (p.cluster_merge) {
        // ---- the splits of this (kv-head, column tile) are one thread-block cluster (cluster rank == split): every CTA merges a
        // slice of the head dims of all columns straight out of its peers' shared memory -- no global round trip, no ticket
        // (the global-memory merge below took 7 us of a 20 us launch at 7B @ 4k, profiles/r02_fattn_trace_timeline.log)
        cluster_sync_all();
        FTRACE(5);
        const int ns = p.n_splits;
        // phase 1: the (M, L) pairs of every split and column into local shared memory, as normalised weights
        float *wts = cpart + (size_t)C * RS;  // [C][FM_MAX_CLUSTER]: exp(M_s - M) / L
        if ((int)threadIdx.x < ncol) {
            const int c = threadIdx.x;
            const float *row = cpart + (size_t)c * RS;
            float ms[FM_MAX_CLUSTER], ls[FM_MAX_CLUSTER];
#pragma unroll
            for (int s2 = 0; s2 < FM_MAX_CLUSTER; s2++) {
                ms[s2] = s2 < ns ? ld_cluster_f32(row + D, s2) : -INFINITY;
                ls[s2] = s2 < ns ? ld_cluster_f32(row + D + 1, s2) : 0.f;
            }
            float M = -INFINITY;
#pragma unroll
            for (int s2 = 0; s2 < FM_MAX_CLUSTER; s2++) M = fmaxf(M, ms[s2]);
            float L = 0.f;
#pragma unroll
            for (int s2 = 0; s2 < FM_MAX_CLUSTER; s2++) {
                ms[s2] = ms[s2] == -INFINITY ? 0.f : __expf(ms[s2] - M);
                L += ms[s2] * ls[s2];
            }
#pragma unroll
            for (int s2 = 0; s2 < FM_MAX_CLUSTER; s2++) wts[c * FM_MAX_CLUSTER + s2] = ms[s2] / L;
        }
        __syncthreads();
        // phase 2: this CTA's slice of the head dims, ns independent remote loads per value
        const int dpc = (D + ns - 1) / ns, d_lo = split * dpc;
        const int nd = min(dpc, D - d_lo);
        for (int idx = threadIdx.x; idx < ncol * nd; idx += FM_NW * 32) {
            const int c = idx / nd, d = d_lo + idx % nd;
            const float *src = cpart + (size_t)c * RS + d;
            float v[FM_MAX_CLUSTER];
#pragma unroll
            for (int s2 = 0; s2 < FM_MAX_CLUSTER; s2++) v[s2] = s2 < ns ? ld_cluster_f32(src, s2) : 0.f;
            float a = 0.f;
#pragma unroll
            for (int s2 = 0; s2 < FM_MAX_CLUSTER; s2++) a += wts[c * FM_MAX_CLUSTER + s2] * v[s2];
            const int col = col0 + c, n = col / p.gqa, h = hk * p.gqa + col % p.gqa;
            p.dst[((size_t)n * p.n_head + h) * D + d] = a;
        }
        cluster_sync_all();  // nobody leaves while a peer may still read its shared memory
        FTRACE(7);
        return;
    }
    FTRACE(5);
    if (p.n_splits == 1) return;

    // ---- split-K merge by the last CTA of this (kv-head, column tile): same scheme as fattn_kernel
    __shared__ int is_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const int tk = atomicAdd(&p.tickets[pair], 1);
        is_last = tk == p.n_splits - 1;
        if (is_last) p.tickets[pair] = 0;  // leave the counter ready for the next launch
    }
    __syncthreads();
    FTRACE(6);
    if (!is_last) return;
    __threadfence();
    const float *all = p.part + (size_t)pair * p.n_splits * C * RS;
    float *wts = reinterpret_cast<float *>(smem);            // [C][FA_MAX_SPLITS]: exp(m_s - M)
    float *lss = wts + FM_MAXC * FA_MAX_SPLITS;               // [C][FA_MAX_SPLITS]: l_s
    const int ns = p.n_splits;
    for (int idx = threadIdx.x; idx < ncol * ns; idx += FM_NW * 32) {
        const int c = idx / ns, s2 = idx - c * ns;
        const float2 ml = __ldcg(reinterpret_cast<const float2 *>(all + ((size_t)s2 * C + c) * RS + D));
        wts[c * FA_MAX_SPLITS + s2] = ml.x;
        lss[c * FA_MAX_SPLITS + s2] = ml.y;
    }
    __syncthreads();
    if ((int)threadIdx.x < ncol) {
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
    for (int idx = threadIdx.x; idx < ncol * (D / 4); idx += FM_NW * 32) {
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
    FTRACE(7);
}

}  // namespace b200
