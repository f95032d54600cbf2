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
constexpr int FM_MAX_CLUSTER = 16;  // splits merged on chip (cluster sizes above 8 are "non-portable": B200 schedules them)

template <int TYPE, int D>
struct MmaGeo {
    static constexpr bool Q8        = TYPE == B200_TYPE_Q8_0 || TYPE == B200_TYPE_Q4_0;  // a quantised cache (q8_0 or q4_0 blocks)
    static constexpr int BLK        = TYPE == B200_TYPE_Q4_0 ? Q4_0_BYTES : Q8_0_BYTES;
    static constexpr int ROW_BYTES  = D * 2;           // f16 tile (what ldmatrix reads)
    static constexpr int ROW_STRIDE = ROW_BYTES + 16;  // ldmatrix: 8 rows x 16 bytes must hit 8 different bank groups
    static constexpr int TILE_BYTES = FM_TP * ROW_STRIDE;
    // q8_0 / q4_0 cache: the ring holds raw rows (D/32 blocks of 34 / 18 bytes, 8- or 4-byte aligned); a warp dequantises the tile it is
    // about to use into ONE f16 tile pair of its own
    static constexpr int RAW_ROW    = D / 32 * BLK;
    static constexpr int RAW_STRIDE = (RAW_ROW + 15) / 16 * 16 + 8;
    static constexpr int RAW_TILE   = (FM_TP * RAW_STRIDE + 16 + 127) / 128 * 128;
    static constexpr int RING_TILE  = Q8 ? RAW_TILE : TILE_BYTES;                                  // one K or V tile in the ring
    static constexpr int WARP_BYTES = FM_NBUF * 2 * RING_TILE + (Q8 ? 2 * TILE_BYTES : 0);         // ring (+ f16 scratch pair)
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
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
// fetch-and-increment of a counter in the shared memory of CTA `rank` of the cluster
__device__ __forceinline__ unsigned atom_inc_cluster(unsigned *local, unsigned rank) {
    unsigned ra, v;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(local)), "r"(rank));
    asm volatile("atom.shared::cluster.add.u32 %0, [%1], 1;" : "=r"(v) : "r"(ra) : "memory");
    return v;
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
    using G = MmaGeo<B200_TYPE_F16, D>;
    constexpr int CPR = G::ROW_BYTES / 16;       // 16-byte chunks per row
    if constexpr (32 % CPR == 0 || CPR % 32 == 0) {
        constexpr int RPP = CPR >= 32 ? 1 : 32 / CPR;  // rows per pass of the warp
        constexpr int CPL = CPR >= 32 ? CPR / 32 : 1;  // chunks per lane and row (D = 256: 1; wider heads: more)
        const int c = lane % CPR, r0 = CPR >= 32 ? 0 : lane / CPR;
        const unsigned char *src = base + (size_t)(row0 + r0) * nb1 + c * 16;
        unsigned char *d = dst + r0 * G::ROW_STRIDE + c * 16;
#pragma unroll
        for (int i = 0; i < FM_TP / RPP; i++) {
#pragma unroll
            for (int k = 0; k < CPL; k++) {
                if (row0 + r0 + i * RPP < n_kv) cp_async16(d + k * 512, src + k * 512);
                else *reinterpret_cast<uint4 *>(d + k * 512) = make_uint4(0, 0, 0, 0);
            }
            src += (size_t)RPP * nb1;
            d += RPP * G::ROW_STRIDE;
        }
    } else {  // head sizes 80 / 96 / 112: 10 / 12 / 14 chunks per row
        for (int i = lane; i < FM_TP * CPR; i += 32) {
            const int r = i / CPR, c = i - r * CPR;
            unsigned char *d = dst + r * G::ROW_STRIDE + c * 16;
            if (row0 + r < n_kv) cp_async16(d, base + (size_t)(row0 + r) * nb1 + c * 16);
            else *reinterpret_cast<uint4 *>(d) = make_uint4(0, 0, 0, 0);
        }
    }
}

// q8_0 / q4_0: copy raw rows [row0, row0+16) of one head; rows >= n_kv are zero-filled (d = 0: dequantises to zeros)
template <int TYPE, int D>
__device__ __forceinline__ void mma_load_tile_q8(unsigned char *dst, const unsigned char *base, size_t nb1, int row0, int n_kv, int lane) {
    using G = MmaGeo<TYPE, D>;
    constexpr int CH = G::RAW_ROW % 8 == 0 ? 8 : 4;  // q8_0: D = 128: 136 = 17 x 8, D = 64: 68 = 17 x 4; q4_0: 72 = 9 x 8, 36 = 9 x 4
    constexpr int CPR = G::RAW_ROW / CH;
    for (int i = lane; i < FM_TP * CPR; i += 32) {
        const int r = i / CPR, c = i - r * CPR;
        unsigned char *d = dst + r * G::RAW_STRIDE + c * CH;
        if (row0 + r < n_kv) {
            if constexpr (CH == 8) cp_async8(d, base + (size_t)(row0 + r) * nb1 + c * CH);
            else cp_async4(d, base + (size_t)(row0 + r) * nb1 + c * CH);
        } else {
            if constexpr (CH == 8) *reinterpret_cast<uint2 *>(d) = make_uint2(0, 0);
            else *reinterpret_cast<unsigned *>(d) = 0u;
        }
    }
}
// four int8 (one word) -> two half2 (exact): (s ^ 0x80) is u = s + 128; 0x6400 | u is the half 1024 + u; minus 1152
__device__ __forceinline__ void i8x4_to_h2x2(unsigned w, __half2 &lo, __half2 &hi) {
    const unsigned u = w ^ 0x80808080u;
    const unsigned a = __byte_perm(u, 0x64646464u, 0x4140), b = __byte_perm(u, 0x64646464u, 0x4342);
    const __half2 off = __floats2half2_rn(1152.0f, 1152.0f);
    lo = __hsub2(*reinterpret_cast<const __half2 *>(&a), off);
    hi = __hsub2(*reinterpret_cast<const __half2 *>(&b), off);
}
// four nibble values 0..15 (one per byte of w) -> two half2 holding value - 8 (exact): 0x6400 | n is the half 1024 + n; minus 1032
__device__ __forceinline__ void u4x4_to_h2x2(unsigned w, __half2 &lo, __half2 &hi) {
    const unsigned a = __byte_perm(w, 0x64646464u, 0x4140), b = __byte_perm(w, 0x64646464u, 0x4342);
    const __half2 off = __floats2half2_rn(1032.0f, 1032.0f);
    lo = __hsub2(*reinterpret_cast<const __half2 *>(&a), off);
    hi = __hsub2(*reinterpret_cast<const __half2 *>(&b), off);
}
// dequantise one raw q8_0 / q4_0 tile (16 rows x D/32 blocks) into an f16 tile: value = fp16(d * q) (one f16 rounding, 2^-11 relative)
template <int TYPE, int D>
__device__ __forceinline__ void mma_dequant_tile_q8(unsigned char *dst, const unsigned char *raw, int lane) {
    using G = MmaGeo<TYPE, D>;
    constexpr int NB = D / 32;
#pragma unroll
    for (int i = 0; i < FM_TP * NB / 32; i++) {
        const int idx = lane + 32 * i;
        const int r = idx / NB, b = idx - r * NB;
        const unsigned char *src = raw + r * G::RAW_STRIDE + b * G::BLK;  // 2-byte aligned: d, then the quants
        const __half dh = *reinterpret_cast<const __half *>(src);
        const __half2 d2 = __half2half2(dh);
        uint4 out[4];
        if constexpr (TYPE == B200_TYPE_Q4_0) {
            // 16 bytes: low nibbles are values 0..15, high nibbles 16..31 (dequantize_row_q4_0, ggml-quants.c:1523-1541)
            unsigned w[4];
            lds_words<4>(src + 2, w);
#pragma unroll
            for (int k = 0; k < 4; k++) {
                __half2 lo, hi;
                u4x4_to_h2x2(w[k] & 0x0f0f0f0fu, lo, hi);
                reinterpret_cast<__half2 *>(out)[2 * k] = __hmul2(lo, d2);
                reinterpret_cast<__half2 *>(out)[2 * k + 1] = __hmul2(hi, d2);
                u4x4_to_h2x2((w[k] >> 4) & 0x0f0f0f0fu, lo, hi);
                reinterpret_cast<__half2 *>(out)[8 + 2 * k] = __hmul2(lo, d2);
                reinterpret_cast<__half2 *>(out)[8 + 2 * k + 1] = __hmul2(hi, d2);
            }
        } else {
            unsigned w[8];
            lds_words<8>(src + 2, w);
#pragma unroll
            for (int k = 0; k < 8; k++) {
                __half2 lo, hi;
                i8x4_to_h2x2(w[k], lo, hi);
                lo = __hmul2(lo, d2);
                hi = __hmul2(hi, d2);
                reinterpret_cast<__half2 *>(out)[2 * k] = lo;
                reinterpret_cast<__half2 *>(out)[2 * k + 1] = hi;
            }
        }
        uint4 *o = reinterpret_cast<uint4 *>(dst + r * G::ROW_STRIDE + b * 64);
        o[0] = out[0]; o[1] = out[1]; o[2] = out[2]; o[3] = out[3];
    }
}

// One (16 keys x K, V) tile against the warp's 16 query columns: S = Q.K^T on the tensor cores, online softmax on the accumulator
// fragments, O += P.V.  kt / vt: shared addresses of the f16 tiles; t: tile number (keys 16 t ..).
template <int D>
__device__ __forceinline__ void mma_tile_step(unsigned kt, unsigned vt, int t, const unsigned (&qa)[D / 16][4], float (&m)[2], float (&l)[2],
                                              float (&o)[D / 8][4], const unsigned char *const (&mrow)[2], const float (&slope)[2], float scale,
                                              float softcap, int n_kv, bool mask_al4, unsigned k_off, unsigned v_off, int t4) {
    constexpr int KS = D / 16;
    constexpr float LOG2E = 1.4426950408889634f;
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
            if (softcap != 0.0f) { s0 = softcap * tanhf(s0); s1 = softcap * tanhf(s1); }
            if (mrow[hh] != nullptr && key < n_kv) {
                float2 mf;
                if (mask_al4) {  // (an odd n_kv reads one half past the row: it is masked out below; buffers carry a tail)
                    mf = __half22float2(*reinterpret_cast<const __half2 *>(mrow[hh] + 2 * (size_t)key));
                } else {
                    mf.x = __half2float(*reinterpret_cast<const __half *>(mrow[hh] + 2 * (size_t)key));
                    mf.y = key + 1 < n_kv ? __half2float(*reinterpret_cast<const __half *>(mrow[hh] + 2 * (size_t)key + 2)) : 0.f;
                }
                s0 += slope[hh] * mf.x;
                s1 += slope[hh] * mf.y;
            }
            s[nb][2 * hh]     = key < n_kv ? s0 : -INFINITY;
            s[nb][2 * hh + 1] = key + 1 < n_kv ? s1 : -INFINITY;
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
}

template <int TYPE, int D>
__global__ void __launch_bounds__(FM_NW * 32, 2) fattn_mma_kernel(const __grid_constant__ FaParams p) {
    using G = MmaGeo<TYPE, D>;
    constexpr bool Q8 = G::Q8;
    constexpr int KS = D / 16;  // k-steps of the K.Q contraction == 16-dim blocks of the P.V output
    constexpr int RS = D + FA_RS_PAD;

    extern __shared__ __align__(128) unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t4 = lane & 3;  // fragment coordinates: row group / thread in group
    const int split = blockIdx.x, hk = blockIdx.y, ct = blockIdx.z;
    const int C = p.cols_per_cta;
    const int col0 = ct * C;
    const int ncol = min(C, p.n_cols - col0);

    unsigned char *my_tiles = smem + (size_t)warp * G::WARP_BYTES;
    unsigned char *my_f16   = my_tiles + FM_NBUF * 2 * G::RING_TILE;  // q8_0: the dequantised (K, V) pair of the current tile
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

    // ---- batches (prompt processing): KV tiles whose mask is -inf for every column of this CTA (the causal upper triangle, the
    // cells of other sequences) are neither loaded nor computed.  Bit i of `skip`: the i-th tile of this warp, t = t_begin + warp
    // + 4 i; lane i checks tile i (two passes cover 64 tiles per warp; later tiles are simply processed).
    // tile walk of this warp: t_first, t_first + t_step, ... < t_hi.  Splits that are one cluster interleave their tiles (warp w of
    // CTA s takes tiles 4 s + w + 4 n_splits k): every CTA sweeps the whole KV range at the same pace, so differences between
    // regions of the cache (DRAM page / channel luck) do not make one CTA finish late; otherwise a split owns a contiguous range
    const bool inter = p.cluster_merge != 0;
    const int nwt = FM_NW * p.n_splits;  // warps of the cluster
    const int t_hi = inter ? n_tiles : t_end;
    const int t_first = inter ? split * FM_NW + warp : t_begin + warp, t_step = inter ? nwt : FM_NW;
    unsigned long long skip = 0;
    if (p.N > 1 && p.mask != nullptr) {
        const bool al16 = (((uintptr_t)p.mask | p.mask_nb1) & 15) == 0;
#pragma unroll 1
        for (int pass = 0; pass < 2; pass++) {
            const int tt = t_first + t_step * (pass * 32 + lane);
            bool masked = false;
            if (tt < t_hi) {
                masked = true;
                const int key = tt * FM_TP;
                for (int c = 0; c < ncol; c++) {
                    const unsigned char *mr = p.mask + (size_t)((col0 + c) / p.gqa) * p.mask_nb1 + 2 * (size_t)key;
                    if (al16 && key + FM_TP <= p.n_kv) {
                        const uint4 u0 = __ldg(reinterpret_cast<const uint4 *>(mr)), u1 = __ldg(reinterpret_cast<const uint4 *>(mr) + 1);
                        masked = masked && (u0.x & u0.y & u0.z & u0.w & u1.x & u1.y & u1.z & u1.w) == 0xfc00fc00u &&
                                 (u0.x | u0.y | u0.z | u0.w | u1.x | u1.y | u1.z | u1.w) == 0xfc00fc00u;
                    } else {
                        for (int e = 0; e < FM_TP; e++)
                            if (key + e < p.n_kv) masked = masked && *reinterpret_cast<const unsigned short *>(mr + 2 * e) == 0xfc00;
                    }
                }
            }
            skip |= (unsigned long long)__ballot_sync(0xffffffffu, masked) << (32 * pass);
        }
    }

    // ---- decode inside a cluster, opt-in (p.dyn_tiles): the KV tiles of the (kv-head, column tile) are handed out DYNAMICALLY
    // to the warps of all its CTAs (a counter in the shared memory of CTA 0, fetched one iteration ahead): with fixed ranges the
    // CTAs of a launch finish up to 4 us apart (they do not get equal shares of the HBM bandwidth) and the merge waits for the
    // last one (profiles/r02_fattn_trace_timeline.log).  Not the default: see fattn.cu.
    const bool dyn = p.dyn_tiles != 0;
    __shared__ unsigned tile_ctr;
    if (dyn) {
        if (threadIdx.x == 0) tile_ctr = (unsigned)(nwt * (FM_NBUF - 1));  // the first FM_NBUF - 1 tiles of every warp are fixed
        cluster_arrive();  // (waited for just before the first fetch)
    }

    int ibuf = 0, n_issued = 0;
    auto issue = [&](int tt) {
        const bool sk = n_issued < 64 && ((skip >> n_issued) & 1ull) != 0;
        n_issued++;
        if (tt < t_hi && !sk) {
            unsigned char *kb = my_tiles + (size_t)ibuf * 2 * G::RING_TILE;
            if constexpr (Q8) {
                mma_load_tile_q8<TYPE, D>(kb, kbase, p.k_nb1, tt * FM_TP, p.n_kv, lane);
                mma_load_tile_q8<TYPE, D>(kb + G::RING_TILE, vbase, p.v_nb1, tt * FM_TP, p.n_kv, lane);
            } else {
                mma_load_tile<D>(kb, kbase, p.k_nb1, tt * FM_TP, p.n_kv, lane);
                mma_load_tile<D>(kb + G::RING_TILE, vbase, p.v_nb1, tt * FM_TP, p.n_kv, lane);
            }
        }
        cp_async_commit();  // always commit (possibly empty) so the group accounting stays uniform
        ibuf = ibuf + 1 == FM_NBUF ? 0 : ibuf + 1;
    };
    static_assert(FM_NBUF == 3, "the tile FIFO below holds FM_NBUF - 1 = 2 entries");
    int f0 = t_first, f1 = t_first + t_step;  // tiles in flight, oldest first
    issue(f0);
    issue(f1);

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
    float m[2] = {-INFINITY, -INFINITY}, l[2] = {0.f, 0.f};  // l: this lane's share (its 4 keys of every tile)
    float o[2 * KS][4];
#pragma unroll
    for (int i = 0; i < 2 * KS; i++) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;

    // ldmatrix lane addresses inside a tile: matrix mi = lane >> 3, row r = lane & 7
    const int mi = lane >> 3, mr = lane & 7;
    const unsigned k_off = (unsigned)(((mi >> 1) * 8 + mr) * G::ROW_STRIDE + (mi & 1) * 16);  // K: keys (mi>>1)*8.., dims (mi&1)*8..
    const unsigned v_off = (unsigned)(((mi & 1) * 8 + mr) * G::ROW_STRIDE + (mi >> 1) * 16);  // V: keys (mi&1)*8.., dims (mi>>1)*8..

    int cbuf = 0, n_done = 0;
    unsigned grabbed = 0;  // dynamic mode: the tile fetched during the previous iteration (lane 0)
    if (dyn) {
        cluster_wait();  // tile_ctr of CTA 0 is initialised
        if (lane == 0) grabbed = atom_inc_cluster(&tile_ctr, 0);
    }
    for (;;) {
        int tnew;
        if (dyn) {
            tnew = (int)__shfl_sync(0xffffffffu, grabbed, 0);
            if (lane == 0 && tnew < t_hi) grabbed = atom_inc_cluster(&tile_ctr, 0);  // for the next iteration: its latency hides here
        } else {
            tnew = f1 + t_step;
        }
        const int t = f0;
        if (t >= t_hi) break;  // tile numbers only grow: nothing is left for this warp
        issue(tnew);
        f0 = f1;
        f1 = tnew;
        cp_async_wait<FM_NBUF - 1>();
        __syncwarp();
        if (n_done == 0 && warp == 0) FTRACE(3);
        const bool sk = n_done < 64 && ((skip >> n_done) & 1ull) != 0;
        n_done++;
        if (sk) {  // every score of the tile is -inf: p = 0, nothing changes
            cbuf = cbuf + 1 == FM_NBUF ? 0 : cbuf + 1;
            continue;
        }
        unsigned kt, vt;
        if constexpr (Q8) {
            const unsigned char *rk = my_tiles + (size_t)cbuf * 2 * G::RING_TILE;
            mma_dequant_tile_q8<TYPE, D>(my_f16, rk, lane);
            mma_dequant_tile_q8<TYPE, D>(my_f16 + G::TILE_BYTES, rk + G::RING_TILE, lane);
            __syncwarp();
            kt = smem_u32(my_f16);
            vt = kt + G::TILE_BYTES;
        } else {
            kt = smem_u32(my_tiles + (size_t)cbuf * 2 * G::RING_TILE);
            vt = kt + G::TILE_BYTES;
        }

        mma_tile_step<D>(kt, vt, t, qa, m, l, o, mrow, slope, scale, p.softcap, p.n_kv, mask_al4, k_off, v_off, t4);
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
        // phase 1: the (M, L) pairs of every split and column into local shared memory, as normalised weights (remote loads in
        // independent batches of 8)
        float *wts = cpart + (size_t)C * RS;  // [C][FM_MAX_CLUSTER]: exp(M_s - M) / L
        if ((int)threadIdx.x < ncol) {
            const int c = threadIdx.x;
            const float *row = cpart + (size_t)c * RS;
            float *w = wts + c * FM_MAX_CLUSTER;
            float M = -INFINITY;
            for (int s0 = 0; s0 < ns; s0 += 8) {
                float ms[8], ls[8];
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    ms[k] = s0 + k < ns ? ld_cluster_f32(row + D, s0 + k) : -INFINITY;
                    ls[k] = s0 + k < ns ? ld_cluster_f32(row + D + 1, s0 + k) : 0.f;
                }
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    M = fmaxf(M, ms[k]);
                    if (s0 + k < ns) { w[s0 + k] = ms[k]; w[FM_MAX_CLUSTER * FM_MAXC + s0 + k] = ls[k]; }  // (second table behind the first)
                }
            }
            float L = 0.f;
            for (int s2 = 0; s2 < ns; s2++) {
                const float f = w[s2] == -INFINITY ? 0.f : __expf(w[s2] - M);
                L += f * w[FM_MAX_CLUSTER * FM_MAXC + s2];
                w[s2] = f;
            }
            const float inv = 1.0f / L;
            for (int s2 = 0; s2 < ns; s2++) w[s2] *= inv;
        }
        __syncthreads();
        // phase 2: this CTA's slice of the head dims, ns independent remote loads per value
        const int dpc = (D + ns - 1) / ns, d_lo = split * dpc;
        const int nd = min(dpc, D - d_lo);
        for (int idx = threadIdx.x; idx < ncol * nd; idx += FM_NW * 32) {
            const int c = idx / nd, d = d_lo + idx % nd;
            const float *src = cpart + (size_t)c * RS + d;
            float a = 0.f;
            for (int s0 = 0; s0 < ns; s0 += 8) {
                float v[8];
#pragma unroll
                for (int k = 0; k < 8; k++) v[k] = s0 + k < ns ? ld_cluster_f32(src, s0 + k) : 0.f;
#pragma unroll
                for (int k = 0; k < 8; k++) a += (s0 + k < ns ? wts[c * FM_MAX_CLUSTER + s0 + k] : 0.f) * v[k];
            }
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

// ------------------------------------------------------------------------------------------------ prompt batches (a12)
// Flash-prefill form of the same arithmetic: a CTA serves 64 query columns of one kv-head (each of its 4 warps owns 16 of them for
// the whole KV range -- no cross-warp merge, no splits) and the 4 warps SHARE every (K, V) tile: one CTA-wide cp.async ring, one
// __syncthreads per tile.  The decode kernel above gives every warp its own tiles instead (right for 1..16 columns); run on a
// 512-token batch it launched 1024 CTAs whose 4 warps saw ~4 tiles each and spent their time in prologue and merge (7B pp512:
// 165 us per layer, 27 % of the step: profiles/r02_launches_pp512_7b.txt).  Causal structure: tile i is skipped by a warp when the
// mask is -inf for all its 16 columns, and not even loaded when that holds for all 64 (the skip words cover the LAST 64 tiles of
// the range: that is where the upper triangle lives).
constexpr int FP_NBUF = 4;   // stages of the CTA ring
constexpr int FP_COLS = 64;  // query columns per CTA

template <int D>
__global__ void __launch_bounds__(FM_NW * 32, 2) fattn_mma_prefill_kernel(const __grid_constant__ FaParams p) {
    using G = MmaGeo<B200_TYPE_F16, D>;
    constexpr int KS = D / 16;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ unsigned long long skip_w[FM_NW];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t4 = lane & 3;
    const int hk = blockIdx.y, ct = blockIdx.x;
    const int col0 = ct * FP_COLS + warp * FM_MAXC;        // this warp's 16 columns
    const int ncol = max(0, min(FM_MAXC, p.n_cols - col0));
    const int n_tiles = (p.n_kv + FM_TP - 1) / FM_TP;
    const unsigned char *kbase = p.k + (size_t)hk * p.k_nb2;
    const unsigned char *vbase = p.v + (size_t)hk * p.v_nb2;

    pdl_wait();
    pdl_trigger();

    // ---- Q fragments, mask rows, slopes of this warp's columns
    unsigned qa[KS][4];
    float slope[2];
    const unsigned char *mrow[2];
#pragma unroll
    for (int hh = 0; hh < 2; hh++) {
        const int c = g + 8 * hh;
        const bool ok = c < ncol;
        const int col = ok ? col0 + c : 0;
        const int n = col / p.gqa, h = hk * p.gqa + col % p.gqa;
        slope[hh] = 1.0f;
        if (p.max_bias > 0.0f) slope[hh] = h < p.n_head_log2 ? powf(p.m0, (float)(h + 1)) : powf(p.m1, (float)(2 * (h - p.n_head_log2) + 1));
        mrow[hh] = p.mask != nullptr ? p.mask + (size_t)n * p.mask_nb1 : nullptr;
        const float *qrow = reinterpret_cast<const float *>(p.q + (size_t)n * p.q_nb1 + (size_t)h * p.q_nb2);
#pragma unroll
        for (int s = 0; s < KS; s++) {
            float2 lo = make_float2(0.f, 0.f), hi = make_float2(0.f, 0.f);
            if (ok) {
                lo = __ldcg(reinterpret_cast<const float2 *>(qrow + 16 * s + 2 * t4));
                hi = __ldcg(reinterpret_cast<const float2 *>(qrow + 16 * s + 8 + 2 * t4));
            }
            qa[s][hh] = pack_h2(lo.x, lo.y);
            qa[s][2 + hh] = pack_h2(hi.x, hi.y);
        }
    }
    // ---- skip words: bit i <-> tile t_skip0 + i is masked for all columns of this warp (lane i checks tile i)
    const int t_skip0 = n_tiles > 64 ? n_tiles - 64 : 0;
    unsigned long long skip = 0;
    if (p.mask != nullptr) {
        const bool al16 = (((uintptr_t)p.mask | p.mask_nb1) & 15) == 0;
#pragma unroll 1
        for (int pass = 0; pass < 2; pass++) {
            const int tt = t_skip0 + pass * 32 + lane;
            bool masked = false;
            if (tt < n_tiles) {
                masked = true;
                const int key = tt * FM_TP;
                for (int c = 0; c < ncol; c++) {
                    const unsigned char *mr = p.mask + (size_t)((col0 + c) / p.gqa) * p.mask_nb1 + 2 * (size_t)key;
                    if (al16 && key + FM_TP <= p.n_kv) {
                        const uint4 u0 = __ldg(reinterpret_cast<const uint4 *>(mr)), u1 = __ldg(reinterpret_cast<const uint4 *>(mr) + 1);
                        masked = masked && (u0.x & u0.y & u0.z & u0.w & u1.x & u1.y & u1.z & u1.w) == 0xfc00fc00u &&
                                 (u0.x | u0.y | u0.z | u0.w | u1.x | u1.y | u1.z | u1.w) == 0xfc00fc00u;
                    } else {
                        for (int e = 0; e < FM_TP; e++)
                            if (key + e < p.n_kv) masked = masked && *reinterpret_cast<const unsigned short *>(mr + 2 * e) == 0xfc00;
                    }
                }
            }
            skip |= (unsigned long long)__ballot_sync(0xffffffffu, masked) << (32 * pass);
        }
    }
    if (ncol == 0) skip = ~0ull;  // a warp without columns never needs a tile
    if (lane == 0) skip_w[warp] = skip;
    __syncthreads();
    const unsigned long long skip_cta = skip_w[0] & skip_w[1] & skip_w[2] & skip_w[3];
    auto is_skipped = [&](unsigned long long w, int tt) { return tt >= t_skip0 && ((w >> (tt - t_skip0)) & 1ull) != 0; };
    // a warp without columns must still not read tiles before t_skip0: give it nothing to do
    const bool idle = ncol == 0;

    // ---- CTA ring: stage = (K tile, V tile); 128 threads copy a tile pair (2 x 16 rows x D/8 chunks of 16 bytes)
    constexpr int CPR = G::ROW_BYTES / 16;
    constexpr int CHUNKS = 2 * FM_TP * CPR;
    auto load_stage = [&](int tt, int st) {
        if (tt < n_tiles && !is_skipped(skip_cta, tt)) {
            unsigned char *base = smem + (size_t)st * 2 * G::TILE_BYTES;
#pragma unroll
            for (int i = threadIdx.x; i < CHUNKS; i += FM_NW * 32) {
                const int kv = i / (FM_TP * CPR), j = i - kv * (FM_TP * CPR), r = j / CPR, c = j - r * CPR;
                const int row = tt * FM_TP + r;
                unsigned char *d = base + kv * G::TILE_BYTES + r * G::ROW_STRIDE + c * 16;
                if (row < p.n_kv) cp_async16(d, (kv ? vbase + (size_t)row * p.v_nb1 : kbase + (size_t)row * p.k_nb1) + c * 16);
                else *reinterpret_cast<uint4 *>(d) = make_uint4(0, 0, 0, 0);
            }
        }
        cp_async_commit();
    };
#pragma unroll
    for (int st = 0; st < FP_NBUF - 1; st++) load_stage(st, st);

    const bool mask_al4 = (((uintptr_t)p.mask | p.mask_nb1) & 3) == 0;
    float scale = p.scale;
    if (p.softcap != 0.0f) scale /= p.softcap;
    float m[2] = {-INFINITY, -INFINITY}, l[2] = {0.f, 0.f};
    float o[2 * KS][4];
#pragma unroll
    for (int i = 0; i < 2 * KS; i++) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    const int mi = lane >> 3, mr = lane & 7;
    const unsigned k_off = (unsigned)(((mi >> 1) * 8 + mr) * G::ROW_STRIDE + (mi & 1) * 16);
    const unsigned v_off = (unsigned)(((mi & 1) * 8 + mr) * G::ROW_STRIDE + (mi >> 1) * 16);

    for (int t = 0; t < n_tiles; t++) {
        cp_async_wait<FP_NBUF - 2>();   // tile t has landed (this thread's copies) ...
        __syncthreads();                // ... everybody's; and everybody is done with tile t - 1, whose stage is refilled now
        load_stage(t + FP_NBUF - 1, (t + FP_NBUF - 1) % FP_NBUF);
        if (idle || is_skipped(skip, t)) continue;
        const unsigned kt = smem_u32(smem + (size_t)(t % FP_NBUF) * 2 * G::TILE_BYTES), vt = kt + G::TILE_BYTES;
        mma_tile_step<D>(kt, vt, t, qa, m, l, o, mrow, slope, scale, p.softcap, p.n_kv, mask_al4, k_off, v_off, t4);
    }
    cp_async_wait<0>();
#pragma unroll
    for (int hh = 0; hh < 2; hh++) {
        l[hh] += __shfl_xor_sync(0xffffffffu, l[hh], 1);
        l[hh] += __shfl_xor_sync(0xffffffffu, l[hh], 2);
    }
#pragma unroll
    for (int hh = 0; hh < 2; hh++) {
        const int c = g + 8 * hh;
        if (c < ncol) {
            const int col = col0 + c, n = col / p.gqa, h = hk * p.gqa + col % p.gqa;
            float *orow = p.dst + ((size_t)n * p.n_head + h) * D;
            const float inv = 1.0f / l[hh];
#pragma unroll
            for (int i = 0; i < 2 * KS; i++)
                *reinterpret_cast<float2 *>(orow + 8 * i + 2 * t4) = make_float2(o[i][2 * hh] * inv, o[i][2 * hh + 1] * inv);
        }
    }
}

}  // namespace b200
