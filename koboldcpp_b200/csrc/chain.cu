// chain.cu -- the persistent decode chain: every dependent phase of a decoded token (GEMVs with their norm / residual / SiLU
// neighbours, ROPE + KV store, flash-decode attention) in ONE cooperative launch (include/b200_ggml.h: b200_decode_chain*).
//
// Replaces the per-node launch sequence of ggml_backend_cuda_graph_compute for batch-1 decode
// (/root/reference/ggml/src/ggml-cuda.cu:2508-2823): quantize_q8_1 + mul_mat_vec_q per matmul (mmvq.cu:50-425), rms_norm, add,
// silu, mul, rope, cpy launches between them and flash_attn_vec_ext + combine (fattn-vec-f16.cuh:8-299, fattn-common.cuh:519-561).
//
// B200 design (what the measurements of the first version said: with 16 warps per SM the kernel is bound by the LATENCY of each
// warp's instruction stream, not by HBM -- 420 warp-instructions per 2.3 KB row at ~9 cycles each):
//   * one CTA per SM for the whole chain; phases separated by grid barriers (a counter in global memory);
//   * a warp streams TWO weight rows per step (consecutive rows of a matrix, or gate row m + up row m) so that the activation
//     registers, the address arithmetic and the pipeline bookkeeping are paid once per pair and the two dot chains interleave;
//   * per-warp byte ring of row steps filled by the TMA engine (cp.async.bulk + mbarrier); the producer cursor of a warp walks
//     ACROSS phase boundaries (weights do not depend on activations), so HBM keeps streaming while a barrier is crossed and the
//     next activation vector is quantised; what a row pair needs (source rows, destination, residual, step sizes) comes from a
//     per-CTA table in shared memory that idle warps build for the phase after next while thread 0 sits in the grid barrier,
//     so the per-pair path is three LDS.128 and no descriptor or address arithmetic;
//   * RMS_NORM folded into the activation quantiser; its sum of squares comes from the producing phase (per-CTA partial sums,
//     summed in a fixed order: deterministic) so the quantiser makes one pass over X;
//   * attention: K/V rows go from HBM straight to registers (the shared memory belongs to the weight ring, which is already
//     holding the wo weights while attention runs).
#include "mmvq_core.cuh"
#include "rope_kv.cuh"

#include <cstdio>

namespace b200 {

constexpr int CH_NW = 15;                 // consumer warps (+ 1 producer warp = 16 warps: 128 registers per thread)
constexpr int CH_THREADS = CH_NW * 32;
constexpr int CH_MAX_SLOT = 8;            // row slots per warp
constexpr int CH_BAR_BYTES = 128 * ((CH_NW * CH_MAX_SLOT * 8 + 127) / 128);
constexpr int CH_SUMSQ_SLOTS = 256;       // >= CTAs of the grid
constexpr int CH_UNIT_CAP = 1 << 20;      // row pairs per CTA and (sub-)phase before a phase is cut (kept for long rows; never hit today)

// ---- the compiled program: one DevPhase per phase, every per-row constant precomputed on the host
struct DevJob {
    const unsigned char *W;
    float       *dst;
    const float *res;
    unsigned stride;      // bytes between rows
    unsigned seg_bytes;   // bytes of 4096 weights of this type
    unsigned tail_bytes;  // bytes of the last step of a row
    int      units;       // row pairs (silu pair: rows)
    int      type;
    int      _pad;
};
struct DevRope {
    const float *q, *k, *v;
    float *q_out, *k_out;
    char *k_cache, *v_cache;
    char *const *k_slot, *const *v_slot;
    const int32_t *pos;
    const float *ff;
    int D, H, HK, kv_type;
    RopeDev rd;
};
struct DevAttn {
    const float *q;
    const unsigned char *k, *v, *mask;
    float *dst;
    int   *tickets;   // [HK]
    float *part;      // [grid][2][gqa][D + 2]
    unsigned k_nb1, k_nb2, v_nb1, v_nb2;
    int D, H, HK, n_kv, kv_type, gqa;
    int tiles_per_head, tiles_q, tiles_rem;  // flattened (kv-head, tile) space cut into one range per CTA
    float scale, max_bias, softcap, m0, m1;
    int n_head_log2;
};
struct DevPhase {
    int kind, n_jobs, K, n_it;
    int silu_pair, units_q, units_rem, unit_lo;  // this (sub-)phase covers units [unit_lo, unit_lo + units_q * grid + units_rem)
    float norm_eps;
    int cont;                                    // continuation of the previous GEMV phase: same X, no barrier in front
    const float *X, *norm_w;
    float *ss_out;
    const float *ss_in;
    union {
        DevJob  jobs[B200_CHAIN_MAX_JOBS];
        DevRope rope;
        DevAttn attn;
    } u;
};

struct ChainParams {
    DevPhase ph[B200_CHAIN_MAX_PHASES];   // used when `dev` is NULL
    const DevPhase *dev;
    int n_phases, ring_bytes, act_bytes, tab_cap;  // tab_cap: row pairs per CTA and phase the tables hold
    unsigned *barrier;
    unsigned long long *trace;  // developer timeline: [CTA][phase % 12][4] globaltimer stamps, or NULL
};

__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define UNIT_TRACE(p, i) do { if (prm.trace && threadIdx.x == 0 && first_unit && (p) < B200_CHAIN_MAX_PHASES) prm.trace[(size_t)gridDim.x * B200_CHAIN_MAX_PHASES * 4 + ((size_t)blockIdx.x * B200_CHAIN_MAX_PHASES + (p)) * 8 + (i)] = gtimer(); } while (0)
#define CHAIN_TRACE(p, i) do { if (prm.trace && threadIdx.x == 0 && (p) < B200_CHAIN_MAX_PHASES) prm.trace[((size_t)blockIdx.x * B200_CHAIN_MAX_PHASES + (p)) * 4 + (i)] = gtimer(); } while (0)

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

// one 4096-weight step of one row against the CTA's activation column 0
__device__ __forceinline__ void slice_dot(int type, const unsigned char *sp, int lane, int it, unsigned char *act_base, int n_it, int nb, float (&acc)[1]) {
    switch (type) {
        case B200_TYPE_Q4_K: SliceDot<B200_TYPE_Q4_K, 1>::run(sp, lane, it, act_base, n_it, nb, acc); break;
        case B200_TYPE_Q5_K: SliceDot<B200_TYPE_Q5_K, 1>::run(sp, lane, it, act_base, n_it, nb, acc); break;
        case B200_TYPE_Q6_K: SliceDot<B200_TYPE_Q6_K, 1>::run(sp, lane, it, act_base, n_it, nb, acc); break;
        case B200_TYPE_Q8_0: SliceDot<B200_TYPE_Q8_0, 1>::run(sp, lane, it, act_base, n_it, nb, acc); break;
        default:             SliceDot<B200_TYPE_Q4_0, 1>::run(sp, lane, it, act_base, n_it, nb, acc); break;
    }
}

// the same step for two rows of one type: one dispatch, and the two dot chains share the activation loads
__device__ __forceinline__ void pair_dot(int type, const unsigned char *spA, const unsigned char *spB, int lane, int it, unsigned char *act_base, int n_it,
                                         int nb, float (&accA)[1], float (&accB)[1]) {
#define B200_PAIR(T) SliceDot<T, 1>::run(spA, lane, it, act_base, n_it, nb, accA); SliceDot<T, 1>::run(spB, lane, it, act_base, n_it, nb, accB); break
    switch (type) {
        case B200_TYPE_Q4_K: B200_PAIR(B200_TYPE_Q4_K);
        case B200_TYPE_Q5_K: B200_PAIR(B200_TYPE_Q5_K);
        case B200_TYPE_Q6_K: B200_PAIR(B200_TYPE_Q6_K);
        case B200_TYPE_Q8_0: B200_PAIR(B200_TYPE_Q8_0);
        default:             B200_PAIR(B200_TYPE_Q4_0);
    }
#undef B200_PAIR
}

// ---- producer -> consumer hand-off, per consumer warp (shared memory)
struct __align__(16) UnitRec {     // what the consumer needs to finish one row pair
    float       *dst;              // pair of outputs (silu pair: one output)
    const float *res;              // residual pair or NULL
    int type, flags;               // flags: 1 = silu pair, 2 = row of jobs[0]
    int _pad[2];
};
constexpr int CH_RECS = 8;         // unit records per consumer warp (>= units in flight)
constexpr int CH_NPROD = 1;        // producer warps: lane = (consumer warp, row of the pair)
constexpr int CH_ALL_THREADS = CH_THREADS + 32 * CH_NPROD;
static_assert(CH_NPROD * 32 >= 2 * CH_NW, "one producer lane per (consumer warp, row)");

struct WarpChan {                  // one per consumer warp
    unsigned long long full[CH_MAX_SLOT];
    unsigned item_off[CH_MAX_SLOT];   // byte offset of the item's first weight byte in the warp's ring
    unsigned item_end[CH_MAX_SLOT];   // end of the item's allocation (producer-private)
    unsigned consumed;                // items the consumer has released
    unsigned _pad[3];
    UnitRec  rec[CH_RECS];
};

__device__ __forceinline__ unsigned ld_acquire_shared(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_shared(unsigned *p, unsigned v) {
    asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(CH_THREADS) : "memory"); }

// FIFO placement in a byte ring of R bytes: the in-use region runs (circularly) from `rd` to `wr`
__device__ __forceinline__ bool ring_place(unsigned wr, unsigned alloc, unsigned rd, bool busy, unsigned R, unsigned &at) {
    at = wr;
    if (busy) {
        if (wr > rd) { if (wr + alloc > R) { if (alloc > rd) return false; at = 0; } }
        else if (wr + alloc > rd) return false;
    } else if (wr + alloc > R) at = 0;
    return true;
}

// The producer side: lane (cw, h) streams row h of every pair of consumer warp cw, across all GEMV phases of the program.
// The loop is CONVERGENT: every pass each lane tests "barriers and ring space for my next step?" and issues it or not; no lane
// ever spins on its own (SIMT would serialise 16 private spin loops and starve the consumers that share the scheduler).
struct ProdState {
    int p, k, it, cnt, n_it, j, base;
    const unsigned char *srcA, *srcB;
    unsigned seg, tail;
    bool done;
};
// position on the unit (p, k): job walk, row addresses, and the unit record for the consumer (written by lane h == 0)
__device__ __forceinline__ void prod_load_unit(ProdState &s, const DevPhase &ph, int bid, WarpChan *ch, unsigned n_units, int h) {
    const int g0 = bid * ph.units_q + (bid < ph.units_rem ? bid : ph.units_rem);
    const int g = ph.unit_lo + g0 + s.k;
    const int silu = ph.silu_pair;
    if (!silu) while (s.j < ph.n_jobs - 1 && g - s.base >= ph.u.jobs[s.j].units) { s.base += ph.u.jobs[s.j].units; s.j++; }
    const DevJob &jb = ph.u.jobs[s.j];
    const int u = g - s.base;
    s.seg = jb.seg_bytes; s.tail = jb.tail_bytes;
    if (silu) { s.srcA = jb.W + (size_t)u * jb.stride; s.srcB = ph.u.jobs[1].W + (size_t)u * jb.stride; }
    else { s.srcA = jb.W + (size_t)(2 * u) * jb.stride; s.srcB = s.srcA + jb.stride; }
    if (h == 0) {
        UnitRec r;
        r.dst = silu ? jb.dst + u : jb.dst + 2 * u;
        r.res = (!silu && jb.res) ? jb.res + 2 * u : nullptr;
        r.type = jb.type; r.flags = (silu ? 1 : 0) | (s.j == 0 ? 2 : 0); r._pad[0] = r._pad[1] = 0;
        ch->rec[n_units & (CH_RECS - 1)] = r;  // (reaches the consumer through the release of the unit's first barrier)
    }
}
// first GEMV phase at or after s.p in which consumer warp cw has a unit
__device__ __forceinline__ void prod_enter_phase(ProdState &s, const DevPhase *phs, int n_phases, int bid, int cw) {
    for (; s.p < n_phases; s.p++) {
        const DevPhase &ph = phs[s.p];
        if (ph.kind != B200_CHAIN_GEMV) continue;
        s.cnt = ph.units_q + (bid < ph.units_rem ? 1 : 0);
        if (cw >= s.cnt) continue;
        s.k = cw; s.it = 0; s.n_it = ph.n_it; s.j = 0; s.base = 0;
        return;
    }
    s.done = true;
}
__device__ __forceinline__ void chain_producer(const DevPhase *phs, int n_phases, WarpChan *chans, unsigned char *rings, unsigned R, int bid, int pl) {
    const int cw = pl >> 1, h = pl & 1;
    WarpChan *ch = chans + (cw < CH_NW ? cw : 0);
    unsigned char *ring = rings + (size_t)cw * R;
    unsigned wr = 0, n_issued = 0, n_units = 0;
    ProdState s;
    s.p = 0; s.done = cw >= CH_NW;
    if (!s.done) prod_enter_phase(s, phs, n_phases, bid, cw);
    if (!s.done) { prod_load_unit(s, phs[s.p], bid, ch, n_units, h); n_units++; }
    for (;;) {
        bool issued = false;
        if (!s.done) {
            const unsigned char *sA = s.srcA + (size_t)s.it * s.seg, *sB = s.srcB + (size_t)s.it * s.seg;
            const unsigned bytes = s.it == s.n_it - 1 ? s.tail : s.seg;
            const unsigned misA = (unsigned)((uintptr_t)sA & 15), misB = (unsigned)((uintptr_t)sB & 15);
            const unsigned lenA = (misA + bytes + 15u) & ~15u, lenB = (misB + bytes + 15u) & ~15u;
            const unsigned allocA = lenA + 16, allocB = lenB + 16;  // + word over-read slack of the unaligned formats
            const unsigned c = ld_acquire_shared(&ch->consumed);
            const unsigned busy = n_issued - c;
            unsigned atA = 0, atB = 0;
            bool ok = busy <= CH_MAX_SLOT - 2;
            if (ok) {
                const unsigned rd = c ? ch->item_end[(c - 1) & (CH_MAX_SLOT - 1)] : 0u;
                ok = ring_place(wr, allocA, rd, busy > 0, R, atA) && ring_place(atA + allocA, allocB, rd, true, R, atB);
            }
            if (ok) {
                const unsigned slot = (n_issued + h) & (CH_MAX_SLOT - 1);
                const unsigned at = h ? atB : atA, mis = h ? misB : misA, len = h ? lenB : lenA;
                ch->item_off[slot] = at + mis;
                ch->item_end[slot] = at + len + 16;
                mbar_expect_tx(&ch->full[slot], len);
                bulk_g2s(ring + at, (h ? sB : sA) - mis, len, &ch->full[slot]);
                wr = atB + allocB;
                n_issued += 2;
                issued = true;
                if (++s.it == s.n_it) {
                    s.it = 0;
                    s.k += CH_NW;
                    if (s.k >= s.cnt) { s.p++; prod_enter_phase(s, phs, n_phases, bid, cw); }
                    if (!s.done) { prod_load_unit(s, phs[s.p], bid, ch, n_units, h); n_units++; }
                }
            }
        }
        __syncwarp();  // also orders the pair's item_end stores before the partner's next read
        if (__all_sync(0xffffffffu, s.done)) break;
        if (!__any_sync(0xffffffffu, issued)) __nanosleep(64);
    }
}

__global__ void __launch_bounds__(CH_ALL_THREADS, 1) chain_kernel(const __grid_constant__ ChainParams prm) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ float red[CH_NW];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, bid = blockIdx.x, nblk = gridDim.x;
    const DevPhase *phs = prm.dev ? prm.dev : prm.ph;
    const int n_phases = prm.n_phases;
    const unsigned R = (unsigned)prm.ring_bytes;  // bytes of one consumer warp's ring
    WarpChan *chans = reinterpret_cast<WarpChan *>(smem);
    unsigned char *act_base = smem + CH_NW * sizeof(WarpChan);
    unsigned char *rings = act_base + prm.act_bytes;
    if (warp < CH_NW && lane == 0) {
        for (int s = 0; s < CH_MAX_SLOT; s++) mbar_init(&chans[warp].full[s], 1);
        chans[warp].consumed = 0;
        mbar_fence_init();
    }
    __syncthreads();

    if (warp >= CH_NW) {  // producer warps: no part in anything below
        chain_producer(phs, n_phases, chans, rings, R, bid, (warp - CH_NW) * 32 + lane);
        return;
    }

    WarpChan *ch = chans + warp;
    unsigned char *my_ring = rings + (size_t)warp * R;
    unsigned n_cons = 0, n_unit = 0;  // items / units this warp has consumed
    unsigned c_par = 0;               // parity bit per barrier
    unsigned n_bar = 0;

    for (int p = 0; p < n_phases; p++) {
        const DevPhase &ph = phs[p];
        const bool gemv = ph.kind == B200_CHAIN_GEMV;
        CHAIN_TRACE(p, 0);
        if (p > 0 && !(gemv && ph.cont)) {  // grid barrier: results of phase p-1 become visible
            consumer_sync();
            if (threadIdx.x == 0) {
                __threadfence();
                atomicAdd(prm.barrier, 1u);
                const unsigned target = (n_bar + 1) * (unsigned)nblk;
                while (ld_acquire_u32(prm.barrier) < target) { }
                __threadfence();
            }
            n_bar++;
            consumer_sync();
        }
        CHAIN_TRACE(p, 1);
        if (ph.kind == B200_CHAIN_ROPE_KV) {
            const DevRope &r = ph.u.rope;
            RopeKvDev a;
            a.q = r.q; a.k = r.k; a.v = r.v; a.q_out = r.q_out; a.k_out = r.k_out; a.k_cache = r.k_cache; a.v_cache = r.v_cache;
            a.k_slot = r.k_slot; a.v_slot = r.v_slot; a.D = r.D; a.H = r.H; a.HK = r.HK; a.T = 1; a.kv_type = r.kv_type;
            const int64_t items = rope_kv_items(a.D, a.H, a.HK, 1);
            for (int64_t t4 = (int64_t)bid * CH_THREADS + threadIdx.x; t4 < items; t4 += (int64_t)nblk * CH_THREADS)
                rope_kv_store_item(a, r.pos, r.ff, r.rd, t4);
            continue;
        }
        if (!gemv) continue;

        // ---- activations: quantise X (RMS-normalised when norm_w) into shared memory
        const int n_it = ph.n_it;
        const int cnt = ph.units_q + (bid < ph.units_rem ? 1 : 0);
        if (cnt > 0 && !ph.cont) {
            float nscale = 1.0f;
            if (ph.norm_w != nullptr) {
                if (ph.ss_in != nullptr) {  // per-CTA partial sums of the producing phase, summed in a fixed order
                    float t = 0.f;
                    for (int i = lane; i < nblk; i += 32) t += __ldcg(ph.ss_in + i);
                    t = warp_sum(t);
                    nscale = rsqrtf(t / (float)ph.K + ph.norm_eps);
                } else {
                    nscale = rms_scale<1, CH_THREADS>(ph.X, ph.K, ph.norm_eps, red);  // (contains consumer barriers)
                }
            }
            quantize_activation_row<CH_THREADS>(act_base, ph.X, ph.K, n_it, ph.norm_w, nscale);
            consumer_sync();
        }
        CHAIN_TRACE(p, 2);
        int last_nblk = ph.K / 32 - 4 * ((n_it - 1) * 32 + lane);
        last_nblk = last_nblk < 0 ? 0 : (last_nblk > 4 ? 4 : last_nblk);
        float ssq = 0.f;  // lane 0: sum of squares of the jobs[0] outputs this warp produced

        for (int k = warp; k < cnt; k += CH_NW) {
            UnitRec rec;
            float2 r2 = make_float2(0.f, 0.f);
            float accA[1] = {0.f}, accB[1] = {0.f};
            for (int it = 0; it < n_it; it++) {
                const int bA = n_cons & (CH_MAX_SLOT - 1), bB = bA + 1;  // items come in pairs: bA is even
                mbar_wait(&ch->full[bA], (c_par >> bA) & 1u);
                if (it == 0) {
                    rec = ch->rec[n_unit & (CH_RECS - 1)];
                    if (lane == 0 && rec.res != nullptr) r2 = __ldcg(reinterpret_cast<const float2 *>(rec.res));  // in flight during the dot
                }
                mbar_wait(&ch->full[bB], (c_par >> bB) & 1u);
                const uint2 off = *reinterpret_cast<const uint2 *>(&ch->item_off[bA]);
                const int nb = it == n_it - 1 ? last_nblk : 4;
                if (nb > 0) pair_dot(rec.type, my_ring + off.x, my_ring + off.y, lane, it, act_base, n_it, nb, accA, accB);
                __syncwarp();
                n_cons += 2;
                if (lane == 0) {
                    fence_proxy_async_smem();  // our generic reads of the items precede the bulk writes that will replace them
                    st_release_shared(&ch->consumed, n_cons);
                }
                c_par ^= 3u << bA;
            }
            n_unit++;
            float va = accA[0], vb = accB[0];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                va += __shfl_xor_sync(0xffffffffu, va, o);
                vb += __shfl_xor_sync(0xffffffffu, vb, o);
            }
            if (lane == 0) {
                if (rec.flags & 1) {
                    rec.dst[0] = (va / (1.0f + __expf(-va))) * vb;
                } else {
                    const float2 o2 = make_float2(va + r2.x, vb + r2.y);
                    *reinterpret_cast<float2 *>(rec.dst) = o2;
                    if (rec.flags & 2) ssq += o2.x * o2.x + o2.y * o2.y;
                }
            }
        }
        if (ph.ss_out != nullptr) {  // deterministic per-CTA partial of sum(out^2)
            consumer_sync();
            if (lane == 0) red[warp] = ssq;
            consumer_sync();
            if (warp == 0) {
                float t = lane < CH_NW ? red[lane] : 0.f;
                t = warp_sum(t);
                if (lane == 0) ph.ss_out[bid] = ph.cont ? ph.ss_out[bid] + t : t;
            }
        }
        CHAIN_TRACE(p, 3);
    }
    // leave the barrier word zero for the next launch: the last CTA to pass resets it
    consumer_sync();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned old = atomicAdd(prm.barrier, 1u);
        if (old == (n_bar + 1) * (unsigned)nblk - 1) atomicExch(prm.barrier, 0u);
    }
}

// ------------------------------------------------------------------------------------------------ host side
static int g_ch_sm = 0;
static size_t g_ch_smem = 0;
static unsigned long long *g_chain_trace = nullptr;

static int chain_props() {
    if (g_ch_sm) return B200_OK;
    int dev = 0, sm = 0, smem = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    B200_CUDA_TRY(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
    B200_CUDA_TRY(cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    if (sm > CH_SUMSQ_SLOTS) sm = CH_SUMSQ_SLOTS;
    g_ch_sm = sm;
    g_ch_smem = (size_t)smem;
    return B200_OK;
}

static bool chain_type_ok(int t) { return t == B200_TYPE_Q4_0 || t == B200_TYPE_Q8_0 || t == B200_TYPE_Q4_K || t == B200_TYPE_Q5_K || t == B200_TYPE_Q6_K; }

// public phase -> DevPhase (validation + every constant the device loop would otherwise derive per row)
static int chain_compile_phase(const b200_chain_phase &in, DevPhase &out, int p, int grid) {
    memset(&out, 0, sizeof(out));
    out.kind = in.kind;
    if (in.kind == B200_CHAIN_GEMV) {
        if (in.n_jobs < 1 || in.n_jobs > B200_CHAIN_MAX_JOBS || in.K <= 0 || !in.X || ((uintptr_t)in.X & 15)) { set_error("b200_decode_chain: phase %d: bad GEMV header", p); return B200_ERR_INVALID; }
        if (in.norm_weight && ((uintptr_t)in.norm_weight & 15)) { set_error("b200_decode_chain: phase %d: norm weight alignment", p); return B200_ERR_UNSUPPORTED; }
        if (in.silu_pair && (in.n_jobs != 2 || in.u.jobs[0].M != in.u.jobs[1].M || in.u.jobs[0].type != in.u.jobs[1].type || in.u.jobs[0].w_row_bytes != in.u.jobs[1].w_row_bytes)) {
            set_error("b200_decode_chain: phase %d: silu_pair needs two jobs of equal M, type and row stride", p);
            return B200_ERR_INVALID;
        }
        out.n_jobs = in.n_jobs; out.K = in.K; out.n_it = (in.K + 4095) / 4096; out.silu_pair = in.silu_pair;
        out.norm_eps = in.norm_eps; out.X = in.X; out.norm_w = in.norm_weight; out.ss_out = in.sumsq_out; out.ss_in = in.norm_weight ? in.sumsq_in : nullptr;
        int64_t units = 0;
        for (int j = 0; j < in.n_jobs; j++) {
            const b200_chain_job &jb = in.u.jobs[j];
            const int t = jb.type;
            if (!chain_type_ok(t)) { set_error("b200_decode_chain: phase %d job %d: weight type %d", p, j, t); return B200_ERR_UNSUPPORTED; }
            const int bv = type_block_values(t);
            if (in.K % bv || !jb.W || !jb.dst || jb.M <= 0 || (jb.w_row_bytes & 1) || ((uintptr_t)jb.W & 1)) { set_error("b200_decode_chain: phase %d job %d: bad job", p, j); return B200_ERR_INVALID; }
            const size_t row_bytes = (size_t)(in.K / bv) * type_block_bytes(t);
            if (jb.w_row_bytes < row_bytes || jb.w_row_bytes > 0xffffffffu) { set_error("b200_decode_chain: phase %d job %d: row stride", p, j); return B200_ERR_INVALID; }
            if (!in.silu_pair && ((jb.M & 1) || ((uintptr_t)jb.dst & 7) || ((uintptr_t)jb.residual & 7))) {
                set_error("b200_decode_chain: phase %d job %d: M must be even and dst / residual 8-byte aligned (rows are streamed in pairs)", p, j);
                return B200_ERR_UNSUPPORTED;
            }
            DevJob &d = out.u.jobs[j];
            d.W = (const unsigned char *)jb.W; d.dst = jb.dst; d.res = jb.residual; d.stride = (unsigned)jb.w_row_bytes;
            d.seg_bytes = (unsigned)(type_slice_bytes(t) * 32);
            d.tail_bytes = (unsigned)(row_bytes - (size_t)(out.n_it - 1) * d.seg_bytes);
            d.units = in.silu_pair ? jb.M : jb.M / 2;
            d.type = t;
            units += d.units;
        }
        if (in.silu_pair) units = in.u.jobs[0].M;
        if (units > 0x3fffffff) { set_error("b200_decode_chain: phase %d: too many rows", p); return B200_ERR_UNSUPPORTED; }
        out.unit_lo = 0;
        out.units_q = (int)(units / grid);
        out.units_rem = (int)(units % grid);
        return B200_OK;
    }
    if (in.kind == B200_CHAIN_ROPE_KV) {
        const auto &r = in.u.rope;
        if (!r.q || !r.k || !r.v || !r.q_out || !r.k_out || !r.pos || (!(r.k_cache || r.k_slot)) || (!(r.v_cache || r.v_slot))) { set_error("b200_decode_chain: phase %d: rope pointers", p); return B200_ERR_INVALID; }
        if ((r.mode & 2) || r.D % 32 || (r.kv_type != B200_TYPE_F16 && r.kv_type != B200_TYPE_Q8_0)) { set_error("b200_decode_chain: phase %d: rope mode / head size / cache type", p); return B200_ERR_UNSUPPORTED; }
        DevRope &d = out.u.rope;
        d.q = r.q; d.k = r.k; d.v = r.v; d.q_out = r.q_out; d.k_out = r.k_out; d.k_cache = (char *)r.k_cache; d.v_cache = (char *)r.v_cache;
        d.k_slot = (char *const *)r.k_slot; d.v_slot = (char *const *)r.v_slot; d.pos = r.pos; d.ff = r.freq_factors;
        d.D = r.D; d.H = r.H; d.HK = r.HK; d.kv_type = r.kv_type;
        // derived constants (ggml_rope_yarn_corr_dims, ggml.c:14262-14270)
        auto corr_dim = [&](float n_rot) { return r.n_dims * logf(r.n_ctx_orig / (n_rot * 2 * 3.14159265358979323846f)) / (2 * logf(r.freq_base)); };
        const float start = floorf(corr_dim(r.beta_fast)), end = ceilf(corr_dim(r.beta_slow));
        d.rd.n_dims = r.n_dims; d.rd.neox = 0; d.rd.freq_scale = r.freq_scale; d.rd.ext_factor = r.ext_factor; d.rd.attn_factor = r.attn_factor;
        d.rd.theta_scale = powf(r.freq_base, -2.0f / r.n_dims);
        d.rd.corr0 = start > 0 ? start : 0;
        d.rd.corr1 = end < r.n_dims - 1 ? end : (float)(r.n_dims - 1);
        return B200_OK;
    }
    set_error("b200_decode_chain: phase %d: unknown kind %d", p, in.kind);
    return B200_ERR_INVALID;
}

// A GEMV phase whose row pairs exceed the per-CTA table is cut into sub-phases that continue each other (no barrier, same X).
// Returns the number of compiled phases written to out (capacity cap), or -1 after set_error.
static int chain_compile_program(const b200_chain_phase *phases, int n_phases, DevPhase *out, int cap, int grid) {
    int n = 0;
    for (int p = 0; p < n_phases; p++) {
        DevPhase d;
        if (chain_compile_phase(phases[p], d, p, grid) != B200_OK) return -1;
        if (d.kind != B200_CHAIN_GEMV) {
            if (n >= cap) { set_error("b200_decode_chain: program too long"); return -1; }
            out[n++] = d;
            continue;
        }
        const int64_t units = (int64_t)d.units_q * grid + d.units_rem;
        const int64_t per = (int64_t)CH_UNIT_CAP * grid;
        const int chunks = (int)((units + per - 1) / per);
        for (int c = 0; c < chunks; c++) {
            if (n >= cap) { set_error("b200_decode_chain: program too long"); return -1; }
            const int64_t lo = units * c / chunks, hi = units * (c + 1) / chunks;
            DevPhase s = d;
            s.unit_lo = (int)lo;
            s.units_q = (int)((hi - lo) / grid);
            s.units_rem = (int)((hi - lo) % grid);
            s.cont = c > 0;
            out[n++] = s;
        }
    }
    return n;
}
static int chain_count_compiled(const b200_chain_phase *phases, int n_phases, int grid) {
    int n = 0;
    for (int p = 0; p < n_phases; p++) {
        if (phases[p].kind != B200_CHAIN_GEMV) { n++; continue; }
        int64_t units = 0;
        for (int j = 0; j < phases[p].n_jobs && j < B200_CHAIN_MAX_JOBS; j++) units += phases[p].silu_pair ? 0 : phases[p].u.jobs[j].M / 2;
        if (phases[p].silu_pair) units = phases[p].u.jobs[0].M;
        const int64_t per = (int64_t)CH_UNIT_CAP * grid;
        n += (int)((units + per - 1) / per > 0 ? (units + per - 1) / per : 1);
    }
    return n;
}

// shared-memory geometry from the compiled phases
static int chain_geometry(const DevPhase *ph, int n, ChainParams &prm) {
    int max_nit = 1, slot = 0, n_gemv = 0;
    for (int p = 0; p < n; p++) {
        if (ph[p].kind != B200_CHAIN_GEMV) continue;
        n_gemv++;
        if (ph[p].n_it > max_nit) max_nit = ph[p].n_it;
        for (int j = 0; j < ph[p].n_jobs; j++) {
            const int sb = (int)((ph[p].u.jobs[j].seg_bytes + 32 + 15) / 16 * 16);  // head misalignment (< 16) + word over-read slack
            if (sb > slot) slot = sb;
        }
    }
    if (n_gemv == 0) { set_error("b200_decode_chain: no GEMV phase"); return B200_ERR_INVALID; }
    prm.tab_cap = 0;
    prm.act_bytes = (int)(((size_t)act_col_bytes(max_nit) + 127) / 128 * 128);
    const size_t budget = g_ch_smem - 256;
    const size_t fixed = CH_NW * sizeof(WarpChan) + (size_t)prm.act_bytes;
    // a step needs its two items in flight whatever the position of the allocation pointer: 3 maximal items per warp
    if (fixed + 3 * (size_t)CH_NW * (slot + 16) > budget) { set_error("b200_decode_chain: activations do not fit in shared memory"); return B200_ERR_UNSUPPORTED; }
    prm.ring_bytes = (int)((budget - fixed) / CH_NW / 16 * 16);
    return B200_OK;
}

static int chain_launch(ChainParams &prm, cudaStream_t st) {
    const size_t smem = CH_NW * sizeof(WarpChan) + (size_t)prm.act_bytes + (size_t)CH_NW * prm.ring_bytes;
    static bool attr_set_dev[16] = {};
    int cur_dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&cur_dev));
    if (!attr_set_dev[cur_dev & 15]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(g_ch_smem - 256)));
        attr_set_dev[cur_dev & 15] = true;
    }
    prm.trace = g_chain_trace;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)g_ch_sm);
    cfg.blockDim = dim3(CH_ALL_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;  // all CTAs co-resident: the grid barrier cannot deadlock
    at[0].val.cooperative = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, chain_kernel, prm));
    return B200_OK;
}

}  // namespace b200

using namespace b200;

// developer hook (not in the public header): device buffer of 148 * B200_CHAIN_MAX_PHASES * 4 u64 time stamps, or NULL
extern "C" __attribute__((visibility("default"))) void b200_decode_chain_trace(unsigned long long *dev) { g_chain_trace = dev; }

extern "C" int b200_decode_chain_sumsq_slots(void) { return CH_SUMSQ_SLOTS; }

extern "C" size_t b200_decode_chain_program_bytes(const b200_chain_phase *phases, int n_phases) {
    if (!phases || n_phases < 1 || chain_props() != B200_OK) return 0;
    return (size_t)chain_count_compiled(phases, n_phases, g_ch_sm) * sizeof(DevPhase) + 64;
}

// blob: [int ring_bytes, tab_cap, act_bytes, n_compiled_phases, pad...(64 bytes)] [DevPhase x n]
extern "C" int b200_decode_chain_compile(const b200_chain_phase *phases, int n_phases, void *program_host) {
    if (!phases || n_phases < 1 || !program_host) { set_error("b200_decode_chain_compile: bad arguments"); return B200_ERR_INVALID; }
    int rc = chain_props();
    if (rc) return rc;
    DevPhase *dp = reinterpret_cast<DevPhase *>((unsigned char *)program_host + 64);
    const int n = chain_compile_program(phases, n_phases, dp, chain_count_compiled(phases, n_phases, g_ch_sm), g_ch_sm);
    if (n < 0) return B200_ERR_INVALID;
    ChainParams prm;
    rc = chain_geometry(dp, n, prm);
    if (rc) return rc;
    int *hdr = (int *)program_host;
    memset(hdr, 0, 64);
    hdr[0] = prm.ring_bytes; hdr[1] = prm.tab_cap; hdr[2] = prm.act_bytes; hdr[3] = n;
    return B200_OK;
}

// program_dev is the device copy of a blob produced by b200_decode_chain_compile, `geometry` the first 16 bytes of the host blob.
extern "C" int b200_decode_chain_run(const void *program_dev, const void *geometry, unsigned *barrier, b200_stream_t stream) {
    if (!program_dev || !geometry || !barrier) { set_error("b200_decode_chain_run: bad arguments"); return B200_ERR_INVALID; }
    int rc = chain_props();
    if (rc) return rc;
    const int *hdr = (const int *)geometry;
    ChainParams prm;
    memset(&prm, 0, sizeof(prm));
    prm.ring_bytes = hdr[0]; prm.tab_cap = hdr[1]; prm.act_bytes = hdr[2]; prm.n_phases = hdr[3];
    prm.dev = reinterpret_cast<const DevPhase *>((const unsigned char *)program_dev + 64);
    prm.barrier = barrier;
    return chain_launch(prm, (cudaStream_t)stream);
}

extern "C" int b200_decode_chain(const b200_chain_phase *phases, int n_phases, unsigned *barrier, b200_stream_t stream) {
    if (!phases || n_phases < 1 || n_phases > B200_CHAIN_MAX_PHASES || !barrier) { set_error("b200_decode_chain: bad arguments"); return B200_ERR_INVALID; }
    int rc = chain_props();
    if (rc) return rc;
    ChainParams prm;
    memset(&prm, 0, sizeof(prm));
    const int n = chain_compile_program(phases, n_phases, prm.ph, B200_CHAIN_MAX_PHASES, g_ch_sm);
    if (n < 0) return B200_ERR_INVALID;
    rc = chain_geometry(prm.ph, n, prm);
    if (rc) return rc;
    prm.n_phases = n;
    prm.barrier = barrier;
    return chain_launch(prm, (cudaStream_t)stream);
}

extern "C" size_t b200_decode_chain_attn_workspace_bytes(int32_t D, int32_t H, int32_t HK) {
    if (D <= 0 || H <= 0 || HK <= 0 || H % HK) return 0;
    return 256 * (((size_t)HK * 4 + 255) / 256) + (size_t)CH_SUMSQ_SLOTS * 2 * (H / HK) * (D + 2) * sizeof(float);
}
