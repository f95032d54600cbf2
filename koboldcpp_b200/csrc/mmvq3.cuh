// mmvq3.cuh -- the decode (one activation column) GEMV: lean skeleton around the per-type row dots of mmvq_core.cuh.
//
// Why a third skeleton (measured on B200, profiles/r02_gemv_chain_hot_cold_dbg.log + the ncu source page of the pair kernel): a
// chain of decode GEMVs ran as fast from L2 as from HBM -- the launches were bound by their own instruction stream, not by memory:
// ~1030 warp instructions of prologue per warp (a per-CTA row table, ring fills issued serially by lane 0 at ~67 instructions
// per copy pair, eight predicated activation loads whatever K is) and ~400 per (two rows x 4096 weights) step, of which 128 are
// the dp4a / nibble-mask pairs that do the work.  This kernel keeps the arithmetic (RowDot<TYPE>: results are bit-identical)
// and removes the rest:
//   * no row table: a row's source / destination is a handful of integer operations (jobs scanned in constant memory), done
//     by the two lanes that need it;
//   * the ring is filled by many lanes at once (lane s / 16+s issue the two rows of stage s), refills by lanes 0 and 16;
//   * a stage is exactly two row segments (no slack for 16-byte aligned types): 3 stages of 2 x 2304 bytes per warp for Q4_K;
//   * K <= 4096: the lane's 128 activations are loaded into registers ONCE per kernel (NIT1), not once per step;
//   * both rows of a step complete on one mbarrier (two arrivals, one per issuing lane).
// Requirements (host dispatch): N == 1, every job reads the same X.  Included by mmvq.cu only.
#pragma once
#include "mmvq_core.cuh"

namespace b200 {

__device__ __forceinline__ float ld_cg_f32(const float *p) {
    float v;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

template <int TYPE>
struct Geo3 {
    static constexpr bool ALIGNED16 = (TYPE == B200_TYPE_Q4_K || TYPE == B200_TYPE_Q5_K);  // host guarantees 16-byte rows
    static constexpr int  SEG       = Geo<TYPE>::SEG_BYTES;
    static constexpr int  STAGE     = ALIGNED16 ? SEG : Geo<TYPE>::STAGE_BYTES;  // one row segment (+ misalignment slack)
    static constexpr int  STAGE2    = 2 * STAGE;
};

template <int TYPE, int TYPE2>
struct Geo3Pair {
    static constexpr int STAGE2 = Geo3<TYPE>::STAGE2 > Geo3<TYPE2>::STAGE2 ? Geo3<TYPE>::STAGE2 : Geo3<TYPE2>::STAGE2;
};

struct RowRef {
    const unsigned char *src;  // first byte of the weight row; NULL: no such row, or an expert id out of range
    float               *dst;  // NULL: no such row
    int                  job;
    int                  t2;   // mixed-type launch: the row belongs to a job of the second weight type
};

// every field a row dot of either type may read
template <int TYPE, int TYPE2>
__device__ __forceinline__ void lane_acts_load2(LaneActs &A, unsigned char *act_base, int n_it, int it, int lane) {
    if constexpr (TYPE == TYPE2) {
        lane_acts_load<TYPE>(A, act_base, n_it, it, lane);
    } else {
        const ActView a = act_view(act_base, n_it, 0);
#pragma unroll
        for (int j = 0; j < 8; j++) A.u[j] = a.q[((it * 8 + j) * 32) + lane];
        A.d8 = a.d[it * 32 + lane];
        A.s8 = a.s[it * 32 + lane];
        A.g = a.g[it * 32 + lane];
    }
}

// TYPE2 != TYPE: a multi-job launch whose jobs are of two weight types (attn_q / attn_k Q4_K + attn_v Q6_K of the *_K_M mixes): bit j
// of p.t2_mask says job j is TYPE2; its rows are contiguous (row stride = row size).
template <int TYPE, int TYPE2, bool MULTI, bool NIT1>
__global__ void __launch_bounds__(MMVQ_THREADS, 1) mmvq3_kernel(const __grid_constant__ MmvqParams p) {
    using G3 = Geo3<TYPE>;
    using G3b = Geo3<TYPE2>;
    constexpr bool MIXED = TYPE != TYPE2;
    static_assert(!MIXED || MULTI, "two weight types need a multi-job launch");
    constexpr bool ALIGNED16 = G3::ALIGNED16 && G3b::ALIGNED16;
    constexpr int SEG = G3::SEG, SEGb = G3b::SEG;
    constexpr int STAGE = G3::STAGE > G3b::STAGE ? G3::STAGE : G3b::STAGE, STAGE2 = 2 * STAGE;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ float red[MMVQ_NW];

    TRACE(0);
    if (!(p.dbg & 4)) pdl_trigger();  // one CTA per SM: a dependent CTA can only become resident when this one retires
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = lane >> 4;  // lanes 0-15 work for row A of a pair, lanes 16-31 for row B
    const int n_stage = p.n_stage, n_it = NIT1 ? 1 : p.n_it;
    // K > 4096: the warps form n_it groups of g; a group owns one 4096-weight slice of K for every row of the CTA (so a lane's
    // activations stay in registers here too) and the per-slice row sums meet in shared memory
    const int g   = NIT1 ? MMVQ_NW : p.grp;
    const int wit = NIT1 ? 0 : warp / g;         // this warp's slice of K
    const int wg  = NIT1 ? warp : warp - wit * g;  // its index inside the group
    const bool active = NIT1 || wit < n_it;
    unsigned long long *my_bars = reinterpret_cast<unsigned long long *>(smem) + warp * MMVQ_MAX_STAGE;
    unsigned char *act_base = smem + MMVQ_BAR_BYTES;
    unsigned char *ring     = act_base + ((size_t)act_col_bytes(n_it) + 127) / 128 * 128;
    unsigned char *my_ring  = ring + (size_t)warp * n_stage * STAGE2;
    float *part = reinterpret_cast<float *>(ring + (size_t)(NIT1 ? MMVQ_NW : g * n_it) * n_stage * STAGE2);  // [n_it][p.part_rows]

    const int bid = blockIdx.x;
    const int r0 = bid * p.rows_q + (bid < p.rows_rem ? bid : p.rows_rem);
    const int np = p.rows_q + (bid < p.rows_rem ? 1 : 0);  // rows (silu: output rows = (gate, up) pairs) of this CTA
    const bool silu = MULTI && p.silu_pair != 0;
    // pair j of this warp: rows wg + 2 g j (A) and wg + 2 g j + g (B) of the CTA; silu: gate and up row of output wg + g j
    int npair = 0;
    if (active && np > wg) npair = silu ? (np - wg + g - 1) / g : ((np - wg + g - 1) / g + 1) >> 1;

    if (lane < n_stage) mbar_init(&my_bars[lane], 2);
    mbar_fence_init();
    __syncwarp();

    // local row lr of the CTA (silu: output lr, hh = 0 gate / 1 up)
    auto resolve_row = [&](int lr, int hh) -> RowRef {
        RowRef r = {nullptr, nullptr, 0, 0};
        if (lr >= np) return r;
        int64_t row = (int64_t)r0 + lr;
        if (silu) {
            r.src = (const unsigned char *)p.jobs[hh].W + (size_t)row * p.w_row_bytes;
            r.dst = p.jobs[0].dst + row;
            return r;
        }
        int jb = 0;
        if constexpr (MULTI) {
            while (row >= p.jobs[jb].M) { row -= p.jobs[jb].M; jb++; }
        }
        const unsigned char *w = (const unsigned char *)p.jobs[jb].W;
        if constexpr (MULTI) {
            if (p.ids != nullptr && p.jobs[jb].id_index >= 0) {
                const int e = p.ids[p.jobs[jb].id_index];
                w = (e >= 0 && e < p.n_expert) ? w + (size_t)e * p.expert_bytes : nullptr;
            }
        }
        if constexpr (MIXED) r.t2 = (int)((p.t2_mask >> jb) & 1u);
        r.src = w != nullptr ? w + (size_t)row * (MIXED && r.t2 ? (size_t)p.row_bytes2 : p.w_row_bytes) : nullptr;
        r.dst = p.jobs[jb].dst + row;
        r.job = jb;
        return r;
    };
    // this lane's row (half h) of pair j
    auto resolve = [&](int j) -> RowRef { return silu ? resolve_row(wg + g * j, h) : resolve_row(wg + 2 * g * j + g * h, 0); };
    const unsigned seg_bytes = wit == n_it - 1 ? p.row_bytes - (unsigned)(n_it - 1) * SEG : (unsigned)SEG;
    const unsigned seg_bytes2 = wit == n_it - 1 ? p.row_bytes2 - (unsigned)(n_it - 1) * SEGb : (unsigned)SEGb;
    // this lane copies the warp's segment of its row of pair j into stage s (or just arrives when there is nothing to copy)
    auto issue_unit = [&](int j, int s) {
        const RowRef r = resolve(j);
        unsigned long long *bar = &my_bars[s];
        if (r.src != nullptr) {
            const bool b2 = MIXED && r.t2;
            const unsigned char *src = r.src + (size_t)wit * (b2 ? SEGb : SEG);
            // aligned-down source, length rounded up to 16: the slack lands in the stage's head / tail padding
            const unsigned mis = ALIGNED16 ? 0u : (unsigned)((uintptr_t)src & 15);
            const unsigned len = (mis + (b2 ? seg_bytes2 : seg_bytes) + 15u) & ~15u;
            mbar_expect_tx(bar, len);
            bulk_g2s(my_ring + (size_t)s * STAGE2 + h * STAGE, src - mis, len, bar);
        } else {
            mbar_arrive(bar);
        }
    };

    // ---- ring fill: lane s (row A) and lane 16 + s (row B) own stage s; p.prefill row segments per warp go out before the wait for the
    // producer kernel (weights do not depend on it), the rest behind the activation loads (see act_issue)
    const int fs = lane & 15;
    const bool filler = fs < n_stage && fs < npair;
    const bool ids_first = MULTI && p.ids != nullptr;  // expert ids are written by the producer kernel
    if (ids_first) pdl_wait();
    if (filler && 2 * fs + h < p.prefill) issue_unit(fs, fs);
    TRACE(1);
    constexpr int RR = NIT1 ? 2 : ACT_RR;  // float4 of the activation row per thread
    const bool acts_first = (NIT1 || p.K <= ACT_RR_MAX) && np > 0;
    ActRegsT<RR> aregs;
    if (!ids_first) pdl_wait();
    TRACE(2);
    if (acts_first && !(p.dbg & 1)) act_issue(aregs, p.jobs[0].X, p.norm_w, p.K);
    if (filler && 2 * fs + h >= p.prefill) issue_unit(fs, fs);
    if (p.dbg & 4) pdl_trigger();  // developer knob: late trigger
    TRACE(3);
    // K-sliced form: the residual of output row threadIdx.x travels from now until the epilogue
    float eres = 0.f;
    if constexpr (!NIT1) {
        if (!silu && (int)threadIdx.x < np) {
            const RowRef rr = resolve_row(threadIdx.x, 0);
            if (p.residual[rr.job]) eres = ld_cg_f32(p.residual[rr.job] + (rr.dst - p.jobs[rr.job].dst));
        }
    }

    // lanes past the end of the row in the last slice (and, for q4_0/q8_0, a partial last 128-weight slice)
    int nblk = 4;
    if (wit == n_it - 1) {
        nblk = (int)(p.K / 32) - 4 * ((n_it - 1) * 32 + lane);
        nblk = nblk < 0 ? 0 : (nblk > 4 ? 4 : nblk);
    }

    LaneActs A;
    if (np > 0) {
        // the RMS_NORM / MUL nodes folded into this prologue keep their values: the first CTA stores them once
        float *n_out = bid == 0 ? p.norm_out : nullptr, *nm_out = bid == 0 ? p.norm_mul_out : nullptr;
        if (p.dbg & 1) {}
        else if (acts_first) act_finish(aregs, act_base, p.K, n_it, p.norm_w != nullptr, p.norm_eps, red, n_out, nm_out);
        else if (p.norm_w == nullptr) {
            // a row above ACT_RR_MAX values (Llama-3-70B ffn_down: K = 28672) goes through the registers in chunks
            for (int64_t k0 = 0; k0 < p.K; k0 += ACT_RR_MAX) {
                const int64_t kc = p.K - k0 < ACT_RR_MAX ? p.K - k0 : ACT_RR_MAX;
                act_issue(aregs, p.jobs[0].X + k0, nullptr, kc);
                act_finish(aregs, act_base, kc, n_it, false, 0.f, red, nullptr, nullptr, (int)(k0 / 4096));
            }
        }
        else load_activations<1>(act_base, p.jobs[0].X, p.x_stride, p.K, n_it, p.norm_w, p.norm_eps, red, n_out, nm_out);
        __syncthreads();
        if (active) lane_acts_load2<TYPE, TYPE2>(A, act_base, n_it, wit, lane);  // once per kernel
        TRACE(4);
    }

    int ij = n_stage;  // next pair to issue (lanes 0 and 16): n_stage pairs are in flight
    int cstage = 0;
    unsigned cphase = 0;
    for (int j = 0; j < npair; j++) {
        const RowRef r = resolve(j);
        const unsigned okmask = __ballot_sync(0xffffffffu, r.src != nullptr);
        const bool vA = (okmask & 1u) != 0, vB = (okmask & 0x10000u) != 0;
        unsigned misA = 0, misB = 0;
        if constexpr (!ALIGNED16) {  // SEG is a multiple of 16: the misalignment is the same for every segment of a row
            const unsigned lo = (unsigned)(uintptr_t)r.src & 15u;
            misA = __shfl_sync(0xffffffffu, lo, 0);
            misB = __shfl_sync(0xffffffffu, lo, 16);
        }
        // the residual value travels while the dot runs
        const float *res = nullptr;
        float *raw = nullptr;
        float resv = 0.f;
        if (NIT1 && (lane & 15) == 0 && r.dst != nullptr) {
            const int64_t o = r.dst - p.jobs[r.job].dst;
            res = p.residual[r.job] ? p.residual[r.job] + o : nullptr;
            raw = p.raw_dst[r.job] ? p.raw_dst[r.job] + o : nullptr;
            if (res != nullptr) resv = ld_cg_f32(res);
        }
        float accA = 0.f, accB = 0.f;
        mbar_wait(&my_bars[cstage], cphase);
        if (j == 0) TRACE(5);
        const unsigned char *st = my_ring + (size_t)cstage * STAGE2;
        unsigned tmask = 0;  // mixed-type launch: which of the two rows are of the second type (all lanes vote: nblk is per lane)
        if constexpr (MIXED) tmask = __ballot_sync(0xffffffffu, r.t2 != 0);
        if (nblk > 0 && !(p.dbg & 2)) {
            if constexpr (!MIXED) {
                if (vA && vB) {
                    RowDot<TYPE>::run(st + misA, lane, A, nblk, accA);
                    RowDot<TYPE>::run(st + STAGE + misB, lane, A, nblk, accB);
                } else if (vA) {
                    RowDot<TYPE>::run(st + misA, lane, A, nblk, accA);
                } else if (vB) {
                    RowDot<TYPE>::run(st + STAGE + misB, lane, A, nblk, accB);
                }
            } else {
                const bool tA = (tmask & 1u) != 0, tB = (tmask & 0x10000u) != 0;
                if (vA && vB && !tA && !tB) {
                    RowDot<TYPE>::run(st + misA, lane, A, nblk, accA);
                    RowDot<TYPE>::run(st + STAGE + misB, lane, A, nblk, accB);
                } else if (vA && vB && tA && tB) {
                    RowDot<TYPE2>::run(st + misA, lane, A, nblk, accA);
                    RowDot<TYPE2>::run(st + STAGE + misB, lane, A, nblk, accB);
                } else {  // a pair across the boundary of the two types, or half a pair
                    if (vA) { if (tA) RowDot<TYPE2>::run(st + misA, lane, A, nblk, accA); else RowDot<TYPE>::run(st + misA, lane, A, nblk, accA); }
                    if (vB) { if (tB) RowDot<TYPE2>::run(st + STAGE + misB, lane, A, nblk, accB); else RowDot<TYPE>::run(st + STAGE + misB, lane, A, nblk, accB); }
                }
            }
        }
        __syncwarp();
        if (ij < npair) {
            if ((lane & 15) == 0) {
                fence_proxy_async_smem();  // the warp's generic reads of the stage precede the next bulk write to it
                issue_unit(ij, cstage);
            }
            ij++;
        }
        cstage = cstage + 1 == n_stage ? 0 : cstage + 1;
        cphase ^= (cstage == 0);
        // both row sums in one butterfly: lanes 0-15 end with sum(accA), lanes 16-31 with sum(accB) (same pairing order as
        // warp_sum)
        float v = h ? accB : accA;
        v += __shfl_xor_sync(0xffffffffu, h ? accA : accB, 16);
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if constexpr (!NIT1) {
            // this slice's share of the row: the epilogue adds the slices in slice order
            if ((lane & 15) == 0 && r.dst != nullptr)
                part[(size_t)wit * p.part_rows + (silu ? 2 * (wg + g * j) + h : wg + 2 * g * j + g * h)] = v;
        } else if (silu) {
            const float up = __shfl_sync(0xffffffffu, v, 16);
            if (lane == 0 && r.dst != nullptr) {
                const float gate_v = v;
                const int64_t row = r.dst - p.jobs[0].dst;
                const float sg = gate_v / (1.0f + expf(-gate_v));  // silu_mul_kernel's expression (ops.cu)
                if (p.gate_out) p.gate_out[row] = gate_v;
                if (p.up_out) p.up_out[row] = up;
                if (p.silu_out) p.silu_out[row] = sg;
                r.dst[0] = sg * up;
            }
        } else if ((lane & 15) == 0 && r.dst != nullptr) {
            if (MULTI && r.src == nullptr) v = __int_as_float(0x7fc00000);
            if (raw) raw[0] = v;
            r.dst[0] = res ? v + resv : v;
        }
        if (j == 0) TRACE(6);
    }
    if constexpr (!NIT1) {
        __syncthreads();
        for (int t = threadIdx.x; t < np; t += MMVQ_THREADS) {
            const RowRef r = resolve_row(t, 0);
            if (silu) {
                float gate_v = 0.f, up = 0.f;
                for (int it = 0; it < n_it; it++) {
                    gate_v += part[(size_t)it * p.part_rows + 2 * t];
                    up += part[(size_t)it * p.part_rows + 2 * t + 1];
                }
                const int64_t row = r.dst - p.jobs[0].dst;
                const float sg = gate_v / (1.0f + expf(-gate_v));
                if (p.gate_out) p.gate_out[row] = gate_v;
                if (p.up_out) p.up_out[row] = up;
                if (p.silu_out) p.silu_out[row] = sg;
                r.dst[0] = sg * up;
            } else {
                float v = 0.f;
                for (int it = 0; it < n_it; it++) v += part[(size_t)it * p.part_rows + t];
                if (MULTI && r.src == nullptr) v = __int_as_float(0x7fc00000);
                const int64_t o = r.dst - p.jobs[r.job].dst;
                if (p.raw_dst[r.job]) p.raw_dst[r.job][o] = v;
                if (p.residual[r.job]) v += t < MMVQ_THREADS ? eres : ld_cg_f32(p.residual[r.job] + o);
                r.dst[0] = v;
            }
        }
    }
    TRACE(7);
}

}  // namespace b200
