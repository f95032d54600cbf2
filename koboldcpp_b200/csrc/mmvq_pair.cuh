// mmvq_pair.cuh -- the decode (one activation column) form of the GEMV: every warp step covers the same 4096-weight slice of TWO
// weight rows (rows w + 2j*NW and w + (2j+1)*NW of the CTA's range; in silu_pair mode the gate row and the up row of one output).
// The lane's 128 activations are loaded into registers once per step and contracted with both rows: the two dot chains interleave
// (twice the instruction-level parallelism of one row per step -- the kernel was bound by the latency of each warp's instruction
// stream, 16 warps per SM), and the mbarrier wait, the proxy fence, the ring bookkeeping and the TMA issue are paid once per two
// rows.  Both rows land in one ring stage under one mbarrier (two bulk copies, one expect_tx).  The two row sums are reduced with
// one 6-shuffle butterfly (lanes 0-15 end with row A, lanes 16-31 with row B; same pairing order as warp_sum: bit-identical), and
// the residual values are fetched at the START of the pair so their L2 round trip hides behind the dot.
//
// Requirements (the host dispatches here only when they hold): N == 1, every job reads the same X.
// Included by mmvq.cu only (uses its TRACE macro and MMVQ_MAX_TAB).
#pragma once
#include "mmvq_core.cuh"

namespace b200 {

__device__ __forceinline__ float ld_cg_f32(const float *p) {
    float v;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}

template <int TYPE, bool MULTI>
__global__ void __launch_bounds__(MMVQ_THREADS, 1) mmvq_pair_kernel(const __grid_constant__ MmvqParams p) {
    using G = Geo<TYPE>;
    constexpr bool ALIGNED16 = (TYPE == B200_TYPE_Q4_K || TYPE == B200_TYPE_Q5_K);  // host guarantees 16-byte rows
    constexpr int STAGE2 = 2 * G::STAGE_BYTES;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ float red[MMVQ_NW];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_stage = p.n_stage, n_it = p.n_it;
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(smem);  // [NW][MAX_STAGE]
    unsigned char *act_base  = smem + MMVQ_BAR_BYTES;
    unsigned char *ring      = act_base + ((size_t)act_col_bytes(n_it) + 127) / 128 * 128;
    unsigned char *my_ring   = ring + (size_t)warp * n_stage * STAGE2;
    unsigned long long *my_bars = bars + warp * MMVQ_MAX_STAGE;
    unsigned char *tab_base = ring + (size_t)MMVQ_NW * n_stage * STAGE2;
    const unsigned char **rsrc = reinterpret_cast<const unsigned char **>(tab_base);
    float **rdst = reinterpret_cast<float **>(tab_base + 8 * (size_t)p.tab_rows);
    signed char *rjob = reinterpret_cast<signed char *>(tab_base + 16 * (size_t)p.tab_rows);

    const int bid = blockIdx.x;
    const int r0 = bid * p.rows_q + (bid < p.rows_rem ? bid : p.rows_rem);
    const bool silu = MULTI && p.silu_pair != 0;
    const int np = p.rows_q + (bid < p.rows_rem ? 1 : 0);  // rows (silu: output rows = (gate, up) pairs) of this CTA
    const int nrows = silu ? 2 * MMVQ_NW * ((np + MMVQ_NW - 1) / MMVQ_NW) : np;
    const int nmy = silu ? 2 * ((np - warp + MMVQ_NW - 1) / MMVQ_NW)
                         : (warp < nrows ? (nrows - warp + MMVQ_NW - 1) / MMVQ_NW : 0);  // rows of this warp
    const int npair = (nmy + 1) >> 1;

    if (lane == 0) {
        for (int s = 0; s < n_stage; s++) mbar_init(&my_bars[s], 1);
        mbar_fence_init();
    }
    if constexpr (MULTI) {
        if (p.ids != nullptr) pdl_wait();  // expert ids are written by the producer kernel
        for (int lr = threadIdx.x; lr < nrows; lr += MMVQ_THREADS) {
            if (silu) {
                const int k = lr / MMVQ_NW, w = lr - k * MMVQ_NW, pl = (k >> 1) * MMVQ_NW + w;
                const bool valid = pl < np;
                const int64_t prow = (int64_t)r0 + pl;
                rsrc[lr] = valid ? (const unsigned char *)p.jobs[k & 1].W + (size_t)prow * p.w_row_bytes : nullptr;
                rdst[lr] = p.jobs[0].dst + prow;
                rjob[lr] = 0;
                continue;
            }
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
    }
    pdl_trigger();  // dependents may start their own (independent) prologue as our CTAs retire

    // ---- producer side of the warp-private ring (lane 0 issues): pair ip, step iit -> stage istage
    const unsigned char *w0 = (const unsigned char *)p.jobs[0].W;
    const unsigned tail_bytes = p.row_bytes - (unsigned)(n_it - 1) * G::SEG_BYTES;
    int ip = 0, iit = 0, istage = 0;
    auto issue = [&]() {
        if (lane == 0) {
            const int lrA = warp + 2 * ip * MMVQ_NW, lrB = lrA + MMVQ_NW;
            const bool hasB = 2 * ip + 1 < nmy;
            const unsigned char *sa, *sb;
            if constexpr (MULTI) { sa = rsrc[lrA]; sb = hasB ? rsrc[lrB] : nullptr; }
            else { sa = w0 + (size_t)(r0 + lrA) * p.w_row_bytes; sb = hasB ? w0 + (size_t)(r0 + lrB) * p.w_row_bytes : nullptr; }
            const unsigned bytes = iit == n_it - 1 ? tail_bytes : (unsigned)G::SEG_BYTES;
            unsigned misa = 0, lena = 0, misb = 0, lenb = 0;
            if (sa != nullptr) { sa += (size_t)iit * G::SEG_BYTES; misa = ALIGNED16 ? 0u : (unsigned)((uintptr_t)sa & 15); lena = (misa + bytes + 15u) & ~15u; }
            if (sb != nullptr) { sb += (size_t)iit * G::SEG_BYTES; misb = ALIGNED16 ? 0u : (unsigned)((uintptr_t)sb & 15); lenb = (misb + bytes + 15u) & ~15u; }
            unsigned char *st = my_ring + (size_t)istage * STAGE2;
            if (lena + lenb != 0) {
                mbar_expect_tx(&my_bars[istage], lena + lenb);
                if (lena) bulk_g2s(st, sa - misa, lena, &my_bars[istage]);
                if (lenb) bulk_g2s(st + G::STAGE_BYTES, sb - misb, lenb, &my_bars[istage]);
            } else {
                // nothing to copy (invalid expert ids / padding): complete the phase so the consumer does not hang
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&my_bars[istage])) : "memory");
            }
        }
        istage = istage + 1 == n_stage ? 0 : istage + 1;
        if (++iit == n_it) { iit = 0; ip++; }
    };
    TRACE(1);
    const bool acts_first = p.K <= ACT_RR_MAX && nrows > 0;
    ActRegs aregs;
    // p.prefill stages of the ring are filled BEFORE the wait for the producer kernel (weights do not depend on it)
    int s_fill = 0;
    if (!MULTI || p.ids == nullptr)
        for (; s_fill < p.prefill && s_fill < n_stage && ip < npair; s_fill++) issue();
    if (acts_first) {
        pdl_wait();
        act_issue(aregs, p.jobs[0].X, p.norm_w, p.K);
    }
    for (; s_fill < n_stage && ip < npair; s_fill++) issue();
    TRACE(2);
    if (!acts_first && (!MULTI || p.ids == nullptr)) pdl_wait();
    TRACE(3);

    // lanes past the end of the row in the last step (and, for q4_0/q8_0, a partial last slice)
    int last_nblk = (int)(p.K / 32) - 4 * ((n_it - 1) * 32 + lane);
    last_nblk = last_nblk < 0 ? 0 : (last_nblk > 4 ? 4 : last_nblk);

    if (nrows > 0) {
        // the RMS_NORM / MUL nodes folded into this prologue keep their values: the first CTA stores them once
        float *n_out = bid == 0 ? p.norm_out : nullptr, *nm_out = bid == 0 ? p.norm_mul_out : nullptr;
        if (acts_first) act_finish(aregs, act_base, p.K, n_it, p.norm_w != nullptr, p.norm_eps, red, n_out, nm_out);
        else load_activations<1>(act_base, p.jobs[0].X, p.x_stride, p.K, n_it, p.norm_w, p.norm_eps, red, n_out, nm_out);
        __syncthreads();
        TRACE(4);
    }

    int cstage = 0;
    unsigned cphase = 0;
    const bool isB = lane >= 16;
    for (int j = 0; j < npair; j++) {
        const int lrA = warp + 2 * j * MMVQ_NW, lrB = lrA + MMVQ_NW;
        const bool hasB = 2 * j + 1 < nmy;
        uintptr_t srcA, srcB;
        if constexpr (MULTI) { srcA = (uintptr_t)rsrc[lrA]; srcB = hasB ? (uintptr_t)rsrc[lrB] : 0; }
        else { srcA = (uintptr_t)w0 + (size_t)(r0 + lrA) * p.w_row_bytes; srcB = hasB ? (uintptr_t)w0 + (size_t)(r0 + lrB) * p.w_row_bytes : 0; }
        // this lane's output row (lanes 0-15: row A, lanes 16-31: row B); only lanes 0 and 16 store
        const int lr = isB ? lrB : lrA;
        const bool row_ok = !isB || hasB;
        float *dst = nullptr, *raw = nullptr;
        const float *res = nullptr;
        if (row_ok) {
            if constexpr (MULTI) {
                const int job = rjob[lr];
                dst = rdst[lr];
                res = p.residual[job] ? p.residual[job] + (dst - p.jobs[job].dst) : nullptr;
                raw = p.raw_dst[job] ? p.raw_dst[job] + (dst - p.jobs[job].dst) : nullptr;
            } else {
                dst = p.jobs[0].dst + (r0 + lr);
                res = p.residual[0] ? p.residual[0] + (r0 + lr) : nullptr;
                raw = p.raw_dst[0] ? p.raw_dst[0] + (r0 + lr) : nullptr;
            }
        }
        // the residual value travels while the dot runs
        float resv = 0.f;
        if ((lane & 15) == 0 && res != nullptr) resv = ld_cg_f32(res);

        float accA = 0.f, accB = 0.f;
        for (int it = 0; it < n_it; it++) {
            mbar_wait(&my_bars[cstage], cphase);
            if (j == 0 && it == 0) TRACE(5);
            const int nblk = it == n_it - 1 ? last_nblk : 4;
            const unsigned char *st = my_ring + (size_t)cstage * STAGE2;
            if (nblk > 0 && !(p.dbg & 2)) {
                LaneActs A;
                lane_acts_load<TYPE>(A, act_base, n_it, it, lane);
                const unsigned misA = ALIGNED16 ? 0u : (unsigned)((srcA + (size_t)it * G::SEG_BYTES) & 15);
                const unsigned misB = ALIGNED16 ? 0u : (unsigned)((srcB + (size_t)it * G::SEG_BYTES) & 15);
                if (srcA != 0 && srcB != 0) {
                    RowDot<TYPE>::run(st + misA, lane, A, nblk, accA);
                    RowDot<TYPE>::run(st + G::STAGE_BYTES + misB, lane, A, nblk, accB);
                } else if (srcA != 0) {
                    RowDot<TYPE>::run(st + misA, lane, A, nblk, accA);
                } else if (srcB != 0) {
                    RowDot<TYPE>::run(st + G::STAGE_BYTES + misB, lane, A, nblk, accB);
                }
            }
            __syncwarp();
            if (lane == 0) fence_proxy_async_smem();  // our generic reads of the stage precede the next bulk write to it
            cstage = cstage + 1 == n_stage ? 0 : cstage + 1;
            cphase ^= (cstage == 0);
            if (ip < npair) issue();
        }
        // both row sums in one butterfly: lanes 0-15 end with sum(accA), lanes 16-31 with sum(accB)
        float v = isB ? accB : accA;
        v += __shfl_xor_sync(0xffffffffu, isB ? accA : accB, 16);
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (silu) {
            const float up = __shfl_sync(0xffffffffu, v, 16);
            if (lane == 0 && srcA != 0) {
                const float gate_v = v;
                const int64_t row = dst - p.jobs[0].dst;
                const float sg = gate_v / (1.0f + expf(-gate_v));  // silu_mul_kernel's expression (ops.cu)
                if (p.gate_out) p.gate_out[row] = gate_v;
                if (p.up_out) p.up_out[row] = up;
                if (p.silu_out) p.silu_out[row] = sg;
                dst[0] = sg * up;
            }
        } else if ((lane & 15) == 0 && row_ok) {
            if (MULTI && (isB ? srcB : srcA) == 0) v = __int_as_float(0x7fc00000);
            if (raw) raw[0] = v;
            dst[0] = res ? v + resv : v;
        }
        if (j == 0) TRACE(6);
    }
    __syncthreads();
    TRACE(7);
}

}  // namespace b200
