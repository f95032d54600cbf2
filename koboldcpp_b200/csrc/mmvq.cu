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

#include <cstdio>
#include <cstdlib>

#include "mmvq_core.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------------ the kernel
// CTA c owns the contiguous row range [r0, r1) of the concatenated jobs; warp w owns local rows w, w+NW, ...
// A row is consumed in n_it steps of 4096 weights; step u of a warp lives in ring stage u % n_stage.
//
// MULTI (several jobs, or expert ids): a per-CTA row table in shared memory (source pointer, destination pointer,
// job) replaces index arithmetic, so the streaming loop is identical and prefetch runs across job boundaries.
constexpr int MMVQ_MAX_TAB = 1024;  // rows per CTA in MULTI mode (17 bytes each)

#ifdef B200_MMVQ_TRACE
// developer timeline: [launch slot][CTA][stamp]; the slot (p.dbg >> 8) is a per-launch counter set by the host, so a captured
// graph keeps one slot per launch and the whole chain can be read back after a replay (tools/trace_mmvq.py)
constexpr int TRACE_SLOTS = 256;
__device__ unsigned long long g_mmvq_trace[TRACE_SLOTS * 148 * 8];
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define TRACE(i) do { if (threadIdx.x == 0 && blockIdx.x < 148) g_mmvq_trace[((size_t)(p.dbg >> 8) * 148 + blockIdx.x) * 8 + (i)] = gtime(); } while (0)
static int g_trace_slot = 0;
#define TRACE_SLOT(p) do { (p).dbg |= (g_trace_slot++ % TRACE_SLOTS) << 8; } while (0)
#else
#define TRACE(i) do {} while (0)
#define TRACE_SLOT(p) do {} while (0)
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
    const bool pair = MULTI && NCOLS == 1 && p.silu_pair != 0;
    // pair mode: the CTA owns `np` (gate, up) row pairs starting at pair r0; warp w streams gate row, up row, gate row, ... of the
    // pairs w, w+NW, ...: table slot lr = w + NW * k holds the gate (k even) or up (k odd) row of pair r0 + (k >> 1) * NW + w
    const int np = p.rows_q + (bid < p.rows_rem ? 1 : 0);
    const int nrows = pair ? 2 * MMVQ_NW * ((np + MMVQ_NW - 1) / MMVQ_NW) : np;
    const int nmy = pair ? 2 * ((np - warp + MMVQ_NW - 1) / MMVQ_NW)
                         : (warp < nrows ? (nrows - warp + MMVQ_NW - 1) / MMVQ_NW : 0);  // rows of this warp

    if (lane == 0) {
        for (int s = 0; s < n_stage; s++) mbar_init(&my_bars[s], 1);
        mbar_fence_init();
    }
    int jfirst = 0, jlast = 0;
    if constexpr (MULTI) {
        if (p.ids != nullptr) pdl_wait();  // expert ids are written by the producer kernel
        // row table: every thread resolves a few rows (<= 16 jobs: linear scan)
        for (int lr = threadIdx.x; lr < nrows; lr += MMVQ_THREADS) {
            if (pair) {
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
    // One column (decode) with a row that fits the register file: the activation loads go out in FRONT of the weight ring
    // (see act_issue); the wait for the producer therefore comes first -- with one CTA per SM this kernel only starts on an SM
    // when the producer's CTA there has retired, so little overlap is lost.
    const bool acts_first = NCOLS == 1 && p.K <= ACT_RR_MAX && nrows > 0;
    ActRegs aregs;
    // p.prefill stages of the ring are filled BEFORE the wait for the producer kernel (weights do not depend on it)
    int s_fill = 0;
    if (!MULTI || p.ids == nullptr)
        for (; s_fill < p.prefill && s_fill < n_stage && ik < nmy; s_fill++) issue();
    if (acts_first) {
        pdl_wait();
        if (!(p.dbg & 1)) act_issue(aregs, p.jobs[jfirst].X, p.norm_w, p.K);
    }
    for (; s_fill < n_stage && ik < nmy; s_fill++) issue();
    TRACE(2);

    // ---- the ring fill above is independent of the producer kernel's output
    if (!acts_first && (!MULTI || p.ids == nullptr)) pdl_wait();
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
    float gate_v = 0.f;
    for (int job = jfirst; job <= jlast; job++) {
        const float *xj = p.jobs[job].X;
        if (loaded_x != xj && nrows > 0) {
            if (loaded_x != nullptr) __syncthreads();  // everyone is done with the previous activations
            // the RMS_NORM / MUL nodes folded into this prologue keep their values: the first CTA stores them once
            float *n_out = bid == 0 && loaded_x == nullptr ? p.norm_out : nullptr, *nm_out = bid == 0 && loaded_x == nullptr ? p.norm_mul_out : nullptr;
            if (p.dbg & 1) {}
            else if (acts_first && loaded_x == nullptr) act_finish(aregs, act_base, p.K, n_it, p.norm_w != nullptr, p.norm_eps, red, n_out, nm_out);
            else load_activations<NCOLS>(act_base, xj, p.x_stride, p.K, n_it, p.norm_w, p.norm_eps, red, n_out, nm_out);
            __syncthreads();
            loaded_x = xj;
            TRACE(4);
        }
        for (; k < nmy; k++) {
            const int lr = warp + k * MMVQ_NW;
            uintptr_t src;
            float *dst, *raw;
            const float *res;
            if constexpr (MULTI) {
                if (rjob[lr] != job) break;
                src = (uintptr_t)rsrc[lr];
                dst = rdst[lr];
                res = p.residual[job] ? p.residual[job] + (dst - p.jobs[job].dst) : nullptr;
                raw = p.raw_dst[job] ? p.raw_dst[job] + (dst - p.jobs[job].dst) : nullptr;
            } else {
                src = (uintptr_t)w0 + (size_t)(r0 + lr) * p.w_row_bytes;
                dst = p.jobs[0].dst + (r0 + lr);
                res = p.residual[0] ? p.residual[0] + (r0 + lr) : nullptr;
                raw = p.raw_dst[0] ? p.raw_dst[0] + (r0 + lr) : nullptr;
            }
            for (int it = 0; it < n_it; it++) {
                mbar_wait(&my_bars[cstage], cphase);
                if (k == 0 && it == 0) TRACE(5);
                const int nblk = it == n_it - 1 ? last_nblk : 4;
                if ((!MULTI || src != 0) && nblk > 0 && !(p.dbg & 2)) {
                    const unsigned mis = ALIGNED16 ? 0u : (unsigned)((src + (size_t)it * G::SEG_BYTES) & 15);
                    SliceDot<TYPE, NCOLS>::run(my_ring + (size_t)cstage * G::STAGE_BYTES + mis, lane, it, act_base, n_it, nblk, acc);
                }
                __syncwarp();
                if (lane == 0) fence_proxy_async_smem();  // our generic reads of the stage precede the next bulk write to it
                cstage = cstage + 1 == n_stage ? 0 : cstage + 1;
                cphase ^= (cstage == 0);
                if (ik < nmy) issue();
            }
            if (pair) {  // (NCOLS == 1)
                const float v = warp_sum(acc[0]);
                acc[0] = 0.f;
                if ((k & 1) == 0) gate_v = v;
                else if (lane == 0) {
                    const int64_t row = dst - p.jobs[0].dst;
                    const float sg = gate_v / (1.0f + expf(-gate_v));  // silu_mul_kernel's expression (ops.cu)
                    if (p.gate_out) p.gate_out[row] = gate_v;
                    if (p.up_out) p.up_out[row] = v;
                    if (p.silu_out) p.silu_out[row] = sg;
                    dst[0] = sg * v;
                }
                continue;
            }
#pragma unroll
            for (int n = 0; n < NCOLS; n++) {
                float v = warp_sum(acc[n]);
                if (MULTI && src == 0) v = __int_as_float(0x7fc00000);
                if (lane == 0) {
                    if (raw) raw[(int64_t)n * p.dst_stride] = v;
                    dst[(int64_t)n * p.dst_stride] = res ? v + res[(int64_t)n * p.dst_stride] : v;
                }
                acc[n] = 0.f;
            }
            if (k == 0) TRACE(6);
        }
    }
    __syncthreads();
    TRACE(7);
}

}  // namespace b200
#include "mmvq3.cuh"
namespace b200 {

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
    const int64_t units = p.silu_pair ? p.total_rows / 2 : p.total_rows;  // rows, or (gate, up) pairs
    const int64_t grid = units < g_sm_count ? units : g_sm_count;
    const int64_t rows_per_cta = p.silu_pair ? 2 * MMVQ_NW * (((units + grid - 1) / grid + MMVQ_NW - 1) / MMVQ_NW) : (p.total_rows + grid - 1) / grid;
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
    static const int prefill_pct = getenv("B200_MMVQ_PREFILL") ? atoi(getenv("B200_MMVQ_PREFILL")) : 0;
    p.prefill = pdl ? (int)(ns * prefill_pct / 100) : 0;
    static const int dbg = getenv("B200_MMVQ_DBG") ? atoi(getenv("B200_MMVQ_DBG")) : 0;
    p.dbg = dbg;
    TRACE_SLOT(p);
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
    p.rows_q = (int)(units / grid);
    p.rows_rem = (int)(units % grid);
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

// decode form, lean skeleton (mmvq3.cuh)
template <int TYPE, bool MULTI, int TYPE2 = TYPE>
static int launch_mmvq3(MmvqParams &p, cudaStream_t st, bool pdl) {
    using G3 = Geo3Pair<TYPE, TYPE2>;  // a stage holds a row segment of either type
    int rc = device_props();
    if (rc) return rc;
    if (p.n_it > 8) return -1000;  // K > 32768: fewer than two warps per slice; the one-row kernel handles it
    const bool nit1 = p.n_it == 1;
    const int grp = MMVQ_NW / p.n_it, n_ring = nit1 ? MMVQ_NW : grp * p.n_it;  // warps per K-slice, warps that stream
    const size_t bar_bytes = MMVQ_BAR_BYTES;
    const size_t act_bytes = ((size_t)act_col_bytes(p.n_it) + 127) / 128 * 128;
    const int64_t units = p.silu_pair ? p.total_rows / 2 : p.total_rows;  // rows, or (gate, up) pairs
    const int64_t grid = units < g_sm_count ? units : g_sm_count;
    const int64_t units_per_cta = (units + grid - 1) / grid;
    static const int onerow_below = getenv("B200_MMVQ_ONEROW_BELOW") ? atoi(getenv("B200_MMVQ_ONEROW_BELOW")) : 0;  // developer knob
    if (units_per_cta <= onerow_below) return -1000;
    const int64_t part_rows = nit1 ? 0 : (p.silu_pair ? 2 : 1) * units_per_cta;
    const size_t part_bytes = (size_t)part_rows * p.n_it * 4;
    const size_t cap = g_smem_optin - 256;
    if (bar_bytes + act_bytes + part_bytes + 2 * (size_t)n_ring * G3::STAGE2 > cap) return -1000;  // caller falls back to the one-row kernel
    size_t ns = (cap - bar_bytes - act_bytes - part_bytes) / ((size_t)n_ring * G3::STAGE2);
    // no point in more stages than a warp has pairs
    const int64_t per_warp = (units_per_cta + grp - 1) / grp;
    const int64_t steps = p.silu_pair ? per_warp : (per_warp + 1) / 2;
    if ((int64_t)ns > steps) ns = (size_t)(steps < 1 ? 1 : steps);
    if (ns > (size_t)MMVQ_MAX_STAGE) ns = MMVQ_MAX_STAGE;
    p.n_stage = (int)ns;
    p.grp = grp;
    p.part_rows = (int)part_rows;
    // row segments per warp that go out before griddepcontrol.wait: the activation loads queue behind them on the SM's ingest
    // path, so small launches (whose critical path is the activation phase) send one, large ones two (measured,
    // profiles/r02_gemv_prefill_sweep.log)
    static const int prefill = getenv("B200_MMVQ_PREFILL") ? atoi(getenv("B200_MMVQ_PREFILL")) : -1;
    p.prefill = !pdl ? 0 : prefill >= 0 ? prefill : (nit1 && steps >= 2 ? 2 : 1);
    static const int dbg = getenv("B200_MMVQ_DBG") ? atoi(getenv("B200_MMVQ_DBG")) : 0;  // developer knobs: 1 no activation phase, 2 no dot
    p.dbg = dbg;
    TRACE_SLOT(p);
    const size_t smem = bar_bytes + act_bytes + ns * n_ring * G3::STAGE2 + part_bytes;
    auto kern = nit1 ? mmvq3_kernel<TYPE, TYPE2, MULTI, true> : mmvq3_kernel<TYPE, TYPE2, MULTI, false>;
    static bool attr_set_dev[16][2] = {};  // cudaFuncSetAttribute is per device (one table per template instantiation)
    int cur_dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&cur_dev));
    bool &attr_set = attr_set_dev[cur_dev & 15][nit1 ? 1 : 0];
    if (!attr_set) {
        B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cap));
        attr_set = true;
    }
    p.rows_q = (int)(units / grid);
    p.rows_rem = (int)(units % grid);
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
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, p));
    return B200_OK;
}

template <int TYPE>
static int dispatch_cols(MmvqParams &p, int N, cudaStream_t st, bool pdl) {
    const bool multi = p.n_jobs > 1 || p.ids != nullptr;
    static const bool no_pair = getenv("B200_MMVQ_NO_MMVQ3") != nullptr;
    if (N == 1 && !no_pair) {
        bool same_x = true;
        for (int j = 1; j < p.n_jobs; j++) same_x = same_x && p.jobs[j].X == p.jobs[0].X;
        if (same_x) {
            const int rc = multi ? launch_mmvq3<TYPE, true>(p, st, pdl) : launch_mmvq3<TYPE, false>(p, st, pdl);
            if (rc != -1000) return rc;
        }
    }
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
    if (!(type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0 || type == B200_TYPE_Q4_K || type == B200_TYPE_Q5_K || type == B200_TYPE_Q6_K ||
          type == B200_TYPE_Q4_1 || type == B200_TYPE_Q5_0 || type == B200_TYPE_Q5_1 || type == B200_TYPE_Q2_K || type == B200_TYPE_Q3_K ||
          type == B200_TYPE_IQ4_NL || type == B200_TYPE_IQ4_XS)) {
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
        for (int j = 0; j < n_jobs; j++) { p.residual[j] = fuse->residual[j]; p.raw_dst[j] = fuse->residual[j] ? fuse->raw_dst[j] : nullptr; }
        if (fuse->norm_weight) {
            p.norm_out = fuse->norm_out;
            p.norm_mul_out = fuse->norm_mul_out;
            if (((uintptr_t)p.norm_out | (uintptr_t)p.norm_mul_out) & 15) { set_error("b200_mul_mat_vec_q_fused: norm_out / norm_mul_out must be 16-byte aligned"); return B200_ERR_UNSUPPORTED; }
        }
        if (fuse->silu_pair) {
            if (n_jobs != 2 || N != 1 || ids || jobs[0].M != jobs[1].M || jobs[0].X != jobs[1].X || fuse->residual[0] || fuse->residual[1]) {
                set_error("b200_mul_mat_vec_q_fused: silu_pair needs two jobs (gate, up) of equal M on the same X, N == 1, no residual");
                return B200_ERR_INVALID;
            }
            p.silu_pair = 1;
            p.gate_out = fuse->gate_out; p.up_out = fuse->up_out; p.silu_out = fuse->silu_out;
        }
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
    // jobs of a second weight type (b200_gemv_job.type2): the mixed-type form of the decode kernel
    int type2 = 0;
    unsigned mask = 0;
    for (int j = 0; j < n_jobs; j++) {
        const int tj = jobs[j]._pad;
        if (tj == 0 || tj == type) continue;
        if (type2 != 0 && tj != type2) { set_error("b200_mul_mat_vec_q_multi: at most two weight types per launch"); return B200_ERR_UNSUPPORTED; }
        type2 = tj;
        mask |= 1u << j;
    }
    if (type2 != 0) {
        bool same_x = true;
        for (int j = 1; j < n_jobs; j++) same_x = same_x && jobs[j].X == jobs[0].X;
        if (N != 1 || ids != nullptr || p.silu_pair || !same_x || K % 256) { set_error("b200_mul_mat_vec_q_multi: two weight types need N == 1, one X, no expert ids"); return B200_ERR_UNSUPPORTED; }
        int ta = type, tb = type2;
        if (ta > tb) { const int t = ta; ta = tb; tb = t; mask = ~mask & ((1u << n_jobs) - 1u); }  // canonical order: TYPE < TYPE2
        const size_t rb_a = (size_t)(K / 256) * type_block_bytes(ta), rb_b = (size_t)(K / 256) * type_block_bytes(tb);
        if (ta != type && w_row_bytes != row_bytes) { set_error("b200_mul_mat_vec_q_multi: mixed-type launches need contiguous rows"); return B200_ERR_UNSUPPORTED; }
        p.w_row_bytes = ta == type ? w_row_bytes : rb_a;
        p.row_bytes = (unsigned)rb_a;
        p.row_bytes2 = (unsigned)rb_b;
        p.t2_mask = mask;
        int rc = -1000;
        if (ta == B200_TYPE_Q4_K && tb == B200_TYPE_Q6_K) rc = launch_mmvq3<B200_TYPE_Q4_K, true, B200_TYPE_Q6_K>(p, st, pdl);
        else if (ta == B200_TYPE_Q5_K && tb == B200_TYPE_Q6_K) rc = launch_mmvq3<B200_TYPE_Q5_K, true, B200_TYPE_Q6_K>(p, st, pdl);
        else if (ta == B200_TYPE_Q4_K && tb == B200_TYPE_Q5_K) rc = launch_mmvq3<B200_TYPE_Q4_K, true, B200_TYPE_Q5_K>(p, st, pdl);
        if (rc == -1000) { set_error("b200_mul_mat_vec_q_multi: weight types %d + %d cannot share a launch", ta, tb); return B200_ERR_UNSUPPORTED; }
        return rc;
    }
    switch (type) {
        case B200_TYPE_Q4_0: return dispatch_cols<B200_TYPE_Q4_0>(p, (int)N, st, pdl);
        case B200_TYPE_Q8_0: return dispatch_cols<B200_TYPE_Q8_0>(p, (int)N, st, pdl);
        case B200_TYPE_Q4_K: return dispatch_cols<B200_TYPE_Q4_K>(p, (int)N, st, pdl);
        case B200_TYPE_Q5_K: return dispatch_cols<B200_TYPE_Q5_K>(p, (int)N, st, pdl);
        case B200_TYPE_Q6_K: return dispatch_cols<B200_TYPE_Q6_K>(p, (int)N, st, pdl);
        case B200_TYPE_Q4_1: return dispatch_cols<B200_TYPE_Q4_1>(p, (int)N, st, pdl);
        case B200_TYPE_Q5_0: return dispatch_cols<B200_TYPE_Q5_0>(p, (int)N, st, pdl);
        case B200_TYPE_Q5_1: return dispatch_cols<B200_TYPE_Q5_1>(p, (int)N, st, pdl);
        case B200_TYPE_Q2_K: return dispatch_cols<B200_TYPE_Q2_K>(p, (int)N, st, pdl);
        case B200_TYPE_Q3_K: return dispatch_cols<B200_TYPE_Q3_K>(p, (int)N, st, pdl);
        case B200_TYPE_IQ4_NL: return dispatch_cols<B200_TYPE_IQ4_NL>(p, (int)N, st, pdl);
        case B200_TYPE_IQ4_XS: return dispatch_cols<B200_TYPE_IQ4_XS>(p, (int)N, st, pdl);
    }
    return B200_ERR_UNSUPPORTED;
}


}  // namespace b200

using namespace b200;

#ifdef B200_MMVQ_TRACE
extern "C" __attribute__((visibility("default"))) int b200_mmvq_trace(unsigned long long *out) {
    B200_CUDA_TRY(cudaMemcpyFromSymbol(out, g_mmvq_trace, sizeof(unsigned long long) * TRACE_SLOTS * 148 * 8));
    g_trace_slot = 0;
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
