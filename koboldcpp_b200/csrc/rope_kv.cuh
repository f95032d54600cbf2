// rope_kv.cuh -- device code of the fused ROPE(Q) + ROPE(K) + KV-cache store (shared by ops.cu and the decode chain kernel).
#pragma once
#include "b200_common.cuh"

namespace b200 {

struct RopeDev {
    int   n_dims, neox;
    float freq_scale, ext_factor, attn_factor, theta_scale, corr0, corr1;
};
// ------------------------------------------------------------------------------------------------ ROPE + KV store
// One launch for the four nodes llm_build_kv emits per layer (src/llama.cpp:9161-9203, 10504-10523): ROPE(Q), ROPE(K),
// CPY(K -> cache view), CPY(V -> cache view).  A thread owns 4 consecutive values (two rotation pairs of the NORM ordering);
// 8 threads own one 32-value q8_0 block.  Q and K are also written back as f32 (they are graph nodes other ops may read).
struct RopeKvDev {
    const float *q, *k, *v;   // contiguous [D, H, T], [D, HK, T], [D*HK, T]
    float *q_out, *k_out;     // contiguous, same shapes
    char *k_cache, *v_cache;  // destination of the T new rows (used when the slot pointers are NULL)
    char *const *k_slot, *const *v_slot;
    int D, H, HK, T, kv_type;
};
__device__ __forceinline__ void store_kv4(char *base, int kv_type, int64_t e, float4 v, unsigned mask) {
    if (kv_type == B200_TYPE_F16) {
        const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
        uint2 pk;
        pk.x = *reinterpret_cast<const unsigned *>(&a);
        pk.y = *reinterpret_cast<const unsigned *>(&b);
        *reinterpret_cast<uint2 *>(base + e * 2) = pk;
    } else {  // q8_0: the 8 lanes of a block agree on amax (quantize_row_q8_0_ref, ggml-quants.c:849-872)
        float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(mask, amax, o));
        const float d  = __fdiv_rn(amax, 127.0f);
        const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
        const int q0 = (int)roundf(__fmul_rn(v.x, id)) & 0xff, q1 = (int)roundf(__fmul_rn(v.y, id)) & 0xff;
        const int q2 = (int)roundf(__fmul_rn(v.z, id)) & 0xff, q3 = (int)roundf(__fmul_rn(v.w, id)) & 0xff;
        unsigned char *out = reinterpret_cast<unsigned char *>(base) + (e >> 5) * Q8_0_BYTES;
        const int part = (int)((e & 31) >> 2);
        unsigned short *o16 = reinterpret_cast<unsigned short *>(out + 2 + 4 * part);
        o16[0] = (unsigned short)(q0 | (q1 << 8));
        o16[1] = (unsigned short)(q2 | (q3 << 8));
        if (part == 0) *reinterpret_cast<__half *>(out) = __float2half_rn(d);
    }
}
__device__ __forceinline__ float4 rope4(float4 v, int ip, int pos, const float *ff, const RopeDev &rp) {
    float r[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int i = ip + h;
        if (2 * i >= rp.n_dims) continue;
        const float theta_extrap = (float)pos * powf(rp.theta_scale, (float)i) / (ff ? ff[i] : 1.0f);
        float theta = rp.freq_scale * theta_extrap;
        float mscale = rp.attn_factor;
        if (rp.ext_factor != 0.0f) {
            const float y = ((float)i - rp.corr0) / fmaxf(0.001f, rp.corr1 - rp.corr0);
            const float ramp_mix = (1.0f - fminf(1.0f, fmaxf(0.0f, y))) * rp.ext_factor;
            theta = theta * (1.0f - ramp_mix) + theta_extrap * ramp_mix;
            mscale *= 1.0f + 0.1f * logf(1.0f / rp.freq_scale);
        }
        float sn, cs;
        sincosf(theta, &sn, &cs);
        sn *= mscale;
        cs *= mscale;
        const float x0 = r[2 * h], x1 = r[2 * h + 1];
        r[2 * h] = x0 * cs - x1 * sn;
        r[2 * h + 1] = x0 * sn + x1 * cs;
    }
    return make_float4(r[0], r[1], r[2], r[3]);
}
// one float4 work item t4 of the [Q | K | V] index space (regions padded to whole warps)
__device__ __forceinline__ void rope_kv_store_item(const RopeKvDev &a, const int32_t *__restrict__ pos, const float *__restrict__ ff, const RopeDev &rp, int64_t t4) {
    const int64_t nq = (int64_t)a.D * a.H * a.T / 4, nk = (int64_t)a.D * a.HK * a.T / 4;
    // whole warps fall in one region (D*H*T/4 and D*HK*T/4 are multiples of 8; regions are padded to 32 by the launcher)
    const int64_t nq_pad = (nq + 31) / 32 * 32, nk_pad = (nk + 31) / 32 * 32;
    if (t4 < nq_pad) {
        if (t4 >= nq) return;
        const int64_t e = t4 * 4;
        const int tok = (int)(e / ((int64_t)a.D * a.H));
        const float4 v = *reinterpret_cast<const float4 *>(a.q + e);
        *reinterpret_cast<float4 *>(a.q_out + e) = rope4(v, (int)(e % a.D) / 2, pos[tok], ff, rp);
        return;
    }
    char *kc = a.k_slot ? *a.k_slot : a.k_cache, *vc = a.v_slot ? *a.v_slot : a.v_cache;
    if (t4 < nq_pad + nk_pad) {
        const int64_t i = t4 - nq_pad;
        const bool active = i < nk;
        float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
        const int64_t e = i * 4;
        if (active) {
            const int tok = (int)(e / ((int64_t)a.D * a.HK));
            r = rope4(*reinterpret_cast<const float4 *>(a.k + e), (int)(e % a.D) / 2, pos[tok], ff, rp);
            *reinterpret_cast<float4 *>(a.k_out + e) = r;
        }
        // 8-lane groups are wholly active or wholly inactive (nk is a multiple of 8), so the shuffles inside stay in `mask`
        const unsigned mask = __ballot_sync(0xffffffffu, active);
        if (active) store_kv4(kc, a.kv_type, e, r, mask);
        return;
    }
    const int64_t i = t4 - nq_pad - nk_pad;
    const bool active = i < nk;
    const int64_t e = i * 4;
    float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
    if (active) r = *reinterpret_cast<const float4 *>(a.v + e);
    const unsigned mask = __ballot_sync(0xffffffffu, active);
    if (active) store_kv4(vc, a.kv_type, e, r, mask);
}
__host__ __device__ inline int64_t rope_kv_items(int D, int H, int HK, int T) {
    const int64_t nq = (int64_t)D * H * T / 4, nk = (int64_t)D * HK * T / 4;
    return (nq + 31) / 32 * 32 + 2 * ((nk + 31) / 32 * 32);
}

}  // namespace b200
