"""GPU parity of the fused launches the graph executor substitutes for node groups: they must equal the unfused kernels
(bit-exact where the arithmetic is the same, else within the activation-quantisation noise)."""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest

import koboldcpp_b200.host as H
import orc as O
from koboldcpp_b200.lib import F16, F32, Q4_0, Q4_K, Q6_K, Q8_0, GemvFuse, GemvJob, RopeParams, Tensor, check, lib, row_size
from koboldcpp_b200.llama import random_blocks

pytestmark = pytest.mark.gpu


def dev(a):
    return H.DeviceBuffer.from_numpy(np.ascontiguousarray(a), pad=256)


@pytest.mark.parametrize("t", [Q4_K, Q6_K])
@pytest.mark.parametrize("K,M,N", [(4096, 512, 1), (2048, 300, 3), (11008, 64, 1)])
def test_gemv_with_norm_prologue_and_residual(t, K, M, N):
    L = lib()
    rng = np.random.default_rng(K + M + t)
    w = random_blocks(rng, t, M, K)
    x = rng.standard_normal((N, K)).astype(np.float32) * 3.0
    nw = (1.0 + 0.2 * rng.standard_normal(K)).astype(np.float32)
    res = rng.standard_normal((N, M)).astype(np.float32)
    eps = 1e-5
    dw, dx, dnw, dres, dout = dev(w), dev(x), dev(nw), dev(res), H.DeviceBuffer(4 * N * M + 256)
    job = (GemvJob * 1)()
    job[0].W, job[0].X, job[0].dst, job[0].M, job[0].id_index = dw.ptr, dx.ptr, dout.ptr, M, -1
    fuse = GemvFuse()
    fuse.norm_weight, fuse.norm_eps = dnw.ptr, eps
    fuse.residual[0] = dres.ptr
    check(L.b200_mul_mat_vec_q_fused(t, job, 1, row_size(t, K), K, M, K, N, C.byref(fuse), None))
    got = dout.to_numpy(np.float32, (N, M))
    # reference: normalise on the host in f32 the way the unfused kernels do, then the exact q8_1 dot of the oracle + residual
    scale = 1.0 / np.sqrt((x.astype(np.float64) ** 2).mean(axis=1, keepdims=True) + eps)
    xn = ((x * scale.astype(np.float32)) * nw).astype(np.float32)
    exp = O.mul_mat_q8_1(t, w, xn, K, M) + res
    err = np.abs(got - exp).max() / np.abs(exp).max()
    # the device scale comes from an f32 tree sum: a last-bit difference can flip a few int8 roundings
    assert err < 2e-3, err
    # residual aliasing the destination (in-place ADD): dst = W.x + dst
    check(L.b200_memcpy_d2d(dout.ptr, dres.ptr, 4 * N * M, None))
    fuse.residual[0] = dout.ptr
    check(L.b200_mul_mat_vec_q_fused(t, job, 1, row_size(t, K), K, M, K, N, C.byref(fuse), None))
    got2 = dout.to_numpy(np.float32, (N, M))
    assert np.array_equal(got, got2)


@pytest.mark.parametrize("t", [Q4_K, Q6_K])
@pytest.mark.parametrize("K,M,norm", [(4096, 11008, True), (2048, 37, False), (4096, 148 * 16 + 5, True), (8192, 3584, False)])
def test_gemv_gate_up_pair_with_silu_epilogue(t, K, M, norm):
    """silu_pair: dst = silu(gate . x) * (up . x) must equal the two-job launch followed by b200_op_silu_mul bit for bit
    (same dot products, same expression), for row counts that leave partial pair groups on the last CTAs."""
    L = lib()
    rng = np.random.default_rng(K + M + t)
    wg, wu = random_blocks(rng, t, M, K), random_blocks(rng, t, M, K)
    x = rng.standard_normal((1, K)).astype(np.float32) * 2.0
    nw = (1.0 + 0.2 * rng.standard_normal(K)).astype(np.float32)
    dg, du, dx, dnw = dev(wg), dev(wu), dev(x), dev(nw)
    og, ou, oref, ofused = (H.DeviceBuffer(4 * M + 256, zero=True) for _ in range(4))
    jobs = (GemvJob * 2)()
    for j, (w, o) in enumerate(((dg, og), (du, ou))):
        jobs[j].W, jobs[j].X, jobs[j].dst, jobs[j].M, jobs[j].id_index = w.ptr, dx.ptr, o.ptr, M, -1
    fuse = GemvFuse()
    if norm:
        fuse.norm_weight, fuse.norm_eps = dnw.ptr, 1e-5
    check(L.b200_mul_mat_vec_q_fused(t, jobs, 2, row_size(t, K), K, M, K, 1, C.byref(fuse), None))
    check(L.b200_op_silu_mul(og.ptr, ou.ptr, None, oref.ptr, M, None))
    exp = oref.to_numpy(np.float32, (M,))
    jobs[0].dst = ofused.ptr
    jobs[1].dst = ofused.ptr
    fuse.silu_pair = 1
    check(L.b200_mul_mat_vec_q_fused(t, jobs, 2, row_size(t, K), K, M, K, 1, C.byref(fuse), None))
    got = ofused.to_numpy(np.float32, (M,))
    assert np.array_equal(got, exp), float(np.abs(got - exp).max())
    # and against the oracle's exact q8_1 dot (no norm: the activation quantisation is then bit-identical)
    if not norm:
        g, u = O.mul_mat_q8_1(t, wg, x, K, M)[0].astype(np.float64), O.mul_mat_q8_1(t, wu, x, K, M)[0].astype(np.float64)
        ref = g / (1.0 + np.exp(-g)) * u
        assert np.abs(got - ref).max() <= 1e-4 * np.abs(ref).max()


@pytest.mark.parametrize("kv_type", [F16, Q8_0])
@pytest.mark.parametrize("D,NH,HK,T", [(128, 32, 32, 1), (128, 32, 8, 3), (64, 32, 4, 5)])
def test_rope_kv_store_equals_separate_nodes(kv_type, D, NH, HK, T):
    L = lib()
    rng = np.random.default_rng(D + NH + T)
    q = rng.standard_normal((T, NH, D)).astype(np.float32)
    k = rng.standard_normal((T, HK, D)).astype(np.float32)
    v = rng.standard_normal((T, HK * D)).astype(np.float32)
    pos = np.arange(37, 37 + T, dtype=np.int32)
    rp = RopeParams(D, 0, 4096, 10000.0, 1.0, 0.0, 1.0, 32.0, 1.0)
    dq, dk, dv, dpos = dev(q), dev(k), dev(v), dev(pos)
    rowb = row_size(kv_type, HK * D)
    outs = []
    for fused in (False, True):
        dqo, dko = H.DeviceBuffer(q.nbytes + 256, zero=True), H.DeviceBuffer(k.nbytes + 256, zero=True)
        kc, vc = H.DeviceBuffer(rowb * T + 256, zero=True), H.DeviceBuffer(rowb * T + 256, zero=True)
        tq, tqo = Tensor.contiguous(dq.ptr, F32, (D, NH, T)), Tensor.contiguous(dqo.ptr, F32, (D, NH, T))
        tk, tko = Tensor.contiguous(dk.ptr, F32, (D, HK, T)), Tensor.contiguous(dko.ptr, F32, (D, HK, T))
        tv = Tensor.contiguous(dv.ptr, F32, (D * HK, T))
        if fused:
            check(L.b200_op_rope_kv_store(tq, tqo, tk, tko, tv, dpos.ptr, None, C.byref(rp), kv_type, kc.ptr, vc.ptr, None, None, None))
        else:
            check(L.b200_op_rope(tq, dpos.ptr, None, tqo, C.byref(rp), None))
            check(L.b200_op_rope(tk, dpos.ptr, None, tko, C.byref(rp), None))
            tkc, tvc = Tensor.contiguous(kc.ptr, kv_type, (D * HK * T,)), Tensor.contiguous(vc.ptr, kv_type, (D * HK * T,))
            tko1 = Tensor.contiguous(dko.ptr, F32, (D * HK * T,))
            tv1 = Tensor.contiguous(dv.ptr, F32, (D * HK * T,))
            check(L.b200_op_cpy(tko1, tkc, None))
            check(L.b200_op_cpy(tv1, tvc, None))
        outs.append((dqo.to_numpy(np.float32, q.shape), dko.to_numpy(np.float32, k.shape), kc.to_numpy(np.uint8, (T, rowb)), vc.to_numpy(np.uint8, (T, rowb))))
    for a, b in zip(*outs):
        assert np.array_equal(a, b)
    # and V against the CPU definition of the cache row (f16 rounding / quantize_row_q8_0_ref)
    exp_v = O.pack_kv(kv_type, v.reshape(1, T, HK * D))[0] if kv_type == F16 else O.quantize_rows(Q8_0, v)
    assert np.array_equal(outs[1][3].reshape(-1), np.asarray(exp_v).view(np.uint8).reshape(-1))


def test_folded_nodes_keep_their_values():
    """The graph executor folds RMS_NORM->MUL into the GEMV prologue, ADD into its epilogue and SILU*up into the pair epilogue.
    The folded nodes are still graph nodes another ggml_backend_sched split may read: with the optional outputs of
    b200_gemv_fuse the launch must store exactly what the unfused kernels would have stored."""
    L = lib()
    K, M = 4096, 1000
    rng = np.random.default_rng(5)
    t = Q4_K
    wg, wu = random_blocks(rng, t, M, K), random_blocks(rng, t, M, K)
    x = rng.standard_normal((1, K)).astype(np.float32) * 2.0
    nw = (1.0 + 0.2 * rng.standard_normal(K)).astype(np.float32)
    res = rng.standard_normal((1, M)).astype(np.float32)
    dg, du, dx, dnw, dres = dev(wg), dev(wu), dev(x), dev(nw), dev(res)
    # unfused reference on the device: rms_norm_mul kernel (stores both values), plain GEMVs, silu_mul kernel
    nrm, nrm_mul = H.DeviceBuffer(4 * K + 256, zero=True), H.DeviceBuffer(4 * K + 256, zero=True)
    tx, tdst = Tensor.contiguous(dx.ptr, F32, (K, 1)), Tensor.contiguous(nrm_mul.ptr, F32, (K, 1))
    check(L.b200_op_rms_norm_mul(tx, dnw.ptr, nrm.ptr, tdst, 1e-5, None))
    og, ou, osilu, opar = (H.DeviceBuffer(4 * M + 256, zero=True) for _ in range(4))
    check(L.b200_mul_mat_vec_q(t, dg.ptr, row_size(t, K), nrm_mul.ptr, K, og.ptr, M, K, M, 1, None))
    check(L.b200_mul_mat_vec_q(t, du.ptr, row_size(t, K), nrm_mul.ptr, K, ou.ptr, M, K, M, 1, None))
    check(L.b200_op_silu_mul(og.ptr, ou.ptr, osilu.ptr, opar.ptr, M, None))
    # fused pair launch with every node value requested
    f_nrm, f_nrm_mul = H.DeviceBuffer(4 * K + 256, zero=True), H.DeviceBuffer(4 * K + 256, zero=True)
    fg, fu, fsilu, fpar = (H.DeviceBuffer(4 * M + 256, zero=True) for _ in range(4))
    jobs = (GemvJob * 2)()
    for j, w in enumerate((dg, du)):
        jobs[j].W, jobs[j].X, jobs[j].dst, jobs[j].M, jobs[j].id_index = w.ptr, dx.ptr, fpar.ptr, M, -1
    fuse = GemvFuse()
    fuse.norm_weight, fuse.norm_eps, fuse.silu_pair = dnw.ptr, 1e-5, 1
    fuse.norm_out, fuse.norm_mul_out, fuse.gate_out, fuse.up_out, fuse.silu_out = f_nrm.ptr, f_nrm_mul.ptr, fg.ptr, fu.ptr, fsilu.ptr
    check(L.b200_mul_mat_vec_q_fused(t, jobs, 2, row_size(t, K), K, M, K, 1, C.byref(fuse), None))
    a_n, b_n = nrm.to_numpy(np.float32, (K,)), f_nrm.to_numpy(np.float32, (K,))
    # the norm scale comes from two different f32 reduction trees (row kernel vs GEMV prologue): compare to f32 rounding
    assert np.abs(a_n - b_n).max() <= 4e-7 * np.abs(a_n).max()
    assert np.abs(nrm_mul.to_numpy(np.float32, (K,)) - f_nrm_mul.to_numpy(np.float32, (K,))).max() <= 4e-7 * np.abs(a_n).max() * 2
    for ref, got in ((og, fg), (ou, fu), (osilu, fsilu), (opar, fpar)):
        r, g = ref.to_numpy(np.float32, (M,)), got.to_numpy(np.float32, (M,))
        assert np.abs(r - g).max() <= 2e-3 * np.abs(r).max()
    # the stored gate / up / silu values are exactly the ones the final product was built from
    g, u, sg, par = (b_.to_numpy(np.float32, (M,)) for b_ in (fg, fu, fsilu, fpar))
    assert np.array_equal(sg * u, par)
    # residual epilogue with the raw product kept
    raw, out = H.DeviceBuffer(4 * M + 256, zero=True), H.DeviceBuffer(4 * M + 256, zero=True)
    job = (GemvJob * 1)()
    job[0].W, job[0].X, job[0].dst, job[0].M, job[0].id_index = dg.ptr, dx.ptr, out.ptr, M, -1
    fuse2 = GemvFuse()
    fuse2.residual[0] = dres.ptr
    fuse2.raw_dst[0] = raw.ptr
    check(L.b200_mul_mat_vec_q_fused(t, job, 1, row_size(t, K), K, M, K, 1, C.byref(fuse2), None))
    plain = H.DeviceBuffer(4 * M + 256, zero=True)
    check(L.b200_mul_mat_vec_q(t, dg.ptr, row_size(t, K), dx.ptr, K, plain.ptr, M, K, M, 1, None))
    r = raw.to_numpy(np.float32, (M,))
    assert np.array_equal(r, plain.to_numpy(np.float32, (M,)))
    assert np.array_equal(out.to_numpy(np.float32, (M,)), r + res[0])


@pytest.mark.parametrize("neox", [0, 2])
def test_rope_f16_in_place_k_shift(neox):
    """ROPE over an f16 cache view in place (build_k_shift, src/llama.cpp:10181-10186): equals the f32 kernel on the
    f16-rounded input, rounded once to f16."""
    L = lib()
    D, HK, T = 128, 4, 37
    rng = np.random.default_rng(3)
    k16 = rng.standard_normal((T, HK, D)).astype(np.float16)
    shift = rng.integers(-20, 20, size=T).astype(np.int32)
    rp = RopeParams(D, neox, 4096, 10000.0, 1.0, 0.0, 1.0, 32.0, 1.0)
    dk16, dpos = dev(k16), dev(shift)
    t16 = Tensor.contiguous(dk16.ptr, F16, (D, HK, T))
    check(L.b200_op_rope(t16, dpos.ptr, None, t16, C.byref(rp), None))
    got = dk16.to_numpy(np.float16, k16.shape)
    k32 = k16.astype(np.float32)
    d32, o32 = dev(k32), H.DeviceBuffer(k32.nbytes + 256, zero=True)
    check(L.b200_op_rope(Tensor.contiguous(d32.ptr, F32, (D, HK, T)), dpos.ptr, None, Tensor.contiguous(o32.ptr, F32, (D, HK, T)), C.byref(rp), None))
    exp = o32.to_numpy(np.float32, k32.shape).astype(np.float16)
    assert np.array_equal(got, exp)


def test_cpy_f32_to_q4_0_is_the_reference_quantiser():
    """KV-cache store with a q4_0 cache (ggml_cpy f32 -> q4_0, src/llama.cpp:9146-9172): bit-exact quantize_row_q4_0_ref, including
    the tie rule (the first value of largest magnitude sets the sign of d) and an all-zero block."""
    L = lib()
    D, HK, T = 128, 4, 19
    rng = np.random.default_rng(9)
    x = rng.standard_normal((T, HK * D)).astype(np.float32)
    x[0, :32] = 0.0                       # d = 0
    x[1, 5] = 3.0; x[1, 20] = -3.0        # tie in magnitude: the first one wins
    x[2, 40] = -2.5; x[2, 33] = 2.5
    dx, out = dev(x), H.DeviceBuffer(T * HK * D // 32 * 18 + 256, zero=True)
    src = Tensor.contiguous(dx.ptr, F32, (D, HK, T))
    dst = Tensor.contiguous(out.ptr, Q4_0, (D, HK, T))
    check(L.b200_op_cpy(src, dst, None))
    got = out.to_numpy(np.uint8, (T, HK * D // 32 * 18))
    exp = np.asarray(O.quantize_rows(Q4_0, x)).view(np.uint8).reshape(T, -1)
    assert np.array_equal(got, exp)


def test_cpy_q8_0_to_f32_is_the_reference_dequantisation():
    """ggml_cast(q8_0 K cache -> f32) of the K-shift graph (src/llama.cpp:10165-10168): bit-exact dequantize_row_q8_0."""
    L = lib()
    D, HK, T = 128, 4, 19
    rng = np.random.default_rng(4)
    x = rng.standard_normal((T, HK * D)).astype(np.float32)
    q = np.asarray(O.quantize_rows(Q8_0, x)).view(np.uint8).reshape(T, -1)
    dq, out = dev(q), H.DeviceBuffer(x.nbytes + 256, zero=True)
    src = Tensor.contiguous(dq.ptr, Q8_0, (D, HK, T))
    dst = Tensor.contiguous(out.ptr, F32, (D, HK, T))
    check(L.b200_op_cpy(src, dst, None))
    got = out.to_numpy(np.float32, x.shape)
    exp = np.stack([O.dequantize(Q8_0, r, HK * D) for r in q])
    assert np.array_equal(got, exp)


@pytest.mark.parametrize("n_expert,n_used,n_tok", [(8, 2, 1), (8, 2, 7), (64, 6, 3), (256, 8, 5), (5, 5, 2)])
def test_moe_router_equals_the_node_chain(n_expert, n_used, n_tok):
    """b200_op_moe_router against soft_max -> argsort(desc)[:k] -> gather -> sum -> div done in numpy (f32)."""
    L = lib()
    rng = np.random.default_rng(n_expert * 31 + n_tok)
    logits = (rng.standard_normal((n_tok, n_expert)) * 2.0).astype(np.float32)
    dl = dev(logits)
    probs, ids = H.DeviceBuffer(4 * n_tok * n_expert + 256, zero=True), H.DeviceBuffer(4 * n_tok * n_expert + 256, zero=True)
    wg, ws, wo = H.DeviceBuffer(4 * n_tok * n_used + 256), H.DeviceBuffer(4 * n_tok + 256), H.DeviceBuffer(4 * n_tok * n_used + 256)
    check(L.b200_op_moe_router(dl.ptr, n_expert, probs.ptr, n_expert, ids.ptr, n_expert, wg.ptr, ws.ptr, wo.ptr, n_expert, n_used, n_tok, None))
    e = np.exp(logits - logits.max(axis=1, keepdims=True))
    p = (e / e.sum(axis=1, keepdims=True)).astype(np.float32)
    order = np.argsort(-p, axis=1, kind="stable")[:, :n_used]
    w = np.take_along_axis(p, order, axis=1)
    got_p = probs.to_numpy(np.float32, (n_tok, n_expert))
    got_ids = ids.to_numpy(np.int32, (n_tok, n_expert))[:, :n_used]
    assert np.allclose(got_p, p, rtol=2e-6, atol=1e-9)
    assert np.array_equal(got_ids, order)
    assert np.array_equal(wg.to_numpy(np.float32, (n_tok, n_used)), np.take_along_axis(got_p, got_ids, axis=1))
    assert np.allclose(ws.to_numpy(np.float32, (n_tok,)), w.sum(axis=1), rtol=1e-6)
    assert np.allclose(wo.to_numpy(np.float32, (n_tok, n_used)), w / w.sum(axis=1, keepdims=True), rtol=1e-5)
