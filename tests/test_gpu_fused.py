"""GPU parity of the fused launches the graph executor substitutes for node groups: they must equal the unfused kernels
(bit-exact where the arithmetic is the same, else within the activation-quantisation noise)."""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest

import koboldcpp_b200.host as H
import orc as O
from koboldcpp_b200.lib import F16, F32, Q4_K, Q6_K, Q8_0, GemvFuse, GemvJob, RopeParams, Tensor, check, lib, row_size
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
