"""GPU drop-in test: the REFERENCE's own host code (compiled ggml.c / ggml-alloc.c / ggml-backend.cpp in oracle/_ref) builds
the graphs, allocates tensors through our buffer type, uploads with ggml_backend_tensor_set and calls graph_compute on the
B200 backend -- then the reference's ggml_backend_compare_graph_backend (ggml-backend.cpp:2566-2603) evaluates every node on
both the B200 backend and the reference CPU backend and hands the pair to the checker below.

This is the method of upstream's test-backend-ops: normalised mean squared error per node.  Thresholds: 1e-6 for float ops;
2e-3 for the quantised matmuls and attention (the CPU path quantises activations to q8_0 / q8_K, ours to q8_1: SURVEY.md 8c)."""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest

import ggml_ref as G
from koboldcpp_b200.lib import lib as b200lib
from koboldcpp_b200.llama import random_blocks

pytestmark = pytest.mark.gpu

P, I, I64, SZ, F = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_float
TP = C.POINTER(G.Tensor)
EVAL_CB = C.CFUNCTYPE(C.c_bool, C.c_int, TP, TP, C.c_void_p)
OP_NAMES = {26: "MUL_MAT", 27: "MUL_MAT_ID", 60: "FLASH_ATTN_EXT"}
LOOSE_OPS = {26, 27, 60}


def bind():
    R = G.load("portable")
    L = R.lib
    for n, res, args in [
        ("ggml_rms_norm", TP, [P, TP, F]), ("ggml_mul", TP, [P, TP, TP]), ("ggml_add", TP, [P, TP, TP]), ("ggml_silu", TP, [P, TP]),
        ("ggml_reshape_2d", TP, [P, TP, I64, I64]), ("ggml_reshape_3d", TP, [P, TP, I64, I64, I64]), ("ggml_cpy", TP, [P, TP, TP]),
        ("ggml_view_1d", TP, [P, TP, I64, SZ]), ("ggml_rope_ext", TP, [P, TP, TP, TP, I, I, I, F, F, F, F, F, F]),
        ("ggml_cast", TP, [P, TP, I]), ("ggml_get_rows", TP, [P, TP, TP]),
        ("ggml_backend_cpu_init", P, []), ("ggml_backend_free", None, [P]), ("ggml_backend_buffer_free", None, [P]),
        ("ggml_backend_alloc_ctx_tensors", P, [P, P]), ("ggml_backend_tensor_set", None, [TP, P, SZ, SZ]),
        ("ggml_backend_tensor_get", None, [TP, P, SZ, SZ]), ("ggml_backend_graph_compute", I, [P, P]),
        ("ggml_backend_compare_graph_backend", C.c_bool, [P, P, P, EVAL_CB, P]),
        ("ggml_backend_get_default_buffer_type", P, [P]), ("ggml_gallocr_new", P, [P]), ("ggml_gallocr_alloc_graph", C.c_bool, [P, P]),
        ("ggml_gallocr_free", None, [P]), ("ggml_backend_synchronize", None, [P]), ("ggml_nelements", I64, [TP]),
        ("ggml_backend_supports_op", C.c_bool, [P, TP]),
    ]:
        f = getattr(L, n)
        f.restype = res
        f.argtypes = args
    return R, L


def fetch(L, t) -> np.ndarray:
    """Tensor contents as f32 (f32 / f16 / i32 tensors) or raw bytes (quantised)."""
    n = L.ggml_nbytes(t)
    raw = np.empty(n, dtype=np.uint8)
    L.ggml_backend_tensor_get(t, raw.ctypes.data, 0, n)
    ty = t.contents.type
    if ty == G.F32:
        return raw.view(np.float32).astype(np.float64)
    if ty == G.F16:
        return raw.view(np.float16).astype(np.float64)
    if ty == G.I32:
        return raw.view(np.int32).astype(np.float64)
    if ty == G.Q8_0:  # cache rows: compare the values they decode to
        return G.load("portable").dequantize(G.Q8_0, raw, n // 34 * 32).astype(np.float64)
    return raw.astype(np.float64)


def layer_graph(R, L, ctx, rng, T, n_embd=512, n_head=8, n_head_kv=4, n_ff=1024, n_kv=256, kv_type=G.F16):
    """One llama layer exactly as build_llama emits it (flash-attention branch), on tensors of `ctx` (no_alloc)."""
    D = n_embd // n_head
    kvw = D * n_head_kv
    t = {}
    def w(name, ty, ne0, ne1):
        t[name] = (L.ggml_new_tensor_2d(ctx, ty, ne0, ne1), random_blocks(rng, ty, ne1, ne0))
        return t[name][0]
    def f32(name, ne0, ne1, data):
        t[name] = (L.ggml_new_tensor_2d(ctx, G.F32, ne0, ne1), np.ascontiguousarray(data, dtype=np.float32).view(np.uint8))
        return t[name][0]
    x = f32("x", n_embd, T, rng.standard_normal((T, n_embd)))
    an = f32("attn_norm", n_embd, 1, 1 + 0.1 * rng.standard_normal(n_embd))
    fn = f32("ffn_norm", n_embd, 1, 1 + 0.1 * rng.standard_normal(n_embd))
    wq, wk = w("wq", G.Q4_K, n_embd, n_embd), w("wk", G.Q4_K, n_embd, kvw)
    wv, wo = w("wv", G.Q6_K, n_embd, kvw), w("wo", G.Q4_K, n_embd, n_embd)
    wg, wu, wd = w("wg", G.Q4_K, n_embd, n_ff), w("wu", G.Q4_K, n_embd, n_ff), w("wd", G.Q6_K, n_ff, n_embd)
    pos = L.ggml_new_tensor_1d(ctx, G.I32, T)
    t["pos"] = (pos, np.arange(n_kv - T, n_kv, dtype=np.int32).view(np.uint8))
    rs = G.row_size(kv_type, kvw)
    kc, vc = L.ggml_new_tensor_1d(ctx, kv_type, kvw * n_kv), L.ggml_new_tensor_1d(ctx, kv_type, kvw * n_kv)
    kvdata = rng.standard_normal((n_kv, kvw)).astype(np.float32)
    packed = kvdata.astype(np.float16).view(np.uint8) if kv_type == G.F16 else R.quantize(kv_type, kvdata)
    t["kc"], t["vc"] = (kc, packed.reshape(-1)), (vc, packed.reshape(-1).copy())
    mrows = (T + 31) // 32 * 32
    mask = np.full((mrows, n_kv), -np.inf, dtype=np.float32)
    for j in range(T):
        mask[j, : n_kv - T + j + 1] = 0
    mk = L.ggml_new_tensor_2d(ctx, G.F32, n_kv, mrows)
    t["mask"] = (mk, mask.view(np.uint8).reshape(-1))

    cur = L.ggml_mul(ctx, L.ggml_rms_norm(ctx, x, 1e-5), an)
    q, k, v = L.ggml_mul_mat(ctx, wq, cur), L.ggml_mul_mat(ctx, wk, cur), L.ggml_mul_mat(ctx, wv, cur)
    rope = lambda a, nh: L.ggml_rope_ext(ctx, L.ggml_reshape_3d(ctx, a, D, nh, T), pos, None, D, 0, 4096, 10000.0, 1.0, 0.0, 1.0, 32.0, 1.0)
    q, k = rope(q, n_head), rope(k, n_head_kv)
    gf = L.ggml_new_graph(ctx)
    for e in (q, k, v):
        L.ggml_build_forward_expand(gf, e)
    L.ggml_build_forward_expand(gf, L.ggml_cpy(ctx, k, L.ggml_view_1d(ctx, kc, T * kvw, rs * (n_kv - T))))
    L.ggml_build_forward_expand(gf, L.ggml_cpy(ctx, v, L.ggml_view_1d(ctx, vc, T * kvw, rs * (n_kv - T))))
    qp = L.ggml_permute(ctx, q, 0, 2, 1, 3)
    k3 = L.ggml_view_3d(ctx, kc, D, n_kv, n_head_kv, rs, G.row_size(kv_type, D), 0)
    v3 = L.ggml_view_3d(ctx, vc, D, n_kv, n_head_kv, rs, G.row_size(kv_type, D), 0)
    att = L.ggml_flash_attn_ext(ctx, qp, k3, v3, L.ggml_cast(ctx, mk, G.F16), 1.0 / np.sqrt(D), 0.0, 0.0)
    a = L.ggml_mul_mat(ctx, wo, L.ggml_reshape_2d(ctx, att, n_embd, T))
    ffn_inp = L.ggml_add(ctx, a, x)
    h = L.ggml_mul(ctx, L.ggml_rms_norm(ctx, ffn_inp, 1e-5), fn)
    up = L.ggml_mul_mat(ctx, wu, h)
    g = L.ggml_silu(ctx, L.ggml_mul_mat(ctx, wg, h))
    out = L.ggml_add(ctx, L.ggml_mul_mat(ctx, wd, L.ggml_mul(ctx, g, up)), ffn_inp)
    L.ggml_build_forward_expand(gf, out)
    return gf, out, t


@pytest.mark.parametrize("T,kv_type", [(1, G.F16), (1, G.Q8_0), (3, G.F16), (40, G.F16)])
def test_reference_host_drives_b200_backend_node_by_node(T, kv_type):
    if not G.available("portable"):
        pytest.skip("oracle/_ref not built")
    R, L = bind()
    B = b200lib()
    B.ggml_backend_b200_init.restype = P
    B.ggml_backend_b200_init.argtypes = [I]
    be = B.ggml_backend_b200_init(0)
    assert be
    cpu = L.ggml_backend_cpu_init()
    ctx = L.ggml_init(G.InitParams(64 << 20, None, True))  # no_alloc: data lives in backend buffers
    try:
        rng = np.random.default_rng(17 + T)
        gf, out, tensors = layer_graph(R, L, ctx, rng, T, kv_type=kv_type)
        buf = L.ggml_backend_alloc_ctx_tensors(ctx, be)  # -> our buft.alloc_buffer / get_alloc_size / buffer.init_tensor
        assert buf
        for name, (tt, data) in tensors.items():
            assert data.nbytes == L.ggml_nbytes(tt), name
            L.ggml_backend_tensor_set(tt, data.ctypes.data, 0, data.nbytes)  # -> our buffer.set_tensor
        worst = {}

        def cb(i, t1, t2, _):
            assert L.ggml_backend_supports_op(be, t1), f"node {i} op {t1.contents.op} would fall back to the CPU"
            a, b = fetch(L, t1), fetch(L, t2)
            fin = np.isfinite(b)
            assert np.array_equal(np.isfinite(a), fin), f"node {i}: non-finite mismatch"
            nmse = float(((a[fin] - b[fin]) ** 2).sum() / max((b[fin] ** 2).sum(), 1e-30))
            worst[i] = (t1.contents.op, nmse, False)
            return True
        cbf = EVAL_CB(cb)
        ok = L.ggml_backend_compare_graph_backend(be, cpu, gf, cbf, None)
        assert ok
        assert len(worst) >= 20
        after_matmul = False  # every backend feeds a node its OWN earlier results: downstream of a matmul the inputs differ
        for i in sorted(worst):
            op, e, is_bytes = worst[i]
            after_matmul |= op in LOOSE_OPS
            lim = 0.08 if is_bytes else (2e-3 if after_matmul else 1e-6)
            assert e < lim, f"node {i} op {op}: {e}"
        L.ggml_backend_buffer_free(buf)
    finally:
        L.ggml_free(ctx)
        L.ggml_backend_free(cpu)
        L.ggml_backend_free(be)


@pytest.mark.parametrize("T", [1, 2])
def test_whole_graph_under_ggml_alloc_matches_cpu(T):
    """The same layer allocated by the reference's graph allocator (ggml_gallocr: tensors share and reuse memory, ops run
    in place) and computed in ONE graph_compute call -- the executor's node fusion and CUDA-graph capture must survive real
    ggml-alloc aliasing.  Compared with ggml_graph_compute_with_ctx on the reference CPU backend."""
    if not G.available("portable"):
        pytest.skip("oracle/_ref not built")
    R, L = bind()
    B = b200lib()
    B.ggml_backend_b200_init.restype = P
    B.ggml_backend_b200_init.argtypes = [I]
    be = B.ggml_backend_b200_init(0)
    ctx = L.ggml_init(G.InitParams(64 << 20, None, True))
    cctx = L.ggml_init(G.InitParams(512 << 20, None, False))
    try:
        gf, out, tensors = layer_graph(R, L, ctx, np.random.default_rng(5), T)
        ga = L.ggml_gallocr_new(L.ggml_backend_get_default_buffer_type(be))
        assert L.ggml_gallocr_alloc_graph(ga, gf)
        for name, (tt, data) in tensors.items():
            L.ggml_backend_tensor_set(tt, data.ctypes.data, 0, data.nbytes)
        for _ in range(3):  # first call captures, the next ones replay the CUDA graph (KV rows are rewritten identically)
            for name, (tt, data) in tensors.items():  # gallocr also places the leafs and may reuse their bytes after the last use
                L.ggml_backend_tensor_set(tt, data.ctypes.data, 0, data.nbytes)
            assert L.ggml_backend_graph_compute(be, gf) == 0
        got = fetch(L, out)
        cgf, cout, ctens = layer_graph(R, L, cctx, np.random.default_rng(5), T)
        for name, (tt, data) in ctens.items():
            C.memmove(tt.contents.data, data.ctypes.data, data.nbytes)
        assert L.ggml_graph_compute_with_ctx(cctx, cgf, 4) == 0
        exp = np.ctypeslib.as_array(C.cast(cout.contents.data, C.POINTER(C.c_float)), (got.size,)).astype(np.float64)
        nmse = float(((got - exp) ** 2).sum() / (exp ** 2).sum())
        assert nmse < 2e-3, nmse
        L.ggml_gallocr_free(ga)
    finally:
        L.ggml_free(ctx)
        L.ggml_free(cctx)
        L.ggml_backend_free(be)


@pytest.mark.parametrize("wt,N", [(G.Q4_K, 1), (G.Q6_K, 3), (G.Q8_0, 20), (G.Q4_0, 5)])
def test_broadcast_mul_mat_with_quantised_weights(wt, N):
    """3-D / 4-D MUL_MAT with quantised weights (ggml.c:12486-12560; the reference loops i02 / i03 in ggml-cuda.cu:1358-1658):
    weights [K, M, 2, 1], activations [K, N, 4, 3] -- every weight slice serves r2 x r3 = 2 x 3 activation slices.  Evaluated on the
    B200 backend and on the reference CPU backend by ggml_backend_compare_graph_backend."""
    if not G.available("portable"):
        pytest.skip("oracle/_ref not built")
    R, L = bind()
    L.ggml_new_tensor_4d.restype = TP
    B = b200lib()
    B.ggml_backend_b200_init.restype = P
    B.ggml_backend_b200_init.argtypes = [I]
    be = B.ggml_backend_b200_init(0)
    cpu = L.ggml_backend_cpu_init()
    ctx = L.ggml_init(G.InitParams(16 << 20, None, True))
    try:
        rng = np.random.default_rng(3 + N)
        K, M = 512, 96
        w = L.ggml_new_tensor_4d(ctx, wt, K, M, 2, 1)
        x = L.ggml_new_tensor_4d(ctx, G.F32, K, N, 4, 3)
        out = L.ggml_mul_mat(ctx, w, x)
        gf = L.ggml_new_graph(ctx)
        L.ggml_build_forward_expand(gf, out)
        buf = L.ggml_backend_alloc_ctx_tensors(ctx, be)
        assert buf
        wd = random_blocks(rng, wt, 2 * M, K)
        xd = rng.standard_normal((3, 4, N, K)).astype(np.float32)
        L.ggml_backend_tensor_set(w, wd.ctypes.data, 0, wd.nbytes)
        L.ggml_backend_tensor_set(x, xd.ctypes.data, 0, xd.nbytes)
        seen = []

        def cb(i, t1, t2, _):
            assert L.ggml_backend_supports_op(be, t1), "the broadcast MUL_MAT would fall back to the CPU"
            a, b = fetch(L, t1), fetch(L, t2)
            seen.append(float(((a - b) ** 2).sum() / max((b ** 2).sum(), 1e-30)))
            return True
        cbf = EVAL_CB(cb)
        assert L.ggml_backend_compare_graph_backend(be, cpu, gf, cbf, None)
        assert len(seen) == 1 and seen[0] < 2e-3, seen
        # and against fp64 on the dequantised weights: slice (i12, i13) uses weight slice i12 // 2
        got = np.empty((3, 4, N, M), dtype=np.float32)
        L.ggml_backend_graph_compute(be, gf)
        L.ggml_backend_tensor_get(out, got.ctypes.data, 0, got.nbytes)
        wf = np.stack([R.dequantize(wt, r, K) for r in wd]).astype(np.float64).reshape(2, M, K)
        for i13 in range(3):
            for i12 in range(4):
                exp = xd[i13, i12].astype(np.float64) @ wf[i12 // 2].T
                assert np.linalg.norm(got[i13, i12] - exp) <= 2e-2 * np.linalg.norm(exp)
        L.ggml_backend_buffer_free(buf)
    finally:
        L.ggml_free(ctx)
        L.ggml_backend_free(cpu)
        L.ggml_backend_free(be)
