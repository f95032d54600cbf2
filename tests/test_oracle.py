"""CPU tests: the oracle (oracle/ggml_oracle.c) against the reference's committed golden vectors and,
when oracle/_ref exists, against the reference itself (method of gguf-py/tests/test_quants.py:144-230)."""
import numpy as np
import pytest

import ggml_ref as R
import orc as O

WT = R.ORACLE_WEIGHT_TYPES
NAME = R.TYPE_NAMES


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


# ---------------------------------------------------------------- against committed golden vectors
@pytest.mark.parametrize("t", WT, ids=[NAME[t] for t in WT])
def test_dequant_bit_exact_golden(golden, t):
    blocks, want = golden[f"w_{NAME[t]}_blocks"], golden[f"w_{NAME[t]}_dequant"]
    got = np.stack([O.dequantize(t, b, want.shape[1]) for b in blocks])
    assert np.array_equal(bits(got), bits(want))  # includes raw random-byte blocks (NaN/Inf scales compare by bits)


@pytest.mark.parametrize("t", (R.Q8_0, R.Q8_1, R.Q8_K), ids=["q8_0", "q8_1", "q8_K"])
def test_activation_quant_bit_exact_golden(golden, t):
    want = golden[f"act_{NAME[t]}"]
    got = np.stack([O.quantize(t, a) for a in golden["act_src"]])
    if t == R.Q8_K:
        # all-zero rows: the reference leaves bsums unspecified (ggml-quants.c:3800-3805); compare d and qs only
        got, want = got.reshape(len(got), -1, 292)[:, :, :260], want.reshape(len(want), -1, 292)[:, :, :260]
    assert np.array_equal(got, want)


def test_q4_0_weight_quant_bit_exact_golden(golden):
    got = np.stack([O.quantize(R.Q4_0, r) for r in golden["w_q4_0_src"]])
    assert np.array_equal(got, golden["w_q4_0_blocks"][:4])


@pytest.mark.parametrize("t", WT, ids=[NAME[t] for t in WT])
def test_vec_dot_and_mul_mat_golden(golden, t):
    w, x = golden[f"mm_{NAME[t]}_w"], golden["mm_x"]
    K, M = x.shape[1], w.shape[0]
    out = O.mul_mat(t, w, x, K, M, nthreads=2)
    assert np.array_equal(bits(out), bits(golden[f"mm_{NAME[t]}_out"]))  # scalar reference: same summation order
    at = R.act_type(t)
    a = O.quantize(at, x[0])
    vd = np.array([O.vec_dot(t, w[i], a, K) for i in range(M)], dtype=np.float32)
    assert np.array_equal(bits(vd), bits(golden[f"vd_{NAME[t]}"]))


@pytest.mark.parametrize("t", WT, ids=[NAME[t] for t in WT])
def test_q8_1_contraction_close_to_cpu_and_fp64(golden, t):
    """The GPU-path arithmetic (q8_1 activations) vs the CPU path (q8_0/q8_K) vs fp64 dequant(W)@x.
    This measures the tolerance the GPU parity tests use."""
    w, x = golden[f"mm_{NAME[t]}_w"], golden["mm_x"]
    K, M = x.shape[1], w.shape[0]
    wd = np.stack([O.dequantize(t, r, K) for r in w]).astype(np.float64)
    exact = x.astype(np.float64) @ wd.T
    cpu = golden[f"mm_{NAME[t]}_out"].astype(np.float64)
    gpu = O.mul_mat_q8_1(t, w, x, K, M)
    nrm = np.linalg.norm(exact)
    err_cpu, err_gpu = np.linalg.norm(cpu - exact) / nrm, np.linalg.norm(gpu - exact) / nrm
    assert err_cpu < 2e-2 and err_gpu < 2e-2
    assert err_gpu < 2.5 * err_cpu + 1e-3  # 32-wide blocks are at least about as accurate as 256-wide q8_K


@pytest.mark.parametrize("t", (R.Q4_K, R.Q6_K, R.Q8_0), ids=["q4_K", "q6_K", "q8_0"])
def test_mul_mat_id_golden(golden, t):
    w, ids = golden[f"mmid_{NAME[t]}_w"], golden["mmid_ids"]
    for tag in ("1", "2"):
        b = golden[f"mmid_b{tag}"]
        out = O.mul_mat_id(t, w, ids, b, b.shape[2], 16, 4, nthreads=1)
        assert np.array_equal(bits(out), bits(golden[f"mmid_{NAME[t]}_out{tag}"]))


def test_mul_mat_id_rejects_bad_expert(golden):
    w = golden["mmid_q8_0_w"]
    b = golden["mmid_b1"]
    with pytest.raises(AssertionError):
        O.mul_mat_id(R.Q8_0, w, np.array([[0, 4]], dtype=np.int32), b[:1], b.shape[2], 16, 4)


@pytest.mark.parametrize("N", (1, 5))
@pytest.mark.parametrize("tkv", (R.F16, R.Q8_0, R.Q4_0), ids=["f16", "q8_0", "q4_0"])
def test_flash_attn_golden(golden, tkv, N):
    q, mask = golden[f"fa_q_N{N}"], golden[f"fa_mask_N{N}"]
    k, v = golden[f"fa_k_{NAME[tkv]}"], golden[f"fa_v_{NAME[tkv]}"]
    D = q.shape[2]
    out = O.flash_attn_ext(q, k, v, mask, tkv, 1.0 / np.sqrt(D), nthreads=2)
    assert np.array_equal(bits(out), bits(golden[f"fa_out_{NAME[tkv]}_N{N}"]))
    f64 = O.flash_attn_f64(q, k, v, mask, tkv, 1.0 / np.sqrt(D))
    assert np.abs(out - f64).max() < 5e-3  # f16 V accumulation noise of the CPU path


@pytest.mark.parametrize("N", (1, 5))
def test_flash_attn_softcap_alibi_golden(golden, N):
    q, mask = golden[f"fa_q_N{N}"], golden[f"fa_mask_N{N}"]
    out = O.flash_attn_ext(q, golden["fa_k_f16"], golden["fa_v_f16"], mask, R.F16, 1.0 / np.sqrt(q.shape[2]), 8.0, 30.0)
    want = golden[f"fa_out_f16_N{N}_softcap_alibi"]
    assert np.abs(out - want).max() <= 2e-6  # powf/tanhf come from libm on both sides; allow 1-ulp libm differences


def test_fp16_roundtrip_all_values():
    L = O.lib()
    h = np.arange(65536, dtype=np.uint16)
    f = h.view(np.float16).astype(np.float32)
    for i in range(0, 65536, 97):
        got = L.orc_fp16_to_fp32(int(h[i]))
        assert (np.isnan(got) and np.isnan(f[i])) or np.float32(got).view(np.uint32) == f[i].view(np.uint32)
    rng = np.random.default_rng(3)
    xs = np.r_[rng.standard_normal(4000) * 10.0 ** rng.integers(-9, 6, 4000), [0.0, -0.0, 65504, 65519.9, 65520, 1e-8, 5.96e-8, 2.98e-8, 2.9802322e-8]]
    for x in xs.astype(np.float32):
        assert L.orc_fp32_to_fp16(float(x)) == int(np.float32(x).astype(np.float16).view(np.uint16)), x


# ---------------------------------------------------------------- against the reference itself (when built)
needs_ref = pytest.mark.skipif(not R.available("scalar"), reason="oracle/_ref not built (needs /root/reference)")


@needs_ref
@pytest.mark.parametrize("variant", ("scalar", "portable"))
@pytest.mark.parametrize("t", WT, ids=[NAME[t] for t in WT])
def test_oracle_vs_reference_fresh_data(t, variant):
    ref = R.load(variant)
    rng = np.random.default_rng(100 + t)
    K, M, N = 2048, 12, 2
    x = (rng.standard_normal((M, K)) * rng.uniform(0.01, 2.0)).astype(np.float32)
    w = ref.quantize(t, x)
    raw = rng.integers(0, 256, size=w.shape, dtype=np.uint8)
    for blk in (w, raw):
        a = np.stack([ref.dequantize(t, r, K) for r in blk]); b = np.stack([O.dequantize(t, r, K) for r in blk])
        assert np.array_equal(bits(a), bits(b))
    act = rng.standard_normal((N, K)).astype(np.float32)
    for at in (R.Q8_0, R.Q8_1, R.Q8_K):
        for r in act:
            assert np.array_equal(ref.quantize_act_ref(at, r), O.quantize(at, r))
    got, want = O.mul_mat(t, w, act, K, M), ref.mul_mat(t, w, act, K, M, nthreads=2)
    if variant == "scalar":
        assert np.array_equal(bits(got), bits(want))
    else:  # SIMD summation order differs
        assert np.abs(got - want).max() <= 1e-5 * max(1.0, np.abs(want).max())


@needs_ref
@pytest.mark.parametrize("tkv", (R.F16, R.Q8_0, R.Q4_0), ids=["f16", "q8_0", "q4_0"])
def test_oracle_flash_attn_vs_reference_fresh_data(tkv):
    ref = R.load("scalar")
    rng = np.random.default_rng(7)
    nh, nhkv, D, nkv, N = 4, 4, 64, 96, 2
    q = rng.standard_normal((nh, N, D)).astype(np.float32)
    k = O.pack_kv(tkv, rng.standard_normal((nhkv, nkv, D))); v = O.pack_kv(tkv, rng.standard_normal((nhkv, nkv, D)))
    for mask in (None, np.where(rng.random((32, nkv)) < 0.3, -np.inf, 0).astype(np.float16)):
        a = ref.flash_attn_ext(q, k, v, mask, tkv, 0.125); b = O.flash_attn_ext(q, k, v, mask, tkv, 0.125)
        assert np.array_equal(bits(a), bits(b))
