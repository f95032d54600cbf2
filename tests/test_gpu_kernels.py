"""GPU parity tests: the CUDA path (through the C ABI, host buffers in/out) against the oracle on identical
seeded inputs, against the reference's committed golden vectors, and -- at BASELINE.json's full sizes --
through size-independent properties (linearity, row permutation, split invariance).

Bars: bit-exact for block dequantisation and q8_1 quantisation; for MUL_MAT / MUL_MAT_ID /
FLASH_ATTN_EXT a stated tolerance against (1) the integer-exact q8_1 contraction, (2) the CPU backend's result
and (3) fp64 ground truth.
"""
import numpy as np
import pytest

import ggml_ref as R
import orc as O

pytestmark = pytest.mark.gpu

WT = R.WEIGHT_TYPES
NAME = R.TYPE_NAMES

# tolerances (measured headroom: see DESIGN.md "Parity")
TOL_Q81 = 3e-5     # |gpu - exact q8_1 contraction| / max|ref|: only fp32 summation order differs
TOL_CPU_NMSE = 2e-2  # ||gpu - cpu|| / ||cpu||: q8_1 (32-wide) vs q8_0/q8_K (CPU) activation rounding
TOL_FA = {1: 2.5e-3, 8: 5e-3, 2: 5e-3}  # max abs error of attention outputs vs fp64 (outputs are O(0.3)); q8_0 KV also rounds Q to int8
                                # (the CPU reference itself is 2.7e-3 / 1.3e-3 away from fp64 on the golden case)


@pytest.fixture(scope="module")
def H():
    import koboldcpp_b200.host as host
    host.require_gpu()
    return host


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def same_f32(a, b):
    """Bit-identical, except that NaNs only have to be NaNs (x86 propagates NaN payloads, the GPU canonicalises them)."""
    a, b = np.ascontiguousarray(a, dtype=np.float32), np.ascontiguousarray(b, dtype=np.float32)
    na, nb = np.isnan(a), np.isnan(b)
    return np.array_equal(na, nb) and np.array_equal(bits(a)[~na], bits(b)[~nb])


def quantize_weights(t, x):
    """Weights as the reference quantiser produces them when oracle/_ref is present, else random valid blocks."""
    if R.available("portable"):
        return R.load("portable").quantize(t, x)
    return random_blocks(t, x.shape[0], x.shape[1], np.random.default_rng(x.shape[0] * 7 + t))


def random_blocks(t, M, K, rng):
    """Random but well-formed blocks: random quant bytes, small finite fp16 scales."""
    bs, by = R.BLOCK[t]
    nb = K // bs
    raw = rng.integers(0, 256, size=(M, nb, by), dtype=np.uint8)

    def put_half(off, lo, hi):
        v = rng.uniform(lo, hi, size=(M, nb)).astype(np.float16).view(np.uint16)
        raw[:, :, off] = v & 0xff
        raw[:, :, off + 1] = v >> 8

    if t in (R.Q4_0, R.Q8_0, R.Q5_0):
        put_half(0, -0.02, 0.02)
    elif t in (R.Q4_1, R.Q5_1):
        put_half(0, -0.02, 0.02); put_half(2, -0.1, 0.1)
    elif t in (R.Q4_K, R.Q5_K):
        put_half(0, 0.0005, 0.002); put_half(2, 0.0005, 0.002)
    elif t == R.Q6_K:
        put_half(208, -0.001, 0.001)
    elif t == R.Q2_K:
        put_half(80, 0.001, 0.004); put_half(82, 0.001, 0.004)
    elif t == R.Q3_K:
        put_half(108, -0.002, 0.002)
    elif t == R.IQ4_NL:
        put_half(0, -0.002, 0.002)
    elif t == R.IQ4_XS:
        put_half(0, -0.0001, 0.0001)
    return raw.reshape(M, nb * by)


# ------------------------------------------------------------------------------ a2: dequantisation (bit-exact)
@pytest.mark.parametrize("t", WT, ids=[NAME[t] for t in WT])
def test_dequantize_bit_exact(H, golden, t):
    blocks, want = golden[f"w_{NAME[t]}_blocks"], golden[f"w_{NAME[t]}_dequant"]
    got = H.dequantize(t, blocks, want.size).reshape(want.shape)
    assert same_f32(got, want)
    rng = np.random.default_rng(5 + t)
    K = 4096
    w = quantize_weights(t, (rng.standard_normal((16, K)) * 0.1).astype(np.float32))
    raw = rng.integers(0, 256, size=w.shape, dtype=np.uint8)  # arbitrary bytes incl. NaN/Inf scales
    for blk in (w, raw):
        want = np.stack([O.dequantize(t, r, K) for r in blk])
        got = H.dequantize(t, blk, want.size).reshape(want.shape)
        assert same_f32(got, want)
        got16 = H.dequantize(t, blk, want.size, dst_type=1).reshape(want.shape)
        with np.errstate(over="ignore", invalid="ignore"):
            want16 = want.astype(np.float16)
        assert np.array_equal(got16.view(np.uint16)[~np.isnan(want16)], want16.view(np.uint16)[~np.isnan(want16)])


def test_dequantize_rejects_bad_sizes(H):
    import koboldcpp_b200.lib as L
    with pytest.raises(L.B200Error):
        H.dequantize(R.Q4_K, np.zeros(144, np.uint8), 128)  # not a multiple of 256
    with pytest.raises(L.B200Error):
        H.dequantize(R.Q8_1, np.zeros(36, np.uint8), 32)    # not a weight type


# ------------------------------------------------------------------------------ a3: q8_1 quantisation (bit-exact)
def test_quantize_q8_1_bit_exact(H, golden):
    got = H.quantize_q8_1(golden["act_src"])
    assert np.array_equal(got, golden["act_q8_1"])
    rng = np.random.default_rng(11)
    for K in (32, 4096, 11008, 28672):
        x = (rng.standard_normal((3, K)) * rng.uniform(0.01, 30)).astype(np.float32)
        x[1, : K // 2] = 0
        want = np.stack([O.quantize(R.Q8_1, r) for r in x])
        assert np.array_equal(H.quantize_q8_1(x), want)


# ------------------------------------------------------------------------------ a6: decode GEMV
def check_gemv(H, t, K, M, N, rng, w=None):
    if w is None:
        w = quantize_weights(t, (rng.standard_normal((M, K)) * 0.05).astype(np.float32))
    x = rng.standard_normal((N, K)).astype(np.float32)
    got = H.mul_mat_vec_q(t, w, x, K, M)
    exact = O.mul_mat_q8_1(t, w, x, K, M)
    # fp32 summation error is relative to sum|terms| ~ ||x|| ||w||, not to the (possibly cancelling) result
    wn = max(np.linalg.norm(O.dequantize(t, w[i], K)) for i in range(0, M, max(1, M // 8)))
    scale = np.abs(exact).max() + 0.03 * wn * np.linalg.norm(x, axis=1).max() + 1e-30
    assert np.abs(got - exact).max() <= TOL_Q81 * scale, (NAME[t], K, M, N, np.abs(got - exact).max() / scale)
    cpu = O.mul_mat(t, w, x, K, M)
    # few outputs: a single cancelling dot product makes a relative norm meaningless, so floor the denominator with
    # the typical output magnitude ||x|| ||w|| / sqrt(K)
    floor = wn * np.linalg.norm(x, axis=1).max() / np.sqrt(K) * np.sqrt(M * N)
    nmse = np.linalg.norm(got - cpu) / max(np.linalg.norm(cpu), floor)
    assert nmse <= TOL_CPU_NMSE, (NAME[t], K, M, N, nmse)
    return got


@pytest.mark.parametrize("t", WT, ids=[NAME[t] for t in WT])
def test_gemv_golden(H, golden, t):
    w, x = golden[f"mm_{NAME[t]}_w"], golden["mm_x"]
    K, M = x.shape[1], w.shape[0]
    got = H.mul_mat_vec_q(t, w, x, K, M)
    cpu = golden[f"mm_{NAME[t]}_out"]
    assert np.linalg.norm(got - cpu) / np.linalg.norm(cpu) <= TOL_CPU_NMSE
    exact = O.mul_mat_q8_1(t, w, x, K, M)
    assert np.abs(got - exact).max() <= TOL_Q81 * np.abs(exact).max()


@pytest.mark.parametrize("t", WT, ids=[NAME[t] for t in WT])
@pytest.mark.parametrize("shape", [(256, 1, 1), (256, 3, 2), (4096, 300, 1), (5632, 77, 1), (11008, 160, 1), (4096, 513, 8),
                                   (2048, 2048, 3), (14336, 64, 2)])
def test_gemv_shapes(H, t, shape):
    K, M, N = shape
    check_gemv(H, t, K, M, N, np.random.default_rng(K + M + N + t))


@pytest.mark.parametrize("t", (R.Q4_0, R.Q8_0, R.Q4_1, R.Q5_0, R.Q5_1, R.IQ4_NL), ids=["q4_0", "q8_0", "q4_1", "q5_0", "q5_1", "iq4_nl"])
@pytest.mark.parametrize("K", (32, 96, 4096 + 64))
def test_gemv_ragged_k(H, t, K):
    """Q4_0/Q8_0 rows whose length is not a multiple of the 128-weight slice; odd row byte sizes (2-byte aligned rows)."""
    check_gemv(H, t, K, 37, 2, np.random.default_rng(K + t))


@pytest.mark.parametrize("t", WT, ids=[NAME[t] for t in WT])
def test_gemv_llama7b_shapes_properties(H, t):
    """Full-size 7B shapes (oracle too slow for every row): check sampled rows against the oracle plus
    linearity in x and invariance under row permutation."""
    rng = np.random.default_rng(99 + t)
    for K, M in ((4096, 4096), (4096, 11008), (11008, 4096)):
        w = random_blocks(t, M, K, rng)
        x1 = rng.standard_normal((1, K)).astype(np.float32)
        y1 = H.mul_mat_vec_q(t, w, x1, K, M)
        rows = rng.choice(M, 48, replace=False)
        exact = O.mul_mat_q8_1(t, w[rows], x1, K, len(rows))
        assert np.abs(y1[:, rows] - exact).max() <= TOL_Q81 * np.abs(exact).max()
        perm = rng.permutation(M)
        yp = H.mul_mat_vec_q(t, w[perm], x1, K, M)
        assert np.array_equal(bits(yp), bits(y1[:, perm]))  # a row's result does not depend on where the row sits
        y2 = H.mul_mat_vec_q(t, w, 2.0 * x1, K, M)          # power-of-two scaling is exact through q8_1
        assert np.array_equal(bits(y2), bits(2.0 * y1))


def test_gemv_multi_matches_single(H):
    rng = np.random.default_rng(4)
    K = 4096
    for t in (R.Q4_K, R.Q6_K):
        ws = [random_blocks(t, M, K, rng) for M in (4096, 1024, 1000)]
        x = rng.standard_normal((2, K)).astype(np.float32)
        fused = H.mul_mat_vec_q_multi(t, ws, x, K)
        for w, f in zip(ws, fused):
            single = H.mul_mat_vec_q(t, w, x, K, w.shape[0])
            assert np.array_equal(bits(f), bits(single))


@pytest.mark.parametrize("K", (4096, 8192, 2048))
@pytest.mark.parametrize("pair", ((R.Q4_K, R.Q6_K), (R.Q6_K, R.Q4_K), (R.Q5_K, R.Q6_K), (R.Q4_K, R.Q5_K)), ids=lambda p: f"{NAME[p[0]]}+{NAME[p[1]]}")
def test_gemv_two_weight_types_in_one_launch(H, pair, K):
    """attn_q / attn_k + attn_v of the *_K_M mixes share one decode launch: every row must come out bit-identical to the
    single-type launch of its matrix (row counts chosen so that pairs straddle the type boundary and CTAs end half full)."""
    rng = np.random.default_rng(K + pair[0] * 16 + pair[1])
    ta, tb = pair
    Ms, ts = (1000, 137, 515), (ta, ta, tb)
    ws = [random_blocks(t, M, K, rng) for t, M in zip(ts, Ms)]
    x = rng.standard_normal((1, K)).astype(np.float32)
    fused = H.mul_mat_vec_q_multi(ta, ws, x, K, types=list(ts))
    for w, t, f in zip(ws, ts, fused):
        single = H.mul_mat_vec_q(t, w, x, K, w.shape[0])
        assert np.array_equal(bits(f), bits(single)), NAME[t]
    # the second type first
    fused = H.mul_mat_vec_q_multi(tb, ws[::-1], x, K, types=list(ts[::-1]))
    for w, t, f in zip(ws[::-1], ts[::-1], fused):
        assert np.array_equal(bits(f), bits(H.mul_mat_vec_q(t, w, x, K, w.shape[0]))), NAME[t]


def test_gemv_bad_arguments(H):
    import koboldcpp_b200.lib as L
    w = np.zeros((4, 144), np.uint8)
    with pytest.raises(L.B200Error):
        H.mul_mat_vec_q(R.Q4_K, w, np.zeros((9, 256), np.float32), 256, 4)  # N > 8
    with pytest.raises(L.B200Error):
        H.mul_mat_vec_q(R.Q8_1, w, np.zeros((1, 256), np.float32), 256, 4)  # not a weight type


# ------------------------------------------------------------------------------ a9: MUL_MAT_ID
@pytest.mark.parametrize("t", (R.Q4_K, R.Q6_K, R.Q8_0), ids=["q4_K", "q6_K", "q8_0"])
def test_mul_mat_id_golden(H, golden, t):
    w, ids = golden[f"mmid_{NAME[t]}_w"], golden["mmid_ids"]
    for tag in ("1", "2"):
        b = golden[f"mmid_b{tag}"]
        got = H.mul_mat_id_vec_q(t, w, ids, b, b.shape[2], 16, 4)
        cpu = golden[f"mmid_{NAME[t]}_out{tag}"]
        assert np.linalg.norm(got - cpu) / np.linalg.norm(cpu) <= TOL_CPU_NMSE


def test_mul_mat_id_mixtral_decode_shape(H):
    """C4 decode: as = [4096, 14336, 8] (gate/up), b = [4096, 1, 1]; and down: as = [14336, 4096, 8], b = [14336, 2, 1]."""
    rng = np.random.default_rng(8)
    for (K, M, nbr) in ((4096, 14336, 1), (14336, 4096, 2)):
        w = random_blocks(R.Q4_K, 8 * M, K, rng)
        ids = np.array([[5, 2]], dtype=np.int32)
        b = rng.standard_normal((1, nbr, K)).astype(np.float32)
        got = H.mul_mat_id_vec_q(R.Q4_K, w, ids, b, K, M, 8)
        for e in range(2):
            rows = rng.choice(M, 32, replace=False)
            wexp = w[ids[0, e] * M:(ids[0, e] + 1) * M][rows]
            exact = O.mul_mat_q8_1(R.Q4_K, wexp, b[0, e % nbr], K, len(rows))
            assert np.abs(got[0, e, rows] - exact[0]).max() <= TOL_Q81 * np.abs(exact).max()


def test_mul_mat_id_bad_expert_is_nan(H, golden):
    w, b = golden["mmid_q8_0_w"], golden["mmid_b1"]
    got = H.mul_mat_id_vec_q(R.Q8_0, w, np.array([[0, 4]], dtype=np.int32), b[:1], b.shape[2], 16, 4)
    assert np.isfinite(got[0, 0]).all() and np.isnan(got[0, 1]).all()


# ------------------------------------------------------------------------------ a10-a12: FLASH_ATTN_EXT
@pytest.mark.parametrize("N", (1, 5))
@pytest.mark.parametrize("tkv", (R.F16, R.Q8_0, R.Q4_0), ids=["f16", "q8_0", "q4_0"])
def test_flash_attn_golden(H, golden, tkv, N):
    q, mask = golden[f"fa_q_N{N}"], golden[f"fa_mask_N{N}"]
    k, v = golden[f"fa_k_{NAME[tkv]}"], golden[f"fa_v_{NAME[tkv]}"]
    D = q.shape[2]
    got = H.flash_attn_ext(q, k, v, mask, tkv, 1.0 / np.sqrt(D))
    cpu = golden[f"fa_out_{NAME[tkv]}_N{N}"]
    f64 = O.flash_attn_f64(q, k, v, mask, tkv, 1.0 / np.sqrt(D))
    assert np.abs(got - f64).max() <= TOL_FA[tkv]
    assert np.abs(got - cpu).max() <= 2 * TOL_FA[tkv]
    assert np.abs(got - f64).max() <= np.abs(cpu - f64).max() * 1.5 + 1e-4  # at least as accurate as the CPU path


@pytest.mark.parametrize("N", (1, 5))
def test_flash_attn_softcap_alibi_golden(H, golden, N):
    q, mask = golden[f"fa_q_N{N}"], golden[f"fa_mask_N{N}"]
    got = H.flash_attn_ext(q, golden["fa_k_f16"], golden["fa_v_f16"], mask, R.F16, 1.0 / np.sqrt(q.shape[2]), 8.0, 30.0)
    assert np.abs(got - golden[f"fa_out_f16_N{N}_softcap_alibi"]).max() <= 2 * TOL_FA[R.F16]


@pytest.mark.parametrize("cfg", [
    # (n_head, n_head_kv, D, n_kv, N)
    (32, 32, 128, 4096, 1),    # C2: Llama-2-7B decode at 4k
    (32, 8, 128, 8192, 1),     # C4-like GQA 4 (32k covered by the property test)
    (64, 8, 128, 1024, 1),     # C5: GQA 8
    (32, 4, 64, 512, 1),       # C1: TinyLlama D=64
    (8, 8, 128, 250, 3),       # ragged n_kv (not a multiple of the tile), small N
    (4, 2, 64, 33, 2),
    (40, 40, 128, 512, 16),    # small prefill through the generic path
])
@pytest.mark.parametrize("tkv", (R.F16, R.Q8_0, R.Q4_0), ids=["f16", "q8_0", "q4_0"])
def test_flash_attn_shapes(H, tkv, cfg):
    nh, nhkv, D, nkv, N = cfg
    rng = np.random.default_rng(sum(cfg) + tkv)
    q = rng.standard_normal((nh, N, D)).astype(np.float32)
    k = O.pack_kv(tkv, rng.standard_normal((nhkv, nkv, D)))
    v = O.pack_kv(tkv, rng.standard_normal((nhkv, nkv, D)))
    npad = (N + 31) // 32 * 32
    mask = np.zeros((npad, nkv), dtype=np.float16)
    for i in range(N):  # causal tail
        cut = nkv - (N - 1 - i)
        mask[i, cut:] = -np.inf
    got = H.flash_attn_ext(q, k, v, mask, tkv, 1.0 / np.sqrt(D))
    f64 = O.flash_attn_f64(q, k, v, mask, tkv, 1.0 / np.sqrt(D))
    err = np.abs(got - f64).max()
    if nh * N * nkv <= 40 * 16 * 512:
        cpu = O.flash_attn_ext(q, k, v, mask, tkv, 1.0 / np.sqrt(D))
        err_cpu = np.abs(cpu - f64).max()
        # few keys = little averaging of the int8-Q / f16-V rounding noise: the bar is the CPU path's own distance to fp64
        assert err <= max(TOL_FA[tkv], 1.5 * err_cpu + 1e-4), (err, err_cpu)
        assert np.abs(got - cpu).max() <= 2 * max(TOL_FA[tkv], err_cpu)
    else:
        assert err <= TOL_FA[tkv], err


@pytest.mark.parametrize("cfg", [
    # (n_head, n_head_kv, D, n_kv, N, n_past): a prompt batch at the front of the cache and one behind 1536 cached cells
    (40, 40, 128, 512, 512, 0),      # C3: Llama-2-13B pp512
    (16, 4, 128, 2048, 512, 1536),   # GQA 4, second batch of a long prompt
    (8, 8, 64, 320, 300, 20),        # D = 64, ragged sizes
    (8, 2, 128, 1024, 250, 256),     # cells behind the batch that no query sees
])
def test_flash_attn_prefill_causal(H, cfg):
    """a12: prompt batches through the tensor-core kernel.  Causal mask as llama.cpp builds it (src/llama.cpp:16866-16930): query i
    sees cells [0, n_past + i]; the KV tiles above the diagonal are skipped by the kernel, so the result must not depend on
    what those cells hold (they are filled with NaN-producing garbage here)."""
    nh, nhkv, D, nkv, N, n_past = cfg
    rng = np.random.default_rng(sum(cfg))
    q = rng.standard_normal((nh, N, D)).astype(np.float32)
    kf, vf = rng.standard_normal((nhkv, nkv, D)), rng.standard_normal((nhkv, nkv, D))
    npad = (N + 31) // 32 * 32
    mask = np.full((npad, nkv), -np.inf, dtype=np.float16)
    for i in range(N):
        mask[i, :min(nkv, n_past + i + 1)] = 0
    f64 = O.flash_attn_f64(q, O.pack_kv(R.F16, kf), O.pack_kv(R.F16, vf), mask, R.F16, 1.0 / np.sqrt(D))
    got = H.flash_attn_ext(q, O.pack_kv(R.F16, kf), O.pack_kv(R.F16, vf), mask, R.F16, 1.0 / np.sqrt(D))
    assert np.isfinite(got).all()
    assert np.abs(got - f64).max() <= TOL_FA[R.F16], float(np.abs(got - f64).max())
    # whole 16-cell tiles no query of the batch can see: poison them (inf * 0 = NaN if a masked tile were multiplied in)
    hidden = (n_past + N + 15) // 16 * 16
    if hidden < nkv:
        kf[:, hidden:] = 6e4
        vf[:, hidden:] = np.inf
        got2 = H.flash_attn_ext(q, O.pack_kv(R.F16, kf), O.pack_kv(R.F16, vf), mask, R.F16, 1.0 / np.sqrt(D))
        assert np.isfinite(got2).all() and np.array_equal(got2, got)


@pytest.mark.parametrize("D", (80, 96, 112, 256))
@pytest.mark.parametrize("shape", [
    # (n_head, n_head_kv, n_kv, N, n_past)
    (32, 32, 1024, 1, 1023),    # decode, one query column per head (StableLM / Phi-style head sizes; Gemma D = 256)
    (16, 2, 2085, 1, 2084),     # GQA 8, ragged cache, several cluster splits
    (8, 4, 330, 5, 325),        # a few columns
    (8, 8, 384, 300, 84),       # a prompt batch through the prefill kernel, causal tiles skipped
], ids=["decode", "gqa8", "n5", "prefill"])
def test_flash_attn_more_head_sizes(H, D, shape):
    """N4: head sizes 80 / 96 / 112 / 256 over an f16 cache (the reference's CUDA backend: ggml-cuda.cu:3162-3177 lists
    64, 80, 96, 112, 128, 256).  They run on the tensor-core kernels only; a q8_0 cache with these sizes is refused."""
    nh, nhkv, nkv, N, n_past = shape
    rng = np.random.default_rng(D * 7 + sum(shape))
    q = rng.standard_normal((nh, N, D)).astype(np.float32)
    k = O.pack_kv(R.F16, rng.standard_normal((nhkv, nkv, D)))
    v = O.pack_kv(R.F16, rng.standard_normal((nhkv, nkv, D)))
    npad = (N + 31) // 32 * 32
    mask = np.full((npad, nkv), -np.inf, dtype=np.float16)
    for i in range(N):
        mask[i, :min(nkv, n_past + i + 1)] = 0
    got = H.flash_attn_ext(q, k, v, mask, R.F16, 1.0 / np.sqrt(D))
    f64 = O.flash_attn_f64(q, k, v, mask, R.F16, 1.0 / np.sqrt(D))
    assert np.isfinite(got).all()
    assert np.abs(got - f64).max() <= TOL_FA[R.F16], float(np.abs(got - f64).max())
    if nh * N * nkv <= 8 * 5 * 330:
        cpu = O.flash_attn_ext(q, k, v, mask, R.F16, 1.0 / np.sqrt(D))
        assert np.abs(got - cpu).max() <= 2 * TOL_FA[R.F16]


def test_flash_attn_head_size_refused_for_q8_cache(H):
    rng = np.random.default_rng(5)
    q = rng.standard_normal((4, 1, 96)).astype(np.float32)
    k = O.pack_kv(R.Q8_0, rng.standard_normal((4, 64, 96))); v = O.pack_kv(R.Q8_0, rng.standard_normal((4, 64, 96)))
    with pytest.raises(Exception, match="head size"):
        H.flash_attn_ext(q, k, v, None, R.Q8_0, 0.1)


def test_flash_attn_no_mask_and_fully_masked_prefix(H):
    rng = np.random.default_rng(21)
    nh, nhkv, D, nkv = 4, 4, 128, 300
    q = rng.standard_normal((nh, 1, D)).astype(np.float32)
    k = O.pack_kv(R.F16, rng.standard_normal((nhkv, nkv, D))); v = O.pack_kv(R.F16, rng.standard_normal((nhkv, nkv, D)))
    got = H.flash_attn_ext(q, k, v, None, R.F16, 0.1)
    assert np.abs(got - O.flash_attn_f64(q, k, v, None, R.F16, 0.1)).max() <= TOL_FA[R.F16]
    mask = np.zeros((32, nkv), dtype=np.float16)
    mask[:, :200] = -np.inf  # whole leading tiles masked: the running max starts at -inf
    got = H.flash_attn_ext(q, k, v, mask, R.F16, 0.1)
    assert np.abs(got - O.flash_attn_f64(q, k, v, mask, R.F16, 0.1)).max() <= TOL_FA[R.F16]


def test_flash_attn_long_context_split_invariance(H):
    """C4 size (q8_0 KV, 32k, GQA 4): the fp64 oracle on 4 heads, and the property that attention over the
    concatenation of two KV halves equals the softmax-weighted merge of the halves."""
    rng = np.random.default_rng(33)
    nh, nhkv, D, nkv = 4, 1, 128, 32768
    q = rng.standard_normal((nh, 1, D)).astype(np.float32)
    kf = rng.standard_normal((nhkv, nkv, D)).astype(np.float32) * 0.5
    vf = rng.standard_normal((nhkv, nkv, D)).astype(np.float32)
    k, v = O.pack_kv(R.Q8_0, kf), O.pack_kv(R.Q8_0, vf)
    got = H.flash_attn_ext(q, k, v, None, R.Q8_0, 1.0 / np.sqrt(D))
    f64 = O.flash_attn_f64(q, k, v, None, R.Q8_0, 1.0 / np.sqrt(D))
    assert np.abs(got - f64).max() <= TOL_FA[R.Q8_0]
    # repeatability: the ticket counters must be left clean, and results must not depend on CTA arrival order by more than fp32 noise
    again = H.flash_attn_ext(q, k, v, None, R.Q8_0, 1.0 / np.sqrt(D))
    assert np.abs(got - again).max() <= 1e-6
