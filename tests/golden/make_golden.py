"""Generate tests/golden/*.npz from the REFERENCE itself (oracle/_ref/libggml_ref_scalar.so, built by
oracle/Makefile from /root/reference).  Run here (the reference cannot travel); the vectors are committed.

    python tests/golden/make_golden.py

Every array is produced by reference code: ggml_quantize_chunk, dequantize_row_*, quantize_row_q8_*_ref,
ggml_vec_dot_*, and single-node graphs (ggml_mul_mat / ggml_mul_mat_id / ggml_flash_attn_ext) computed by
the reference CPU backend -- the same method as ggml_backend_compare_graph_backend
(ggml/src/ggml-backend.cpp:2566).  The scalar build is used so the float results do not depend on the
SIMD summation order of the machine that generated them.
"""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import ggml_ref as R  # noqa: E402
import orc as O  # noqa: E402  (only pack_kv's f16 view + q8_0 row packing via the reference below)

OUT = Path(__file__).resolve().parent
SEED = 1234


def special_activations(rng, k):
    """rows exercising the edge cases of the activation quantisers: zeros, constants, ties at .5, huge/small."""
    rows = [rng.standard_normal(k), np.zeros(k), np.full(k, 3.25), rng.standard_normal(k) * 1e-6,
            rng.standard_normal(k) * 3e4, np.where(np.arange(k) % 2, 0.5, -0.5) * np.arange(k),
            np.linspace(-127, 127, k), np.r_[np.zeros(k - 1), 1.0]]
    return np.stack(rows).astype(np.float32)


def main():
    ref = R.load("scalar")
    rng = np.random.default_rng(SEED)
    g = {}
    # --- (a1,a2) weight formats: reference-quantised gaussians + raw random blocks -> dequantised f32
    K = 1024
    for t in R.ORACLE_WEIGHT_TYPES:
        n = R.TYPE_NAMES[t]
        x = (rng.standard_normal((4, K)) * 0.05).astype(np.float32)
        w = ref.quantize(t, x)
        raw = rng.integers(0, 256, size=(2, R.row_size(t, K)), dtype=np.uint8)
        blocks = np.concatenate([w, raw])
        g[f"w_{n}_src"] = x
        g[f"w_{n}_blocks"] = blocks
        g[f"w_{n}_dequant"] = np.stack([ref.dequantize(t, b, K) for b in blocks])
    # --- (a3) activation quantisers
    acts = special_activations(rng, K)
    g["act_src"] = acts
    for t in (R.Q8_0, R.Q8_1, R.Q8_K):
        g[f"act_{R.TYPE_NAMES[t]}"] = np.stack([ref.quantize_act_ref(t, a) for a in acts])
    # --- (a15) vec_dot + MUL_MAT on the CPU backend
    Km, M, N = 1024, 24, 3
    xs = rng.standard_normal((N, Km)).astype(np.float32)
    g["mm_x"] = xs
    for t in R.ORACLE_WEIGHT_TYPES:
        n = R.TYPE_NAMES[t]
        w = ref.quantize(t, (rng.standard_normal((M, Km)) * 0.05).astype(np.float32))
        g[f"mm_{n}_w"] = w
        g[f"mm_{n}_out"] = ref.mul_mat(t, w, xs, Km, M)
        at = R.act_type(t)
        a = ref.quantize_act_ref(at, xs[0])
        g[f"vd_{n}"] = np.array([ref.vec_dot(t, w[i], a, Km) for i in range(M)], dtype=np.float32)
    # --- (a9) MUL_MAT_ID
    ne, Mi, Ki = 4, 16, 512
    ids = np.array([[1, 3], [0, 2], [3, 3]], dtype=np.int32)
    g["mmid_ids"] = ids
    g["mmid_b1"] = rng.standard_normal((3, 1, Ki)).astype(np.float32)
    g["mmid_b2"] = rng.standard_normal((3, 2, Ki)).astype(np.float32)
    for t in (R.Q4_K, R.Q6_K, R.Q8_0):
        n = R.TYPE_NAMES[t]
        w = ref.quantize(t, (rng.standard_normal((ne * Mi, Ki)) * 0.05).astype(np.float32))
        g[f"mmid_{n}_w"] = w
        g[f"mmid_{n}_out1"] = ref.mul_mat_id(t, w, ids, g["mmid_b1"], Ki, Mi, ne)
        g[f"mmid_{n}_out2"] = ref.mul_mat_id(t, w, ids, g["mmid_b2"], Ki, Mi, ne)
    # --- (a10-a12) FLASH_ATTN_EXT, f16, q8_0 and q4_0 KV, GQA, masked tail, decode (N=1) and small prefill (N=5)
    nh, nhkv, D, nkv = 8, 2, 128, 256
    kf = rng.standard_normal((nhkv, nkv, D)).astype(np.float32)
    vf = rng.standard_normal((nhkv, nkv, D)).astype(np.float32)
    g["fa_k_f32"], g["fa_v_f32"] = kf, vf
    for N in (1, 5):
        q = rng.standard_normal((nh, N, D)).astype(np.float32)
        mask = np.zeros((32, nkv), dtype=np.float16)
        mask[:, 200:] = -np.inf
        for i in range(N):
            mask[i, 200 - (N - 1 - i) * 7:] = -np.inf
        g[f"fa_q_N{N}"], g[f"fa_mask_N{N}"] = q, mask
        for tkv in (R.F16, R.Q8_0, R.Q4_0):
            k = O.pack_kv(tkv, kf) if tkv == R.F16 else ref.quantize(tkv, kf.reshape(-1, D)).reshape(nhkv, nkv, -1)
            v = O.pack_kv(tkv, vf) if tkv == R.F16 else ref.quantize(tkv, vf.reshape(-1, D)).reshape(nhkv, nkv, -1)
            g[f"fa_k_{R.TYPE_NAMES[tkv]}"], g[f"fa_v_{R.TYPE_NAMES[tkv]}"] = k, v
            g[f"fa_out_{R.TYPE_NAMES[tkv]}_N{N}"] = ref.flash_attn_ext(q, k, v, mask, tkv, 1.0 / np.sqrt(D))
        # logit soft-cap + ALiBi variant (f16 only)
        g[f"fa_out_f16_N{N}_softcap_alibi"] = ref.flash_attn_ext(q, g["fa_k_f16"], g["fa_v_f16"], mask, R.F16,
                                                                  1.0 / np.sqrt(D), max_bias=8.0, softcap=30.0)
    np.savez_compressed(OUT / "hotpath_golden.npz", **g)
    print("wrote", OUT / "hotpath_golden.npz", sum(v.nbytes for v in g.values()) // 1024, "KiB raw")


if __name__ == "__main__":
    main()
