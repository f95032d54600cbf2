"""GPU parity of the tcgen05 prefill GEMM (b200_mul_mat_q) against fp64 dequant(W).x (oracle dequantisation) and the
reference CPU backend's MUL_MAT where oracle/_ref is present.  Operands are bf16 on the tensor cores (8 mantissa bits):
the stated tolerance is 4e-3 relative L2 against fp64 -- tighter than the ~1e-2 of the q8 integer paths."""
from __future__ import annotations

import numpy as np
import pytest

import koboldcpp_b200.host as H
import orc as O
from koboldcpp_b200.lib import IQ4_NL, IQ4_XS, Q2_K, Q3_K, Q4_0, Q4_1, Q4_K, Q5_0, Q5_1, Q5_K, Q6_K, Q8_0
from koboldcpp_b200.llama import random_blocks

pytestmark = pytest.mark.gpu
TYPES = [Q4_0, Q8_0, Q4_K, Q5_K, Q6_K, Q4_1, Q5_0, Q5_1, Q2_K, Q3_K, IQ4_NL, IQ4_XS]


def ref64(t, w, x, K, M):
    W = np.stack([O.dequantize(t, w[m], K) for m in range(M)]).astype(np.float64)
    return x.astype(np.float64) @ W.T


def rel_l2(a, b):
    return float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30))


@pytest.mark.parametrize("t", TYPES)
@pytest.mark.parametrize("K,M,N", [(256, 128, 32), (512, 256, 9), (1024, 200, 70), (2048, 384, 130), (4096, 128, 512), (768, 130, 300)])
def test_mmq_matches_fp64(t, K, M, N):
    rng = np.random.default_rng(K * 7 + M + N + t)
    w = random_blocks(rng, t, M, K)
    x = rng.standard_normal((N, K)).astype(np.float32)
    got = H.mul_mat_q(t, w, x, K, M)
    exp = ref64(t, w, x, K, M)
    assert np.isfinite(got).all()
    assert rel_l2(got, exp) < 4e-3, rel_l2(got, exp)


@pytest.mark.parametrize("t", [Q4_K, Q6_K])
def test_mmq_agrees_with_decode_gemv(t):
    """The two matmul paths of the backend (GEMV for N <= 8, tensor-core GEMM above) must agree within their quantisation noise."""
    K, M, N = 2048, 256, 8
    rng = np.random.default_rng(3)
    w = random_blocks(rng, t, M, K)
    x = rng.standard_normal((N, K)).astype(np.float32)
    a = H.mul_mat_q(t, w, x, K, M)
    b = np.stack([H.mul_mat_vec_q(t, w, x[n:n + 1], K, M)[0] for n in range(N)])
    assert rel_l2(a, b) < 1.5e-2, rel_l2(a, b)


# ------------------------------------------------------------------------------ a9: MUL_MAT_ID for batches (grouped tensor-core launch)
def ref64_id(t, w, ids, b, K, M, n_expert):
    """dst[t, s, :] = dequant(as[ids[t, s]]) . b[t, s % n_b_rows]  in fp64 (ggml.c:12711-12916 semantics)."""
    T, n_used = ids.shape
    W = [np.stack([O.dequantize(t, w[e * M + m], K) for m in range(M)]).astype(np.float64) for e in range(n_expert)]
    out = np.zeros((T, n_used, M))
    for tok in range(T):
        for s in range(n_used):
            out[tok, s] = W[ids[tok, s]] @ b[tok, s % b.shape[1]].astype(np.float64)
    return out


@pytest.mark.parametrize("t", [Q4_K, Q6_K, Q8_0])
@pytest.mark.parametrize("K,M,n_expert,n_used,T,nbr", [(512, 256, 8, 2, 40, 1), (768, 130, 4, 2, 70, 2), (1024, 128, 8, 3, 9, 1), (256, 384, 16, 4, 300, 4)])
def test_mul_mat_id_batched_matches_fp64(t, K, M, n_expert, n_used, T, nbr):
    rng = np.random.default_rng(K + M + T + t)
    w = random_blocks(rng, t, n_expert * M, K)
    ids = np.stack([rng.choice(n_expert, n_used, replace=False) for _ in range(T)]).astype(np.int32)
    b = rng.standard_normal((T, nbr, K)).astype(np.float32)
    got = H.mul_mat_id_q(t, w, ids, b, K, M, n_expert)
    exp = ref64_id(t, w, ids, b, K, M, n_expert)
    assert np.isfinite(got).all()
    assert rel_l2(got, exp) < 4e-3, rel_l2(got, exp)


def test_mul_mat_id_batched_skewed_routing_and_oracle():
    """All tokens on one expert, experts with no token, and the oracle's own mul_mat_id (CPU algorithm, q8_K activations)."""
    t, K, M, n_expert, n_used, T = Q4_K, 512, 256, 8, 2, 64
    rng = np.random.default_rng(77)
    w = random_blocks(rng, t, n_expert * M, K)
    b = rng.standard_normal((T, 1, K)).astype(np.float32)
    ids = np.tile(np.array([[3, 5]], dtype=np.int32), (T, 1))
    ids[T // 2:, 1] = 0
    got = H.mul_mat_id_q(t, w, ids, b, K, M, n_expert)
    exp = ref64_id(t, w, ids, b, K, M, n_expert)
    assert rel_l2(got, exp) < 4e-3
    cpu = O.mul_mat_id(t, w, ids, b, K, M, n_expert)
    assert rel_l2(got, cpu.reshape(got.shape)) < 2e-2
    # decode-regime GEMV path and the grouped path agree within their quantisation noise
    gemv = H.mul_mat_id_vec_q(t, w, ids[:8], b[:8], K, M, n_expert)
    assert rel_l2(got[:8], gemv) < 1.5e-2


def test_mul_mat_id_batched_invalid_ids_are_skipped():
    t, K, M, n_expert, n_used, T = Q8_0, 256, 128, 4, 2, 20
    rng = np.random.default_rng(78)
    w = random_blocks(rng, t, n_expert * M, K)
    b = rng.standard_normal((T, 1, K)).astype(np.float32)
    ids = np.stack([rng.choice(n_expert, n_used, replace=False) for _ in range(T)]).astype(np.int32)
    ids[3, 1] = 9
    ids[7, 0] = -1
    got = H.mul_mat_id_q(t, w, ids, b, K, M, n_expert)
    bad = np.zeros((T, n_used), dtype=bool)
    bad[3, 1] = bad[7, 0] = True
    assert np.isnan(got[bad]).all() and np.isfinite(got[~bad]).all()
    ids_ok = np.where(bad, 0, ids)
    exp = ref64_id(t, w, ids_ok, b, K, M, n_expert)
    assert rel_l2(got[~bad], exp[~bad]) < 4e-3


@pytest.mark.parametrize("K,M,N", [(5120, 13824, 512), (13824, 5120, 512), (5120, 5120, 512)])
def test_mmq_c3_shapes_sampled_rows(K, M, N):
    """BASELINE configs[2] (13B pp512) shapes; fp64 check on a sample of output rows (the full fp64 product is 36 G flop)."""
    t = Q5_K if K != 13824 else Q6_K
    rng = np.random.default_rng(K + M)
    w = random_blocks(rng, t, M, K)
    x = rng.standard_normal((N, K)).astype(np.float32)
    got = H.mul_mat_q(t, w, x, K, M)
    rows = rng.choice(M, 96, replace=False)
    W = np.stack([O.dequantize(t, w[m], K) for m in rows]).astype(np.float64)
    exp = x.astype(np.float64) @ W.T
    assert rel_l2(got[:, rows], exp) < 4e-3
