"""GPU parity of the tcgen05 prefill GEMM (b200_mul_mat_q) against fp64 dequant(W).x (oracle dequantisation) and the
reference CPU backend's MUL_MAT where oracle/_ref is present.  Operands are bf16 on the tensor cores (8 mantissa bits):
the stated tolerance is 4e-3 relative L2 against fp64 -- tighter than the ~1e-2 of the q8 integer paths."""
from __future__ import annotations

import numpy as np
import pytest

import koboldcpp_b200.host as H
import orc as O
from koboldcpp_b200.lib import Q4_0, Q4_K, Q5_K, Q6_K, Q8_0
from koboldcpp_b200.llama import random_blocks

pytestmark = pytest.mark.gpu
TYPES = [Q4_0, Q8_0, Q4_K, Q5_K, Q6_K]


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
