"""b200_flash_attn_ext_static: the attention entry point that allocates nothing (scratch in a static device array).

No counterpart in the reference.  Collected last on purpose:
a failure here must not hide the parity tests.  The check is the new entry point against the existing
b200_flash_attn_ext through the C-ABI, on two calls in a row (the second sees the scratch the first one left behind)."""
import ctypes as C

import numpy as np
import pytest

import ggml_ref as R
import orc as O
from koboldcpp_b200 import host
from koboldcpp_b200.host import DeviceBuffer
from koboldcpp_b200.lib import FattnArgs, check, lib

pytestmark = pytest.mark.gpu

TOL = 2.5e-3  # TOL_FA[f16] of tests/test_gpu_kernels.py; the two entry points launch the same kernels, so they agree bit for bit


def _problem(seed, nh, nhkv, D, nkv, N):
    rng = np.random.default_rng(seed)
    q = rng.standard_normal((nh, N, D)).astype(np.float32)
    k = O.pack_kv(R.F16, rng.standard_normal((nhkv, nkv, D)))
    v = O.pack_kv(R.F16, rng.standard_normal((nhkv, nkv, D)))
    mask = np.zeros(((N + 31) // 32 * 32, nkv), dtype=np.float16)
    for i in range(N):
        mask[i, nkv - (N - 1 - i):] = -np.inf
    return q, k, v, mask


def _static_call(q, k, v, mask, scale):
    """The new entry point through ctypes: inputs and output are the caller's, no workspace argument."""
    n_head, N, D = q.shape
    n_head_kv, n_kv, rb = k.shape
    dq, dk, dv = DeviceBuffer.from_numpy(q), DeviceBuffer.from_numpy(k, pad=64), DeviceBuffer.from_numpy(v, pad=64)
    dm = DeviceBuffer.from_numpy(mask)
    dd = DeviceBuffer(4 * N * n_head * D)
    a = FattnArgs(dq.ptr, dk.ptr, dv.ptr, dm.ptr, dd.ptr, D, N, n_head, n_head_kv, n_kv, R.F16, R.F16,
                  4 * D, 4 * D * N, rb, rb * n_kv, rb, rb * n_kv, 2 * n_kv, scale, 0.0, 0.0)
    need = lib().b200_flash_attn_workspace_bytes(C.byref(a))
    check(lib().b200_flash_attn_ext_static(C.byref(a), None), "b200_flash_attn_ext_static")
    return dd.to_numpy(np.float32, (N, n_head, D)), need


def test_static_scratch_matches_workspace_entry_point_twice():
    host.require_gpu()
    assert lib().b200_flash_attn_static_workspace_max() == 8 << 20
    # two different shapes back to back, both split over the cache (they need the scratch), the second smaller than the first so
    # that it runs over the partials and counters the first one left in the array; then the first again
    # (8 heads over a 4 MB cache go to the scalar kernel, whose splits always merge through the global partials and counters)
    shapes = [(11, 32, 8, 128, 4096, 1), (13, 8, 8, 128, 1024, 1), (12, 16, 16, 64, 1500, 2), (11, 32, 8, 128, 4096, 1),
              (13, 8, 8, 128, 1024, 1)]
    for seed, nh, nhkv, D, nkv, N in shapes:
        q, k, v, mask = _problem(seed, nh, nhkv, D, nkv, N)
        scale = 1.0 / np.sqrt(D)
        want = host.flash_attn_ext(q, k, v, mask, R.F16, scale)          # existing entry point, caller-owned workspace
        got, need = _static_call(q, k, v, mask, scale)
        assert 0 < need <= 8 << 20, need                                  # the shape really uses scratch memory
        assert np.isfinite(got).all()
        assert np.abs(got - want).max() <= TOL, float(np.abs(got - want).max())
        f64 = O.flash_attn_f64(q, k, v, mask, R.F16, scale)
        assert np.abs(got - f64).max() <= TOL


def test_static_scratch_rejects_shapes_beyond_the_array():
    host.require_gpu()
    # 64 heads x 64 query columns over a 32k cache: the split partials alone exceed 8 MiB.  Nothing is launched, so the buffers
    # only have to be addresses (never dereferenced): tiny allocations stand in for them.
    nh, nhkv, D, nkv, N = 64, 64, 128, 32768, 24
    tiny = DeviceBuffer(256)
    a = FattnArgs(tiny.ptr, tiny.ptr, tiny.ptr, None, tiny.ptr, D, N, nh, nhkv, nkv, R.F16, R.F16,
                  4 * D, 4 * D * N, 2 * D, 2 * D * nkv, 2 * D, 2 * D * nkv, 0, 0.1, 0.0, 0.0)
    need = lib().b200_flash_attn_workspace_bytes(C.byref(a))
    if need <= 8 << 20:
        pytest.skip(f"plan for this shape needs only {need} bytes")
    rc = lib().b200_flash_attn_ext_static(C.byref(a), None)
    assert rc != 0
    assert b"static array holds" in lib().b200_last_error()
