"""Host-buffer convenience layer over the C ABI (numpy in, numpy out).

Every call allocates device memory, copies the host inputs in, runs the CUDA path and copies the result back --
this is the end-to-end shape the tests use.  bench.py keeps buffers resident and times the kernels itself.
There is no CPU fallback anywhere in this module.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import lib as L
from .lib import F16, F32, Q8_0, Q8_1, FattnArgs, check, lib, row_size


class DeviceBuffer:
    """cudaMalloc'ed bytes with explicit lifetime."""

    def __init__(self, nbytes: int, zero: bool = False):
        self.nbytes = int(nbytes)
        p = C.c_void_p()
        check(lib().b200_malloc(C.byref(p), max(self.nbytes, 1)), "b200_malloc")
        self.ptr = p.value
        if zero:
            check(lib().b200_memset(self.ptr, 0, max(self.nbytes, 1), None), "b200_memset")

    @classmethod
    def from_numpy(cls, a: np.ndarray, pad: int = 0) -> "DeviceBuffer":
        a = np.ascontiguousarray(a)
        b = cls(a.nbytes + pad)
        if a.nbytes:
            check(lib().b200_memcpy_h2d(b.ptr, a.ctypes.data, a.nbytes, None), "h2d")
        return b

    def to_numpy(self, dtype, shape) -> np.ndarray:
        out = np.empty(shape, dtype=dtype)
        assert out.nbytes <= self.nbytes, (out.nbytes, self.nbytes)
        sync()
        if out.nbytes:
            check(lib().b200_memcpy_d2h(out.ctypes.data, self.ptr, out.nbytes, None), "d2h")
            sync()
        return out

    def free(self):
        if self.ptr:
            lib().b200_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def sync():
    check(lib().b200_device_synchronize(), "sync")


def require_gpu():
    if lib().b200_device_count() < 1:
        raise L.B200Error("no CUDA device: the B200 path has no CPU fallback")


def dequantize(t: int, blocks: np.ndarray, n: int, dst_type: int = F32) -> np.ndarray:
    require_gpu()
    src = DeviceBuffer.from_numpy(np.ascontiguousarray(blocks, dtype=np.uint8), pad=16)
    dst = DeviceBuffer(n * (4 if dst_type == F32 else 2))
    check(lib().b200_dequantize(t, src.ptr, dst.ptr, dst_type, n, None), "b200_dequantize")
    return dst.to_numpy(np.float32 if dst_type == F32 else np.float16, (n,))


def quantize_q8_1(x: np.ndarray) -> np.ndarray:
    """x f32 [N, K] -> uint8 [N, K/32*36] (block_q8_1 rows)."""
    require_gpu()
    x = np.ascontiguousarray(x, dtype=np.float32)
    x2 = x.reshape(-1, x.shape[-1])
    N, K = x2.shape
    src = DeviceBuffer.from_numpy(x2)
    dst = DeviceBuffer(N * row_size(Q8_1, K))
    check(lib().b200_quantize_q8_1(src.ptr, K, dst.ptr, K, N, None), "b200_quantize_q8_1")
    return dst.to_numpy(np.uint8, (N, row_size(Q8_1, K)))


def mul_mat_vec_q(t: int, w: np.ndarray, x: np.ndarray, K: int, M: int) -> np.ndarray:
    """w: packed rows uint8 [M, row_size]; x f32 [N, K] -> f32 [N, M]."""
    require_gpu()
    w = np.ascontiguousarray(w, dtype=np.uint8)
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, K)
    N = x.shape[0]
    dw, dx, dd = DeviceBuffer.from_numpy(w, pad=16), DeviceBuffer.from_numpy(x), DeviceBuffer(4 * N * M)
    check(lib().b200_mul_mat_vec_q(t, dw.ptr, row_size(t, K), dx.ptr, K, dd.ptr, M, K, M, N, None), "b200_mul_mat_vec_q")
    return dd.to_numpy(np.float32, (N, M))


def mul_mat_vec_q_multi(t: int, ws: list[np.ndarray], x: np.ndarray, K: int, types: list[int] | None = None) -> list[np.ndarray]:
    """Several weight matrices against the same activations in one launch (fused Q/K/V, gate/up).  types: per-matrix weight type
    when the launch mixes two K-quant types (b200_gemv_job.type2); t is then the type of ws[0]."""
    require_gpu()
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, K)
    N = x.shape[0]
    dx = DeviceBuffer.from_numpy(x)
    dws = [DeviceBuffer.from_numpy(np.ascontiguousarray(w, dtype=np.uint8), pad=16) for w in ws]
    Ms = [w.shape[0] for w in ws]
    # all outputs share one row stride so a single dst_stride describes them
    dd = DeviceBuffer(4 * N * sum(Ms))
    jobs = (L.GemvJob * len(ws))()
    off = 0
    for i, (dw, M) in enumerate(zip(dws, Ms)):
        jobs[i].W, jobs[i].X, jobs[i].dst, jobs[i].M, jobs[i].id_index = dw.ptr, dx.ptr, dd.ptr + 4 * off, M, -1
        jobs[i]._pad = types[i] if types is not None and types[i] != t else 0
        off += M
    tot = sum(Ms)
    check(lib().b200_mul_mat_vec_q_multi(t, jobs, len(ws), row_size(t, K), K, tot, K, N, None, 0, 0, None), "b200_mul_mat_vec_q_multi")
    out = dd.to_numpy(np.float32, (N, tot))
    res, off = [], 0
    for M in Ms:
        res.append(out[:, off:off + M].copy())
        off += M
    return res


def mul_mat_id_vec_q(t: int, w: np.ndarray, ids: np.ndarray, b: np.ndarray, K: int, M: int, n_expert: int) -> np.ndarray:
    """as: uint8 [n_expert*M, row_size]; ids i32 [T, n_used]; b f32 [T, n_b_rows, K] -> f32 [T, n_used, M]."""
    require_gpu()
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    T, n_used = ids.shape
    b = np.ascontiguousarray(b, dtype=np.float32).reshape(T, -1, K)
    nbr = b.shape[1]
    rs = row_size(t, K)
    dw, di, db = DeviceBuffer.from_numpy(np.ascontiguousarray(w, dtype=np.uint8), pad=16), DeviceBuffer.from_numpy(ids), DeviceBuffer.from_numpy(b)
    dd = DeviceBuffer(4 * T * n_used * M)
    check(lib().b200_mul_mat_id_vec_q(t, dw.ptr, rs, rs * M, n_expert, di.ptr, n_used, db.ptr, K, nbr * K,
                                      dd.ptr, M, n_used * M, K, M, n_used, nbr, T, None), "b200_mul_mat_id_vec_q")
    return dd.to_numpy(np.float32, (T, n_used, M))


def mul_mat_id_q(t: int, w: np.ndarray, ids: np.ndarray, b: np.ndarray, K: int, M: int, n_expert: int) -> np.ndarray:
    """Batched MUL_MAT_ID on the tensor cores (b200_mul_mat_id_q); same layouts as mul_mat_id_vec_q.  dst is pre-filled with NaN so
    that rows the kernel must not touch (invalid expert ids) can be told apart."""
    require_gpu()
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    T, n_used = ids.shape
    b = np.ascontiguousarray(b, dtype=np.float32).reshape(T, -1, K)
    nbr = b.shape[1]
    rs = row_size(t, K)
    dw, di, db = DeviceBuffer.from_numpy(np.ascontiguousarray(w, dtype=np.uint8), pad=16), DeviceBuffer.from_numpy(ids), DeviceBuffer.from_numpy(b)
    dd = DeviceBuffer.from_numpy(np.full((T, n_used, M), np.nan, dtype=np.float32))
    ws_bytes = lib().b200_mul_mat_id_q_workspace_bytes(t, K, M, n_used, T, n_expert)
    ws = DeviceBuffer(ws_bytes + 256)
    check(lib().b200_mul_mat_id_q(t, dw.ptr, rs, rs * M, n_expert, di.ptr, n_used, db.ptr, K, nbr * K, dd.ptr, M, n_used * M, K, M, n_used, nbr, T,
                                  ws.ptr, ws_bytes, None), "b200_mul_mat_id_q")
    return dd.to_numpy(np.float32, (T, n_used, M))


def flash_attn_ext(q: np.ndarray, k: np.ndarray, v: np.ndarray, mask: np.ndarray | None, type_kv: int, scale: float,
                   max_bias: float = 0.0, softcap: float = 0.0) -> np.ndarray:
    """q f32 [n_head, N, D]; k, v packed rows uint8 [n_head_kv, n_kv, row_bytes]; mask f16 [N_pad, n_kv] -> f32 [N, n_head, D]."""
    require_gpu()
    q = np.ascontiguousarray(q, dtype=np.float32)
    n_head, N, D = q.shape
    n_head_kv, n_kv, rb = k.shape
    dq, dk, dv = DeviceBuffer.from_numpy(q), DeviceBuffer.from_numpy(np.ascontiguousarray(k), pad=64), DeviceBuffer.from_numpy(np.ascontiguousarray(v), pad=64)
    dm = None if mask is None else DeviceBuffer.from_numpy(np.ascontiguousarray(mask, dtype=np.float16))
    dd = DeviceBuffer(4 * N * n_head * D)
    a = FattnArgs(dq.ptr, dk.ptr, dv.ptr, None if dm is None else dm.ptr, dd.ptr, D, N, n_head, n_head_kv, n_kv, type_kv, type_kv,
                  4 * D, 4 * D * N, rb, rb * n_kv, rb, rb * n_kv, 0 if mask is None else 2 * n_kv, scale, max_bias, softcap)
    ws_bytes = lib().b200_flash_attn_workspace_bytes(C.byref(a))
    ws = DeviceBuffer(max(ws_bytes, 256), zero=True)
    check(lib().b200_flash_attn_ext(C.byref(a), ws.ptr, ws.nbytes, None), "b200_flash_attn_ext")
    out = dd.to_numpy(np.float32, (N, n_head, D))
    # the kernel must leave its ticket counters zeroed for the next launch
    return out


def mul_mat_q(t: int, w: np.ndarray, x: np.ndarray, K: int, M: int) -> np.ndarray:
    """Prefill GEMM on the tensor cores (b200_mul_mat_q): w packed rows [M, row_bytes], x f32 [N, K] -> f32 [N, M]."""
    require_gpu()
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, K)
    N = x.shape[0]
    rs = row_size(t, K)
    dw = DeviceBuffer.from_numpy(np.ascontiguousarray(w).reshape(-1), pad=256)
    dx = DeviceBuffer.from_numpy(x)
    dd = DeviceBuffer(4 * N * M + 256)
    ws_bytes = lib().b200_mul_mat_q_workspace_bytes(t, K, M, N)
    ws = DeviceBuffer(ws_bytes + 16)
    check(lib().b200_mul_mat_q(t, dw.ptr, rs, dx.ptr, K, dd.ptr, M, K, M, N, ws.ptr, ws_bytes, None), "b200_mul_mat_q")
    return dd.to_numpy(np.float32, (N, M))
