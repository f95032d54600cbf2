"""ctypes binding for oracle/libggml_oracle.so (the C restatement; test infrastructure only)."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
ORACLE_DIR = ROOT / "oracle"
F32, F16, Q4_0, Q8_0, Q8_1, Q4_K, Q5_K, Q6_K, Q8_K = 0, 1, 2, 8, 9, 12, 13, 14, 15
Q4_1, Q5_0, Q5_1 = 3, 6, 7
Q2_K, Q3_K = 10, 11
IQ4_NL, IQ4_XS = 20, 23


class FattnParams(C.Structure):
    _fields_ = [("D", C.c_int64), ("N", C.c_int64), ("n_head", C.c_int64), ("n_head_kv", C.c_int64), ("n_kv", C.c_int64),
                ("type_k", C.c_int), ("type_v", C.c_int),
                ("q_nb1", C.c_size_t), ("q_nb2", C.c_size_t), ("k_nb1", C.c_size_t), ("k_nb2", C.c_size_t),
                ("v_nb1", C.c_size_t), ("v_nb2", C.c_size_t), ("mask_nb1", C.c_size_t),
                ("scale", C.c_float), ("max_bias", C.c_float), ("logit_softcap", C.c_float)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        so = ORACLE_DIR / "libggml_oracle.so"
        src = ORACLE_DIR / "ggml_oracle.c"
        if not so.exists() or so.stat().st_mtime < src.stat().st_mtime:
            subprocess.run(["make", "-C", str(ORACLE_DIR), "oracle"], check=True, capture_output=True)
        L = C.CDLL(str(so))
        P, I, I64, SZ = C.c_void_p, C.c_int, C.c_int64, C.c_size_t
        L.orc_blck_size.argtypes = [I]
        L.orc_type_size.restype = SZ; L.orc_type_size.argtypes = [I]
        L.orc_row_size.restype = SZ; L.orc_row_size.argtypes = [I, I64]
        L.orc_fp16_to_fp32.restype = C.c_float; L.orc_fp16_to_fp32.argtypes = [C.c_uint16]
        L.orc_fp32_to_fp16.restype = C.c_uint16; L.orc_fp32_to_fp16.argtypes = [C.c_float]
        L.orc_dequantize_row.argtypes = [I, P, P, I64]
        L.orc_quantize_row.argtypes = [I, P, P, I64]
        L.orc_vec_dot.restype = C.c_float; L.orc_vec_dot.argtypes = [I, I64, P, P]
        L.orc_vec_dot_q8_1.restype = C.c_double; L.orc_vec_dot_q8_1.argtypes = [I, I64, P, P]
        L.orc_mul_mat.argtypes = [I, P, SZ, P, P, I64, I64, I64, I]
        L.orc_mul_mat_q8_1.argtypes = [I, P, SZ, P, P, I64, I64, I64, I]
        L.orc_mul_mat_id.argtypes = [I, P, SZ, SZ, P, P, P, I64, I64, I64, I64, I64, I64, I]
        L.orc_flash_attn_ext.argtypes = [C.POINTER(FattnParams), P, P, P, P, P, I]
        L.orc_flash_attn_f64.argtypes = [C.POINTER(FattnParams), P, P, P, P, P, I]
        _lib = L
    return _lib


def row_size(t: int, ne: int) -> int:
    return lib().orc_row_size(t, ne)


def dequantize(t: int, blocks: np.ndarray, k: int) -> np.ndarray:
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
    out = np.empty(k, dtype=np.float32)
    assert lib().orc_dequantize_row(t, blocks.ctypes.data, out.ctypes.data, k) == 0
    return out


def quantize(t: int, x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.zeros(row_size(t, x.size), dtype=np.uint8)
    assert lib().orc_quantize_row(t, x.ctypes.data, out.ctypes.data, x.size) == 0
    return out


def quantize_rows(t: int, x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32)
    return np.stack([quantize(t, r) for r in x.reshape(-1, x.shape[-1])])


def vec_dot(wt: int, w: np.ndarray, act: np.ndarray, n: int) -> float:
    return lib().orc_vec_dot(wt, n, w.ctypes.data, act.ctypes.data)


def vec_dot_q8_1(wt: int, w: np.ndarray, act: np.ndarray, n: int) -> float:
    return lib().orc_vec_dot_q8_1(wt, n, w.ctypes.data, act.ctypes.data)


def mul_mat(wt: int, w: np.ndarray, x: np.ndarray, K: int, M: int, nthreads: int = 0) -> np.ndarray:
    w = np.ascontiguousarray(w, dtype=np.uint8)
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, K)
    out = np.empty((x.shape[0], M), dtype=np.float32)
    assert lib().orc_mul_mat(wt, w.ctypes.data, row_size(wt, K), x.ctypes.data, out.ctypes.data, K, M, x.shape[0], nthreads) == 0
    return out


def mul_mat_q8_1(wt: int, w: np.ndarray, x: np.ndarray, K: int, M: int, nthreads: int = 0) -> np.ndarray:
    w = np.ascontiguousarray(w, dtype=np.uint8)
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, K)
    out = np.empty((x.shape[0], M), dtype=np.float64)
    assert lib().orc_mul_mat_q8_1(wt, w.ctypes.data, row_size(wt, K), x.ctypes.data, out.ctypes.data, K, M, x.shape[0], nthreads) == 0
    return out


def mul_mat_id(wt: int, w: np.ndarray, ids: np.ndarray, b: np.ndarray, K: int, M: int, n_expert: int, nthreads: int = 0) -> np.ndarray:
    w = np.ascontiguousarray(w, dtype=np.uint8)
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    T, n_used = ids.shape
    b = np.ascontiguousarray(b, dtype=np.float32).reshape(T, -1, K)
    out = np.empty((T, n_used, M), dtype=np.float32)
    rs = row_size(wt, K)
    rc = lib().orc_mul_mat_id(wt, w.ctypes.data, rs, rs * M, ids.ctypes.data, b.ctypes.data, out.ctypes.data,
                              K, M, n_expert, n_used, b.shape[1], T, nthreads)
    assert rc == 0, rc
    return out


def _fa_params(q, k, v, mask, type_kv, scale, max_bias, softcap) -> FattnParams:
    n_head, N, D = q.shape
    n_head_kv, n_kv, rb = k.shape
    return FattnParams(D, N, n_head, n_head_kv, n_kv, type_kv, type_kv,
                       4 * D, 4 * D * N, rb, rb * n_kv, rb, rb * n_kv, 0 if mask is None else 2 * n_kv,
                       scale, max_bias, softcap)


def flash_attn_ext(q, k, v, mask, type_kv, scale, max_bias=0.0, softcap=0.0, nthreads=0) -> np.ndarray:
    """Same layouts as tests/ggml_ref.py::Ref.flash_attn_ext."""
    q = np.ascontiguousarray(q, dtype=np.float32); k = np.ascontiguousarray(k); v = np.ascontiguousarray(v)
    m = None if mask is None else np.ascontiguousarray(mask, dtype=np.float16)
    p = _fa_params(q, k, v, m, type_kv, scale, max_bias, softcap)
    out = np.empty((q.shape[1], q.shape[0], q.shape[2]), dtype=np.float32)
    rc = lib().orc_flash_attn_ext(C.byref(p), q.ctypes.data, k.ctypes.data, v.ctypes.data,
                                  None if m is None else m.ctypes.data, out.ctypes.data, nthreads)
    assert rc == 0, rc
    return out


def flash_attn_f64(q, k, v, mask, type_kv, scale, max_bias=0.0, softcap=0.0, nthreads=0) -> np.ndarray:
    q = np.ascontiguousarray(q, dtype=np.float32); k = np.ascontiguousarray(k); v = np.ascontiguousarray(v)
    m = None if mask is None else np.ascontiguousarray(mask, dtype=np.float16)
    p = _fa_params(q, k, v, m, type_kv, scale, max_bias, softcap)
    out = np.empty((q.shape[1], q.shape[0], q.shape[2]), dtype=np.float64)
    rc = lib().orc_flash_attn_f64(C.byref(p), q.ctypes.data, k.ctypes.data, v.ctypes.data,
                                  None if m is None else m.ctypes.data, out.ctypes.data, nthreads)
    assert rc == 0, rc
    return out


def pack_kv(type_kv: int, x: np.ndarray) -> np.ndarray:
    """f32 [n_head_kv, n_kv, D] -> packed rows uint8 [n_head_kv, n_kv, row_bytes] (f16 or q8_0 KV cache rows)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    if type_kv == F16:
        return x.astype(np.float16).view(np.uint8).reshape(x.shape[0], x.shape[1], -1)
    return quantize_rows(type_kv, x).reshape(x.shape[0], x.shape[1], -1)
