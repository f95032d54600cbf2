"""ctypes binding for the REFERENCE ggml CPU backend built by oracle/Makefile into oracle/_ref/.

Test infrastructure only.  Mirrors the method of the reference's own quant tests
(gguf-py/tests/test_quants.py:41-101: load libggml via ctypes and compare bit-for-bit) and of
ggml_backend_compare_graph_backend (ggml/src/ggml-backend.cpp:2566-2603: run one node on the CPU
backend and hand the output to a checker).

Nothing here reads /root/reference at run time -- only the prebuilt shared objects.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
REF_DIR = ROOT / "oracle" / "_ref"

# ggml/include/ggml.h:361-398
F32, F16, Q4_0, Q8_0, Q8_1, Q4_K, Q5_K, Q6_K, Q8_K, I32 = 0, 1, 2, 8, 9, 12, 13, 14, 15, 26
Q4_1, Q5_0, Q5_1 = 3, 6, 7
Q2_K, Q3_K = 10, 11
IQ4_NL, IQ4_XS = 20, 23
TYPE_NAMES = {F32: "f32", F16: "f16", Q4_0: "q4_0", Q4_1: "q4_1", Q5_0: "q5_0", Q5_1: "q5_1", Q8_0: "q8_0", Q8_1: "q8_1", Q2_K: "q2_K", Q3_K: "q3_K", Q4_K: "q4_K", Q5_K: "q5_K", Q6_K: "q6_K", Q8_K: "q8_K", IQ4_NL: "iq4_nl", IQ4_XS: "iq4_xs"}
# (values per block, bytes per block)  ggml/src/ggml-common.h:144-335
BLOCK = {F32: (1, 4), F16: (1, 2), Q4_0: (32, 18), Q4_1: (32, 20), Q5_0: (32, 22), Q5_1: (32, 24), Q8_0: (32, 34), Q8_1: (32, 36), Q2_K: (256, 84), Q3_K: (256, 110), Q4_K: (256, 144), Q5_K: (256, 176),
         Q6_K: (256, 210), Q8_K: (256, 292), I32: (1, 4), IQ4_NL: (32, 18), IQ4_XS: (256, 136)}
WEIGHT_TYPES = (Q4_0, Q8_0, Q4_K, Q5_K, Q6_K, Q4_1, Q5_0, Q5_1, Q2_K, Q3_K, IQ4_NL, IQ4_XS)
ORACLE_WEIGHT_TYPES = WEIGHT_TYPES  # what the oracle restates (golden vectors); kept apart in case the CUDA path lags behind


def act_type(t: int) -> int:
    """vec_dot_type of a weight type on the CPU backend (ggml.c type_traits)."""
    return Q8_0 if t in (Q4_0, Q8_0, Q5_0, IQ4_NL) else Q8_1 if t in (Q4_1, Q5_1) else Q8_K


def row_size(t: int, ne: int) -> int:
    bs, by = BLOCK[t]
    assert ne % bs == 0
    return ne // bs * by


class Tensor(C.Structure):
    """struct ggml_tensor, ggml/include/ggml.h:584-619 (GGML_MAX_NAME = 128 in this fork, :225-227)."""
    _fields_ = [
        ("type", C.c_int), ("backend", C.c_int), ("buffer", C.c_void_p),
        ("ne", C.c_int64 * 4), ("nb", C.c_size_t * 4),
        ("op", C.c_int), ("op_params", C.c_int32 * 16), ("flags", C.c_int32),
        ("grad", C.c_void_p), ("src", C.c_void_p * 10),
        ("view_src", C.c_void_p), ("view_offs", C.c_size_t),
        ("data", C.c_void_p), ("name", C.c_char * 128), ("extra", C.c_void_p),
    ]


class InitParams(C.Structure):
    _fields_ = [("mem_size", C.c_size_t), ("mem_buffer", C.c_void_p), ("no_alloc", C.c_bool)]


def _cpu_flags() -> set[str]:
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def available(variant: str = "portable") -> bool:
    return _path(variant).exists()


def _path(variant: str) -> Path:
    return REF_DIR / {"portable": "libggml_ref.so", "avx512": "libggml_ref_avx512.so", "scalar": "libggml_ref_scalar.so"}[variant]


def best_variant() -> str:
    """Fastest variant this host can run (for the timed CPU baseline)."""
    fl = _cpu_flags()
    if {"avx512f", "avx512bw", "avx512vl", "avx512dq", "avx512cd", "avx512_vnni"} <= fl and available("avx512"):
        return "avx512"
    return "portable"


_LIBS: dict[str, "Ref"] = {}


def load(variant: str = "portable") -> "Ref":
    if variant not in _LIBS:
        _LIBS[variant] = Ref(variant)
    return _LIBS[variant]


class Ref:
    def __init__(self, variant: str):
        p = _path(variant)
        if not p.exists():
            raise FileNotFoundError(f"{p} missing: run `make -C oracle ref` where /root/reference exists")
        self.variant = variant
        L = self.lib = C.CDLL(str(p), mode=os.RTLD_LOCAL)
        P, I, I64, SZ, F = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_float
        TP = C.POINTER(Tensor)
        L.ggml_init.restype = P; L.ggml_init.argtypes = [InitParams]
        L.ggml_free.argtypes = [P]
        for n, k in (("ggml_new_tensor_1d", 1), ("ggml_new_tensor_2d", 2), ("ggml_new_tensor_3d", 3), ("ggml_new_tensor_4d", 4)):
            f = getattr(L, n); f.restype = TP; f.argtypes = [P, I] + [I64] * k
        L.ggml_mul_mat.restype = TP; L.ggml_mul_mat.argtypes = [P, TP, TP]
        L.ggml_mul_mat_id.restype = TP; L.ggml_mul_mat_id.argtypes = [P, TP, TP, TP]
        L.ggml_flash_attn_ext.restype = TP; L.ggml_flash_attn_ext.argtypes = [P, TP, TP, TP, TP, F, F, F]
        L.ggml_permute.restype = TP; L.ggml_permute.argtypes = [P, TP, I, I, I, I]
        L.ggml_view_3d.restype = TP; L.ggml_view_3d.argtypes = [P, TP, I64, I64, I64, SZ, SZ, SZ]
        L.ggml_new_graph.restype = P; L.ggml_new_graph.argtypes = [P]
        L.ggml_build_forward_expand.argtypes = [P, TP]
        L.ggml_graph_compute_with_ctx.restype = I; L.ggml_graph_compute_with_ctx.argtypes = [P, P, I]
        L.ggml_quantize_chunk.restype = SZ; L.ggml_quantize_chunk.argtypes = [I, P, P, I64, I64, I64, P]
        L.ggml_nbytes.restype = SZ; L.ggml_nbytes.argtypes = [TP]
        L.ggml_tensor_overhead.restype = SZ
        L.ggml_type_size.restype = SZ; L.ggml_type_size.argtypes = [I]
        L.ggml_blck_size.restype = I64; L.ggml_blck_size.argtypes = [I]
        for n in ("q4_0", "q4_1", "q5_0", "q5_1", "q8_0", "q2_K", "q3_K", "q4_K", "q5_K", "q6_K", "iq4_nl", "iq4_xs"):
            f = getattr(L, f"dequantize_row_{n}"); f.restype = None; f.argtypes = [P, P, I64]
        for n in ("q4_0", "q8_0", "q8_1", "q8_K"):
            f = getattr(L, f"quantize_row_{n}_ref"); f.restype = None; f.argtypes = [P, P, I64]
        for n in ("q4_0_q8_0", "q4_1_q8_1", "q5_0_q8_0", "q5_1_q8_1", "q8_0_q8_0", "q2_K_q8_K", "q3_K_q8_K", "q4_K_q8_K", "q5_K_q8_K",
                  "q6_K_q8_K", "iq4_nl_q8_0", "iq4_xs_q8_K"):
            f = getattr(L, f"ggml_vec_dot_{n}"); f.restype = None; f.argtypes = [I, P, SZ, P, SZ, P, SZ, I]
        L.ggml_fp32_to_fp16_row.argtypes = [P, P, I64]
        L.ggml_fp16_to_fp32_row.argtypes = [P, P, I64]
        # our struct mirror must match the library's sizeof(struct ggml_tensor) (+ 32-byte ggml_object)
        assert L.ggml_tensor_overhead() == C.sizeof(Tensor) + 32, (L.ggml_tensor_overhead(), C.sizeof(Tensor))
        # the fp16 lookup table used by GGML_FP16_TO_FP32 is filled by the first ggml_init (ggml.c:3719-3760)
        L.ggml_free(L.ggml_init(InitParams(1 << 20, None, False)))

    # ------------------------------------------------------------- flat helpers
    def quantize(self, t: int, x: np.ndarray) -> np.ndarray:
        """ggml_quantize_chunk(type, rows...) -> packed bytes [nrows, row_size] (ggml.c:22148)."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        nrows, k = (1, x.shape[0]) if x.ndim == 1 else (int(np.prod(x.shape[:-1])), x.shape[-1])
        out = np.empty((nrows, row_size(t, k)), dtype=np.uint8)
        n = self.lib.ggml_quantize_chunk(t, x.ctypes.data, out.ctypes.data, 0, nrows, k, None)
        assert n == out.nbytes
        return out

    def dequantize(self, t: int, blocks: np.ndarray, k: int) -> np.ndarray:
        blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
        out = np.empty(k, dtype=np.float32)
        getattr(self.lib, f"dequantize_row_{TYPE_NAMES[t]}")(blocks.ctypes.data, out.ctypes.data, k)
        return out

    def quantize_act_ref(self, t: int, x: np.ndarray) -> np.ndarray:
        """quantize_row_{q8_0,q8_1,q8_K,q4_0}_ref on one row."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        out = np.zeros(row_size(t, x.size), dtype=np.uint8)
        getattr(self.lib, f"quantize_row_{TYPE_NAMES[t]}_ref")(x.ctypes.data, out.ctypes.data, x.size)
        return out

    def vec_dot(self, wt: int, w: np.ndarray, act: np.ndarray, n: int) -> float:
        at = TYPE_NAMES[act_type(wt)]
        s = C.c_float(0)
        getattr(self.lib, f"ggml_vec_dot_{TYPE_NAMES[wt]}_{at}")(n, C.byref(s), 0, w.ctypes.data, 0, act.ctypes.data, 0, 1)
        return s.value

    # ------------------------------------------------------------- graph helpers
    def _ctx(self, nbytes: int):
        return self.lib.ggml_init(InitParams(nbytes, None, False))

    @staticmethod
    def _fill(t, arr: np.ndarray):
        arr = np.ascontiguousarray(arr)
        C.memmove(t.contents.data, arr.ctypes.data, arr.nbytes)

    def _run(self, ctx, out, nthreads: int) -> None:
        gf = self.lib.ggml_new_graph(ctx)
        self.lib.ggml_build_forward_expand(gf, out)
        st = self.lib.ggml_graph_compute_with_ctx(ctx, gf, nthreads)
        assert st == 0, st

    def mul_mat(self, wt: int, w: np.ndarray, x: np.ndarray, K: int, M: int, nthreads: int = 1) -> np.ndarray:
        """dst[N, M] = ggml_mul_mat(W[K,M] quantised, X[K,N] f32) on the CPU backend (ggml.c:12486)."""
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, K)
        N = x.shape[0]
        ctx = self._ctx(w.nbytes + x.nbytes + 4 * M * N + (64 << 20))
        try:
            a = self.lib.ggml_new_tensor_2d(ctx, wt, K, M)
            b = self.lib.ggml_new_tensor_2d(ctx, F32, K, N)
            assert self.lib.ggml_nbytes(a) == w.nbytes
            self._fill(a, w); self._fill(b, x)
            o = self.lib.ggml_mul_mat(ctx, a, b)
            self._run(ctx, o, nthreads)
            return np.ctypeslib.as_array(C.cast(o.contents.data, C.POINTER(C.c_float)), (N, M)).copy()
        finally:
            self.lib.ggml_free(ctx)

    def mul_mat_id(self, wt: int, w: np.ndarray, ids: np.ndarray, b: np.ndarray, K: int, M: int, n_expert: int, nthreads: int = 1) -> np.ndarray:
        """dst[T, n_used, M] = ggml_mul_mat_id(as[K,M,n_expert], b[K, n_b_rows, T], ids[n_used, T]) (ggml.c:5693-5729)."""
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        T, n_used = ids.shape
        b = np.ascontiguousarray(b, dtype=np.float32).reshape(T, -1, K)
        nbr = b.shape[1]
        ctx = self._ctx(w.nbytes + b.nbytes + 4 * M * n_used * T + (64 << 20))
        try:
            a = self.lib.ggml_new_tensor_3d(ctx, wt, K, M, n_expert)
            bb = self.lib.ggml_new_tensor_3d(ctx, F32, K, nbr, T)
            ii = self.lib.ggml_new_tensor_2d(ctx, I32, n_used, T)
            self._fill(a, w); self._fill(bb, b); self._fill(ii, ids)
            o = self.lib.ggml_mul_mat_id(ctx, a, bb, ii)
            self._run(ctx, o, nthreads)
            return np.ctypeslib.as_array(C.cast(o.contents.data, C.POINTER(C.c_float)), (T, n_used, M)).copy()
        finally:
            self.lib.ggml_free(ctx)

    def flash_attn_ext(self, q: np.ndarray, k: np.ndarray, v: np.ndarray, mask: np.ndarray | None, type_kv: int,
                       scale: float, max_bias: float = 0.0, softcap: float = 0.0, nthreads: int = 1) -> np.ndarray:
        """q f32 [n_head, N, D]; k, v: packed rows [n_head_kv, n_kv, row_bytes] of `type_kv`; mask f16 [N_pad, n_kv].
        Returns dst f32 [N, n_head, D] (ggml.c:15667-15877; constructor ggml.c:7107-7160)."""
        n_head, N, D = q.shape
        n_head_kv, n_kv = k.shape[0], k.shape[1]
        ctx = self._ctx(q.nbytes + k.nbytes + v.nbytes + (mask.nbytes if mask is not None else 0) + 4 * q.size + (64 << 20))
        try:
            tq = self.lib.ggml_new_tensor_3d(ctx, F32, D, N, n_head)
            tk = self.lib.ggml_new_tensor_3d(ctx, type_kv, D, n_kv, n_head_kv)
            tv = self.lib.ggml_new_tensor_3d(ctx, type_kv, D, n_kv, n_head_kv)
            self._fill(tq, q.astype(np.float32)); self._fill(tk, k); self._fill(tv, v)
            tm = None
            if mask is not None:
                tm = self.lib.ggml_new_tensor_2d(ctx, F16, n_kv, mask.shape[0])
                self._fill(tm, mask.astype(np.float16))
            o = self.lib.ggml_flash_attn_ext(ctx, tq, tk, tv, tm, scale, max_bias, softcap)
            self._run(ctx, o, nthreads)
            return np.ctypeslib.as_array(C.cast(o.contents.data, C.POINTER(C.c_float)), (N, n_head, D)).copy()
        finally:
            self.lib.ggml_free(ctx)
