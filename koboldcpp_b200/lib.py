"""ctypes loader for libkcpp_b200.so (include/b200_ggml.h)."""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "libkcpp_b200.so"

# enum ggml_type values (reference ggml/include/ggml.h:361-398)
F32, F16, Q4_0, Q8_0, Q8_1, Q4_K, Q5_K, Q6_K, I32 = 0, 1, 2, 8, 9, 12, 13, 14, 26
Q4_1, Q5_0, Q5_1 = 3, 6, 7
Q2_K, Q3_K = 10, 11
IQ4_NL, IQ4_XS = 20, 23
BLOCK = {F32: (1, 4), F16: (1, 2), Q4_0: (32, 18), Q4_1: (32, 20), Q5_0: (32, 22), Q5_1: (32, 24), Q8_0: (32, 34), Q8_1: (32, 36), Q2_K: (256, 84), Q3_K: (256, 110), Q4_K: (256, 144), Q5_K: (256, 176), Q6_K: (256, 210), I32: (1, 4), IQ4_NL: (32, 18), IQ4_XS: (256, 136)}
TYPE_NAMES = {F32: "f32", F16: "f16", Q4_0: "q4_0", Q4_1: "q4_1", Q5_0: "q5_0", Q5_1: "q5_1", Q8_0: "q8_0", Q8_1: "q8_1", Q2_K: "q2_K", Q3_K: "q3_K", Q4_K: "q4_K", Q5_K: "q5_K", Q6_K: "q6_K", I32: "i32", IQ4_NL: "iq4_nl", IQ4_XS: "iq4_xs"}


def row_size(t: int, ne: int) -> int:
    bs, by = BLOCK[t]
    assert ne % bs == 0, (t, ne)
    return ne // bs * by


class B200Error(RuntimeError):
    pass


class GemvJob(C.Structure):
    _fields_ = [("W", C.c_void_p), ("X", C.c_void_p), ("dst", C.c_void_p), ("M", C.c_int64), ("id_index", C.c_int32), ("_pad", C.c_int32)]


class GemvFuse(C.Structure):
    _fields_ = [("norm_weight", C.c_void_p), ("norm_eps", C.c_float), ("silu_pair", C.c_int32), ("residual", C.c_void_p * 16),
                ("raw_dst", C.c_void_p * 16), ("norm_out", C.c_void_p), ("norm_mul_out", C.c_void_p), ("gate_out", C.c_void_p), ("up_out", C.c_void_p),
                ("silu_out", C.c_void_p)]


class Tensor(C.Structure):
    """b200_tensor (include/b200_ggml_ops.h): ggml-style element counts and byte strides."""
    _fields_ = [("data", C.c_void_p), ("type", C.c_int32), ("_pad", C.c_int32), ("ne", C.c_int64 * 4), ("nb", C.c_size_t * 4)]

    @classmethod
    def contiguous(cls, ptr: int, t: int, ne) -> "Tensor":
        ne = list(ne) + [1] * (4 - len(ne))
        bs, by = BLOCK[t]
        nb = [by, by * (ne[0] // bs), by * (ne[0] // bs) * ne[1], by * (ne[0] // bs) * ne[1] * ne[2]]
        return cls(ptr, t, 0, (C.c_int64 * 4)(*ne), (C.c_size_t * 4)(*nb))


class RopeParams(C.Structure):
    _fields_ = [("n_dims", C.c_int32), ("mode", C.c_int32), ("n_ctx_orig", C.c_int32), ("freq_base", C.c_float), ("freq_scale", C.c_float),
                ("ext_factor", C.c_float), ("attn_factor", C.c_float), ("beta_fast", C.c_float), ("beta_slow", C.c_float)]


class FattnArgs(C.Structure):
    _fields_ = [("q", C.c_void_p), ("k", C.c_void_p), ("v", C.c_void_p), ("mask", C.c_void_p), ("dst", C.c_void_p),
                ("D", C.c_int64), ("N", C.c_int64), ("n_head", C.c_int64), ("n_head_kv", C.c_int64), ("n_kv", C.c_int64),
                ("type_k", C.c_int32), ("type_v", C.c_int32),
                ("q_nb1", C.c_size_t), ("q_nb2", C.c_size_t), ("k_nb1", C.c_size_t), ("k_nb2", C.c_size_t),
                ("v_nb1", C.c_size_t), ("v_nb2", C.c_size_t), ("mask_nb1", C.c_size_t),
                ("scale", C.c_float), ("max_bias", C.c_float), ("logit_softcap", C.c_float)]


_SIGS = {
    # name: (restype, argtypes)
    "b200_last_error": (C.c_char_p, []),
    "b200_version": (C.c_char_p, []),
    "b200_set_pdl": (None, [C.c_int]),
    "b200_device_count": (C.c_int, []),
    "b200_set_device": (C.c_int, [C.c_int]),
    "b200_device_info": (C.c_int, [C.c_int, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "b200_malloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "b200_free": (C.c_int, [C.c_void_p]),
    "b200_malloc_host": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "b200_free_host": (C.c_int, [C.c_void_p]),
    "b200_memcpy_h2d": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200_memcpy_d2h": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200_memcpy_d2d": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200_memset": (C.c_int, [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p]),
    "b200_stream_create": (C.c_int, [C.POINTER(C.c_void_p)]),
    "b200_stream_destroy": (C.c_int, [C.c_void_p]),
    "b200_stream_synchronize": (C.c_int, [C.c_void_p]),
    "b200_device_synchronize": (C.c_int, []),
    "b200_event_create": (C.c_int, [C.POINTER(C.c_void_p)]),
    "b200_event_destroy": (C.c_int, [C.c_void_p]),
    "b200_event_record": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200_event_elapsed_ms": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_float)]),
    "b200_graph_begin_capture": (C.c_int, [C.c_void_p]),
    "b200_graph_end_capture": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "b200_graph_launch": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200_graph_destroy": (C.c_int, [C.c_void_p]),
    "b200_dequantize": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_void_p]),
    "b200_quantize_q8_1": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]),
    "b200_mul_mat_vec_q": (C.c_int, [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]),
    "b200_mul_mat_vec_q_multi": (C.c_int, [C.c_int, C.POINTER(GemvJob), C.c_int, C.c_size_t, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_size_t, C.c_int32, C.c_void_p]),
    "b200_mul_mat_vec_q_fused": (C.c_int, [C.c_int, C.POINTER(GemvJob), C.c_int, C.c_size_t, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.POINTER(GemvFuse), C.c_void_p]),
    "b200_mul_mat_id_vec_q": (C.c_int, [C.c_int, C.c_void_p, C.c_size_t, C.c_size_t, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64,
                                        C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]),
    "b200_mul_mat_id_q_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int32]),
    "b200_mul_mat_id_q": (C.c_int, [C.c_int, C.c_void_p, C.c_size_t, C.c_size_t, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64,
                                    C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200_mul_mat_q_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int64, C.c_int64, C.c_int64]),
    "b200_mul_mat_q": (C.c_int, [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200_flash_attn_workspace_bytes": (C.c_size_t, [C.POINTER(FattnArgs)]),
    "b200_flash_attn_ext": (C.c_int, [C.POINTER(FattnArgs), C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200_flash_attn_static_workspace_max": (C.c_size_t, []),
    "b200_flash_attn_ext_static": (C.c_int, [C.POINTER(FattnArgs), C.c_void_p]),
}

# include/b200_ggml_ops.h (the subset the tests drive directly; the rest is reached through the backend)
_OPS_SIGS = {
    "b200_op_rope": (C.c_int, [C.POINTER(Tensor), C.c_void_p, C.c_void_p, C.POINTER(Tensor), C.POINTER(RopeParams), C.c_void_p]),
    "b200_op_cpy": (C.c_int, [C.POINTER(Tensor), C.POINTER(Tensor), C.c_void_p]),
    "b200_op_rope_kv_store": (C.c_int, [C.POINTER(Tensor), C.POINTER(Tensor), C.POINTER(Tensor), C.POINTER(Tensor), C.POINTER(Tensor), C.c_void_p, C.c_void_p,
                                        C.POINTER(RopeParams), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200_op_moe_router": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                                      C.c_int64, C.c_void_p]),
    "b200_op_silu_mul": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "b200_op_rms_norm_mul": (C.c_int, [C.POINTER(Tensor), C.c_void_p, C.c_void_p, C.POINTER(Tensor), C.c_float, C.c_void_p]),
}

_lib = None


def load(path: Path | None = None):
    """Load the shared library and declare every symbol of include/b200_ggml.h.  Raises if it is missing:
    the product has no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    p = Path(path) if path else Path(os.environ.get("B200_LIB", LIB_PATH))  # B200_LIB: developer A/B of two builds
    if not p.exists():
        raise B200Error(f"{p} not found: build it with `python -m koboldcpp_b200.build` (nvcc, sm_100a)")
    L = C.CDLL(str(p))
    for name, (res, args) in list(_SIGS.items()) + list(_OPS_SIGS.items()):
        f = getattr(L, name)  # AttributeError if the library does not export a declared symbol
        f.restype = res
        f.argtypes = args
    _lib = L
    return L


def lib():
    return load()


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise B200Error(f"{what or 'b200 call'} failed with status {rc}: {lib().b200_last_error().decode()}")
