"""Python mirror of include/b200_llama.h: the llama / mixtral decoder that drives the B200 backend through the
ggml_backend vtables, plus synthetic (random-init) GGUF-shaped models for the configurations BASELINE.json names.

The type mix per tensor follows llama_tensor_get_type for Q4_0 / Q4_K_M / Q5_K_M (/root/reference/src/llama.cpp:17979-18200).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from .lib import BLOCK, F16, F32, IQ4_NL, IQ4_XS, Q2_K, Q3_K, Q4_0, Q4_1, Q4_K, Q5_0, Q5_1, Q5_K, Q6_K, Q8_0, B200Error, check, lib, row_size


class LlamaParams(C.Structure):
    _fields_ = [("n_embd", C.c_int32), ("n_layer", C.c_int32), ("n_head", C.c_int32), ("n_head_kv", C.c_int32), ("n_embd_head", C.c_int32),
                ("n_ff", C.c_int32), ("n_vocab", C.c_int32), ("n_expert", C.c_int32), ("n_expert_used", C.c_int32), ("n_ctx", C.c_int32),
                ("type_k", C.c_int32), ("type_v", C.c_int32), ("rope_neox", C.c_int32), ("n_ctx_orig", C.c_int32),
                ("rms_eps", C.c_float), ("rope_freq_base", C.c_float), ("rope_freq_scale", C.c_float),
                ("main_device", C.c_int32), ("n_split_devices", C.c_int32), ("tensor_split", C.c_float * 16),
                ("tp_rank", C.c_int32), ("tp_world", C.c_int32)]


_SIGS = {
    "b200_llama_new": (C.c_int, [C.POINTER(LlamaParams), C.POINTER(C.c_void_p)]),
    "b200_llama_free": (None, [C.c_void_p]),
    "b200_llama_set_tensor": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]),
    "b200_llama_clone_tensor": (C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p]),
    "b200_llama_finalize": (C.c_int, [C.c_void_p, C.c_int32]),
    "b200_llama_kv_write": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "b200_llama_kv_read": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "b200_llama_decode": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32]),
    "b200_llama_decode_async": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32]),
    "b200_llama_synchronize": (C.c_int, [C.c_void_p]),
    "b200_llama_replay": (C.c_int, [C.c_void_p]),
    "b200_llama_stream": (C.c_void_p, [C.c_void_p]),
    "b200_llama_stats": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "ggml_backend_b200_set_graphs": (None, [C.c_int]),
    "ggml_backend_b200_set_fusion": (None, [C.c_int]),
    "ggml_backend_b200_get_stats": (None, [C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "ggml_backend_b200_reset_stats": (None, []),
    "ggml_backend_b200_get_device_count": (C.c_int, []),
}
_declared = False


def _l():
    global _declared
    L = lib()
    if not _declared:
        for name, (res, args) in _SIGS.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _declared = True
    return L


@dataclass
class ModelShape:
    """Hyper-parameters + quantisation mix of one BASELINE.json configuration."""
    name: str
    n_embd: int
    n_layer: int
    n_head: int
    n_head_kv: int
    n_ff: int
    n_vocab: int
    ftype: str = "Q4_K_M"  # Q4_0 | Q4_1 | Q5_0 | Q5_1 | Q4_K_M | Q5_K_M
    n_expert: int = 0
    n_expert_used: int = 0
    rope_freq_base: float = 10000.0
    rms_eps: float = 1e-5

    @property
    def head_dim(self) -> int:
        return self.n_embd // self.n_head


SHAPES = {
    # SURVEY.md section 8 table (standard public hyper-parameters)
    "tinyllama-1.1b-q4_0": ModelShape("tinyllama-1.1b-q4_0", 2048, 22, 32, 4, 5632, 32000, "Q4_0"),
    "llama2-7b-q4_k_m": ModelShape("llama2-7b-q4_k_m", 4096, 32, 32, 32, 11008, 32000, "Q4_K_M"),
    "llama2-13b-q5_k_m": ModelShape("llama2-13b-q5_k_m", 5120, 40, 40, 40, 13824, 32000, "Q5_K_M"),
    "mixtral-8x7b-q4_k_m": ModelShape("mixtral-8x7b-q4_k_m", 4096, 32, 32, 8, 14336, 32000, "Q4_K_M", 8, 2, 1e6),
    "llama3-70b-q4_k_m": ModelShape("llama3-70b-q4_k_m", 8192, 80, 64, 8, 28672, 128256, "Q4_K_M", 0, 0, 5e5),
    # small shapes for tests
    "test-dense": ModelShape("test-dense", 512, 3, 8, 4, 1024, 1024, "Q4_K_M"),
    "test-moe": ModelShape("test-moe", 512, 2, 8, 2, 768, 512, "Q4_K_M", 4, 2),
}


def use_more_bits(i: int, n: int) -> bool:
    """src/llama.cpp:17986-17988"""
    return i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2


def tensor_types(s: ModelShape) -> dict[str, int]:
    """GGUF tensor name -> ggml type, the mix llama_tensor_get_type produces (src/llama.cpp:17979-18200)."""
    t: dict[str, int] = {}
    base = {"Q2_K": Q2_K, "Q3_K": Q3_K, "Q4_0": Q4_0, "Q4_1": Q4_1, "Q5_0": Q5_0, "Q5_1": Q5_1, "Q4_K_M": Q4_K, "Q5_K_M": Q5_K, "IQ4_NL": IQ4_NL, "IQ4_XS": IQ4_XS}[s.ftype]
    kquant_m = s.ftype in ("Q4_K_M", "Q5_K_M")
    t["token_embd.weight"] = base
    t["output_norm.weight"] = F32
    t["output.weight"] = Q6_K  # :18008-18025 (output.weight is Q6_K unless the ftype is below Q4 territory; also for Q4_0)
    is_70b = s.n_layer == 80 and s.n_head != s.n_head_kv
    for il in range(s.n_layer):
        p = f"blk.{il}."
        t[p + "attn_norm.weight"] = F32
        t[p + "ffn_norm.weight"] = F32
        t[p + "attn_q.weight"] = base
        k = v = o = base
        down = base
        if kquant_m:
            if use_more_bits(il, s.n_layer):
                v = Q6_K  # :18094-18095
                down = Q6_K  # :18151-18156
            if s.ftype == "Q4_K_M" and is_70b and v == Q4_K:
                v = Q5_K  # 70B rule :18097-18102
            if s.n_expert == 8:
                k = v = Q8_0  # :18103-18114
                o = Q5_K if s.ftype == "Q4_K_M" else o  # :18175-18180
        t[p + "attn_k.weight"] = k
        t[p + "attn_v.weight"] = v
        t[p + "attn_output.weight"] = o
        if s.n_expert:
            t[p + "ffn_gate_inp.weight"] = F32
            t[p + "ffn_gate_exps.weight"] = base
            t[p + "ffn_up_exps.weight"] = base
            t[p + "ffn_down_exps.weight"] = down
        else:
            t[p + "ffn_gate.weight"] = base
            t[p + "ffn_up.weight"] = base
            t[p + "ffn_down.weight"] = down
    return t


def tensor_shape(s: ModelShape, name: str) -> tuple[int, int, int]:
    """(ne0, ne1, ne2) in ggml order: ne0 = input features (contiguous)."""
    kv = s.head_dim * s.n_head_kv
    leaf = name.split(".", 2)[2] if name.startswith("blk.") else name
    return {
        "token_embd.weight": (s.n_embd, s.n_vocab, 1), "output.weight": (s.n_embd, s.n_vocab, 1), "output_norm.weight": (s.n_embd, 1, 1),
        "attn_norm.weight": (s.n_embd, 1, 1), "ffn_norm.weight": (s.n_embd, 1, 1),
        "attn_q.weight": (s.n_embd, s.n_embd, 1), "attn_k.weight": (s.n_embd, kv, 1), "attn_v.weight": (s.n_embd, kv, 1),
        "attn_output.weight": (s.n_embd, s.n_embd, 1),
        "ffn_gate.weight": (s.n_embd, s.n_ff, 1), "ffn_up.weight": (s.n_embd, s.n_ff, 1), "ffn_down.weight": (s.n_ff, s.n_embd, 1),
        "ffn_gate_inp.weight": (s.n_embd, s.n_expert, 1),
        "ffn_gate_exps.weight": (s.n_embd, s.n_ff, s.n_expert), "ffn_up_exps.weight": (s.n_embd, s.n_ff, s.n_expert),
        "ffn_down_exps.weight": (s.n_ff, s.n_embd, s.n_expert),
    }[leaf]


def weight_bytes_per_token(s: ModelShape) -> int:
    """Bytes of weights one decoded token must stream: every matrix once, experts: n_expert_used of n_expert;
    the token_embd row and the f32 norm vectors are not counted."""
    total = 0
    for name, ty in tensor_types(s).items():
        ne0, ne1, ne2 = tensor_shape(s, name)
        if name == "token_embd.weight" or ty == F32 and not name.endswith("ffn_gate_inp.weight"):
            continue
        n = row_size(ty, ne0) * ne1
        total += n * (s.n_expert_used if ne2 > 1 else 1)
    return total


def kv_bytes_per_token(s: ModelShape, n_kv: int, type_kv: int) -> int:
    return 2 * s.n_layer * n_kv * row_size(type_kv, s.head_dim * s.n_head_kv)


# ------------------------------------------------------------------------------------------------ synthetic weights
def random_blocks(rng: np.random.Generator, ty: int, n_rows: int, ne0: int) -> np.ndarray:
    """Random-init quantised rows: uniformly random quants / sub-block scales with fp16 super-block scales chosen so the
    dequantised weights are roughly zero-mean with std ~0.02 (no NaN/Inf bit patterns in any scale field).
    Block layouts: ggml/src/ggml-common.h:144-149 (q4_0), 186-191 (q8_0), 286-327 (q4_K, q5_K, q6_K)."""
    if ty == F32:
        return (rng.standard_normal((n_rows, ne0), dtype=np.float32) * 0.02).view(np.uint8).reshape(n_rows, -1)
    if ty == F16:
        return (rng.standard_normal((n_rows, ne0), dtype=np.float32) * 0.02).astype(np.float16).view(np.uint8).reshape(n_rows, -1)
    bs, by = BLOCK[ty]
    nb = ne0 // bs
    raw = rng.integers(0, 256, size=(n_rows, nb, by), dtype=np.uint8)

    def f16(lo, hi, shape):
        return rng.uniform(lo, hi, size=shape).astype(np.float16).view(np.uint8).reshape(*shape, 2)

    if ty == Q4_0:
        raw[:, :, 0:2] = f16(3e-3, 5.5e-3, (n_rows, nb))
    elif ty == Q8_0:
        raw[:, :, 0:2] = f16(2e-4, 3.4e-4, (n_rows, nb))
    elif ty in (Q4_1, Q5_1):  # x d + m with x in [0, 15] / [0, 31]: m = -(mid x) d centres the values
        d = rng.uniform(3e-3, 5.5e-3, size=(n_rows, nb)) / (1 if ty == Q4_1 else 2)
        raw[:, :, 0:2] = d.astype(np.float16).view(np.uint8).reshape(n_rows, nb, 2)
        raw[:, :, 2:4] = (-d * (7.5 if ty == Q4_1 else 15.5)).astype(np.float16).view(np.uint8).reshape(n_rows, nb, 2)
    elif ty == Q5_0:
        raw[:, :, 0:2] = f16(1.5e-3, 2.7e-3, (n_rows, nb))
    elif ty == Q4_K:
        d = rng.uniform(1.1e-4, 1.9e-4, size=(n_rows, nb))
        raw[:, :, 0:2] = d.astype(np.float16).view(np.uint8).reshape(n_rows, nb, 2)
        raw[:, :, 2:4] = (d * 7.5).astype(np.float16).view(np.uint8).reshape(n_rows, nb, 2)
    elif ty == Q5_K:
        d = rng.uniform(4.5e-5, 7.5e-5, size=(n_rows, nb))
        raw[:, :, 0:2] = d.astype(np.float16).view(np.uint8).reshape(n_rows, nb, 2)
        raw[:, :, 2:4] = (d * 15.5).astype(np.float16).view(np.uint8).reshape(n_rows, nb, 2)
    elif ty == Q6_K:
        raw[:, :, 208:210] = f16(1.1e-5, 1.9e-5, (n_rows, nb))
    elif ty == Q2_K:  # x = d sc q - dmin m, sc, m in [0, 15], q in [0, 3]
        d = rng.uniform(1.2e-3, 2e-3, size=(n_rows, nb))
        raw[:, :, 80:82] = d.astype(np.float16).view(np.uint8).reshape(n_rows, nb, 2)
        raw[:, :, 82:84] = (d * 1.5).astype(np.float16).view(np.uint8).reshape(n_rows, nb, 2)
    elif ty == Q3_K:  # x = d (sc - 32) q, q in [-4, 3]
        raw[:, :, 108:110] = f16(3e-4, 5e-4, (n_rows, nb))
    elif ty == IQ4_NL:  # x = d level, levels in [-127, 113]
        raw[:, :, 0:2] = f16(2.2e-4, 3.8e-4, (n_rows, nb))
    elif ty == IQ4_XS:  # x = d (ls - 32) level
        raw[:, :, 0:2] = f16(1.2e-5, 2e-5, (n_rows, nb))
    else:
        raise B200Error(f"random_blocks: unsupported type {ty}")
    return raw.reshape(n_rows, nb * by)


class LlamaModel:
    """A model resident on the device(s), driven through b200_llama_* (include/b200_llama.h)."""

    def __init__(self, shape: ModelShape, n_ctx: int, type_kv: int = F16, n_split_devices: int = 0, main_device: int = 0,
                 tp_rank: int = 0, tp_world: int = 1):
        L = _l()
        if L.ggml_backend_b200_get_device_count() <= 0:
            raise B200Error("no CUDA device: the B200 backend has no CPU fallback")
        self.shape = shape
        self.n_ctx = n_ctx
        self.type_kv = type_kv
        self.tp_rank, self.tp_world = tp_rank, max(1, tp_world)
        if self.tp_world > 1:
            from . import tp
            tp.check_divisible(shape.n_head, shape.n_head_kv, shape.n_ff, shape.n_vocab, shape.head_dim, self.tp_world)
        p = LlamaParams(n_embd=shape.n_embd, n_layer=shape.n_layer, n_head=shape.n_head, n_head_kv=shape.n_head_kv, n_embd_head=shape.head_dim,
                        n_ff=shape.n_ff, n_vocab=shape.n_vocab, n_expert=shape.n_expert, n_expert_used=shape.n_expert_used, n_ctx=n_ctx,
                        type_k=type_kv, type_v=type_kv, rope_neox=0, n_ctx_orig=n_ctx, rms_eps=shape.rms_eps, rope_freq_base=shape.rope_freq_base,
                        rope_freq_scale=1.0, main_device=main_device, n_split_devices=n_split_devices, tp_rank=tp_rank, tp_world=self.tp_world)
        h = C.c_void_p()
        check(L.b200_llama_new(C.byref(p), C.byref(h)), "b200_llama_new")
        self.h = h
        self._logits = None
        self._tok = None

    def set_tensor(self, name: str, ty: int, ne: tuple[int, int, int], data: np.ndarray) -> None:
        data = np.ascontiguousarray(data)
        expect = row_size(ty, ne[0]) * ne[1] * ne[2]
        if data.nbytes != expect:
            raise B200Error(f"{name}: {data.nbytes} bytes given, {expect} expected")
        check(_l().b200_llama_set_tensor(self.h, name.encode(), ty, ne[0], ne[1], ne[2], data.ctypes.data), f"set_tensor {name}")

    def clone_tensor(self, name: str, src: str) -> None:
        check(_l().b200_llama_clone_tensor(self.h, name.encode(), src.encode()), f"clone_tensor {name}")

    def finalize(self, max_batch: int = 1) -> None:
        check(_l().b200_llama_finalize(self.h, max_batch), "b200_llama_finalize")
        self.max_batch = max_batch
        # pinned host staging for the timed path
        n = self.shape.n_vocab * max_batch * 4
        p = C.c_void_p()
        check(lib().b200_malloc_host(C.byref(p), n), "malloc_host")
        self._logits_ptr = p
        self._logits = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(max_batch, self.shape.n_vocab))
        self._tok = np.zeros(max_batch, dtype=np.int32)

    def init_random(self, seed: int = 1234, replicate_layers: bool = True) -> None:
        """Random-init weights of the configuration's architecture and type mix.  With replicate_layers the layers share
        the contents of the first layer that has the same tensor type (device-to-device clone into distinct memory): the
        bytes streamed per token are identical to an all-unique model, only host generation time is saved."""
        rng = np.random.default_rng(seed)
        s = self.shape
        seen: dict[tuple[str, int], str] = {}
        for name, ty in tensor_types(s).items():
            ne = tensor_shape(s, name)
            leaf = name.split(".", 2)[2] if name.startswith("blk.") else name
            if leaf.endswith("norm.weight"):
                self.set_tensor(name, F32, ne, np.ones(ne[0], dtype=np.float32))
                continue
            key = (leaf, ty)
            if replicate_layers and key in seen:
                self.clone_tensor(name, seen[key])
                continue
            data = random_blocks(rng, ty, ne[1] * ne[2], ne[0])  # every rank draws the FULL tensor from the same stream ...
            if self.tp_world > 1:                                 # ... and keeps its shard
                from . import tp
                data = tp.shard_tensor(name, ty, ne, data, self.tp_rank, self.tp_world)
                ne = tp.shard_shape(name, ne, self.tp_world)
            self.set_tensor(name, ty, ne, data)
            seen[key] = name

    def kv_fill_random(self, n_cells: int, seed: int = 7) -> None:
        """Fill KV cells [0, n_cells) of every layer with N(0,1) values stored in the cache type."""
        rng = np.random.default_rng(seed)
        s = self.shape
        width = s.head_dim * s.n_head_kv
        x = rng.standard_normal((n_cells, width), dtype=np.float32)
        if self.type_kv == F16:
            rows = x.astype(np.float16).view(np.uint8).reshape(n_cells, -1)
        elif self.type_kv != Q8_0:
            raise B200Error("fill_kv: f16 and q8_0 caches only")
        else:
            xb = x.reshape(n_cells, width // 32, 32)
            amax = np.abs(xb).max(axis=2)
            d = (amax / 127.0).astype(np.float32)
            q = np.where(d[..., None] > 0, np.round(xb / np.where(d[..., None] > 0, d[..., None], 1)), 0).astype(np.int8)
            rows = np.concatenate([d.astype(np.float16).view(np.uint8).reshape(n_cells, width // 32, 2), q.view(np.uint8)], axis=2).reshape(n_cells, -1)
        if self.tp_world > 1:
            from . import tp
            rows = tp.shard_kv_rows(rows, self.type_kv, s.n_head_kv, s.head_dim, self.tp_rank, self.tp_world)
        rows = np.ascontiguousarray(rows)
        for il in range(s.n_layer):
            for is_v in (0, 1):
                check(_l().b200_llama_kv_write(self.h, il, is_v, 0, n_cells, rows.ctypes.data), "kv_write")

    def kv_read(self, il: int, is_v: int, cell0: int, n: int) -> np.ndarray:
        row = row_size(self.type_kv, self.shape.head_dim * self.shape.n_head_kv // self.tp_world)
        out = np.empty((n, row), dtype=np.uint8)
        check(_l().b200_llama_kv_read(self.h, il, is_v, cell0, n, out.ctypes.data), "kv_read")
        return out

    def decode(self, tokens, n_past: int, logits_all: bool = False) -> np.ndarray:
        """llama_decode + get_logits: host tokens in, host logits out (synchronous)."""
        n = len(tokens)
        self._tok[:n] = tokens
        check(_l().b200_llama_decode(self.h, self._tok.ctypes.data, n, n_past, self._logits_ptr, 1 if logits_all else 0), "b200_llama_decode")
        return self._logits[: (n if logits_all else 1)]

    def decode_async(self, tokens, n_past: int, want_logits: bool = True) -> None:
        n = len(tokens)
        self._tok[:n] = tokens
        check(_l().b200_llama_decode_async(self.h, self._tok.ctypes.data, n, n_past, self._logits_ptr if want_logits else None, 0), "decode_async")

    def synchronize(self) -> None:
        check(_l().b200_llama_synchronize(self.h), "synchronize")

    def replay(self) -> None:
        check(_l().b200_llama_replay(self.h), "replay")

    @property
    def stream(self):
        return _l().b200_llama_stream(self.h)

    def stats(self) -> dict:
        a, b, c, d = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        check(_l().b200_llama_stats(self.h, C.byref(a), C.byref(b), C.byref(c), C.byref(d)), "stats")
        return {"n_nodes": a.value, "n_kernels": b.value, "weight_bytes": c.value, "kv_bytes": d.value}

    def close(self) -> None:
        if getattr(self, "h", None):
            _l().b200_llama_free(self.h)
            self.h = None
            if self._logits is not None:
                lib().b200_free_host(self._logits_ptr)
                self._logits = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
