"""TEST / BENCH INFRASTRUCTURE: the reference's own CPU implementation of the hot path, timed on this box's host cores.

Runs one transformer layer of the configuration (the ops of build_llama, /root/reference/src/llama.cpp:10453-10617: RMS_NORM,
MUL, MUL_MAT x7, ROPE, CPY into an f16/q8_0 cache, FLASH_ATTN_EXT over n_kv cells, SILU, ADD) and the output head through the
compiled reference library oracle/_ref/libggml_ref*.so (ggml_graph_compute_with_ctx, all host threads), and scales the
measured layer times to whole-token decode: t_token = sum over the layer type-mix variants + t_head.  This is a bounded
sample of the same workload (the full model would need its 4-42 GB of weights generated and streamed on the host for the
same per-layer arithmetic).  Only bench.py's CPU legs and tests may import this module."""
from __future__ import annotations

import ctypes as C
import os
import time

import numpy as np

import ggml_ref as G
from koboldcpp_b200.lib import F16, F32, I32, row_size
from koboldcpp_b200.llama import ModelShape, random_blocks, tensor_shape, tensor_types


def _bind(L):
    P, I, I64, SZ, F = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_float
    TP = C.POINTER(G.Tensor)
    for n, args in {
        "ggml_rms_norm": [P, TP, F], "ggml_mul": [P, TP, TP], "ggml_add": [P, TP, TP], "ggml_silu": [P, TP],
        "ggml_reshape_2d": [P, TP, I64, I64], "ggml_reshape_3d": [P, TP, I64, I64, I64], "ggml_cpy": [P, TP, TP],
        "ggml_view_1d": [P, TP, I64, SZ], "ggml_get_rows": [P, TP, TP],
        "ggml_rope_ext": [P, TP, TP, TP, I, I, I, F, F, F, F, F, F],
    }.items():
        f = getattr(L, n)
        f.restype = TP
        f.argtypes = args


def host_threads() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def _layer_variants(s: ModelShape):
    """Distinct per-layer type mixes and how many layers have each."""
    types = tensor_types(s)
    variants: dict[tuple, list[int]] = {}
    for il in range(s.n_layer):
        key = tuple(sorted((k.split(".", 2)[2], v) for k, v in types.items() if k.startswith(f"blk.{il}.")))
        variants.setdefault(key, []).append(il)
    return [(dict(k), len(v)) for k, v in variants.items()]


def time_decode(s: ModelShape, n_kv: int, type_kv: int = F16, threads: int | None = None, budget_s: float = 20.0, seed: int = 0) -> dict:
    """Returns {"tok_s", "ms_per_token", "cores", "variant", "sample"} for batch-1 decode at n_kv cached cells."""
    if s.n_expert:
        raise NotImplementedError("CPU baseline sample covers the dense configurations")
    variant = G.best_variant()
    R = G.load(variant)
    L = R.lib
    _bind(L)
    threads = threads or host_threads()
    rng = np.random.default_rng(seed)
    D, H, HK = s.head_dim, s.n_head, s.n_head_kv
    kvw = D * HK
    variants = _layer_variants(s)
    per_variant_budget = budget_s / (len(variants) + 1)
    t_token = 0.0
    detail = []

    def run_timed(ctx, out, budget):
        gf = L.ggml_new_graph(ctx)
        L.ggml_build_forward_expand(gf, out)
        L.ggml_graph_compute_with_ctx(ctx, gf, threads)  # warm-up (page in, thread pool)
        n, t0 = 0, time.perf_counter()
        while True:
            L.ggml_graph_compute_with_ctx(ctx, gf, threads)
            n += 1
            dt = time.perf_counter() - t0
            if dt > budget or n >= 200:
                return dt / n, n

    for types, count in variants:
        wbytes = sum(row_size(ty, tensor_shape(s, "blk.0." + leaf)[0]) * tensor_shape(s, "blk.0." + leaf)[1] for leaf, ty in types.items())
        kv_bytes = 2 * n_kv * row_size(type_kv, kvw)
        ctx = R._ctx(wbytes + kv_bytes + (256 << 20))
        try:
            w = {}
            for leaf, ty in types.items():
                ne = tensor_shape(s, "blk.0." + leaf)
                t = L.ggml_new_tensor_2d(ctx, ty, ne[0], ne[1])
                data = np.ones(ne[0], np.float32).view(np.uint8) if ty == F32 else random_blocks(rng, ty, ne[1], ne[0])
                R._fill(t, data)
                w[leaf] = t
            kc = L.ggml_new_tensor_1d(ctx, type_kv, kvw * n_kv)
            vc = L.ggml_new_tensor_1d(ctx, type_kv, kvw * n_kv)
            kvrow = rng.standard_normal((n_kv, kvw), dtype=np.float32).astype(np.float16)
            if type_kv == F16:
                R._fill(kc, kvrow); R._fill(vc, kvrow)
            else:
                q = R.quantize(type_kv, kvrow.astype(np.float32))
                R._fill(kc, q); R._fill(vc, q)
            x = L.ggml_new_tensor_2d(ctx, F32, s.n_embd, 1)
            R._fill(x, rng.standard_normal(s.n_embd).astype(np.float32))
            pos = L.ggml_new_tensor_1d(ctx, I32, 1)
            R._fill(pos, np.array([n_kv - 1], np.int32))
            mask = L.ggml_new_tensor_2d(ctx, F16, n_kv, 32)
            R._fill(mask, np.zeros((32, n_kv), np.float16))
            cur = L.ggml_mul(ctx, L.ggml_rms_norm(ctx, x, s.rms_eps), w["attn_norm.weight"])
            qc = L.ggml_mul_mat(ctx, w["attn_q.weight"], cur)
            kcur = L.ggml_mul_mat(ctx, w["attn_k.weight"], cur)
            vcur = L.ggml_mul_mat(ctx, w["attn_v.weight"], cur)
            rope = lambda t, nh: L.ggml_rope_ext(ctx, L.ggml_reshape_3d(ctx, t, D, nh, 1), pos, None, D, 0, n_kv, s.rope_freq_base, 1.0, 0.0, 1.0, 32.0, 1.0)
            qc, kcur = rope(qc, H), rope(kcur, HK)
            rs = row_size(type_kv, kvw)
            kst = L.ggml_cpy(ctx, kcur, L.ggml_view_1d(ctx, kc, kvw, rs * (n_kv - 1)))
            vst = L.ggml_cpy(ctx, vcur, L.ggml_view_1d(ctx, vc, kvw, rs * (n_kv - 1)))
            qp = L.ggml_permute(ctx, qc, 0, 2, 1, 3)
            k3 = L.ggml_view_3d(ctx, kc, D, n_kv, HK, rs, row_size(type_kv, D), 0)
            v3 = L.ggml_view_3d(ctx, vc, D, n_kv, HK, rs, row_size(type_kv, D), 0)
            att = L.ggml_flash_attn_ext(ctx, qp, k3, v3, mask, 1.0 / np.sqrt(D), 0.0, 0.0)
            a = L.ggml_mul_mat(ctx, w["attn_output.weight"], L.ggml_reshape_2d(ctx, att, D * H, 1))
            ffn_inp = L.ggml_add(ctx, a, x)
            h = L.ggml_mul(ctx, L.ggml_rms_norm(ctx, ffn_inp, s.rms_eps), w["ffn_norm.weight"])
            up = L.ggml_mul_mat(ctx, w["ffn_up.weight"], h)
            g = L.ggml_silu(ctx, L.ggml_mul_mat(ctx, w["ffn_gate.weight"], h))
            out = L.ggml_add(ctx, L.ggml_mul_mat(ctx, w["ffn_down.weight"], L.ggml_mul(ctx, g, up)), ffn_inp)
            # one graph: cache stores first (they do not feed `out`, so expand them explicitly), then the layer
            gf = L.ggml_new_graph(ctx)
            L.ggml_build_forward_expand(gf, kst)
            L.ggml_build_forward_expand(gf, vst)
            L.ggml_build_forward_expand(gf, out)
            L.ggml_graph_compute_with_ctx(ctx, gf, threads)
            n, t0 = 0, time.perf_counter()
            while True:
                L.ggml_graph_compute_with_ctx(ctx, gf, threads)
                n += 1
                dt = time.perf_counter() - t0
                if dt > per_variant_budget or n >= 200:
                    break
            t_layer = dt / n
            t_token += t_layer * count
            detail.append(f"{count}x layer[{'/'.join(sorted(set(G.TYPE_NAMES[t] for t in types.values() if t != F32)))}] {t_layer * 1e3:.2f} ms ({n} runs)")
        finally:
            L.ggml_free(ctx)
    # output head: final norm + output.weight GEMV
    ty = tensor_types(s)["output.weight"]
    ne = tensor_shape(s, "output.weight")
    ctx = R._ctx(row_size(ty, ne[0]) * ne[1] + (64 << 20))
    try:
        wo = L.ggml_new_tensor_2d(ctx, ty, ne[0], ne[1])
        R._fill(wo, random_blocks(rng, ty, ne[1], ne[0]))
        wn = L.ggml_new_tensor_1d(ctx, F32, s.n_embd)
        R._fill(wn, np.ones(s.n_embd, np.float32))
        x = L.ggml_new_tensor_2d(ctx, F32, s.n_embd, 1)
        R._fill(x, rng.standard_normal(s.n_embd).astype(np.float32))
        out = L.ggml_mul_mat(ctx, wo, L.ggml_mul(ctx, L.ggml_rms_norm(ctx, x, s.rms_eps), wn))
        t_head, n = run_timed(ctx, out, per_variant_budget)
        t_token += t_head
        detail.append(f"head {t_head * 1e3:.2f} ms ({n} runs)")
    finally:
        L.ggml_free(ctx)
    return {"tok_s": 1.0 / t_token, "ms_per_token": t_token * 1e3, "cores": threads, "variant": variant,
            "sample": f"reference ggml CPU backend ({variant} build), one layer per type mix + output head timed and scaled to "
                      f"{s.n_layer} layers at n_kv={n_kv}: " + "; ".join(detail)}
