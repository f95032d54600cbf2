"""GGUF fixtures for the llama.cpp-level tests (TEST INFRASTRUCTURE: binds oracle/_ref).

reference_quantizer() hands tools/make_gguf.py the reference's own ggml_quantize_chunk (ggml/src/ggml.c:21944) from
oracle/_ref/libggml_ref.so, run over row chunks on all host threads (ctypes releases the GIL).  build_fixture() caches the
files under $B200_GGUF_CACHE (default /tmp/b200_gguf): nothing large is committed, the generating code is.
"""
from __future__ import annotations

import concurrent.futures as cf
import dataclasses
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))

import ggml_ref as R  # noqa: E402


def reference_quantizer(n_threads: int | None = None):
    ref = R.load("portable")
    n_threads = n_threads or os.cpu_count() or 1
    pool = cf.ThreadPoolExecutor(max_workers=n_threads)

    def quantize(ty: int, x: np.ndarray) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        rows = x.shape[0]
        step = max(1, (rows + 4 * n_threads - 1) // (4 * n_threads))
        parts = list(pool.map(lambda r0: ref.quantize(ty, x[r0:r0 + step]), range(0, rows, step)))
        return np.concatenate(parts, axis=0)

    return quantize


def cache_dir() -> Path:
    d = Path(os.environ.get("B200_GGUF_CACHE", "/tmp/b200_gguf"))
    d.mkdir(parents=True, exist_ok=True)
    return d


def build_fixture(shape_name: str, seed: int = 1234, quantizer: str = "reference", **over) -> Path:
    """Return the path of a cached synthetic GGUF of SHAPES[shape_name] (fields overridden by **over, e.g. n_layer=2)."""
    import make_gguf
    from koboldcpp_b200.llama import SHAPES

    layer_gain = over.pop("layer_gain", 1.0)
    s = dataclasses.replace(SHAPES[shape_name], **over) if over else SHAPES[shape_name]
    tag = "_".join(f"{k}{v}" for k, v in sorted(over.items()))
    out = cache_dir() / f"{shape_name}{'_' + tag if tag else ''}_s{seed}_g{layer_gain}_{quantizer}.gguf"
    if not out.exists():
        tmp = out.with_suffix(".tmp")
        make_gguf.build(s, tmp, reference_quantizer() if quantizer == "reference" else None, seed, layer_gain=layer_gain)
        tmp.rename(out)
    return out
