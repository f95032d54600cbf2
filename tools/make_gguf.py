"""Synthetic (random-init) GGUF models of the shapes BASELINE.json names.

    python tools/make_gguf.py --shape tinyllama-1.1b-q4_0 --out /tmp/tiny.gguf [--quantizer reference|blocks] [--layers N]

The file is written with gguf-py's GGUFWriter (the reference's own writer: gguf-py/gguf/gguf_writer.py:62 `GGUFWriter`,
:368 `add_tensor`; the same package is installed in this image) and loads with the reference's unmodified
llama_load_model_from_file (src/llama.cpp:5396, 6196: tokenizer model "no_vocab", token ids are fed directly).

Tensor types follow llama_tensor_get_type for the ftype (koboldcpp_b200.llama.tensor_types; src/llama.cpp:17979-18200).

Two quantisers:
  * "reference": f32 weights drawn from numpy.random.default_rng(seed) and quantised BY THE REFERENCE's
    ggml_quantize_chunk (ggml/src/ggml.c:21944) through oracle/_ref/libggml_ref.so.  This is test-fixture generation: the
    callable is injected by the caller (tests/gguf_fixtures.py), this module itself never touches oracle/.
  * "blocks": random quantised blocks with sane scales (koboldcpp_b200.llama.random_blocks) -- throughput runs only.

Greedy-parity design ("margin model"): random-init transformers produce near-flat logits whose argmax flips under 1e-3
relative noise, which would make "identical greedy tokens" a coin toss between ANY two correct implementations (the
reference's own AVX2 / AVX-512 / scalar builds disagree).  The generator therefore ties the output head to the embedding
through a fixed random permutation sigma (output.weight[j] = token_embd[sigma(j)] + noise) and scales the layer weights so
that the residual stream keeps the current token's embedding as one component among the layers' contributions
(`--layer-gain`): the top-1 logit then leads by a margin far above quantisation noise while all layers still shape the
logits that the parity tests compare numerically.
"""
from __future__ import annotations

import argparse
import sys
import time
from pathlib import Path
from typing import Callable

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from koboldcpp_b200.llama import SHAPES, ModelShape, random_blocks, tensor_shape, tensor_types  # noqa: E402
from koboldcpp_b200.lib import F16, F32, row_size  # noqa: E402

# ggml type id -> gguf enum
_LLAMA_FTYPE = {"Q2_K": 10, "Q3_K": 12, "Q4_0": 2, "Q4_1": 3, "Q5_0": 8, "Q5_1": 9, "Q4_K_M": 15, "Q5_K_M": 17, "IQ4_NL": 25, "IQ4_XS": 30}  # include/llama.h:150-190 (LLAMA_FTYPE_MOSTLY_*)

Quantizer = Callable[[int, np.ndarray], np.ndarray]  # (ggml type, f32 [rows, ne0]) -> uint8 [rows, row_bytes]


def _std_for(leaf: str, s: ModelShape, layer_gain: float) -> float:
    """Weight std per tensor kind: attention / ffn products are scaled so that one layer adds ~layer_gain / sqrt(n_layer)
    of the embedding's RMS to the residual stream (the embedding itself has unit RMS)."""
    d, ff = s.n_embd, s.n_ff
    g = layer_gain / np.sqrt(max(1, s.n_layer))
    if leaf in ("attn_q.weight", "attn_k.weight"):
        return 1.0 / np.sqrt(d)                 # q.k ~ N(0, D): softmax is neither flat nor one-hot after the 1/sqrt(D) scale
    if leaf == "attn_v.weight":
        return 1.0 / np.sqrt(d)                 # v ~ unit RMS
    if leaf == "attn_output.weight":
        return g / np.sqrt(d)
    if leaf in ("ffn_gate.weight", "ffn_up.weight", "ffn_gate_exps.weight", "ffn_up_exps.weight"):
        return 1.0 / np.sqrt(d)
    if leaf in ("ffn_down.weight", "ffn_down_exps.weight"):
        return 2.5 * g / np.sqrt(ff)            # silu(g)*u of unit-variance g, u has RMS ~0.4
    if leaf == "ffn_gate_inp.weight":
        return 1.0 / np.sqrt(d)
    raise KeyError(leaf)


def build(shape: ModelShape, out: Path, quantize: Quantizer | None, seed: int = 1234, n_ctx_train: int = 4096, layer_gain: float = 1.0,
          head_noise: float = 0.25, verbose: bool = False, replicate_layers: bool = False) -> dict:
    """Write the GGUF.  quantize == None selects random blocks (throughput only).  Returns a small summary dict.
    replicate_layers (random blocks only): layers reuse the bytes of the first layer with the same tensor type -- the file and
    the bytes streamed per token are those of an all-unique model, only host generation time is saved."""
    import gguf

    s = shape
    t0 = time.time()
    w = gguf.GGUFWriter(str(out), "llama")
    w.add_name(s.name)
    w.add_context_length(n_ctx_train)
    w.add_embedding_length(s.n_embd)
    w.add_block_count(s.n_layer)
    w.add_feed_forward_length(s.n_ff)
    w.add_head_count(s.n_head)
    w.add_head_count_kv(s.n_head_kv)
    w.add_layer_norm_rms_eps(s.rms_eps)
    w.add_rope_dimension_count(s.head_dim)
    w.add_rope_freq_base(s.rope_freq_base)
    w.add_vocab_size(s.n_vocab)
    w.add_file_type(_LLAMA_FTYPE[s.ftype])
    if s.n_expert:
        w.add_expert_count(s.n_expert)
        w.add_expert_used_count(s.n_expert_used)
    w.add_tokenizer_model("no_vocab")

    rng = np.random.default_rng(seed)
    types = tensor_types(s)
    total = 0
    emb_f32 = None
    cache: dict = {}
    sigma = np.random.default_rng(seed + 1).permutation(s.n_vocab)
    for name, ty in types.items():
        ne0, ne1, ne2 = tensor_shape(s, name)
        leaf = name.split(".", 2)[2] if name.startswith("blk.") else name
        rows = ne1 * ne2
        if leaf.endswith("norm.weight"):
            w.add_tensor(name, np.ones(ne0, dtype=np.float32))
            continue
        if quantize is None:
            key = (leaf, ty)
            if replicate_layers and key in cache:
                data = cache[key]
            else:
                data = random_blocks(rng, ty, rows, ne0)
                if replicate_layers:
                    cache[key] = data
        else:
            if leaf == "token_embd.weight":
                x = rng.standard_normal((rows, ne0), dtype=np.float32)
                emb_f32 = x
            elif leaf == "output.weight":
                x = emb_f32[sigma] + head_noise * rng.standard_normal((rows, ne0), dtype=np.float32)
            else:
                x = rng.standard_normal((rows, ne0), dtype=np.float32) * np.float32(_std_for(leaf, s, layer_gain))
            data = x.view(np.uint8).reshape(rows, -1) if ty == F32 else quantize(ty, x)
        assert data.shape == (rows, row_size(ty, ne0)), (name, data.shape)
        total += data.nbytes
        if ty == F32:
            arr = data.view(np.float32).reshape((ne2, ne1, ne0) if ne2 > 1 else (ne1, ne0))
            w.add_tensor(name, arr)
        elif ty == F16:
            arr = data.view(np.float16).reshape((ne2, ne1, ne0) if ne2 > 1 else (ne1, ne0))
            w.add_tensor(name, arr)
        else:
            # quantised payload: byte array whose last dim is the row size; GGUFWriter derives ne from raw_dtype
            arr = data.reshape((ne2, ne1, -1) if ne2 > 1 else (ne1, -1))
            w.add_tensor(name, arr, raw_dtype=gguf.GGMLQuantizationType(ty))
        if verbose:
            print(f"  {name:32s} type {ty:2d} [{ne0}, {ne1}, {ne2}]", file=sys.stderr)
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    return {"path": str(out), "bytes": total, "seconds": round(time.time() - t0, 2), "n_tensors": len(types), "sigma_seed": seed + 1}


def shape_from_args(a) -> ModelShape:
    import dataclasses
    s = SHAPES[a.shape]
    over = {}
    if a.layers:
        over["n_layer"] = a.layers
    if a.vocab:
        over["n_vocab"] = a.vocab
    return dataclasses.replace(s, **over) if over else s


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", required=True, choices=sorted(SHAPES))
    ap.add_argument("--out", required=True)
    ap.add_argument("--quantizer", default="blocks", choices=["blocks", "reference"])
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--vocab", type=int, default=0)
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--layer-gain", type=float, default=1.0)
    ap.add_argument("-v", action="store_true")
    a = ap.parse_args()
    q = None
    if a.quantizer == "reference":
        sys.path.insert(0, str(ROOT / "tests"))
        import gguf_fixtures  # test infrastructure: binds oracle/_ref
        q = gguf_fixtures.reference_quantizer()
    info = build(shape_from_args(a), Path(a.out), q, a.seed, layer_gain=a.layer_gain, verbose=a.v)
    print(info)


if __name__ == "__main__":
    main()
