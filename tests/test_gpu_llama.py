"""GPU parity of whole-token decode through the ggml_backend boundary (b200_llama_* -> graph_compute) against the fp64
restatement in llama_ref.py.  Tolerances: matmuls run on q8_1-quantised activations (like the reference CUDA path), which
puts ~1e-2 relative noise on each matmul; logits of a 2-3 layer model must agree to 4e-2 of their
RMS and the executor variants (graphs / fusion on or off) must agree with each other far tighter than that."""
from __future__ import annotations

import numpy as np
import pytest

from koboldcpp_b200.lib import F16, F32, Q4_0, Q8_0
from koboldcpp_b200.llama import SHAPES, LlamaModel, _l, random_blocks, tensor_shape, tensor_types

pytestmark = pytest.mark.gpu


def make_tensors(shape, seed):
    rng = np.random.default_rng(seed)
    out = {}
    for name, ty in tensor_types(shape).items():
        ne = tensor_shape(shape, name)
        if name.endswith("norm.weight"):
            out[name] = (F32, (1.0 + 0.1 * rng.standard_normal(ne[0])).astype(np.float32).view(np.uint8))
        else:
            out[name] = (ty, random_blocks(rng, ty, ne[1] * ne[2], ne[0]).reshape(-1))
    # a wider spread on the embeddings and the router so that tokens / experts differ visibly
    return out


def build(shape, tensors, n_ctx, type_kv, max_batch=8):
    m = LlamaModel(shape, n_ctx, type_kv)
    for name, (ty, raw) in tensors.items():
        m.set_tensor(name, ty, tensor_shape(shape, name), raw)
    m.finalize(max_batch)
    return m


def rel_rms(a, b):
    return float(np.sqrt(((a - b) ** 2).mean()) / (np.sqrt((b ** 2).mean()) + 1e-30))


@pytest.mark.parametrize("cfg,type_kv", [("test-dense", F16), ("test-dense", Q8_0), ("test-dense", Q4_0), ("test-moe", F16)])
def test_decode_matches_fp64_reference(cfg, type_kv):
    from llama_ref import RefModel
    shape = SHAPES[cfg]
    tensors = make_tensors(shape, 11)
    ref = RefModel(shape, tensors, 512, type_kv)
    m = build(shape, tensors, 512, type_kv)
    # q4_0 cache: both sides store bit-identical blocks for identical inputs, but the inputs differ in the last f32 bits (q8_1 vs
    # fp64 matmuls), and a value that sits on a rounding boundary of the 4-bit quantiser then lands on the neighbouring level -- one
    # level is 1/8 of the block's largest value.  The bar is wider there; greedy-token parity with a q4_0 cache is checked against the
    # reference itself in test_gpu_llamacpp.py.
    tol = 1e-1 if type_kv == Q4_0 else 4e-2
    try:
        prompt = [(i * 7919 + 3) % shape.n_vocab for i in range(5)]
        got = m.decode(prompt, 0, logits_all=True).copy()
        exp = ref.decode(prompt, 0)
        assert np.isfinite(got).all()
        assert rel_rms(got, exp) < tol, rel_rms(got, exp)
        n_past = len(prompt)
        tok = int(np.argmax(exp[-1]))
        for _ in range(6):
            got = m.decode([tok], n_past).copy()
            exp = ref.decode([tok], n_past)
            assert rel_rms(got[0], exp[0]) < tol, rel_rms(got[0], exp[0])
            n_past += 1
            tok = int(np.argmax(exp[0]))
        # the cache rows written by the CPY nodes: f16 rounding of nearly equal values
        if type_kv == F16:
            k = m.kv_read(0, 0, 0, n_past).view(np.float16).astype(np.float64)
            assert rel_rms(k, ref.k[0, :n_past]) < 2e-2
    finally:
        m.close()


@pytest.mark.parametrize("cfg,type_kv,n_prompt", [("test-dense", F16, 70), ("test-dense", Q8_0, 33)])
def test_prefill_batch_matches_fp64_reference(cfg, type_kv, n_prompt):
    """A prompt processed as ONE batch (N > 8: tensor-core GEMM + multi-column flash attention), then decode on top of the
    cache it wrote.  bf16 tensor-core operands: logits within 4e-2 of their RMS like the decode path."""
    from llama_ref import RefModel
    shape = SHAPES[cfg]
    tensors = make_tensors(shape, 21)
    ref = RefModel(shape, tensors, 512, type_kv)
    m = build(shape, tensors, 512, type_kv, max_batch=128)
    try:
        prompt = [(i * 7919 + 11) % shape.n_vocab for i in range(n_prompt)]
        got = m.decode(prompt, 0, logits_all=True).copy()
        exp = ref.decode(prompt, 0)
        assert np.isfinite(got).all()
        assert rel_rms(got, exp) < 4e-2, rel_rms(got, exp)
        got1 = m.decode([5], n_prompt).copy()
        exp1 = ref.decode([5], n_prompt)
        assert rel_rms(got1[0], exp1[0]) < 4e-2
        # last-token-only logits (inp_out_ids path, src/llama.cpp:10541-10546) equal the last row of logits_all
        m2 = build(shape, tensors, 512, type_kv, max_batch=128)
        try:
            last = m2.decode(prompt, 0, logits_all=False).copy()
            assert rel_rms(last[0], got[-1]) < 2e-2  # last layer runs as GEMV (q8_1) here, as GEMM (bf16) above
        finally:
            m2.close()
    finally:
        m.close()


def test_executor_variants_agree():
    """CUDA-graph replay and node fusion must not change results: graphs on/off bit-identical, fusion on/off within
    float reassociation (the fused kernels do the same arithmetic per element)."""
    shape = SHAPES["test-dense"]
    tensors = make_tensors(shape, 5)
    outs = {}
    L = _l()
    try:
        for graphs in (1, 0):
            for fusion in (1, 0):
                L.ggml_backend_b200_set_graphs(graphs)
                L.ggml_backend_b200_set_fusion(fusion)
                m = build(shape, tensors, 256, F16)
                try:
                    seq = []
                    n_past = 0
                    for t in range(20):
                        seq.append(m.decode([(t * 31 + 1) % shape.n_vocab], n_past).copy())
                        n_past += 1
                    outs[(graphs, fusion)] = np.concatenate(seq)
                    if graphs and fusion:
                        st = m.stats()
                        assert st["n_kernels"] < st["n_nodes"], st  # fusion + view elision: fewer launches than nodes
                finally:
                    m.close()
    finally:
        L.ggml_backend_b200_set_graphs(1)
        L.ggml_backend_b200_set_fusion(1)
    assert np.array_equal(outs[(1, 1)], outs[(0, 1)])
    assert np.array_equal(outs[(1, 0)], outs[(0, 0)])
    # the fused GEMV prologue derives the RMS scale with a different reduction tree: a last-bit change can flip int8 roundings
    assert rel_rms(outs[(1, 1)], outs[(1, 0)]) < 5e-3


def test_graph_replay_counts():
    """After the first token of a steady-state decode every token is one cudaGraphLaunch (the KV write position moves
    through the CPY slot table, not through a re-capture)."""
    import ctypes as C
    shape = SHAPES["test-dense"]
    m = build(shape, make_tensors(shape, 3), 512, F16)
    L = _l()
    try:
        m.decode([1], 0)
        L.ggml_backend_b200_reset_stats()
        for t in range(1, 40):
            m.decode([t], t)
        a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
        L.ggml_backend_b200_get_stats(C.byref(a), C.byref(b), C.byref(c))
        assert b.value == 39 and c.value == 0, (a.value, b.value, c.value)
    finally:
        m.close()


@pytest.mark.parametrize("type_kv", [F16, Q8_0])
def test_decoder_input_shortcuts_do_not_change_results(type_kv):
    """The host decoder reuses the previous graph (patched KV-store views) and sends ONE mask element when a single token follows
    the previous one.  A sequence decoded that way must equal, bit for bit, the same sequence decoded by a model that is forced
    onto the general path before every call (a 2-token call in between resets both shortcuts), and jumping back to an earlier
    position must reproduce the logits first seen there (the mask hides the cells behind it again)."""
    shape = SHAPES["test-dense"]
    tensors = make_tensors(shape, 11)
    toks = [(t * 37 + 3) % shape.n_vocab for t in range(24)]
    fast = build(shape, tensors, 512, type_kv)
    slow = build(shape, tensors, 512, type_kv)
    try:
        a, b = [], []
        for t, tok in enumerate(toks):
            a.append(fast.decode([tok], t).copy())
            slow.decode([tok, tok], 300)               # different token count + far-away cells: full rebuild next call
            b.append(slow.decode([tok], t).copy())
        assert np.array_equal(np.concatenate(a), np.concatenate(b))
        # rewind: position 10 again (cells 10.. still hold newer rows; the rebuilt mask must hide them)
        again = fast.decode([toks[10]], 10).copy()
        assert np.array_equal(again, a[10])
        nxt = fast.decode([toks[11]], 11).copy()      # and the one-element update continues from the rewound state
        assert np.array_equal(nxt, a[11])
    finally:
        fast.close()
        slow.close()
