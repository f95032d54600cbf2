"""GPU (>= 2 devices) parity of the row-split buffer type (a13) and the multi-device matmul (a14): a model whose weight
matrices live in the split buffer type must produce the logits of the same model held on one device."""
from __future__ import annotations

import numpy as np
import pytest

from koboldcpp_b200.lib import F16, lib
from koboldcpp_b200.llama import SHAPES, LlamaModel, tensor_shape
from test_gpu_llama import make_tensors, rel_rms

pytestmark = pytest.mark.gpu


def build(shape, tensors, n_split, max_batch=64):
    m = LlamaModel(shape, 512, F16, n_split_devices=n_split)
    for name, (ty, raw) in tensors.items():
        m.set_tensor(name, ty, tensor_shape(shape, name), raw)
    m.finalize(max_batch)
    return m


@pytest.mark.parametrize("n_split", [2, 4, 8])
def test_row_split_model_matches_single_device(n_split):
    if lib().b200_device_count() < n_split:
        pytest.skip(f"needs {n_split} GPUs")
    from llama_ref import RefModel
    shape = SHAPES["test-dense"]
    tensors = make_tensors(shape, 31)
    one, many = build(shape, tensors, 0), build(shape, tensors, n_split)
    ref = RefModel(shape, tensors, 512, F16)
    try:
        prompt = [(i * 7919 + 5) % shape.n_vocab for i in range(40)]  # N > 8: tensor-core GEMM on every device
        a, b = one.decode(prompt, 0, logits_all=True).copy(), many.decode(prompt, 0, logits_all=True).copy()
        exp = ref.decode(prompt, 0)
        assert rel_rms(b, a) < 5e-3, rel_rms(b, a)
        assert rel_rms(b, exp) < 4e-2
        n_past = len(prompt)
        for t in range(5):  # decode GEMVs on every device
            tok = int(np.argmax(exp[-1])) if t == 0 else t
            a, b = one.decode([tok], n_past).copy(), many.decode([tok], n_past).copy()
            assert rel_rms(b[0], a[0]) < 5e-3, rel_rms(b[0], a[0])
            n_past += 1
    finally:
        one.close()
        many.close()
