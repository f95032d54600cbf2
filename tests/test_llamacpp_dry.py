"""Drop-in boundary under the reference's REAL llama.cpp + ggml_backend_sched, on a GPU-less host (B200_DRY_RUN).

oracle/_ref/llama_harness links the unmodified reference (src/llama.cpp, ggml-backend.cpp) and attaches libkcpp_b200.so with
ggml_backend_register (Route 2).  In a dry run the backend emulates its devices in host memory: model load, buffer placement,
scheduler split assignment (ggml-backend.cpp:1615-1975), supports_op coverage and the graph planner all run for real; kernels
do not (outputs are zeros).  These tests pin WHICH nodes the scheduler leaves on the CPU backend: only the token-embedding
GET_ROWS, whose weight llama.cpp itself keeps in a host buffer (src/llama.cpp: buft_input = CPU).  The numeric side of the same
runs is tests/test_gpu_llamacpp.py.
"""
from __future__ import annotations

import pytest

import gguf_fixtures as G
import llamacpp as LC

pytestmark = pytest.mark.skipif(not LC.available(), reason="oracle/_ref/llama_harness not built (needs /root/reference at build time)")

ONLY_EMBD = [("GET_ROWS", "inp_embd")]


@pytest.fixture(scope="module")
def dense():
    return G.build_fixture("test-dense", layer_gain=1.5)


@pytest.fixture(scope="module")
def moe():
    return G.build_fixture("test-moe", layer_gain=1.5)


@pytest.mark.parametrize("opts", [dict(), dict(type_kv="q8_0"), dict(type_kv="q4_0"), dict(fa=0), dict(n_prompt=200, n_batch=64, n_ubatch=64)])
def test_dense_graph_stays_on_the_backend(dense, opts):
    o = LC.run(dense, "b200", n_prompt=opts.pop("n_prompt", 40), n_gen=4, sched_debug=True, dry_devices=1, **opts)
    assert o["sched"]["cpu_nodes"] == ONLY_EMBD, o["sched"]
    assert o["sched"]["nodes_per_backend"].get("B200_", 0) > 100


@pytest.mark.parametrize("ftype", ["Q4_1", "Q5_0", "Q5_1", "Q2_K", "Q3_K", "IQ4_NL", "IQ4_XS"])
def test_other_weight_types_stay_on_the_backend(ftype):
    """N4: every matmul of a model quantised to one of the further formats is accepted by supports_op and planned on the backend."""
    model = G.build_fixture("test-dense", layer_gain=1.5, ftype=ftype)
    for n_prompt in (1, 40):
        o = LC.run(model, "b200", n_prompt=n_prompt, n_gen=2, sched_debug=True, dry_devices=1)
        assert o["sched"]["cpu_nodes"] == ONLY_EMBD, o["sched"]


@pytest.mark.parametrize("type_kv", ["f16", "q8_0"])
def test_k_shift_graph_stays_on_the_backend(dense, type_kv):
    """koboldcpp's ContextShift: llama_kv_cache_seq_rm + seq_add -> build_k_shift (src/llama.cpp:10144-10192): ROPE in place on the
    f16 cache, or CPY(q8_0 -> f32) -> ROPE -> CPY(f32 -> q8_0) pinned to the KV backend (used to abort / fall to the CPU)."""
    o = LC.run(dense, "b200", n_prompt=40, n_gen=8, sched_debug=True, dry_devices=1, type_kv=type_kv, shift_at=4, shift_by=8)
    assert o["sched"]["cpu_nodes"] == ONLY_EMBD, o["sched"]


@pytest.mark.parametrize("n_prompt", [1, 40])
def test_moe_graph_stays_on_the_backend(moe, n_prompt):
    """MUL_MAT_ID for decode (expert GEMV) AND for prompt batches (grouped tensor-core GEMM): no expert matmul on the CPU."""
    o = LC.run(moe, "b200", n_prompt=n_prompt, n_gen=3, sched_debug=True, dry_devices=1)
    assert o["sched"]["cpu_nodes"] == ONLY_EMBD, o["sched"]


@pytest.mark.parametrize("mode,n_dev", [("row", 2), ("row", 8), ("layer", 2)])
def test_multi_device_split_modes_load_and_schedule(dense, mode, n_dev):
    """LLAMA_SPLIT_MODE_ROW finds the split buffer type through get_proc_address("ggml_backend_split_buffer_type")
    (src/llama.cpp:3498-3510, 7039-7062); LAYER creates one backend per device."""
    o = LC.run(dense, "b200", n_prompt=24, n_gen=3, sched_debug=True, dry_devices=n_dev, split_mode=mode)
    assert o["n_devices"] == n_dev
    assert o["sched"]["cpu_nodes"] == ONLY_EMBD, o["sched"]


def test_reference_cpu_run_is_deterministic(dense):
    a = LC.run(dense, "cpu", n_prompt=32, n_gen=16, threads=4)
    b = LC.run(dense, "cpu", n_prompt=32, n_gen=16, threads=2)
    assert a["tokens"] == b["tokens"] and len(a["tokens"]) == 17
