"""Auxiliary measurements of the BASELINE.json configurations that are not the round bench line (bench.py is configs[1]):
  C3  Llama-2-13B-shaped Q5_K_M (Q6_K mix): pp512 prefill through the tensor-core GEMM + multi-column attention
  C4  Mixtral-8x7B-shaped Q4_K_M: batch-1 decode with MUL_MAT_ID expert GEMVs and a q8_0 KV cache at n_kv = 32768
  C1  TinyLlama-1.1B-shaped Q4_0: pp128 + tg128 (the reference's CPU-runnable case, here on the GPU)
One JSON line per configuration; random-init weights, synthetic KV; timed like bench.py (CUDA events on the backend stream
for the resident replay, wall clock around the public call for e2e / prefill).

    python tools/bench_configs.py [c1] [c3] [c4]
"""
import ctypes as C
import json
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import koboldcpp_b200.llama as LL  # noqa: E402
from koboldcpp_b200.lib import F16, Q8_0, check, lib  # noqa: E402


def decode_leg(m, L, n_prefilled, n_ctx, steps=64, warmup=3):
    span = n_ctx - n_prefilled
    tok = 1
    for i in range(warmup):
        tok = int(np.argmax(m.decode([tok], n_prefilled + i % span)[0]))
    m.synchronize()
    t0 = time.perf_counter()
    for i in range(steps):
        tok = int(np.argmax(m.decode([tok], n_prefilled + (warmup + i) % span)[0]))
    e2e_ms = (time.perf_counter() - t0) * 1e3 / steps
    e0, e1 = C.c_void_p(), C.c_void_p()
    check(L.b200_event_create(C.byref(e0))); check(L.b200_event_create(C.byref(e1)))
    for _ in range(warmup):
        m.replay()
    m.synchronize()
    check(L.b200_event_record(e0, m.stream))
    for _ in range(steps):
        m.replay()
    check(L.b200_event_record(e1, m.stream))
    ms = C.c_float()
    check(L.b200_event_elapsed_ms(e0, e1, C.byref(ms)))
    return ms.value / steps, e2e_ms


def prefill_leg(m, shape, n_tokens, reps=4):
    toks = [(i * 7919) % shape.n_vocab for i in range(n_tokens)]
    t = []
    for _ in range(reps):
        m.synchronize()
        t0 = time.perf_counter()
        m.decode(toks, 0)
        t.append((time.perf_counter() - t0) * 1e3)
    return min(t[1:])


def main():
    which = [a.lower() for a in sys.argv[1:]] or ["c1", "c3", "c4"]
    L = lib()
    check(L.b200_set_device(0))
    peak = 6458.1
    try:
        peak = float(json.loads((Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
    except Exception:
        pass
    if "c1" in which:
        s = LL.SHAPES["tinyllama-1.1b-q4_0"]
        m = LL.LlamaModel(s, 2048, F16)
        m.init_random(seed=1); m.finalize(max_batch=128); m.kv_fill_random(128, seed=2)
        pp = prefill_leg(m, s, 128)
        res, e2e = decode_leg(m, L, 128, 256, steps=128)
        alg = LL.weight_bytes_per_token(s) + LL.kv_bytes_per_token(s, 256, F16)
        print(json.dumps({"config": "C1 tinyllama-1.1b-q4_0, f16 KV, pp128 + tg128 at n_kv=256", "pp128_tok_s": 128 / (pp / 1e3), "decode_tok_s": 1e3 / res,
                          "decode_e2e_tok_s": 1e3 / e2e, "ms_per_token": res, "hbm_frac": alg / res / 1e6 / peak, "launches_per_token": m.stats()["n_kernels"]}), flush=True)
        m.close()
    if "c3" in which:
        s = LL.SHAPES["llama2-13b-q5_k_m"]
        m = LL.LlamaModel(s, 1024, F16)
        m.init_random(seed=3); m.finalize(max_batch=512)
        pp = prefill_leg(m, s, 512)
        # GEMM flops of one 512-token batch: 2 * 512 * (weights per token, in elements)
        n_w = sum(np.prod(LL.tensor_shape(s, n)[:2]) for n, ty in LL.tensor_types(s).items() if n != "token_embd.weight" and not n.endswith("norm.weight"))
        # only the last token's logits are computed (inp_out_ids): the head runs on 1 column
        n_w_batch = n_w - s.n_embd * s.n_vocab
        tflop = 2.0 * 512 * n_w_batch / 1e12
        print(json.dumps({"config": "C3 llama2-13b-q5_k_m (Q6_K mix), pp512 at n_past=0, f16 KV", "pp512_tok_s": 512 / (pp / 1e3), "ms": pp, "gemm_tflop": tflop,
                          "tflop_s_whole_step": tflop / (pp / 1e3), "note": "wall clock around b200_llama_decode (host tokens in, last-token logits out), best of 3"}), flush=True)
        m.close()
    if "c4" in which:
        s = LL.SHAPES["mixtral-8x7b-q4_k_m"]
        n_ctx = 32768
        m = LL.LlamaModel(s, n_ctx, Q8_0)
        m.init_random(seed=4); m.finalize(max_batch=8)
        m.kv_fill_random(n_ctx - 128, seed=5)
        res, e2e = decode_leg(m, L, n_ctx - 128, n_ctx, steps=48)
        alg = LL.weight_bytes_per_token(s) + LL.kv_bytes_per_token(s, n_ctx, Q8_0)
        print(json.dumps({"config": "C4 mixtral-8x7b-q4_k_m, MUL_MAT_ID experts (2 of 8), q8_0 KV, decode at n_kv=32768", "decode_tok_s": 1e3 / res, "decode_e2e_tok_s": 1e3 / e2e,
                          "ms_per_token": res, "alg_bytes_per_token": alg, "hbm_frac": alg / res / 1e6 / peak, "launches_per_token": m.stats()["n_kernels"]}), flush=True)
        m.close()


if __name__ == "__main__":
    main()
