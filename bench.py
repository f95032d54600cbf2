#!/usr/bin/env python
"""Round benchmark: batch-1 decode of a Llama-2-7B-shaped Q4_K_M model at 4k context on B200 (BASELINE.json configs[1]).

One "step" = one decoded token = one pass of the hot path (all MUL_MATs as decode GEMVs + FLASH_ATTN_EXT over the f16 KV
cache + the small ops of the graph) through the ggml_backend boundary of libkcpp_b200.so.

Legs (all printed in ONE JSON line by rank 0):
  value      tokens/s with everything resident in HBM: graph_compute replayed K times, CUDA events on the backend stream.
  e2e        tokens/s through the public call b200_llama_decode with HOST buffers: per step host token id -> H2D (token,
             position, KQ mask), graph_compute, D2H of the logits (pinned), host argmax feeds the next token (greedy).
  roofline   the decode GEMV family: the token's GEMV launch sequence (same types, shapes and order, rotating weight sets
             larger than L2) timed alone with CUDA events; achieved = algorithmic bytes / time against the measured HBM peak.
  e2e_llamacpp  the same metric with the REFERENCE's unmodified llama.cpp as the host: llama_decode under ggml_backend_sched with
             libkcpp_b200.so attached by ggml_backend_register, on a GGUF file of the workload (tools/make_gguf.py).
  cpu_baseline  the reference's own llama.cpp + CPU backend (oracle/_ref) decoding the same GGUF on this box's host cores
             (bounded sample: 16 tokens at n_kv = 4096).
  tp_70b     auxiliary: BASELINE.json configs[4], ONE Llama-3-70B-shaped Q4_K_M model sharded over all N ranks (strong scaling,
             NCCL all-reduce / all-gather in the captured graph); at N = 1 the same model on one GPU, so 1 -> N speed-up is
             the ratio of the tp_70b values of two runs.
`--impl reference` prints the CPU arm only.   N > 1: the headline legs run one 7B replica per rank (weak scaling, no data-path
collective: the 7B model does not need sharding); the sharded path is the tp_70b leg.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOAD = "llama2-7b-q4_k_m"
N_CTX = 4096
N_PREFILLED = 3968  # tg128 at 4k context: cells [0, 3968) pre-filled, positions 3968..4095 decoded (n_kv = 4096 throughout)
PP_BATCH = 512     # auxiliary leg: pp512 prefill of the same model (tensor-core GEMM + multi-column flash attention)
METRIC = "decode_tokens_per_s"
UNIT = "tok/s"


def peaks() -> tuple[float, str]:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.rows: list[list[str]] = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ llama.cpp-level legs
# The reference's own, unmodified llama.cpp (oracle/_ref/libllama_ref*.so + llama_harness, compiled from /root/reference by
# oracle/Makefile; the binaries travel to the GPU box) decodes the SAME GGUF file: on its CPU backend (reference arm,
# cpu_baseline) and with libkcpp_b200.so attached through ggml_backend_register (e2e_llamacpp).
def _cpu_flags() -> set:
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def harness_path():
    ref = ROOT / "oracle" / "_ref"
    if {"avx512f", "avx512bw", "avx512vl", "avx512dq", "avx512_vnni"} <= _cpu_flags() and (ref / "llama_harness_avx512").exists():
        return ref / "llama_harness_avx512", "avx512-vnni build"
    return (ref / "llama_harness"), "avx2 build"


def host_threads() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def bench_gguf(shape) -> Path:
    """The workload as a GGUF file (gguf-py writer, tools/make_gguf.py): the configuration's shapes and llama_tensor_get_type mix
    with random-init quantised blocks (throughput runs need no trained or even reference-quantised values)."""
    d = Path(os.environ.get("B200_GGUF_CACHE", "/tmp/b200_gguf"))
    d.mkdir(parents=True, exist_ok=True)
    out = d / f"bench_{shape.name}.gguf"
    if not out.exists():
        sys.path.insert(0, str(ROOT / "tools"))
        import make_gguf
        tmp = out.with_suffix(".tmp")
        make_gguf.build(shape, tmp, None, seed=1234, replicate_layers=True)
        tmp.rename(out)
    return out


def run_harness(model: Path, backend: str, extra: list, timeout: float = 900.0) -> dict:
    exe, build = harness_path()
    if not exe.exists():
        raise FileNotFoundError(f"{exe} missing (built by oracle/Makefile where /root/reference exists)")
    cmd = [str(exe), "--model", str(model), "--quiet", "--n-ctx", str(N_CTX)] + [str(x) for x in extra]
    if backend == "b200":
        cmd += ["--backend-lib", str(ROOT / "koboldcpp_b200" / "libkcpp_b200.so")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    if r.returncode != 0:
        raise RuntimeError(f"llama_harness rc={r.returncode}: {r.stderr[-600:]}")
    j = json.loads(r.stdout.strip().splitlines()[-1])
    j["build"] = build
    return j


def llamacpp_cpu_decode(shape, n_gen: int, n_warm: int) -> dict:
    """tg at n_kv = 4096 through llama_decode on the reference CPU backend with all host threads.  The KV cache is brought to
    N_PREFILLED occupied cells by the harness (--kv-fill: the prompt's rows repeated through the reference's own session-state
    format) instead of a 3968-token CPU prompt (~15 tok/s: minutes); the timed decode steps are the real thing."""
    thr = host_threads()
    j = run_harness(bench_gguf(shape), "cpu", ["--n-prompt", 32, "--kv-fill", N_PREFILLED, "--n-gen", n_gen, "--n-warm", n_warm, "--threads", thr])
    return {"tok_s": j["tg_tok_s"], "ms_per_token": 1e3 / j["tg_tok_s"], "cores": j["threads"],
            "sample": f"reference llama.cpp + ggml CPU backend ({j['build']}), llama_decode on the same GGUF, {j['n_gen']} greedy tokens timed after {n_warm} warm-up "
                      f"at n_kv={N_CTX} (KV filled to {N_PREFILLED} cells via llama_state_set_data), {j['threads']} threads"}


def reference_arm(args, rank: int) -> int:
    """The reference's own CPU implementation of the path on this box's host cores: its llama.cpp + ggml CPU backend."""
    if rank != 0:
        return 0
    from koboldcpp_b200.llama import SHAPES
    shape = SHAPES[WORKLOAD]
    try:
        r = llamacpp_cpu_decode(shape, max(4, min(args.steps, 160)), max(1, min(args.warmup, 8)))
    except Exception as ex:  # the harness is prebuilt in oracle/_ref; fall back to the per-layer sample of the same backend
        sys.path.insert(0, str(ROOT / "tests"))
        import cpu_baseline as B
        r = B.time_decode(shape, N_CTX, budget_s=min(150.0, max(10.0, 1.0 * (args.steps + args.warmup))))
        r["sample"] += f" [llama.cpp harness unavailable: {ex}]"
    line = {"impl": "reference", "metric": METRIC, "value": r["tok_s"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": r["ms_per_token"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8 dot / f32 accumulate",
            "data": "synthetic", "config": config_dict(shape),
            "cpu_baseline": {"value": r["tok_s"], "unit": UNIT, "cores": r["cores"], "kind": "reference", "sample": r["sample"]},
            "e2e": {"value": r["tok_s"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


def config_dict(shape) -> dict:
    return {"workload": f"{shape.name} random-init GGUF-shaped model, batch-1 greedy decode (tg128) at n_kv={N_CTX} (KV pre-filled to {N_PREFILLED}), f16 KV, flash attention",
            "n_embd": shape.n_embd, "n_layer": shape.n_layer, "n_head": shape.n_head, "n_ff": shape.n_ff, "n_vocab": shape.n_vocab,
            "l2_policy": "inputs larger than L2: 4.0 GB of weights + 2.1 GB of KV streamed per step vs 126 MB L2"}


# ------------------------------------------------------------------------------------------------ GEMV roofline leg
def gemv_launch_plan(shape):
    """The decode GEMV launches of one token as graph_compute issues them (backend_graph.cu: siblings reading the same
    activation are one multi-job launch; two K-quant weight types may share it).  Returns per layer a list of
    (K, [(type, M), ...])."""
    from koboldcpp_b200.llama import tensor_shape, tensor_types
    types = tensor_types(shape)
    kq = {12, 13, 14} if os.environ.get("B200_MIXED_TYPES") else set()  # the planner's opt-in rule (backend_graph.cu: g_mixed_types)
    layers = []
    for il in range(shape.n_layer):
        p = f"blk.{il}."
        launches = []
        for group in (["attn_q.weight", "attn_k.weight", "attn_v.weight"], ["attn_output.weight"], ["ffn_gate.weight", "ffn_up.weight"], ["ffn_down.weight"]):
            K = tensor_shape(shape, p + group[0])[0]
            members = [(types[p + leaf], tensor_shape(shape, p + leaf)[1]) for leaf in group]
            tys = []
            for ty, _ in members:
                if ty not in tys:
                    tys.append(ty)
            if len(tys) == 1 or (len(tys) == 2 and set(tys) <= kq):
                launches.append((K, members))
            else:
                for ty in tys:
                    launches.append((K, [(t, M) for (t, M) in members if t == ty]))
        layers.append(launches)
    head = (shape.n_embd, [(types["output.weight"], shape.n_vocab)])
    return layers, head


def gemv_leg(shape, steps: int, warmup: int) -> dict:
    """Time the token's GEMV sequence alone: weights rotate through 4 layer-sized sets (> L2 between reuses)."""
    import koboldcpp_b200.host as H
    from koboldcpp_b200.lib import GemvJob, check, lib, row_size
    L = lib()
    layers, head = gemv_launch_plan(shape)
    n_sets = 4
    rng = np.random.default_rng(0)
    kinds = sorted({(K, tuple(mem)) for lay in layers for (K, mem) in lay} | {(head[0], tuple(head[1]))})
    pool = {}
    for (K, mem) in kinds:
        sets = []
        for _ in range(1 if (K, list(mem)) == (head[0], list(head[1])) else n_sets):
            sets.append([H.DeviceBuffer.from_numpy(rng.integers(0, 256, size=row_size(ty, K) * M, dtype=np.uint8), pad=256) for (ty, M) in mem])
        pool[(K, mem)] = sets
    kmax = max(k for k, _ in kinds)
    mmax = max(sum(M for _, M in mem) for _, mem in kinds)
    x = H.DeviceBuffer.from_numpy(rng.standard_normal(kmax).astype(np.float32))
    y = H.DeviceBuffer(4 * mmax + 256)
    st = C.c_void_p()
    check(L.b200_stream_create(C.byref(st)))
    L.b200_set_pdl(1)
    alg_bytes = 0
    n_launch = 0

    def issue():
        nonlocal alg_bytes, n_launch
        alg_bytes = 0
        n_launch = 0
        seq = [(il, l) for il, lay in enumerate(layers) for l in lay] + [(0, head)]
        for il, (K, mem) in seq:
            sets = pool[(K, tuple(mem))]
            bufs = sets[il % len(sets)]
            ty0 = mem[0][0]
            rs = row_size(ty0, K)
            if len(mem) == 1:
                check(L.b200_mul_mat_vec_q(ty0, bufs[0].ptr, rs, x.ptr, K, y.ptr, mem[0][1], K, mem[0][1], 1, st))
            else:
                jobs = (GemvJob * len(mem))()
                off = 0
                for j, (ty, M) in enumerate(mem):
                    jobs[j].W = bufs[j].ptr; jobs[j].X = x.ptr; jobs[j].dst = y.ptr + 4 * off; jobs[j].M = M; jobs[j].id_index = -1
                    jobs[j]._pad = ty if ty != ty0 else 0
                    off += M
                check(L.b200_mul_mat_vec_q_multi(ty0, jobs, len(mem), rs, K, off, K, 1, None, 0, 0, st))
            alg_bytes += sum(row_size(ty, K) * M + 4 * M for (ty, M) in mem) + 4 * K
            n_launch += 1

    check(L.b200_graph_begin_capture(st))
    issue()
    g = C.c_void_p()
    check(L.b200_graph_end_capture(st, C.byref(g)))
    for _ in range(max(3, warmup)):
        check(L.b200_graph_launch(g, st))
    check(L.b200_stream_synchronize(st))
    e0, e1 = C.c_void_p(), C.c_void_p()
    check(L.b200_event_create(C.byref(e0))); check(L.b200_event_create(C.byref(e1)))
    check(L.b200_event_record(e0, st))
    for _ in range(steps):
        check(L.b200_graph_launch(g, st))
    check(L.b200_event_record(e1, st))
    ms = C.c_float()
    check(L.b200_event_elapsed_ms(e0, e1, C.byref(ms)))
    L.b200_graph_destroy(g)
    L.b200_set_pdl(0)
    per_token_ms = ms.value / steps
    for sets in pool.values():
        for bufs in sets:
            for b in bufs:
                b.free()
    x.free(); y.free()
    L.b200_stream_destroy(st)
    return {"ms_per_token": per_token_ms, "launches_per_token": n_launch, "alg_bytes_per_token": alg_bytes,
            "avg_launch_us": per_token_ms * 1e3 / n_launch, "gbs": alg_bytes / per_token_ms / 1e6}


# ------------------------------------------------------------------------------------------------ 70B tensor-parallel leg
TP_WORKLOAD = "llama3-70b-q4_k_m"


def tp_leg(rank: int, world: int, local_rank: int, dist, steps: int, warmup: int) -> dict:
    """BASELINE.json configs[4]: Llama-3-70B-shaped Q4_K_M, ONE model sharded over all `world` ranks (strong scaling): q/k/v, gate/up
    and the head split by rows (whole heads per rank), attn_output / ffn_down split along K with one NCCL all-reduce each, the
    logits all-gathered (koboldcpp_b200/tp.py, csrc/comm.cu).  world == 1 is the same model on one GPU (42 GB of weights).
    Timed like the main legs: resident replay with CUDA events on the backend stream, and the public call with host tokens in /
    host logits out; max over ranks."""
    import koboldcpp_b200.llama as LL
    from koboldcpp_b200 import tp
    from koboldcpp_b200.lib import F16, check, lib
    L = lib()
    shape = LL.SHAPES[TP_WORKLOAD]
    t0 = time.time()
    if world > 1:
        tp.comm_init_from_torch(rank, world)
    m = LL.LlamaModel(shape, N_CTX, F16, main_device=local_rank, tp_rank=rank, tp_world=world)
    m.init_random(seed=4321)            # same stream on every rank: each keeps its shard of the same model
    m.finalize(max_batch=PP_BATCH)
    m.kv_fill_random(N_PREFILLED, seed=11)
    setup_s = time.time() - t0
    span = N_CTX - N_PREFILLED
    stream = m.stream
    ev = [C.c_void_p() for _ in range(4)]
    for e in ev:
        check(L.b200_event_create(C.byref(e)))

    def barrier():
        if dist is not None:
            import torch
            torch.cuda.synchronize()
            dist.barrier()
        m.synchronize()

    tok = 1
    for i in range(warmup):
        tok = int(np.argmax(m.decode([tok], N_PREFILLED + i % span)[0]))
    barrier()
    check(L.b200_event_record(ev[0], stream))
    t1 = time.perf_counter()
    for i in range(steps):
        tok = int(np.argmax(m.decode([tok], N_PREFILLED + (warmup + i) % span)[0]))
    check(L.b200_event_record(ev[1], stream))
    ms = C.c_float()
    check(L.b200_event_elapsed_ms(ev[0], ev[1], C.byref(ms)))
    e2e_ms = max(ms.value, (time.perf_counter() - t1) * 1e3)
    st = m.stats()
    barrier()
    for _ in range(warmup):
        m.replay()
    barrier()
    check(L.b200_event_record(ev[2], stream))
    for _ in range(steps):
        m.replay()
    check(L.b200_event_record(ev[3], stream))
    check(L.b200_event_elapsed_ms(ev[2], ev[3], C.byref(ms)))
    res_ms = ms.value
    barrier()
    pp_tokens = [(i * 7919) % shape.n_vocab for i in range(PP_BATCH)]
    pp_ms = []
    for _ in range(3):
        barrier()
        t2 = time.perf_counter()
        m.decode(pp_tokens, 0)
        pp_ms.append((time.perf_counter() - t2) * 1e3)
    pp_best = min(pp_ms[1:])
    if dist is not None:
        import torch
        t = torch.tensor([res_ms, e2e_ms, pp_best], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res_ms, e2e_ms, pp_best = float(t[0]), float(t[1]), float(t[2])
    m.close()
    if world > 1:
        tp.comm_destroy()
    peak, _ = peaks()
    alg = LL.weight_bytes_per_token(shape) + LL.kv_bytes_per_token(shape, N_CTX, F16)
    return {"workload": f"{shape.name} random-init, batch-1 greedy decode at n_kv={N_CTX}, f16 KV; ONE model over {world} GPU(s)",
            "parallelism": f"tp{world}" if world > 1 else "single", "scaling": "strong", "steps": steps, "warmup": warmup,
            "value": steps / (res_ms / 1e3), "unit": UNIT, "ms_per_step": res_ms / steps,
            "e2e": {"value": steps / (e2e_ms / 1e3), "unit": UNIT, "ms_per_step": e2e_ms / steps},
            "pp512_tokens_per_s": PP_BATCH / (pp_best / 1e3), "pp512_ms": pp_best,
            "alg_bytes_per_step": alg, "hbm_frac_of_aggregate_peak": alg / (res_ms / steps) / 1e6 / (peak * world),
            "launches_per_step": int(st["n_kernels"]), "collectives_per_step": (2 * shape.n_layer + 1) if world > 1 else 0, "setup_s": setup_s}


# ------------------------------------------------------------------------------------------------ product arm
def product_arm(args, rank: int, world: int, local_rank: int) -> int:
    import koboldcpp_b200.llama as LL
    from koboldcpp_b200.lib import F16, check, lib
    L = lib()
    if L.b200_device_count() <= 0:
        print("bench.py: no CUDA device and the product has no CPU path", file=sys.stderr)
        return 2
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist_
        torch.cuda.set_device(local_rank)
        dist_.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist = dist_
    check(L.b200_set_device(local_rank), "set_device")
    if os.environ.get("B200_BENCH_ONLY_TP70"):  # developer shortcut: the sharded 70B leg alone (not a round bench line)
        tp70 = tp_leg(rank, world, local_rank, dist, min(args.steps, 32), 3)
        if rank == 0:
            print(json.dumps({"tp_70b_only": tp70}))
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return 0
    shape = LL.SHAPES[WORKLOAD]
    steps, warmup = args.steps, max(3, args.warmup)
    t_setup = time.time()
    m = LL.LlamaModel(shape, N_CTX, F16, main_device=local_rank)
    m.init_random(seed=1234 + rank)
    m.finalize(max_batch=PP_BATCH)
    m.kv_fill_random(N_PREFILLED, seed=7 + rank)
    setup_s = time.time() - t_setup
    span = N_CTX - N_PREFILLED  # positions cycle inside [3968, 4096): n_kv stays 4096 for any K

    def pos(i):
        return N_PREFILLED + (i % span)

    ev = [C.c_void_p() for _ in range(4)]
    for e in ev:
        check(L.b200_event_create(C.byref(e)))
    stream = m.stream

    def barrier():
        if dist is not None:
            import torch
            torch.cuda.synchronize()
            dist.barrier()
        m.synchronize()

    # ---- e2e leg: the public call with host buffers, greedy token feedback
    tok = 1
    for i in range(warmup):
        tok = int(np.argmax(m.decode([tok], pos(i))[0]))
    barrier()
    clocks = ClockSampler(local_rank)
    check(L.b200_event_record(ev[0], stream))
    t0 = time.perf_counter()
    for i in range(steps):
        tok = int(np.argmax(m.decode([tok], pos(warmup + i))[0]))
    check(L.b200_event_record(ev[1], stream))
    ms = C.c_float()
    check(L.b200_event_elapsed_ms(ev[0], ev[1], C.byref(ms)))
    e2e_wall_ms = (time.perf_counter() - t0) * 1e3
    e2e_ms = max(ms.value, e2e_wall_ms)  # host-synchronous loop: the wall clock bounds the device span from above
    st = m.stats()
    barrier()

    # ---- resident leg: graph_compute replayed with inputs in HBM
    for _ in range(warmup):
        m.replay()
    barrier()
    check(L.b200_event_record(ev[2], stream))
    for _ in range(steps):
        m.replay()
    check(L.b200_event_record(ev[3], stream))
    check(L.b200_event_elapsed_ms(ev[2], ev[3], C.byref(ms)))
    res_ms = ms.value
    clk = clocks.stop()
    barrier()

    # ---- auxiliary leg: pp512 through the same public call (host tokens in, last-token logits out)
    pp_tokens = [(i * 7919) % shape.n_vocab for i in range(PP_BATCH)]
    pp_ms = []
    for _ in range(4):
        t1 = time.perf_counter()
        m.decode(pp_tokens, 0)
        pp_ms.append((time.perf_counter() - t1) * 1e3)
    pp_best = min(pp_ms[1:])
    barrier()

    if dist is not None:
        import torch
        t = torch.tensor([res_ms, e2e_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res_ms, e2e_ms = float(t[0]), float(t[1])
    m.close()

    tp70 = None
    if not args.no_tp70:
        try:
            tp70 = tp_leg(rank, world, local_rank, dist, min(steps, 32), 3)
        except Exception as ex:  # the auxiliary leg must not hide the headline numbers
            tp70 = {"error": f"{type(ex).__name__}: {ex}"}

    if rank == 0:
        peak, peak_src = peaks()
        g = gemv_leg(shape, steps, warmup)
        # DRAM bytes per launch from the committed ncu --set full capture (measured / algorithmic ratio of the same kernels)
        traffic, traffic_src = None, None
        try:
            tj = json.loads((ROOT / "profiles" / "ncu_traffic.json").read_text())
            traffic = float(tj["gemv_dram_bytes_over_algorithmic_bytes"]) * g["alg_bytes_per_token"] / g["launches_per_token"]
            traffic_src = "profiles/ncu_traffic.json"
        except Exception:
            pass
        wbytes = LL.weight_bytes_per_token(shape)
        kvbytes = LL.kv_bytes_per_token(shape, N_CTX, F16)
        tok_s = world * steps / (res_ms / 1e3)
        e2e_tok_s = world * steps / (e2e_ms / 1e3)
        # per step: token id + position + ONE mask element (the cell that becomes visible); the decoder re-sends the whole
        # 4 * n_kv * 32-byte mask only when the position does not follow the previous one, i.e. when pos() wraps (every `span` steps)
        wraps = sum(1 for i in range(steps) if pos(warmup + i) != pos(warmup + i - 1) + 1)
        mask_bytes = 4 + (4 * N_CTX * 32 * wraps) // steps
        line = {
            "metric": METRIC, "value": tok_s, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": res_ms / steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8 dot (dp4a) / f32 accumulate; f16 KV",
            "data": "synthetic", "config": config_dict(shape),
            "e2e": {"value": e2e_tok_s, "unit": UNIT, "h2d_bytes_per_step": 4 + 4 + mask_bytes, "d2h_bytes_per_step": 4 * shape.n_vocab,
                    "ms_per_step": e2e_ms / steps},
            "gpu_launches": int(st["n_kernels"]) * steps * 2,
            "launches_per_step": int(st["n_kernels"]), "graph_nodes_per_step": int(st["n_nodes"]),
            "roofline": {"bound": "hbm", "kernel": "decode GEMV family (mmvq_kernel<TYPE,1,MULTI>), the token's launch sequence timed alone",
                         "achieved": g["gbs"], "peak": peak, "unit": "GB/s", "frac": g["gbs"] / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src,
                         "alg_bytes_per_launch": g["alg_bytes_per_token"] / g["launches_per_token"], "avg_launch_us": g["avg_launch_us"],
                         "launches_per_step": g["launches_per_token"], "share_of_step": g["ms_per_token"] / (res_ms / steps)},
            "step_roofline": {"alg_bytes_per_step": wbytes + kvbytes, "achieved": (wbytes + kvbytes) / (res_ms / steps) / 1e6, "peak": peak, "unit": "GB/s",
                              "frac": (wbytes + kvbytes) / (res_ms / steps) / 1e6 / peak, "frac_of_8TBs_nominal": (wbytes + kvbytes) / (res_ms / steps) / 1e6 / 8000.0},
            "prefill": {"metric": "pp512_tokens_per_s", "value": PP_BATCH / (pp_best / 1e3), "ms": pp_best, "n_tokens": PP_BATCH,
                        "note": "same model, one 512-token batch at n_past=0 through b200_llama_decode (host sync timing, best of 3 after 1 warm-up)"},
            "clocks": clk, "setup_s": setup_s,
        }
        if tp70 is not None:
            line["tp_70b"] = tp70
        if world == 1 and not args.no_llamacpp:
            # the same metric through the REFERENCE's host code: unmodified llama.cpp + ggml_backend_sched drive libkcpp_b200.so
            # (Route 2 attach), host token ids in, host logits out, greedy; real 3968-token prompt in 512-token batches first
            try:
                j = run_harness(bench_gguf(shape), "b200", ["--n-prompt", N_PREFILLED, "--n-gen", min(steps, N_CTX - N_PREFILLED - 8), "--n-warm", 3, "--n-batch", PP_BATCH, "--n-ubatch", PP_BATCH])
                line["e2e_llamacpp"] = {"value": j["tg_tok_s"], "unit": UNIT, "pp_tokens_per_s": j["pp_tok_s"], "n_prompt": j["n_prompt"], "n_gen": j["n_gen"],
                                        "host": "reference llama.cpp (llama_decode, ggml_backend_sched) with the B200 backend attached by ggml_backend_register",
                                        "h2d_bytes_per_step": 4 + 4 + 4 * N_CTX * 32 + 4, "d2h_bytes_per_step": 4 * shape.n_vocab}
            except Exception as ex:
                line["e2e_llamacpp"] = {"value": None, "unit": UNIT, "error": f"{type(ex).__name__}: {ex}"[:300]}
        if world == 1 and not args.no_cpu:
            try:
                r = llamacpp_cpu_decode(shape, 16, 2)
                line["cpu_baseline"] = {"value": r["tok_s"], "unit": UNIT, "cores": r["cores"], "kind": "reference", "sample": r["sample"]}
            except Exception as ex:
                try:  # per-layer sample of the same compiled reference backend
                    sys.path.insert(0, str(ROOT / "tests"))
                    import cpu_baseline as B
                    r = B.time_decode(shape, N_CTX, budget_s=args.cpu_budget)
                    line["cpu_baseline"] = {"value": r["tok_s"], "unit": UNIT, "cores": r["cores"], "kind": "reference", "sample": r["sample"] + f" [llama.cpp harness unavailable: {ex}]"[:200]}
                except Exception as ex2:  # the checker library is prebuilt in oracle/_ref; its absence must not hide the GPU numbers
                    line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"unavailable: {ex2}"}
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=125)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-llamacpp", action="store_true", help="skip the e2e_llamacpp leg (reference llama.cpp driving the B200 backend)")
    ap.add_argument("--cpu-budget", type=float, default=18.0)
    ap.add_argument("--no-tp70", action="store_true", default=bool(os.environ.get("B200_BENCH_NO_TP70")),
                    help="skip the auxiliary 70B tensor-parallel leg (BASELINE.json configs[4])")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return reference_arm(args, rank)
    return product_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    sys.exit(main())
