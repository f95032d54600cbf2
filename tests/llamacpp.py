"""Run the reference's unmodified llama.cpp through oracle/_ref/llama_harness (TEST INFRASTRUCTURE).

run() returns the harness JSON; with sched_debug=True it also parses GGML_SCHED_DEBUG output
(ggml-backend.cpp:1552-1584) into per-backend node counts and the list of nodes that were scheduled on the CPU backend.
"""
from __future__ import annotations

import json
import os
import re
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
REF = ROOT / "oracle" / "_ref"
B200_LIB = ROOT / "koboldcpp_b200" / "libkcpp_b200.so"

_NODE = re.compile(r"^node #\s*(\d+) \(\s*([A-Z_0-9]+)\):\s+(.{0,20}?) \(\s*\S+\) \[\s*(\S*)")


def harness_path() -> Path:
    import ggml_ref as R
    return REF / ("llama_harness_avx512" if R.best_variant() == "avx512" else "llama_harness")


def available() -> bool:
    return (REF / "llama_harness").exists()


def run(model: Path | str, backend: str = "cpu", n_prompt: int = 128, n_gen: int = 128, sched_debug: bool = False, dry_devices: int = 0,
        exe: Path | None = None, env_extra: dict | None = None, timeout: float = 1800, **opts) -> dict:
    cmd = [str(exe or (REF / "llama_harness")), "--model", str(model), "--n-prompt", str(n_prompt), "--n-gen", str(n_gen), "--quiet"]
    if backend == "b200":
        cmd += ["--backend-lib", str(B200_LIB)]
    for k, v in opts.items():
        cmd += ["--" + k.replace("_", "-"), str(v)]
    env = dict(os.environ)
    if sched_debug:
        env["GGML_SCHED_DEBUG"] = "1"
    if dry_devices:
        env["B200_DRY_RUN"] = str(dry_devices)
    env.update(env_extra or {})
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=timeout)
    if r.returncode != 0:
        raise RuntimeError(f"llama_harness rc={r.returncode}\ncmd: {' '.join(cmd)}\nstderr tail:\n{r.stderr[-4000:]}")
    out = json.loads(r.stdout.strip().splitlines()[-1])
    out["stderr"] = r.stderr
    if sched_debug:
        per_backend: dict[str, int] = {}
        cpu_nodes: list[tuple[str, str]] = []
        graphs = 0
        for line in r.stderr.splitlines():
            if line.startswith("## SPLIT #0:"):
                graphs += 1
            m = _NODE.match(line)
            if not m:
                continue
            be = m.group(4) or "?"
            per_backend[be] = per_backend.get(be, 0) + 1
            if be.startswith("CPU"):
                cpu_nodes.append((m.group(2), m.group(3).strip()))
        out["sched"] = {"graphs": graphs, "nodes_per_backend": per_backend, "cpu_nodes": sorted(set(cpu_nodes))}
    return out
