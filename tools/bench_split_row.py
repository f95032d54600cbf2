"""Developer measurement: the reference's unmodified llama.cpp (oracle/_ref/llama_harness) driving this backend with
LLAMA_SPLIT_MODE_ROW over the row-split buffer type (a13 / a14) on all visible GPUs, next to the same run on one GPU.
7B-shaped Q4_K_M GGUF of bench.py, 512-token prompt, 64 greedy tokens."""
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
from koboldcpp_b200.llama import SHAPES  # noqa: E402

shape = SHAPES[bench.WORKLOAD]
model = bench.bench_gguf(shape)
common = ["--n-prompt", 512, "--n-gen", 64, "--n-warm", 3, "--n-batch", 512, "--n-ubatch", 512]
out = {}
os.environ["CUDA_VISIBLE_DEVICES"] = os.environ.get("CUDA_VISIBLE_DEVICES", "0,1")
j = bench.run_harness(model, "b200", common + ["--split-mode", "row"])
out["split_mode_row"] = {"n_devices": j.get("n_devices"), "tg_tok_s": j["tg_tok_s"], "pp_tok_s": j["pp_tok_s"]}
os.environ["CUDA_VISIBLE_DEVICES"] = "0"
j = bench.run_harness(model, "b200", common)
out["one_gpu"] = {"n_devices": j.get("n_devices"), "tg_tok_s": j["tg_tok_s"], "pp_tok_s": j["pp_tok_s"]}
print(json.dumps(out))
