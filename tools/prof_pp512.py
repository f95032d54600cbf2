"""Developer tool: ONE pp512 step of a model shape for an ncu launch list (python tools/prof_pp512.py [shape] [n_layer])."""
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import koboldcpp_b200.llama as LL  # noqa: E402
from koboldcpp_b200.lib import F16  # noqa: E402
import dataclasses  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "llama2-7b-q4_k_m"
shape = LL.SHAPES[name]
if len(sys.argv) > 2:
    shape = dataclasses.replace(shape, n_layer=int(sys.argv[2]))
m = LL.LlamaModel(shape, 1024, F16)
m.init_random(seed=1)
m.finalize(max_batch=512)
toks = [(i * 7919) % shape.n_vocab for i in range(512)]
for _ in range(2):
    t0 = time.perf_counter()
    m.decode(toks, 0)
    ms = (time.perf_counter() - t0) * 1e3
print(f"{name} n_layer={shape.n_layer}: pp512 {ms:.2f} ms")
m.close()
