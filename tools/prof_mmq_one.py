"""Developer tool: a few launches of the prefill GEMM at one shape for ncu (python tools/prof_mmq_one.py TYPE K M N)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent))
import bench_mmq  # noqa: E402

t, K, M, N = (int(a) for a in sys.argv[1:5])
ms, tf = bench_mmq.bench(t, K, M, N, iters=4)
print(f"type {t} K {K} M {M} N {N}: {ms:.3f} ms {tf:.1f} TFLOP/s")
