"""Developer tool: a short chain of one decode GEMV shape for ncu (python tools/prof_gemv_one.py TYPE K M [hot])."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent))
import gemv_chain  # noqa: E402

t, K, M = (int(a) for a in sys.argv[1:4])
hot = len(sys.argv) > 4 and sys.argv[4] == "hot"
us, gbs = gemv_chain.chain(t, K, M, hot, n=6, reps=1)
print(f"type {t} K {K} M {M} {'hot' if hot else 'cold'}: {us:.2f} us {gbs:.0f} GB/s")
