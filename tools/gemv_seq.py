"""Developer tool: the 7B token's GEMV launch sequence alone (bench.py's roofline leg), for kernel tuning."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from koboldcpp_b200.llama import SHAPES  # noqa: E402

r = bench.gemv_leg(SHAPES[sys.argv[2] if len(sys.argv) > 2 else "llama2-7b-q4_k_m"], int(sys.argv[1]) if len(sys.argv) > 1 else 30, 3)
print("gemv_seq: %.3f ms/token, %d launches, %.2f us/launch, %.0f GB/s" % (r["ms_per_token"], r["launches_per_token"], r["avg_launch_us"], r["gbs"]))
