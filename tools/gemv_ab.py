"""Developer tool: A/B the decode GEMV chain between two builds of the library (python tools/gemv_ab.py path/to/lib.so [shapes...]).
Shapes: TYPE:K:M triples; default = the per-rank slices of a Llama-3-70B Q4_K_M layer under 8-way tensor parallelism."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import koboldcpp_b200.lib as L  # noqa: E402

L.load(Path(sys.argv[1]))
sys.path.insert(0, str(ROOT / "tools"))
import gemv_chain  # noqa: E402

shapes = [tuple(int(x) for x in a.split(":")) for a in sys.argv[2:]] or [
    (12, 8192, 1280), (12, 1024, 8192), (12, 8192, 7168), (12, 3584, 8192), (14, 3584, 8192), (14, 8192, 16032),
    (12, 8192, 10240), (12, 8192, 8192), (12, 28672, 8192)]
for (t, K, M) in shapes:
    uc, gc = gemv_chain.chain(t, K, M, False)
    uh, gh = gemv_chain.chain(t, K, M, True)
    print("type %2d K=%6d M=%6d %6.2f MB  cold %6.2f us %5.0f GB/s   hot(L2) %6.2f us" % (t, K, M, L.row_size(t, K) * M / 1e6, uc, gc, uh))
