"""Developer tool: per-CTA timeline of the decode GEMV inside a CHAIN of launches (CUDA graph, PDL on).
Needs build/libtrace.so (tools/build_trace.sh: the library compiled with -DB200_MMVQ_TRACE).
    python tools/trace_mmvq.py TYPE K M [hot]
Stamps per launch and CTA (globaltimer ns): start, prologue, issued, pdl_wait, acts, first_data, first_row, end.
Printed: for every stamp, the median over CTAs and launches of (stamp - end of the previous launch = max over its CTAs)."""
import ctypes as C
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import koboldcpp_b200.lib as L  # noqa: E402

lib = L.load(ROOT / "build" / "libtrace.so")
import koboldcpp_b200.host as H  # noqa: E402

t, K, M = (int(a) for a in sys.argv[1:4])
hot = len(sys.argv) > 4 and sys.argv[4] == "hot"
N = 32
rs = L.row_size(t, K)
rng = np.random.default_rng(0)
host = rng.integers(0, 256, size=rs * M, dtype=np.uint8)
ncopy = 1 if hot else max(2, int(500e6 // host.nbytes) + 1)
bufs = [H.DeviceBuffer.from_numpy(host, pad=256) for _ in range(ncopy)]
x = H.DeviceBuffer.from_numpy(rng.standard_normal((1, K)).astype(np.float32))
d = H.DeviceBuffer(4 * M + 256)
st = C.c_void_p()
L.check(lib.b200_stream_create(C.byref(st)))
lib.b200_set_pdl(1)
out = np.zeros(256 * 148 * 8, dtype=np.uint64)
lib.b200_mmvq_trace.argtypes = [C.c_void_p]
L.check(lib.b200_mmvq_trace(out.ctypes.data))  # resets the slot counter
L.check(lib.b200_graph_begin_capture(st))
for i in range(N):
    L.check(lib.b200_mul_mat_vec_q(t, bufs[i % ncopy].ptr, rs, x.ptr, K, d.ptr, M, K, M, 1, st))
g = C.c_void_p()
L.check(lib.b200_graph_end_capture(st, C.byref(g)))
for _ in range(4):
    L.check(lib.b200_graph_launch(g, st))
L.check(lib.b200_stream_synchronize(st))
L.check(lib.b200_mmvq_trace(out.ctypes.data))
tr = out.reshape(256, 148, 8)[:N].astype(np.int64)
grid = min(148, M)
tr = tr[:, :grid, :]
names = ["start", "prefilled", "pdl_wait", "issued", "acts", "first_data", "first_row", "end"]
ends = tr[:, :, 7].max(axis=1)
print(f"type {t} K {K} M {M} {'hot' if hot else 'cold'}: launch period {np.median(np.diff(ends[8:])):.0f} ns; ns relative to the END of the previous launch")
print(f"  {'stamp':10s} {'min':>7s} {'median':>7s} {'max':>7s}   (median over launches 8..{N - 1} of the per-launch min / median / max over CTAs)")
for i, n in enumerate(names):
    rel = tr[8:, :, i] - ends[7:-1, None]
    print(f"  {n:10s} {int(np.median(rel.min(axis=1))):7d} {int(np.median(np.median(rel, axis=1))):7d} {int(np.median(rel.max(axis=1))):7d}")
