"""Developer tool: per-CTA timeline of the GEMV kernel (needs build/libtrace.so compiled with -DB200_MMVQ_TRACE)."""
import ctypes as C
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import koboldcpp_b200.lib as L  # noqa: E402

lib = L.load(ROOT / "build" / "libtrace.so")
import koboldcpp_b200.host as H  # noqa: E402

t, K, M, pdl = (int(a) for a in sys.argv[1:5])
rs = L.row_size(t, K)
rng = np.random.default_rng(0)
host = rng.integers(0, 256, size=rs * M, dtype=np.uint8)
ncopy = max(2, int(400e6 // host.nbytes) + 1)
bufs = [H.DeviceBuffer.from_numpy(host, pad=16) for _ in range(ncopy)]
x = H.DeviceBuffer.from_numpy(rng.standard_normal((1, K)).astype(np.float32))
d = H.DeviceBuffer(4 * M)
lib.b200_set_pdl(pdl)
for i in range(30):
    L.check(lib.b200_mul_mat_vec_q(t, bufs[i % ncopy].ptr, rs, x.ptr, K, d.ptr, M, K, M, 1, None))
H.sync()
out = np.zeros(148 * 8, dtype=np.uint64)
lib.b200_mmvq_trace.argtypes = [C.c_void_p]
L.check(lib.b200_mmvq_trace(out.ctypes.data))
tr = out.reshape(148, 8).astype(np.int64)
t0 = tr[:, 0].min()
rel = tr - t0
names = ["start", "prologue", "issued", "pdl_wait", "acts", "first_data", "first_row", "end"]
print(f"type {t} K {K} M {M} pdl {pdl}: ns relative to the first CTA start (min / median / max over CTAs)")
for i, n in enumerate(names):
    c = rel[:, i]
    print(f"  {n:10s} {c.min():7d} {int(np.median(c)):7d} {c.max():7d}")
