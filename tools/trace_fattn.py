"""Developer tool: per-CTA timeline of the flash-decode kernel inside a CHAIN of launches (CUDA graph, PDL on).
Needs build/libtrace.so (tools/build_trace.sh).   python tools/trace_fattn.py TKV N_HEAD N_HEAD_KV D N_KV
Stamps (globaltimer ns, relative to the END of the previous launch): start, pdl_wait, prologue (Q fragments), first_tile, loop_end,
partial (cross-warp merge + partial written), ticket (arrival counter known), end (merging CTA only)."""
import ctypes as C
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import koboldcpp_b200.lib as L  # noqa: E402

lib = L.load(ROOT / "build" / "libtrace.so")
import koboldcpp_b200.host as H  # noqa: E402
from koboldcpp_b200.lib import FattnArgs  # noqa: E402

tkv, nh, nhkv, D, nkv = (int(a) for a in sys.argv[1:6])
N = 24
rb = D * 2 if tkv == 1 else D // 32 * 34
nbytes = nhkv * nkv * rb
ncopy = max(2, int(400e6 // (2 * nbytes)) + 1)
rng = np.random.default_rng(0)
host = rng.integers(0, 64, size=nbytes, dtype=np.uint8)
ks = [H.DeviceBuffer.from_numpy(host, pad=64) for _ in range(ncopy)]
vs = [H.DeviceBuffer.from_numpy(host, pad=64) for _ in range(ncopy)]
q = H.DeviceBuffer.from_numpy(rng.standard_normal((nh, 1, D)).astype(np.float32))
d = H.DeviceBuffer(4 * nh * D)
args = [FattnArgs(q.ptr, ks[i].ptr, vs[i].ptr, None, d.ptr, D, 1, nh, nhkv, nkv, tkv, tkv, 4 * D, 4 * D, rb * nhkv, rb, rb * nhkv, rb, 0, 0.088, 0.0, 0.0)
        for i in range(ncopy)]
ws = H.DeviceBuffer(max(lib.b200_flash_attn_workspace_bytes(C.byref(args[0])), 256), zero=True)
st = C.c_void_p()
L.check(lib.b200_stream_create(C.byref(st)))
lib.b200_set_pdl(1)
out = np.zeros(64 * 320 * 8, dtype=np.uint64)
lib.b200_fattn_trace.argtypes = [C.c_void_p]
L.check(lib.b200_fattn_trace(out.ctypes.data))
L.check(lib.b200_graph_begin_capture(st))
for i in range(N):
    L.check(lib.b200_flash_attn_ext(C.byref(args[i % ncopy]), ws.ptr, ws.nbytes, st))
g = C.c_void_p()
L.check(lib.b200_graph_end_capture(st, C.byref(g)))
for _ in range(4):
    L.check(lib.b200_graph_launch(g, st))
L.check(lib.b200_stream_synchronize(st))
L.check(lib.b200_fattn_trace(out.ctypes.data))
tr = out.reshape(64, 320, 8)[:N].astype(np.int64)
used = (tr[8, :, 0] != 0)
tr = tr[:, used, :]
names = ["start", "pdl_wait", "prologue", "first_tile", "loop_end", "partial", "ticket", "end"]
ends = np.maximum(tr[:, :, 6].max(axis=1), tr[:, :, 7].max(axis=1))
print(f"tkv {tkv} heads {nh}/{nhkv} D {D} n_kv {nkv}: {used.sum()} CTAs, launch period {np.median(np.diff(ends[8:])):.0f} ns; ns relative to the END of the previous launch")
for i, n in enumerate(names):
    rel = tr[8:, :, i] - ends[7:-1, None]
    if i == 7:
        rel = np.where(tr[8:, :, 7] >= tr[8:, :, 6], rel, -10**9)  # only the merging CTAs stamp it in this launch
        print(f"  {n:10s} max over CTAs (merging CTA) {int(np.median(rel.max(axis=1))):7d}")
        continue
    print(f"  {n:10s} {int(np.median(rel.min(axis=1))):7d} {int(np.median(np.median(rel, axis=1))):7d} {int(np.median(rel.max(axis=1))):7d}")
