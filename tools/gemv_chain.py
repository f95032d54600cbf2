"""Developer tool: a chain of identical decode GEMVs captured in one CUDA graph (PDL on), weights either rotating through > L2
(cold: every launch streams from HBM) or one buffer (hot: L2-resident, so the time is the kernel's fixed + compute cost)."""
import ctypes as C
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import koboldcpp_b200.host as H  # noqa: E402
from koboldcpp_b200.lib import check, lib, row_size  # noqa: E402

L = lib()


def chain(t, K, M, hot, n=40, reps=20):
    rs = row_size(t, K)
    nbytes = rs * M
    ncopy = 1 if hot else max(2, int(500e6 // nbytes) + 1)
    rng = np.random.default_rng(0)
    host = rng.integers(0, 256, size=nbytes, dtype=np.uint8)
    bufs = [H.DeviceBuffer.from_numpy(host, pad=256) for _ in range(ncopy)]
    x = H.DeviceBuffer.from_numpy(rng.standard_normal(K).astype(np.float32))
    d = H.DeviceBuffer(4 * M + 256)
    st = C.c_void_p()
    check(L.b200_stream_create(C.byref(st)))
    L.b200_set_pdl(1)
    check(L.b200_graph_begin_capture(st))
    for i in range(n):
        check(L.b200_mul_mat_vec_q(t, bufs[i % ncopy].ptr, rs, x.ptr, K, d.ptr, M, K, M, 1, st))
    g = C.c_void_p()
    check(L.b200_graph_end_capture(st, C.byref(g)))
    for _ in range(3):
        check(L.b200_graph_launch(g, st))
    check(L.b200_stream_synchronize(st))
    e0, e1 = C.c_void_p(), C.c_void_p()
    check(L.b200_event_create(C.byref(e0))); check(L.b200_event_create(C.byref(e1)))
    check(L.b200_event_record(e0, st))
    for _ in range(reps):
        check(L.b200_graph_launch(g, st))
    check(L.b200_event_record(e1, st))
    ms = C.c_float()
    check(L.b200_event_elapsed_ms(e0, e1, C.byref(ms)))
    L.b200_graph_destroy(g)
    L.b200_set_pdl(0)
    for b in bufs:
        b.free()
    x.free(); d.free()
    us = ms.value * 1e3 / reps / n
    return us, nbytes / us / 1e3


if __name__ == "__main__":
    names = {2: "q4_0", 8: "q8_0", 12: "q4_K", 13: "q5_K", 14: "q6_K"}
    shapes = [(12, 4096, 4096), (12, 4096, 12288), (12, 4096, 22016), (12, 11008, 4096), (14, 11008, 4096), (14, 4096, 4096), (14, 4096, 32000), (12, 8192, 8192), (12, 28672, 8192)]
    for (t, K, M) in shapes:
        uc, gc = chain(t, K, M, False)
        uh, gh = chain(t, K, M, True)
        print("%-5s K=%6d M=%6d %6.1f MB  cold %6.2f us %5.0f GB/s   hot(L2) %6.2f us %5.0f GB/s" % (names[t], K, M, row_size(t, K) * M / 1e6, uc, gc, uh, gh))
