"""Developer micro-benchmark: tensor-pipe throughput of the tcgen05 prefill GEMM (b200_mul_mat_q) at the 13B / 7B pp512 shapes.
Prints TFLOP/s (2*M*N*K / time) per shape; weights rotate through > L2-sized sets.  Not the round bench (bench.py)."""
import ctypes as C
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import koboldcpp_b200.host as H  # noqa: E402
from koboldcpp_b200.lib import Q4_K, Q5_K, Q6_K, TYPE_NAMES, check, lib, row_size  # noqa: E402

L = lib()


def bench(t, K, M, N, iters=30):
    rs = row_size(t, K)
    rng = np.random.default_rng(0)
    from koboldcpp_b200.llama import random_blocks
    host = random_blocks(rng, t, min(M, 512), K)
    host = np.tile(host, ((M + 511) // 512, 1))[:M]
    ncopy = max(2, int(300e6 // host.nbytes) + 1)
    bufs = [H.DeviceBuffer.from_numpy(host, pad=256) for _ in range(ncopy)]
    x = H.DeviceBuffer.from_numpy(rng.standard_normal((N, K)).astype(np.float32))
    d = H.DeviceBuffer(4 * N * M + 256)
    wsb = L.b200_mul_mat_q_workspace_bytes(t, K, M, N)
    ws = H.DeviceBuffer(wsb + 16)

    def fn(i):
        check(L.b200_mul_mat_q(t, bufs[i % ncopy].ptr, rs, x.ptr, K, d.ptr, M, K, M, N, ws.ptr, wsb, None))
    for i in range(3):
        fn(i)
    H.sync()
    e0, e1 = C.c_void_p(), C.c_void_p()
    check(L.b200_event_create(C.byref(e0))); check(L.b200_event_create(C.byref(e1)))
    L.b200_event_record(e0, None)
    for i in range(iters):
        fn(i)
    L.b200_event_record(e1, None)
    ms = C.c_float()
    check(L.b200_event_elapsed_ms(e0, e1, C.byref(ms)))
    t_ms = ms.value / iters
    for b in bufs:
        b.free()
    return t_ms, 2.0 * M * N * K / (t_ms * 1e-3) / 1e12


if __name__ == "__main__":
    print("type,K,M,N,ms,TFLOPs")
    for (t, K, M) in [(Q5_K, 5120, 5120), (Q5_K, 5120, 13824), (Q6_K, 13824, 5120), (Q5_K, 13824, 5120), (Q4_K, 4096, 4096), (Q4_K, 4096, 11008), (Q6_K, 4096, 32000)]:
        for N in (512, 128, 32):
            ms, tf = bench(t, K, M, N)
            print(f"{TYPE_NAMES[t]},{K},{M},{N},{ms:.3f},{tf:.1f}", flush=True)
