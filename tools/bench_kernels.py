"""Developer micro-benchmark: achieved HBM GB/s of the decode GEMV and flash-decode kernels at the 7B shapes.
Weights rotate through > L2-size working sets so every launch streams from HBM.  Not the round bench (bench.py)."""
import ctypes as C
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import koboldcpp_b200.host as H  # noqa: E402
from koboldcpp_b200.lib import FattnArgs, check, lib, row_size  # noqa: E402

L = lib()


def ev():
    e = C.c_void_p()
    check(L.b200_event_create(C.byref(e)))
    return e


def time_loop(fn, iters, warm=5):
    for i in range(warm):
        fn(i)
    H.sync()
    e0, e1 = ev(), ev()
    L.b200_event_record(e0, None)
    for i in range(iters):
        fn(i)
    L.b200_event_record(e1, None)
    ms = C.c_float()
    check(L.b200_event_elapsed_ms(e0, e1, C.byref(ms)))
    return ms.value / iters


def bench_gemv(t, K, M, N=1, pdl=0, iters=200):
    rs = row_size(t, K)
    nbytes = rs * M
    ncopy = max(2, int(400e6 // nbytes) + 1)
    rng = np.random.default_rng(0)
    host = rng.integers(0, 256, size=nbytes, dtype=np.uint8)
    # finite scales are irrelevant for timing; keep bytes random
    bufs = [H.DeviceBuffer.from_numpy(host, pad=16) for _ in range(ncopy)]
    x = H.DeviceBuffer.from_numpy(rng.standard_normal((N, K)).astype(np.float32))
    d = H.DeviceBuffer(4 * N * M)
    L.b200_set_pdl(pdl)

    def fn(i):
        check(L.b200_mul_mat_vec_q(t, bufs[i % ncopy].ptr, rs, x.ptr, K, d.ptr, M, K, M, N, None))
    ms = time_loop(fn, iters)
    L.b200_set_pdl(0)
    alg = nbytes + N * K * 4 + N * M * 4
    return ms * 1e3, alg / ms / 1e6


def bench_fattn(tkv, n_head, n_head_kv, D, n_kv, iters=100):
    rb = D * 2 if tkv == 1 else D // 32 * 34
    nbytes = n_head_kv * n_kv * rb
    ncopy = max(2, int(400e6 // (2 * nbytes)) + 1)
    rng = np.random.default_rng(0)
    host = rng.integers(0, 64, size=nbytes, dtype=np.uint8)
    ks = [H.DeviceBuffer.from_numpy(host, pad=64) for _ in range(ncopy)]
    vs = [H.DeviceBuffer.from_numpy(host, pad=64) for _ in range(ncopy)]
    q = H.DeviceBuffer.from_numpy(rng.standard_normal((n_head, 1, D)).astype(np.float32))
    d = H.DeviceBuffer(4 * n_head * D)
    args = []
    for i in range(ncopy):
        # cache layout of llama.cpp: row stride = n_head_kv * rb, head stride = rb
        args.append(FattnArgs(q.ptr, ks[i].ptr, vs[i].ptr, None, d.ptr, D, 1, n_head, n_head_kv, n_kv, tkv, tkv,
                              4 * D, 4 * D, rb * n_head_kv, rb, rb * n_head_kv, rb, 0, 0.088, 0.0, 0.0))
    wsb = L.b200_flash_attn_workspace_bytes(C.byref(args[0]))
    ws = H.DeviceBuffer(max(wsb, 256), zero=True)

    def fn(i):
        check(L.b200_flash_attn_ext(C.byref(args[i % ncopy]), ws.ptr, ws.nbytes, None))
    ms = time_loop(fn, iters)
    return ms * 1e3, 2 * nbytes / ms / 1e6


if __name__ == "__main__":
    H.require_gpu()
    names = {2: "q4_0", 8: "q8_0", 12: "q4_K", 13: "q5_K", 14: "q6_K"}
    if len(sys.argv) > 1 and sys.argv[1] == "gemv":  # gemv <type> <K> <M> <N> <pdl> <iters>
        t, K, M, N, pdl, iters = (int(a) for a in sys.argv[2:8])
        us, gbs = bench_gemv(t, K, M, N, pdl, iters)
        print(f"gemv,{names[t]},{K},{M},{N},{pdl},{us:.2f},{gbs:.0f}")
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "fattn":  # fattn <tkv> <n_head> <n_head_kv> <D> <n_kv> <iters>
        tkv, nh, nhkv, D, nkv, iters = (int(a) for a in sys.argv[2:8])
        us, gbs = bench_fattn(tkv, nh, nhkv, D, nkv, iters)
        print(f"fattn,{tkv},{nh}/{nhkv},{D},{nkv},0,{us:.2f},{gbs:.0f}")
        sys.exit(0)
    print("kernel,type,K,M,N,pdl,us,GBps")
    for t in (12, 14, 13, 8, 2):
        for (K, M) in ((4096, 4096), (4096, 11008), (11008, 4096), (4096, 32000), (8192, 28672)):
            for pdl in (0, 1):
                us, gbs = bench_gemv(t, K, M, 1, pdl)
                print(f"gemv,{names[t]},{K},{M},1,{pdl},{us:.2f},{gbs:.0f}", flush=True)
    for N in (2, 4, 8):
        us, gbs = bench_gemv(12, 4096, 11008, N)
        print(f"gemv,q4_K,4096,11008,{N},0,{us:.2f},{gbs:.0f}", flush=True)
    for (tkv, nh, nhkv, D, nkv) in ((1, 32, 32, 128, 4096), (1, 32, 32, 128, 512), (8, 32, 8, 128, 32768), (1, 64, 8, 128, 4096), (8, 32, 32, 128, 4096)):
        us, gbs = bench_fattn(tkv, nh, nhkv, D, nkv)
        print(f"fattn,{'f16' if tkv == 1 else 'q8_0'},{nh}/{nhkv},{D},{nkv},0,{us:.2f},{gbs:.0f}", flush=True)
