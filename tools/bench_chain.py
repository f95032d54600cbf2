"""Developer micro-benchmark: achieved HBM GB/s of the decode chain kernel on the GEMV phases of 7B-shaped layers
(QKV with norm prologue -> rope+KV store -> wo + residual -> gate/up with SiLU epilogue -> down + residual), two layers per
launch, weight sets rotating through > L2.  Not the round bench (bench.py)."""
import ctypes as C
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import koboldcpp_b200.host as H  # noqa: E402
from koboldcpp_b200.lib import ChainPhase, Q4_K, Q6_K, check, lib, row_size  # noqa: E402

L = lib()


def ev():
    e = C.c_void_p()
    check(L.b200_event_create(C.byref(e)))
    return e


def main():
    H.require_gpu()
    E, NH, D, FF = 4096, 32, 128, 11008
    rng = np.random.default_rng(0)
    n_sets = 3

    def wbuf(t, K, M):
        host = rng.integers(0, 256, size=row_size(t, K) * M, dtype=np.uint8)
        return H.DeviceBuffer.from_numpy(host, pad=16)

    sets = []
    for _ in range(n_sets):
        layers = []
        for li in range(2):
            down_t = Q6_K if li == 0 else Q4_K
            layers.append(dict(wq=wbuf(Q4_K, E, E), wk=wbuf(Q4_K, E, E), wv=wbuf(down_t, E, E), wo=wbuf(Q4_K, E, E), wg=wbuf(Q4_K, E, FF), wu=wbuf(Q4_K, E, FF),
                               wd=wbuf(down_t, FF, E), down_t=down_t))
        sets.append(layers)
    f = lambda n: H.DeviceBuffer(4 * n + 256, zero=True)
    x, nw = H.DeviceBuffer.from_numpy(rng.standard_normal(E).astype(np.float32)), H.DeviceBuffer.from_numpy(np.ones(E, np.float32))
    q, k, v, qo, ko, h, act, out = f(E), f(E), f(E), f(E), f(E), f(E), f(FF), f(E)
    kc, vc = H.DeviceBuffer(2 * E + 256, zero=True), H.DeviceBuffer(2 * E + 256, zero=True)
    pos = H.DeviceBuffer.from_numpy(np.array([77], np.int32))
    bar = H.DeviceBuffer(256, zero=True)

    ss1, ss2 = H.DeviceBuffer(4 * 256, zero=True), H.DeviceBuffer(4 * 256, zero=True)
    use_ss = "--no-ss" not in sys.argv

    def gemv_phase(X, K, jobs, norm=None, pair=0, ss_out=None, ss_in=None):
        ph = ChainPhase()
        ph.kind, ph.n_jobs, ph.K, ph.silu_pair, ph.norm_eps = 0, len(jobs), K, pair, 1e-5
        ph.X, ph.norm_weight = X, norm
        if use_ss:
            ph.sumsq_out, ph.sumsq_in = ss_out, ss_in
        for i, (W, dst, res, M, t) in enumerate(jobs):
            j = ph.u.jobs[i]
            j.W, j.dst, j.residual, j.w_row_bytes, j.M, j.type = W, dst, res, row_size(t, K), M, t
        return ph

    progs, alg = [], 0
    for layers in sets:
        ph = (ChainPhase * 10)()
        i = 0
        alg = 0
        for ly in layers:
            dt = ly["down_t"]
            ph[i] = gemv_phase(x.ptr, E, [(ly["wq"].ptr, q.ptr, None, E, Q4_K), (ly["wk"].ptr, k.ptr, None, E, Q4_K), (ly["wv"].ptr, v.ptr, None, E, dt)], nw.ptr, ss_in=ss2.ptr if i else None)
            r = ph[i + 1]
            r.kind = 1
            rr = r.u.rope
            rr.q, rr.k, rr.v, rr.q_out, rr.k_out = q.ptr, k.ptr, v.ptr, qo.ptr, ko.ptr
            rr.k_cache, rr.v_cache, rr.pos = kc.ptr, vc.ptr, pos.ptr
            rr.D, rr.H, rr.HK, rr.kv_type, rr.n_dims, rr.mode, rr.n_ctx_orig = D, NH, NH, 1, D, 0, 4096
            rr.freq_base, rr.freq_scale, rr.ext_factor, rr.attn_factor, rr.beta_fast, rr.beta_slow = 10000.0, 1.0, 0.0, 1.0, 32.0, 1.0
            ph[i + 2] = gemv_phase(qo.ptr, E, [(ly["wo"].ptr, h.ptr, x.ptr, E, Q4_K)], ss_out=ss1.ptr)
            ph[i + 3] = gemv_phase(h.ptr, E, [(ly["wg"].ptr, act.ptr, None, FF, Q4_K), (ly["wu"].ptr, act.ptr, None, FF, Q4_K)], nw.ptr, pair=1, ss_in=ss1.ptr)
            ph[i + 4] = gemv_phase(act.ptr, FF, [(ly["wd"].ptr, out.ptr, h.ptr, E, dt)], ss_out=ss2.ptr)
            i += 5
            alg += sum(ly[n].nbytes for n in ("wq", "wk", "wv", "wo", "wg", "wu", "wd"))
        progs.append(ph)

    def run(i):
        check(L.b200_decode_chain(progs[i % n_sets], 10, bar.ptr, None))
    for i in range(6):
        run(i)
    H.sync()
    iters = 60
    e0, e1 = ev(), ev()
    L.b200_event_record(e0, None)
    for i in range(iters):
        run(i)
    L.b200_event_record(e1, None)
    ms = C.c_float()
    check(L.b200_event_elapsed_ms(e0, e1, C.byref(ms)))
    us = ms.value / iters * 1e3
    if "--trace" in sys.argv:
        NP = 12
        tb = H.DeviceBuffer(148 * NP * 12 * 8, zero=True)
        L.b200_decode_chain_trace.argtypes = [C.c_void_p]
        L.b200_decode_chain_trace(tb.ptr)
        run(0)
        H.sync()
        L.b200_decode_chain_trace(None)
        raw = tb.to_numpy(np.uint64, (148 * NP * 12,)).astype(np.int64)
        t = raw[:148 * NP * 4].reshape(148, NP, 4)
        t2 = raw[148 * NP * 4:].reshape(148, NP, 8)
        t0 = t[:, 0, 0].min()
        np.save('gpurun_out/chain_trace.npy', t - t0)
        names = ["qkv", "rope", "wo", "gate/up", "down"] * 2
        print("phase: [min..max over CTAs, ns since kernel start] arrive | barrier passed | activations ready | warp0 done")
        for p_ in range(10):
            r = t[:, p_, :] - t0
            print(f"{p_:2d} {names[p_]:8s} " + " | ".join(f"{r[:, i].min():7d}..{r[:, i].max():7d}" for i in range(4)))
            if names[p_] != "rope":
                u = t2[:, p_, :6] - t[:, p_, 2:3]
                print("     first unit of warp 0 (ns after act ready, median over CTAs): enter %d | data %d | dot %d | issued %d | stored %d | cursor %d" % tuple(int(np.median(u[:, i])) for i in range(6)))
    print(f"chain 2 layers (10 phases): {us:.1f} us per launch, {alg / 1e6:.1f} MB -> {alg / us / 1e3:.0f} GB/s; ideal at 6540 GB/s = {alg / 6540e3:.1f} us")


if __name__ == "__main__":
    main()
