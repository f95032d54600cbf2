"""Developer micro-benchmark (torchrun, one rank per GPU): latency of the one-kernel peer-memory all-reduce (+ residual) against
the NCCL all-reduce + ADD it replaces, for decode-size messages.  Not the round bench."""
import ctypes as C
import os
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def main():
    import torch
    import torch.distributed as dist

    import koboldcpp_b200.host as H
    from koboldcpp_b200 import tp
    from koboldcpp_b200.lib import check, lib
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = lib()
    check(L.b200_set_device(local))
    tp.comm_init_from_torch(rank, world)
    T = tp._l()
    st = C.c_void_p()
    check(L.b200_stream_create(C.byref(st)))
    for n in (4096, 8192, 16384):
        src = H.DeviceBuffer.from_numpy(np.ones(n, np.float32))
        res = H.DeviceBuffer.from_numpy(np.ones(n, np.float32))
        dst = H.DeviceBuffer(4 * n)
        for name, fn in (("peer", lambda: T.b200_all_reduce_sum_add_f32(src.ptr, res.ptr, dst.ptr, n, st)),
                         ("nccl", lambda: T.b200_all_reduce_sum_f32(dst.ptr, n, st))):
            if name == "peer" and not T.b200_comm_peer_ready():
                continue
            iters = 200
            check(L.b200_graph_begin_capture(st))
            for _ in range(iters):
                check(fn())
            g = C.c_void_p()
            check(L.b200_graph_end_capture(st, C.byref(g)))
            for _ in range(3):
                check(L.b200_graph_launch(g, st))
            check(L.b200_stream_synchronize(st))
            dist.barrier()
            e0, e1 = C.c_void_p(), C.c_void_p()
            check(L.b200_event_create(C.byref(e0))); check(L.b200_event_create(C.byref(e1)))
            check(L.b200_event_record(e0, st))
            for _ in range(5):
                check(L.b200_graph_launch(g, st))
            check(L.b200_event_record(e1, st))
            check(L.b200_stream_synchronize(st))
            ms = C.c_float()
            check(L.b200_event_elapsed_ms(e0, e1, C.byref(ms)))
            if rank == 0:
                print(f"world {world} n {n} {name}: {ms.value / (5 * iters) * 1e3:.2f} us per all-reduce", flush=True)
            L.b200_graph_destroy(g)
    tp.comm_destroy()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
