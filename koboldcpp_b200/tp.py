"""Tensor-parallel sharding of a llama model's tensors, one process per GPU (include/b200_comm.h, b200_llama.h tp_world > 1).

The reference shards every weight matrix by output rows over the devices of one process and moves activations / results
through the main GPU for every matmul (/root/reference/ggml/src/ggml-cuda.cu:623-953 split buffer type, :1519-1657 matmul
loop).  Here the rows are cut so that whole heads and whole ffn columns stay on one rank, and the two matrices that close a
residual branch (attn_output, ffn_down) are cut along K instead, which leaves two all-reduces per layer as the only
exchange.  Everything in this file is host logic on numpy arrays of raw GGUF block bytes: it runs (and is tested) without a GPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .lib import BLOCK, B200Error, check, lib

ROW_SPLIT = ("attn_q.weight", "attn_k.weight", "attn_v.weight", "ffn_gate.weight", "ffn_up.weight", "output.weight")
K_SPLIT = ("attn_output.weight", "ffn_down.weight")


def leaf_of(name: str) -> str:
    return name.split(".", 2)[2] if name.startswith("blk.") else name


def split_kind(name: str) -> str:
    """'rows' | 'k' | 'replicate' for a GGUF tensor name."""
    leaf = leaf_of(name)
    if leaf in ROW_SPLIT:
        return "rows"
    if leaf in K_SPLIT:
        return "k"
    return "replicate"


def check_divisible(n_head: int, n_head_kv: int, n_ff: int, n_vocab: int, head_dim: int, world: int) -> None:
    if n_head % world or n_head_kv % world or n_ff % world or n_vocab % world:
        raise B200Error(f"heads {n_head}/{n_head_kv}, ffn {n_ff} or vocabulary {n_vocab} do not divide over {world} ranks")
    if (n_ff // world) % 256 or (n_head // world * head_dim) % 256:
        raise B200Error("a K-split shard must be a whole number of 256-value super-blocks")


def shard_shape(name: str, ne: tuple[int, int, int], world: int) -> tuple[int, int, int]:
    kind = split_kind(name)
    if kind == "rows":
        return (ne[0], ne[1] // world, ne[2])
    if kind == "k":
        return (ne[0] // world, ne[1], ne[2])
    return ne


def shard_tensor(name: str, ty: int, ne: tuple[int, int, int], data: np.ndarray, rank: int, world: int) -> np.ndarray:
    """The bytes of rank `rank`'s shard of a tensor given as raw rows (uint8 [ne1, row_bytes])."""
    kind = split_kind(name)
    bs, by = BLOCK[ty]
    rows = np.asarray(data, dtype=np.uint8).reshape(ne[1] * ne[2], ne[0] // bs * by)
    if world == 1 or kind == "replicate":
        return rows
    if kind == "rows":
        m = ne[1] // world
        return np.ascontiguousarray(rows[rank * m:(rank + 1) * m])
    nb = ne[0] // bs
    if nb % world:
        raise B200Error(f"{name}: {nb} blocks per row do not divide over {world} ranks")
    k = nb // world
    return np.ascontiguousarray(rows.reshape(rows.shape[0], nb, by)[:, rank * k:(rank + 1) * k, :]).reshape(rows.shape[0], k * by)


def shard_kv_rows(rows: np.ndarray, type_kv: int, n_head_kv: int, head_dim: int, rank: int, world: int) -> np.ndarray:
    """KV cache rows [n_cells, row_bytes(n_head_kv * head_dim)] -> the columns of this rank's KV heads."""
    if world == 1:
        return rows
    bs, by = BLOCK[type_kv]
    per_head = head_dim // bs * by
    lkv = n_head_kv // world
    r = rows.reshape(rows.shape[0], n_head_kv, per_head)
    return np.ascontiguousarray(r[:, rank * lkv:(rank + 1) * lkv, :]).reshape(rows.shape[0], lkv * per_head)


# ------------------------------------------------------------------------------------------------ communicator
_declared = False


def _l():
    global _declared
    L = lib()
    if not _declared:
        for name, res, args in (("b200_comm_unique_id", C.c_int, [C.c_void_p]), ("b200_comm_init", C.c_int, [C.c_void_p, C.c_int, C.c_int]),
                                ("b200_comm_destroy", C.c_int, []), ("b200_comm_rank", C.c_int, []), ("b200_comm_world", C.c_int, []),
                                ("b200_all_reduce_sum_f32", C.c_int, [C.c_void_p, C.c_int64, C.c_void_p]),
                                ("b200_all_gather_f32", C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
                                ("b200_comm_peer_window_create", C.c_int, [C.c_void_p]), ("b200_comm_peer_window_attach", C.c_int, [C.c_void_p]),
                                ("b200_comm_peer_window_destroy", C.c_int, []), ("b200_comm_peer_ready", C.c_int, []), ("b200_comm_peer_status", C.c_int, []),
                                ("b200_all_reduce_sum_add_f32", C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p])):
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _declared = True
    return L


def comm_init_from_torch(rank: int, world: int, peer_window: bool = True) -> None:
    """Create the library's NCCL communicator; the 128-byte id travels through the caller's torch.distributed group."""
    import torch
    import torch.distributed as dist
    L = _l()
    buf = (C.c_ubyte * 128)()
    if rank == 0:
        check(L.b200_comm_unique_id(buf), "b200_comm_unique_id")
    t = torch.tensor(list(bytes(buf)), dtype=torch.uint8)
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.broadcast(t, 0)
    raw = bytes(t.cpu().tolist())
    check(L.b200_comm_init(raw, rank, world), "b200_comm_init")
    if peer_window and 2 <= world <= 8:
        peer_window_init(rank, world)


IPC_BYTES = 64


def peer_window_init(rank: int, world: int) -> bool:
    """Map every rank's exchange window into every process (CUDA IPC handles travel through torch.distributed).  Either all
    ranks end up attached or none does (then the backend keeps the NCCL all-reduce); returns which."""
    import torch
    import torch.distributed as dist
    L = _l()
    on_gpu = dist.get_backend() == "nccl"
    h = (C.c_ubyte * IPC_BYTES)()
    ok = L.b200_comm_peer_window_create(h) == 0
    mine = torch.tensor(list(bytes(h)) + [1 if ok else 0], dtype=torch.uint8)
    if on_gpu:
        mine = mine.cuda()
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)
    rows = [bytes(p.cpu().tolist()) for p in parts]
    ok = all(r[IPC_BYTES] == 1 for r in rows)
    if ok:
        ok = L.b200_comm_peer_window_attach(b"".join(r[:IPC_BYTES] for r in rows)) == 0
    flag = torch.tensor([1 if ok else 0], dtype=torch.int32)
    if on_gpu:
        flag = flag.cuda()
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if int(flag.item()) != 1:
        L.b200_comm_peer_window_destroy()
        return False
    return True


def comm_destroy() -> None:
    _l().b200_comm_destroy()
