"""CPU, world_size 2 over gloo: the host logic of the tensor-parallel path (koboldcpp_b200/tp.py).  Each rank cuts its
shards out of the same raw GGUF block bytes; row-split shards must concatenate (all-gather) and K-split shards must sum
(all-reduce) to the unsharded matrix-vector product, computed here on oracle-dequantised weights in fp64."""
from __future__ import annotations

import os
import socket

import numpy as np
import pytest

from koboldcpp_b200 import tp
from koboldcpp_b200.lib import F16, Q4_K, Q6_K, Q8_0, row_size
from koboldcpp_b200.llama import random_blocks


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank: int, world: int, port: int, q) -> None:
    import torch
    import torch.distributed as dist

    import orc as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(5)  # the same stream on every rank: everybody draws the full tensors
        K, M = 1024, 512
        x = rng.standard_normal(K)
        out = {}
        for ty in (Q4_K, Q6_K, Q8_0):
            raw = random_blocks(rng, ty, M, K)
            full = O.dequantize(ty, raw.reshape(-1), K * M).reshape(M, K).astype(np.float64) @ x
            # rows: ffn_gate-like
            sh = tp.shard_tensor("blk.0.ffn_gate.weight", ty, (K, M, 1), raw, rank, world)
            ne = tp.shard_shape("blk.0.ffn_gate.weight", (K, M, 1), world)
            assert sh.shape == (ne[1], row_size(ty, ne[0]))
            part = torch.from_numpy(O.dequantize(ty, sh.reshape(-1), ne[0] * ne[1]).reshape(ne[1], ne[0]).astype(np.float64) @ x)
            parts = [torch.empty_like(part) for _ in range(world)]
            dist.all_gather(parts, part)
            out[("rows", ty)] = float(np.abs(torch.cat(parts).numpy() - full).max())
            # K: ffn_down-like
            sh = tp.shard_tensor("blk.0.ffn_down.weight", ty, (K, M, 1), raw, rank, world)
            ne = tp.shard_shape("blk.0.ffn_down.weight", (K, M, 1), world)
            assert sh.shape == (M, row_size(ty, ne[0]))
            xs = x[rank * ne[0]:(rank + 1) * ne[0]]
            part = torch.from_numpy(O.dequantize(ty, sh.reshape(-1), ne[0] * M).reshape(M, ne[0]).astype(np.float64) @ xs)
            dist.all_reduce(part)
            out[("k", ty)] = float(np.abs(part.numpy() - full).max() / np.abs(full).max())
        if rank == 0:
            q.put(out)
    finally:
        dist.destroy_process_group()


def test_shards_reassemble_over_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for (kind, ty), err in out.items():
        assert err < 1e-9, (kind, ty, err)


def test_split_kinds_and_shapes():
    assert tp.split_kind("blk.3.attn_q.weight") == "rows" and tp.split_kind("output.weight") == "rows"
    assert tp.split_kind("blk.3.attn_output.weight") == "k" and tp.split_kind("blk.0.ffn_down.weight") == "k"
    assert tp.split_kind("token_embd.weight") == "replicate" and tp.split_kind("blk.1.ffn_norm.weight") == "replicate"
    assert tp.shard_shape("blk.0.ffn_down.weight", (28672, 8192, 1), 8) == (3584, 8192, 1)
    assert tp.shard_shape("blk.0.attn_k.weight", (8192, 1024, 1), 8) == (8192, 128, 1)
    with pytest.raises(Exception):
        tp.check_divisible(32, 32, 11008, 32000, 128, 2)  # 11008 / 2 is not a whole number of super-blocks
    tp.check_divisible(64, 8, 28672, 128256, 128, 8)


@pytest.mark.parametrize("type_kv", [F16, Q8_0])
def test_kv_row_shards_are_the_ranks_heads(type_kv):
    n_head_kv, D, cells = 4, 64, 5
    rb = row_size(type_kv, D)
    rows = np.arange(cells * n_head_kv * rb, dtype=np.uint32).astype(np.uint8).reshape(cells, n_head_kv * rb)
    got = [tp.shard_kv_rows(rows, type_kv, n_head_kv, D, r, 2) for r in range(2)]
    assert np.array_equal(np.concatenate(got, axis=1), rows)
    assert got[0].shape == (cells, 2 * rb)
