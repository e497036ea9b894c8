"""CPU tests of the multi-GPU host logic: the tile partition wavefront.cu and multigpu.py agree on, and the one
collective of the data path (all_reduce SUM of disjoint accumulators) over gloo with world_size 2."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lighthouse2_b200 import multigpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("size", [(1920, 1080), (200, 100), (33, 9), (3840, 2160)])
@pytest.mark.parametrize("shards", [1, 2, 3, 4, 8])
def test_tile_partition(size, shards):
    w, h = size
    owner = multigpu.tile_owner_map(w, h, shards)
    assert owner.shape == (h, w) and owner.min() == 0 and owner.max() == min(shards, np.prod(multigpu.tile_grid(w, h))) - 1
    counts = multigpu.shard_pixel_counts(w, h, shards)
    assert counts.sum() == w * h
    if w * h >= 1920 * 1080:
        assert counts.max() / counts.min() < 1.02, "interleaved tiles balance the shards"
    # restate the kernel's mapping (generateEyeRaysKernel + tileCoords of wavefront.cu): local tile lt of shard i is global tile
    # lt * shards + i, and tile t sits in row t // tx at column (t % tx - row * skew) mod tx
    tx, ty = multigpu.tile_grid(w, h)
    skew = multigpu.tile_skew(shards)
    seen = np.zeros((ty, tx), np.int32)
    for i in range(shards):
        own_tiles = (tx * ty - i + shards - 1) // shards if tx * ty > i else 0
        lt = np.arange(own_tiles)
        t = lt * shards + i
        assert (t < tx * ty).all()
        row = t // tx
        col = (t % tx - row * skew) % tx
        seen[row, col] += 1
        x0, y0 = col * multigpu.TILE_W, row * multigpu.TILE_H
        assert (owner[np.minimum(y0, h - 1), np.minimum(x0, w - 1)] == i).all()
    assert (seen == 1).all(), "every tile belongs to exactly one shard"
    if shards > 1 and ty >= 2 * shards and tx >= shards:
        # no shard owns a whole column of tiles (what the row-by-row numbering did whenever shards divides the column count)
        tile_owner = owner[::multigpu.TILE_H, ::multigpu.TILE_W]
        assert all(len(np.unique(tile_owner[:, c])) > 1 for c in range(tile_owner.shape[1]))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, w, h, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(5)
    full = rng.random((h, w, 4), dtype=np.float32)  # what a single core would have accumulated (same on every rank)
    owner = multigpu.tile_owner_map(w, h, world)
    mine = torch.from_numpy(np.where((owner == rank)[..., None], full, np.float32(0)))
    dist.all_reduce(mine, op=dist.ReduceOp.SUM)
    ok = np.array_equal(mine.numpy().view(np.uint32), full.view(np.uint32))
    # the bench's reduction: slowest rank's time, all ranks' rays
    t = torch.tensor([1.0 + rank], dtype=torch.float64)
    n = torch.tensor([100.0 * (rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(n, op=dist.ReduceOp.SUM)
    # the peer-handle table of the peer-memory gather: entry r = rank r's 128-byte blob, on every rank; one failing rank
    # sends every rank down the NCCL path
    blob = bytes([rank * 16 + (i % 16) for i in range(128)])
    table = multigpu.exchange_peer_handles(blob, world)
    ok = ok and len(table) == 128 * world and all(table[r * 128:(r + 1) * 128] == bytes([r * 16 + (i % 16) for i in range(128)]) for r in range(world))
    ok = ok and multigpu.all_ranks_ok(True) and not multigpu.all_ranks_ok(rank != 1)
    out[rank] = (ok, float(t), float(n))
    dist.barrier()
    dist.destroy_process_group()


def test_gather_is_bit_exact_over_gloo_world2():
    world, port = 2, _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, 200, 100, out), nprocs=world, join=True)
    for r in range(world):
        ok, t, n = out[r]
        assert ok, "sum of disjoint tile accumulators must equal the single-core accumulator bit for bit"
        assert t == 2.0 and n == 300.0
