"""One process per GPU: tile- or sample-sharded frames; the accumulator is gathered over peer memory (tiles) or NCCL (samples).

The reference has no multi-GPU path (single device picked by CUDATools::FastestDevice, lib/CUDA/shared_host_code/
cudatools.h:62-87).  Here the scene is replicated, every rank renders its interleaved 32x8 tiles (or its own samples)
into a torch-owned float4[w*h] accumulator that the core writes directly (b200_set_accumulator_buffer), and ONE
collective per frame -- ``all_reduce(SUM)`` over NCCL/NVLink; disjoint pixels, so the sum is the union, bit-identical to
a single-GPU frame -- makes the full accumulator resident on every rank.  Tile-sharded frames skip the collective: the cores
map each other's accumulators (CUDA IPC over NVLink) and ONE kernel per frame pulls the foreign tiles and finalizes the frame
(gather="p2p", wavefront.cu::gatherFinalizeKernel); torch.distributed only carries the 128-byte handles once.  torch /
torch.distributed are plumbing only (device memory, streams, the handle exchange / the collective); all kernels are the core's own.
"""
from __future__ import annotations

import numpy as np

TILE_W, TILE_H = 32, 8  # must match wavefront.cu


def tile_grid(width: int, height: int):
    return (width + TILE_W - 1) // TILE_W, (height + TILE_H - 1) // TILE_H


def tile_skew(shard_count: int) -> int:
    """Row ty of the tile grid is rotated by ty * skew columns before the tiles are numbered (wavefront.cu::tileIndex): the smallest odd
    number >= 3 that is coprime with the shard count, so that the shards' tiles form diagonal, not vertical, stripes."""
    import math
    s = 3
    while math.gcd(s, max(shard_count, 1)) != 1:
        s += 2
    return s


def tile_owner_map(width: int, height: int, shard_count: int) -> np.ndarray:
    """int32[h, w]: which shard renders each pixel (tile t -> shard t % shard_count), as wavefront.cu assigns them."""
    tx, ty = tile_grid(width, height)
    col, row = np.meshgrid(np.arange(tx, dtype=np.int64), np.arange(ty, dtype=np.int64))
    tiles = (row * tx + (col + row * tile_skew(shard_count)) % tx) % shard_count
    return np.repeat(np.repeat(tiles, TILE_H, axis=0), TILE_W, axis=1)[:height, :width].astype(np.int32)


def shard_pixel_counts(width: int, height: int, shard_count: int) -> np.ndarray:
    return np.bincount(tile_owner_map(width, height, shard_count).ravel(), minlength=shard_count)


def exchange_peer_handles(blob: bytes, world: int, device=None) -> bytes:
    """All-gather of the ranks' 128-byte peer handles (b200_peer_handle) -> the table b200_open_peers takes (entry r = rank r).
    Works over gloo (CPU tensors) and NCCL (pass the rank's CUDA device)."""
    import torch
    import torch.distributed as dist
    assert len(blob) == 128
    mine = torch.frombuffer(bytearray(blob), dtype=torch.uint8)
    if device is not None:
        mine = mine.to(device)
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)
    return b"".join(bytes(p.cpu().numpy().tobytes()) for p in parts)


def all_ranks_ok(ok: bool, device=None) -> bool:
    """True iff `ok` on every rank (so that the ranks take the same gather path)."""
    import torch
    import torch.distributed as dist
    t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item())


class ShardedRenderer:
    """CoreAPI-shaped wrapper for rank ``rank`` of ``world``: same calls as RenderCoreB200, Render() returns after the gather."""

    def __init__(self, rank: int, world: int, local_device: int, mode: str = "tiles", existing_core=None, gather: str = "auto"):
        """gather: "p2p" = fused pull + finalize over peer memory (tile mode only), "nccl" = all-reduce, "auto" = p2p where it
        can be set up on every rank (tile mode, world > 1), else nccl."""
        import torch
        from . import core
        assert mode in ("tiles", "samples") and gather in ("auto", "p2p", "nccl")
        self.torch, self.rank, self.world, self.mode = torch, rank, world, mode
        self.gather_request, self.gather = gather, "nccl"
        self.device = torch.device("cuda", local_device)
        self.core = existing_core if existing_core is not None else core.RenderCoreB200(local_device)
        self.core.SetStream(torch.cuda.current_stream(self.device).cuda_stream)
        self.acc = None
        self.spp = 1
        self.samples_taken = 0

    def __getattr__(self, name):  # SetGeometry, SetInstance, UpdateToplevel, SetPointLight, ... go straight to the core
        return getattr(self.core, name)

    def SetTarget(self, width: int, height: int, spp: int = 1):
        torch = self.torch
        self.width, self.height, self.spp = width, height, spp
        self.gather = "nccl"
        if self.mode == "tiles" and self.world > 1 and self.gather_request in ("auto", "p2p"):
            # the core keeps its own accumulator and the ranks map each other's (CUDA IPC)
            self.core.ClosePeers()
            self.core.SetAccumulatorBuffer(0)
            self.core.SetTarget(width, height, spp)
            err = None
            try:
                blob = self.core.PeerHandle()
            except RuntimeError as e:
                blob, err = bytes(128), e
            table = exchange_peer_handles(blob, self.world, self.device)
            if err is None:
                try:
                    self.core.OpenPeers(table, self.world, self.rank)
                except RuntimeError as e:
                    err = e
            if all_ranks_ok(err is None, self.device):
                self.gather = "p2p"
                self.acc = None
            else:
                self.core.ClosePeers()
                if self.gather_request == "p2p":
                    raise RuntimeError("peer-memory gather could not be set up on every rank: %s" % (err,))
        if self.gather == "nccl":
            self.acc = torch.zeros((height, width, 4), dtype=torch.float32, device=self.device)
            self.core.SetAccumulatorBuffer(self.acc.data_ptr())
            self.core.SetTarget(width, height, spp)
        if self.mode == "tiles":
            self.core.Setting("shardCount", self.world)
            self.core.Setting("shardIndex", self.rank)
        self.samples_taken = 0

    def Render(self, view: dict, converge: int = 1):
        """Render this rank's share, then gather (peer-memory pull + finalize, or all-reduce).  Returns (render_ms, gather_ms) device times."""
        import torch.distributed as dist
        torch = self.torch
        if converge == 1:
            self.samples_taken = 0
        if self.mode == "samples":
            # this frame's samples are [taken, taken + world*spp); rank r renders [taken + r*spp, taken + (r+1)*spp)
            self.core.Setting("sampleOffset", self.samples_taken + self.rank * self.spp)
        elif self.samples_taken and self.gather != "p2p":  # (p2p: the core itself keeps converging in its own accumulator)
            self.core.Setting("sampleOffset", self.samples_taken)
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        stream = torch.cuda.current_stream(self.device)
        if self.gather == "p2p":
            # every rank owns its tiles: Converge keeps adding to them in place, the pull overwrites the foreign ones.  With open peers the
            # core's Render() ends with the gather (arrival, pull + finalize, arrival: the tail of the frame's enqueue)
            self.core.Render(view, converge)
            self.samples_taken += self.spp
            st = self.core.GetCoreStats()
            return st["renderTime"] - st["gatherMs"], st["gatherMs"]
        e0.record(stream)
        # each rank's accumulator holds only its own contribution of THIS frame; earlier frames are already global sums
        prev = self.acc.clone() if (converge == 0 and self.samples_taken) else None
        self.core.Render(view, 1)
        e1.record(stream)
        if self.world > 1:
            dist.all_reduce(self.acc, op=dist.ReduceOp.SUM)
        if prev is not None:
            self.acc += prev
        e2.record(stream)
        self.samples_taken += self.spp * (self.world if self.mode == "samples" else 1)
        self.core.Finalize(self.samples_taken)
        torch.cuda.synchronize(self.device)
        return e0.elapsed_time(e1), e1.elapsed_time(e2)

    def ReadAccumulator(self) -> np.ndarray:
        return self.core.ReadAccumulator() if self.acc is None else self.acc.cpu().numpy()

    def ReadFrame(self) -> np.ndarray:
        return self.core.ReadFrame()

    def Shutdown(self):
        if self.gather == "p2p":
            # nobody may unmap / free an accumulator a peer could still be reading
            import torch.distributed as dist
            self.torch.cuda.synchronize(self.device)
            dist.barrier()
            self.core.ClosePeers()
            dist.barrier()
        self.core.Shutdown()
