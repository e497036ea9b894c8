#!/usr/bin/env python
"""BASELINE.json configs[4]: 3840x2160, 16 spp full wavefront frame on N GPUs (torchrun, one process per GPU), C2 scene replicated,
interleaved-tile sharding, accumulator gathered over NVLink by the fused peer-memory pull + finalize kernel (--gather nccl: one NCCL all-reduce).  Rank 0 prints one JSON line.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29577 tools/bench_c5.py [--spp 16] [--frames 3]"""
import json
import os
import sys
import time

os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lighthouse2_b200 import multigpu, scenes  # noqa: E402


def arg(name, default):
    return int(sys.argv[sys.argv.index(name) + 1]) if name in sys.argv else default


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    w, h, spp, frames = arg("--width", 3840), arg("--height", 2160), arg("--spp", 16), arg("--frames", 3)
    gather = sys.argv[sys.argv.index("--gather") + 1] if "--gather" in sys.argv else "auto"
    sr = multigpu.ShardedRenderer(rank, world, local, mode="tiles", gather=gather)
    sr.SetGeometry(0, scenes.sphere_field(1_000_000)); sr.SetInstance(0, 0); sr.SetInstance(1, -1); sr.UpdateToplevel()
    sr.SetTarget(w, h, spp)
    sr.SetPointLight(scenes.C2_LIGHT, (4000.0, 4000.0, 3600.0)); sr.SetSky((0.4, 0.5, 0.7)); sr.Setting("maxPathLength", 5); sr.Setting("russianRoulette", 0)
    view = scenes.view_pyramid(width=w, height=h, **scenes.C2_CAMERA)
    sr.Render(view, 1)
    sync = lambda: (torch.cuda.synchronize(), dist.barrier() if world > 1 else None, torch.cuda.synchronize())
    sync()
    t0 = time.perf_counter()
    r_ms = g_ms = 0.0
    rays = 0
    for _ in range(frames):
        a, b = sr.Render(view, 1)
        r_ms += a; g_ms += b
        rays += sr.GetCoreStats()["totalRays"]
    sync()
    wall = (time.perf_counter() - t0) / frames * 1e3
    v = torch.tensor([wall, r_ms / frames, g_ms / frames], device=dev, dtype=torch.float64)
    n = torch.tensor([rays / frames], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(v, op=dist.ReduceOp.MAX); dist.all_reduce(n, op=dist.ReduceOp.SUM)
    if rank == 0:
        v = v.tolist()
        gbytes = w * h * 16
        print(json.dumps({"config": f"C5 {w}x{h} x {spp} spp, C2 scene replicated, {world} GPUs tile-sharded", "n_gpus": world, "frames": frames,
                          "ms_per_frame_wall": v[0], "render_ms_max_rank": v[1], "gather_ms": v[2], "rays_per_frame": n.item(),
                          "mrays_s_aggregate": n.item() / (v[0] * 1e-3) / 1e6, "gather_bytes": gbytes,
                          "gather": sr.gather if world > 1 else "none",
                          # bytes each GPU moves over NVLink: the pull reads (G-1)/G of the image once, the all-reduce moves twice that
                          "gather_busbw_gbs": ((1 if sr.gather == "p2p" else 2) * (world - 1) / world * gbytes / (v[2] * 1e-3) / 1e9) if world > 1 and v[2] > 0 else None,
                          "frame_mean": float(torch.from_numpy(sr.ReadFrame())[..., :3].mean())}), flush=True)
    sr.Shutdown()
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
