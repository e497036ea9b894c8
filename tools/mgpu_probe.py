#!/usr/bin/env python
"""Multi-GPU frame-time probe (torchrun): wall and device time per frame of the tile-sharded C5-size frame for the frame-loop settings.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29541 tools/mgpu_probe.py [--width ..]"""
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
    w, h, spp, frames = arg("--width", 3840), arg("--height", 2160), arg("--spp", 16), arg("--frames", 6)
    sr = multigpu.ShardedRenderer(rank, world, local, mode="tiles")
    sr.SetGeometry(0, scenes.sphere_field(1_000_000)); sr.SetInstance(0, 0); sr.SetInstance(1, -1); sr.UpdateToplevel()
    sr.SetTarget(w, h, spp)
    sr.SetPointLight(scenes.C2_LIGHT, (4000.0, 4000.0, 3600.0)); sr.SetSky((0.4, 0.5, 0.7)); sr.Setting("maxPathLength", 5); sr.Setting("russianRoulette", 0)
    view = scenes.view_pyramid(width=w, height=h, **scenes.C2_CAMERA)

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    for cfg in (dict(frameGraph=0, waveTimers=1, chains=0), dict(frameGraph=1, waveTimers=1, chains=0),
                dict(frameGraph=1, waveTimers=0, chains=0), dict(frameGraph=0, waveTimers=0, chains=1), dict(frameGraph=0, waveTimers=0, chains=2)):
        for k, v in cfg.items():
            sr.Setting(k, v)
        for _ in range(3):
            sr.Render(view, 1)
        sync()
        per = []
        t0 = time.perf_counter()
        dev_ms = g_ms = 0.0
        for _ in range(frames):
            t1 = time.perf_counter()
            a, b = sr.Render(view, 1)
            per.append((time.perf_counter() - t1) * 1e3)
            dev_ms += a; g_ms += b
        sync()
        wall = (time.perf_counter() - t0) / frames * 1e3
        st = sr.GetCoreStats()
        v = torch.tensor([wall, dev_ms / frames, g_ms / frames], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(v, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(json.dumps(dict(cfg, n_gpus=world, size=f"{w}x{h}x{spp}", ms_wall=round(v[0].item(), 3), render_ms=round(v[1].item(), 3), gather_ms=round(v[2].item(), 3),
                                  rank0_call_ms=[round(x, 2) for x in per], chains_used=st["frameChains"], passes=st["framePasses"], graph_replays=st["frameGraphLaunches"],
                                  gather_kernel_ms=round(st["gatherKernelMs"], 4), host_enqueue_wait_total_ms=[round(st[k], 3) for k in ("hostEnqueueMs", "hostWaitMs", "hostTotalMs")],
                                  device_render_ms_last=round(st["renderTime"], 3))), flush=True)
    sr.Shutdown()
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
