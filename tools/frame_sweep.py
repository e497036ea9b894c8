#!/usr/bin/env python
"""Frame time of Render() on the C2 scene for the frame-loop settings (chains, frameGraph, waveTimers, overlapConnect).
One JSON line per combination: wall ms per frame (host clock around Render(), which synchronises) and CoreStats.renderTime.
  python tools/frame_sweep.py [--width 1920 --height 1080 --spp 1] [--frames 20]"""
import itertools
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lighthouse2_b200 import core, scenes  # noqa: E402


def arg(name, default):
    return int(sys.argv[sys.argv.index(name) + 1]) if name in sys.argv else default


def main():
    w, h, spp, frames, shards = arg("--width", 1920), arg("--height", 1080), arg("--spp", 1), arg("--frames", 20), arg("--shards", 1)
    rc = core.RenderCoreB200(0)
    rc.SetGeometry(0, scenes.sphere_field(1_000_000)); rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
    rc.SetTarget(w, h, spp)
    rc.Setting("shardCount", shards)  # one rank's share of a tile-sharded frame (rank 0 of `shards`)
    rc.SetPointLight(scenes.C2_LIGHT, (4000.0, 4000.0, 3600.0)); rc.SetSky((0.4, 0.5, 0.7)); rc.Setting("maxPathLength", 5); rc.Setting("russianRoulette", 0)
    view = scenes.view_pyramid(width=w, height=h, **scenes.C2_CAMERA)
    chain_set = tuple(int(x) for x in sys.argv[sys.argv.index("--chains") + 1].split(",")) if "--chains" in sys.argv else (1, 2, 3, 4)
    combos = [dict(chains=c, frameGraph=g, waveTimers=t, overlapConnect=o)
              for c, g, t, o in itertools.product(chain_set, (0, 1), (0, 1), (1,))] + [dict(chains=1, frameGraph=0, waveTimers=1, overlapConnect=0)]
    ref = None
    for cfg in combos:
        for k, v in cfg.items():
            rc.Setting(k, v)
        for _ in range(3):
            rc.Render(view, 1)
        t0 = time.perf_counter()
        dev = 0.0
        for _ in range(frames):
            rc.Render(view, 1)
            dev += rc.GetCoreStats()["renderTime"]
        wall = (time.perf_counter() - t0) / frames * 1e3
        st = rc.GetCoreStats()
        acc = rc.ReadAccumulator()
        if ref is None:
            ref = acc
        print(json.dumps(dict(cfg, size=f"{w}x{h}x{spp}/{shards}", ms_wall=round(wall, 4), ms_device=round(dev / frames, 4), rays=st["totalRays"],
                              mrays_s=round(st["totalRays"] / wall / 1e3, 1), passes=st["framePasses"], graph_launches=st["frameGraphLaunches"],
                              pixels_differing_from_first=int((acc.view("uint32") != ref.view("uint32")).any(axis=2).sum()))), flush=True)
    rc.Shutdown()


if __name__ == "__main__":
    main()
