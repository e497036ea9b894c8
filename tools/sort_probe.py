#!/usr/bin/env python
"""What would re-sorting the rays between bounces buy?  The bounce waves of the C2 pass are traced as the shade stage would emit them and
again sorted by a 33-bit key (direction octant | 30-bit Morton code of the origin) -- an upper bound on what a device-side counting sort
could achieve, for free (the sort itself is not timed).  Prints Mrays/s of both orders per wave; run under
  ncu --metrics smsp__thread_inst_executed_per_inst_executed.ratio,gpu__time_duration.sum -k regex:traceKernel --csv
the launches come in the order (wave 1 as is, wave 1 sorted, wave 2 as is, ...) once, after the timing loop (--ncu: only that part)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lighthouse2_b200 import core, scenes, workload  # noqa: E402


def spread10(x):
    x = x.astype(np.uint32) & 0x3FF
    x = (x | (x << 16)) & 0x030000FF
    x = (x | (x << 8)) & 0x0300F00F
    x = (x | (x << 4)) & 0x030C30C3
    x = (x | (x << 2)) & 0x09249249
    return x


def sort_key(O4, D4):
    O = O4[:, :3]
    lo, hi = O.min(axis=0), O.max(axis=0)
    q = np.clip(((O - lo) / np.maximum(hi - lo, 1e-20) * 1023.0), 0, 1023).astype(np.uint32)
    morton = (spread10(q[:, 0]) << 2) | (spread10(q[:, 1]) << 1) | spread10(q[:, 2])
    octant = ((D4[:, 0] < 0).astype(np.uint64) << 2) | ((D4[:, 1] < 0).astype(np.uint64) << 1) | (D4[:, 2] < 0).astype(np.uint64)
    return (octant << np.uint64(30)) | morton.astype(np.uint64)


def main():
    import torch
    only_ncu = "--ncu" in sys.argv
    rc = core.RenderCoreB200(0)
    verts = scenes.sphere_field(1_000_000)
    rc.SetGeometry(0, verts); rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
    view = scenes.view_pyramid(width=1920, height=1080, **scenes.C2_CAMERA)
    O, D = scenes.primary_rays(view, 1920, 1080)

    def tr(o, d):
        h = core.unpack_hits(rc.TraceClosest(*core.pack_rays(o, d, tmin=workload.EPS)))
        return h["t"], h["prim"]
    waves = [w for w in workload.make_pass(tr, verts, O, D, scenes.C2_LIGHT, 4, seed=1) if w["cls"] == "diffuse"]
    dev = torch.device("cuda", 0)
    rc.SetStream(torch.cuda.current_stream().cuda_stream)
    sets = []
    for w in waves:
        order = np.argsort(sort_key(w["O4"], w["D4"]), kind="stable")
        for name, idx in (("as emitted", None), ("sorted", order)):
            O4 = w["O4"] if idx is None else np.ascontiguousarray(w["O4"][idx])
            D4 = w["D4"] if idx is None else np.ascontiguousarray(w["D4"][idx])
            sets.append(dict(bounce=w["bounce"], order=name, n=len(O4), dO=torch.from_numpy(O4).to(dev), dD=torch.from_numpy(D4).to(dev),
                             dH=torch.empty((len(O4), 4), dtype=torch.float32, device=dev), idx=idx))
    torch.cuda.synchronize()
    launch = lambda s: rc.TraceClosestDevice(s["dO"].data_ptr(), s["dD"].data_ptr(), s["n"], s["dH"].data_ptr())
    if not only_ncu:
        for s in sets:
            for _ in range(3):
                launch(s)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ms = []
            for _ in range(20):
                a.record(); launch(s); b.record(); b.synchronize()
                ms.append(a.elapsed_time(b))
            s["ms"] = float(np.median(ms))
        for k in range(0, len(sets), 2):
            u, v = sets[k], sets[k + 1]
            hu, hv = u["dH"].cpu().numpy(), v["dH"].cpu().numpy()
            same = np.array_equal(hu[v["idx"]].view(np.uint32), hv.view(np.uint32))
            print(json.dumps(dict(bounce=u["bounce"], rays=u["n"], mrays_s_as_emitted=round(u["n"] / u["ms"] / 1e3, 1), mrays_s_sorted=round(v["n"] / v["ms"] / 1e3, 1),
                                  gain=round(u["ms"] / v["ms"], 4), same_hits=bool(same))), flush=True)
    for s in sets:  # one launch each, in order, for ncu
        launch(s)
    torch.cuda.synchronize()
    rc.Shutdown()


if __name__ == "__main__":
    main()
