#!/usr/bin/env python
"""bench.py -- Mrays/s of the wavefront extend/shadow path (BASELINE.json metric) on B200.

One STEP = one full wavefront frame through the core's public call (Render: ray generation, then per path length the extend wave,
the shade stage and the fused shadow/connect wave; accumulate + finalize) on the C2 scene (BASELINE.json configs[1]: tessellated-
sphere field, 1 000 000 triangles) at the size of BASELINE.json configs[4]: 3840x2160 x 16 spp, primary + next-event shadow rays
+ 4 diffuse bounces.  That is the largest configuration that fits one GPU and is also the 8-GPU configuration, so the SAME frame is
timed at every N (strong scaling): with N ranks the frame is sharded by interleaved 32x8 tiles (scene replicated), and every frame ends
with the gather of the accumulator over NVLink (one fused peer-memory pull + finalize kernel, wavefront.cu).
  value  = rays of the frame (all ranks) / time of the slowest rank (CUDA events on the core's stream, barrier + synchronize on both sides)
  e2e    = the same frames through the host-facing calls b200_render (which ends with the gather) + b200_read_frame_async: the view pyramid goes in
           from host memory, the finished RGBA32F frame comes back into pinned host memory, both inside the timed region
  check  = every rank renders the frame once unsharded and compares its gathered accumulator with it bit for bit
Sub-blocks (explanation, not the headline):
  waves  = the raw extend / shadow waves of a 1080p pass over resident rays, one launch timed at a time (per-class rates: primary, shadow,
           incoherent diffuse -- the >= 1.5 Grays/s target class); `roofline` is computed from its extend launches (the dominant kernel,
           timed alone on its launch stream); with N ranks every rank traces its own sample (weak scaling, no collective)
  e2e_waves = the same waves through the host-buffer calls (pinned rays in, hits out), PCIe-bound
  cpu_baseline = the reference's own CPU traversal on the host cores over a bounded sample of the wave pass (N = 1 only)

`--impl reference` times the reference's own CPU traversal (oracle/_ref = unmodified lib/RenderCore_Bart/bvh.cpp compiled in place; else
the plain-C port) on the box's host cores: every step is the wave pass (primary + shadow + diffuse bounces) of a bounded pixel sample of
the workload's frame.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # stdout carries exactly one JSON line (NCCL prints its version banner there otherwise)

from lighthouse2_b200 import scenes, workload  # noqa: E402

EXTEND_BYTES_PER_RAY = 48.0      # 32 B ray in + 16 B hit out
SHADOW_BYTES_PER_RAY = 32.125    # 32 B ray in + 1 bit out
METRIC = "Mrays/s (wavefront extend + shadow + diffuse-bounce waves, C2 1M-triangle scene)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--tris", type=int, default=1_000_000)
    ap.add_argument("--width", type=int, default=3840, help="frame workload")
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--wave-width", type=int, default=1920, help="wave leg (per-class rates, roofline)")
    ap.add_argument("--wave-height", type=int, default=1080)
    ap.add_argument("--wave-steps", type=int, default=30)
    ap.add_argument("--bounces", type=int, default=4)
    ap.add_argument("--isect", type=int, default=0, help="0 = reference arithmetic (bit-exact parity), 1 = watertight")
    ap.add_argument("--cpu-sample", type=int, default=65536, help="pixels in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-waves", action="store_true")
    ap.add_argument("--no-refit", action="store_true")
    ap.add_argument("--no-filter", action="store_true")
    ap.add_argument("--no-check", action="store_true", help="skip the bit-identity check of the gathered frame")
    ap.add_argument("--gather", default="auto", choices=["auto", "p2p", "nccl"])
    ap.add_argument("--setting", action="append", default=[], help="name=value, applied to the core before the scene is built")
    ap.add_argument("--stats", action="store_true", help="also report nodes visited / triangles tested per ray (untimed counting run)")
    ap.add_argument("--check", action="store_true", help="also compare a sample of every wave's GPU hits with the CPU checker")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_traffic(key="extend_dram_bytes_per_launch"):
    """dram bytes per launch of the extend kernel (or another entry) from the committed ncu --set full capture (profiles/), or None."""
    p = os.path.join(ROOT, "profiles", "roofline.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get(key)
        except Exception:
            return None
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "20"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, pw, reasons = [], [], [], set()
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 8:
                continue
            try:
                sm.append(float(c[0])); mx.append(float(c[1])); pw.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw) if pw else None}


def frame_rays(args, sample_index=0):
    view = scenes.view_pyramid(width=args.wave_width, height=args.wave_height, **scenes.C2_CAMERA)
    # sub-pixel sample position: rank/sample r uses a different (irrational) offset so pixel grids never coincide
    ox = (0.5 + 0.6180339887 * sample_index) % 1.0
    oy = (0.5 + 0.7548776662 * sample_index) % 1.0
    return scenes.primary_rays(view, args.wave_width, args.wave_height, offset=(ox, oy))


def workload_name(args):
    return (f"C2 sphere field {args.tris} tris at the C5 frame size {args.width}x{args.height} x {args.spp} spp: full wavefront frame, primary + 1 NEE shadow ray "
            f"per lit vertex + {args.bounces} diffuse bounces, fixed depth")


# ------------------------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's CPU traversal on host cores over a bounded pixel sample
# ------------------------------------------------------------------------------------------------------------------
def cpu_pass(args, verts, n_pixels, steps, warmup, seed=1):
    import oracle  # test infrastructure: allowed here only as the timed CPU baseline (never on the product path)
    from oracle import Oracle
    kind = "ref" if oracle.have_ref() else "port"
    if not os.path.exists(oracle.PORT_SO):
        oracle.build()
    cores = os.cpu_count() or 1
    o = Oracle(kind)
    o.SetGeometry(0, verts)
    o.SetInstance(0, 0)
    o.UpdateToplevel()
    O, D = frame_rays(args)
    rng = np.random.default_rng(1234)
    pix = np.sort(rng.choice(len(O), min(n_pixels, len(O)), replace=False))
    Os, Ds = O[pix], D[pix]
    tr = lambda o_, d_: o.trace(o_, d_, nthreads=cores)[:2]
    waves = workload.make_pass(tr, verts, Os, Ds, scenes.C2_LIGHT, args.bounces, seed)
    counts = workload.pass_ray_counts(waves)
    total = sum(counts.values())
    times = []
    per_class = {k: 0.0 for k in counts}
    for it in range(warmup + steps):
        t_step = 0.0
        for w in waves:
            Ow, Dw = np.ascontiguousarray(w["O4"][:, :3]), np.ascontiguousarray(w["D4"][:, :3])
            t0 = time.perf_counter()
            o.trace(Ow, Dw, nthreads=cores)  # any-hit == closest t < tmax in the reference (raytracer.cpp:135-139)
            dt = time.perf_counter() - t0
            t_step += dt
            if it >= warmup:
                per_class[w["cls"]] += dt
        if it >= warmup:
            times.append(t_step)
    sec = float(np.mean(times))
    res = dict(value=total / sec / 1e6, unit="Mrays/s", cores=cores, kind="reference" if kind == "ref" else "port",
               sample=f"the waves of {len(pix)} random pixels of a 1 spp pass over the same scene (primary + shadow + {args.bounces} diffuse bounces: {total} rays/step, "
                      f"{counts}), {steps} steps, {cores} host threads; reference BVH build {o.blas_build_seconds:.1f} s not included",
               per_class={k: counts[k] * steps / per_class[k] / 1e6 for k in counts if per_class[k] > 0},
               ms_per_step=sec * 1e3, bvh_build_s=o.blas_build_seconds)
    o.close()
    return res


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    verts = scenes.sphere_field(args.tris)
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    r = cpu_pass(args, verts, args.cpu_sample, steps, warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": "Mrays/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args), "sampled": r["sample"]},
        "rays": {k: {"mrays_s": v} for k, v in r["per_class"].items()},
        "cpu_baseline": {"value": r["value"], "unit": "Mrays/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
        "e2e": {"value": r["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------------
# the B200 arm
# ------------------------------------------------------------------------------------------------------------------
def wave_leg(args, rc, core, verts, torch, dist, dev, rank, world, barrier):
    """Raw waves of a 1080p pass over rays resident in HBM, one launch timed at a time; then the same through the host-buffer calls."""
    O, D = frame_rays(args, sample_index=rank)

    def gpu_tracer(o_, d_):
        h = core.unpack_hits(rc.TraceClosest(*core.pack_rays(o_, d_, tmin=workload.EPS)))
        return h["t"], h["prim"]

    waves = workload.make_pass(gpu_tracer, verts, O, D, scenes.C2_LIGHT, args.bounces, seed=1 + rank)
    counts = workload.pass_ray_counts(waves)
    total_rays = sum(counts.values())
    stream = torch.cuda.current_stream()
    for w in waves:
        n = w["O4"].shape[0]
        w["n"] = n
        w["dO"] = torch.from_numpy(w["O4"]).to(dev)
        w["dD"] = torch.from_numpy(w["D4"]).to(dev)
        w["dOut"] = torch.empty((n, 4), dtype=torch.float32, device=dev) if w["kind"] == "closest" else \
            torch.empty(((n + 31) // 32,), dtype=torch.int32, device=dev)
        w["hO"] = torch.from_numpy(w["O4"]).pin_memory()
        w["hD"] = torch.from_numpy(w["D4"]).pin_memory()
        w["hOut"] = torch.empty(w["dOut"].shape, dtype=w["dOut"].dtype).pin_memory()
    torch.cuda.synchronize()

    def launch(w):
        if w["kind"] == "closest":
            rc.TraceClosestDevice(w["dO"].data_ptr(), w["dD"].data_ptr(), w["n"], w["dOut"].data_ptr())
        else:
            rc.TraceAnyDevice(w["dO"].data_ptr(), w["dD"].data_ptr(), w["n"], w["dOut"].data_ptr())

    steps = args.wave_steps
    for _ in range(max(3, args.warmup)):
        for w in waves:
            launch(w)
    ev = [[(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in waves] for _ in range(steps)]
    e_begin, e_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e_begin.record(stream)
    for s in range(steps):
        for i, w in enumerate(waves):
            ev[s][i][0].record(stream)
            launch(w)
            ev[s][i][1].record(stream)
    e_end.record(stream)
    barrier()
    total_ms = e_begin.elapsed_time(e_end)
    cls_ms, cls_rays = {}, {}
    ext_ms, ext_rays = 0.0, 0
    for s in range(steps):
        for i, w in enumerate(waves):
            ms = ev[s][i][0].elapsed_time(ev[s][i][1])
            cls_ms[w["cls"]] = cls_ms.get(w["cls"], 0.0) + ms
            cls_rays[w["cls"]] = cls_rays.get(w["cls"], 0) + w["n"]
            if w["kind"] == "closest":
                ext_ms += ms
                ext_rays += w["n"]

    # ---- the same waves through the host-buffer C ABI (H2D + wave + D2H inside each call) --------------------
    e2e_ms = e2e_sync_ms = None
    pcie_h2d = 0.0
    if not args.no_e2e:
        lib, h = rc.lib, rc.h

        def e2e_pass(queued=True):
            # the six waves of a step are independent inputs: they are queued back to back (b200_trace_*_async) and collected with
            # one b200_wait, so that the upload of a wave overlaps the trace and the download of the one before it
            for w in waves:
                if queued:
                    fn = lib.b200_trace_closest_async if w["kind"] == "closest" else lib.b200_trace_any_async
                else:
                    fn = lib.b200_trace_closest if w["kind"] == "closest" else lib.b200_trace_any
                if fn(h, w["hO"].data_ptr(), w["hD"].data_ptr(), w["n"], w["hOut"].data_ptr()):
                    raise RuntimeError(lib.b200_last_error().decode())
            if queued and lib.b200_wait(h):
                raise RuntimeError(lib.b200_last_error().decode())
        e2e_steps = 10
        ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e2e_modes = {}
        for queued in (False, True):
            for _ in range(2):
                e2e_pass(queued)
            barrier()
            ea.record(stream)
            for _ in range(e2e_steps):
                e2e_pass(queued)
            eb.record(stream)
            barrier()
            e2e_modes[queued] = ea.elapsed_time(eb) / e2e_steps
        e2e_ms, e2e_sync_ms = e2e_modes[True], e2e_modes[False]
        # the queued results against the device-resident pass
        for w in waves:
            launch(w)
        torch.cuda.synchronize()
        for w in waves:
            assert torch.equal(w["hOut"].view(torch.int32), w["dOut"].cpu().view(torch.int32)), "queued host-buffer wave differs from the device wave"
        # context: what a plain pinned-memory copy achieves on this box (the floor of the host-buffer pass is h2d_bytes / this)
        big = max(waves, key=lambda w_: w_["n"])
        torch.cuda.synchronize()
        ea.record(stream)
        for _ in range(5):
            big["dO"].copy_(big["hO"], non_blocking=True)
        eb.record(stream)
        torch.cuda.synchronize()
        pcie_h2d = 5 * big["n"] * 16 / (ea.elapsed_time(eb) * 1e-3) / 1e9

    vec = torch.tensor([total_ms, e2e_ms or 0.0, ext_ms] + [cls_ms.get(c, 0.0) for c in ("primary", "shadow", "diffuse")], device=dev, dtype=torch.float64)
    cnt = torch.tensor([total_rays, ext_rays] + [cls_rays.get(c, 0) for c in ("primary", "shadow", "diffuse")], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(vec, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    vec, cnt = vec.tolist(), cnt.tolist()
    out = None
    if rank == 0:
        ms_per_step = vec[0] / steps
        rays = {}
        for k, c in enumerate(("primary", "shadow", "diffuse")):
            if cnt[2 + k] > 0 and vec[3 + k] > 0:
                rays[c] = {"rays_per_step": cnt[2 + k] / steps, "mrays_s": cnt[2 + k] / (vec[3 + k] * 1e-3) / 1e6,
                           "mrays_s_per_gpu": cnt[2 + k] / world / (vec[3 + k] * 1e-3) / 1e6}
        my_ext_ms = sum(ev[s][i][0].elapsed_time(ev[s][i][1]) for s in range(steps) for i, w in enumerate(waves) if w["kind"] == "closest")
        my_ext_rays = sum(w["n"] for w in waves if w["kind"] == "closest") * steps
        my_ext_launches = sum(1 for w in waves if w["kind"] == "closest") * steps
        out = {"block": {"workload": f"raw waves of a {args.wave_width}x{args.wave_height} x 1 spp pass per GPU over rays resident in HBM: primary extend + 1 shadow ray per lit hit + "
                                     f"{args.bounces} diffuse-bounce extend waves, every launch timed alone (CUDA events on the launch stream)",
                         "scaling": "weak (rank r traces sub-pixel sample r; no collective)", "steps": steps, "ms_per_step": ms_per_step,
                         "mrays_s": cnt[0] / (ms_per_step * 1e-3) / 1e6, "rays": rays},
               "rays": rays,
               "extend": {"ms": my_ext_ms, "rays": my_ext_rays, "launches": my_ext_launches}}
        if vec[1]:
            h2d = sum(w["n"] * 32 for w in waves)
            d2h = sum(w["n"] * 16 if w["kind"] == "closest" else ((w["n"] + 31) // 32) * 4 for w in waves)
            out["e2e_waves"] = {"value": cnt[0] / (vec[1] * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                                "ms_per_step": vec[1], "pcie_h2d_gbs_measured": pcie_h2d, "pcie_floor_ms": h2d / (pcie_h2d * 1e9) * 1e3,
                                "api": "b200_trace_closest_async / b200_trace_any_async + one b200_wait per step, pinned host buffers (the six waves of a step queued back to back)",
                                "sync_calls": {"ms_per_step": e2e_sync_ms, "value": (cnt[0] / world) / (e2e_sync_ms * 1e-3) / 1e6,
                                               "note": "rank 0, b200_trace_closest / b200_trace_any: every wave synchronises before the next one is submitted"}}
    return out, waves


def run_b200(args):
    import torch
    import torch.distributed as dist
    from lighthouse2_b200 import core, multigpu

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- RenderCore_B200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    rc = core.RenderCoreB200(local)
    rc.Setting("isectMode", args.isect)
    for kv in args.setting:
        rc.Setting(kv.split("=")[0], float(kv.split("=")[1]))
    verts = scenes.sphere_field(args.tris)
    t0 = time.perf_counter()
    rc.SetGeometry(0, verts)
    rc.SetInstance(0, 0)
    rc.SetInstance(1, -1)
    rc.UpdateToplevel()
    build_s = time.perf_counter() - t0
    stream = torch.cuda.current_stream()
    rc.SetStream(stream.cuda_stream)

    # ---- sub-block: the raw waves (per-class rates, roofline of the dominant kernel) -----------------------------
    waves_out, waves = (None, None)
    if not args.no_waves:
        waves_out, waves = wave_leg(args, rc, core, verts, torch, dist, dev, rank, world, barrier)

    # ---- the frame workload: tile-sharded Render() + gather, strong scaling -----------------------------------------
    W, H, spp = args.width, args.height, args.spp
    view = scenes.view_pyramid(width=W, height=H, **scenes.C2_CAMERA)
    sr = multigpu.ShardedRenderer(rank, world, local, mode="tiles", existing_core=rc, gather=args.gather)
    sr.SetTarget(W, H, spp)
    rc.SetPointLight(scenes.C2_LIGHT, (4000.0, 4000.0, 3600.0))
    rc.SetSky((0.4, 0.5, 0.7))
    rc.Setting("maxPathLength", args.bounces + 1)
    rc.Setting("russianRoulette", 0)  # fixed depth: every frame traces the same waves
    lib, h = rc.lib, rc.h
    vp = core.ViewPyramid()
    for k in ("pos", "p1", "p2", "p3"):
        getattr(vp, k)[:] = [float(x) for x in view[k]]
    vp.aperture, vp.spreadAngle, vp.imagePlane, vp.focalDistance, vp.distortion = 0.0, 0.01, 0.01, 1.0, 0.0
    import ctypes as C
    vp_ref = C.byref(vp)

    def chk(code):
        if code:
            raise RuntimeError(lib.b200_last_error().decode())

    p2p = world > 1 and sr.gather == "p2p"

    def frame():
        # one step: this rank's tiles, then the gather (N > 1).  With open peers b200_render ends with the gather itself (arrival barrier,
        # pull + finalize kernel, second barrier: the tail of the frame's enqueue); the call synchronises its stream before it returns.
        chk(lib.b200_render(h, vp_ref, 1))
        if world > 1 and not p2p:
            dist.all_reduce(sr.acc, op=dist.ReduceOp.SUM)
            chk(lib.b200_finalize(h, spp))

    def digest():
        a = sr.ReadAccumulator()
        return hashlib.sha256(np.ascontiguousarray(a).view(np.uint8)).hexdigest()

    # ---- check: the gathered accumulator against this rank's own unsharded render of the same frame -----------------
    frame_check = None
    if not args.no_check:
        if world > 1:
            rc.Setting("shardCount", 1)
            chk(lib.b200_render(h, vp_ref, 1))
            full = digest()
            rays_full = rc.GetCoreStats()["totalRays"]
            rc.Setting("shardCount", world)
            rc.Setting("shardIndex", rank)
            barrier()
            frame()
            ok = digest() == full
        else:
            frame()
            full = digest()
            rays_full = rc.GetCoreStats()["totalRays"]
            frame()
            ok = digest() == full
        okt = torch.tensor([1 if ok else 0], device=dev, dtype=torch.int32)
        if world > 1:
            dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        frame_check = {"result": "bit-identical" if okt.item() else "DIFFERS", "ranks_checked": world, "sha256_16": full[:16],
                       "what": ("every rank's gathered float4 accumulator == its own unsharded single-GPU render of the frame" if world > 1 else
                                "two consecutive renders of the frame (run-to-run determinism; with N > 1 the gathered accumulator is compared with an unsharded render)"),
                       "rays_unsharded": rays_full}

    for _ in range(args.warmup):
        frame()
    e_begin, e_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = rc.GetCoreStats()["kernelLaunches"]
    barrier()
    clocks = ClockSampler(local) if rank == 0 else None
    wall0 = time.perf_counter()
    e_begin.record(stream)
    render_ms = gather_ms = gather_kernel_ms = 0.0
    rays_mine = 0
    for _ in range(args.steps):
        chk(lib.b200_render(h, vp_ref, 1))
        st = rc.GetCoreStats()
        render_ms += st["renderTime"] - st["gatherMs"]
        rays_mine += st["totalRays"]
        if p2p:
            gather_ms += st["gatherMs"]
            gather_kernel_ms += st["gatherKernelMs"]
        elif world > 1:
            g0.record(stream)
            dist.all_reduce(sr.acc, op=dist.ReduceOp.SUM)
            chk(lib.b200_finalize(h, spp))
            g1.record(stream)
            g1.synchronize()
            gather_ms += g0.elapsed_time(g1)
    e_end.record(stream)
    barrier()
    wall_ms = (time.perf_counter() - wall0) * 1e3
    total_ms = e_begin.elapsed_time(e_end)
    clock_info = clocks.stop() if clocks else None
    st = rc.GetCoreStats()
    launches = st["kernelLaunches"] - launches0

    # ---- e2e: the same frames with the finished frame read back into pinned host memory every step -------------------
    e2e_ms = e2e_blocking_ms = None
    frame_bytes = W * H * 16
    if not args.no_e2e:
        host_frames = [torch.empty((H, W, 4), dtype=torch.float32).pin_memory() for _ in range(2)]
        e2e_steps = max(3, min(args.steps, 10))

        # every step: Render() (the view pyramid and the pass parameters go up) and the finished frame into page-locked host memory on rank 0, where
        # the application (RenderSystem) lives.  b200_read_frame_async lets frame k travel while frame k + 1 renders (two host buffers, as a
        # display loop would hold them); the timed region ends when the last frame has arrived.  The blocking call is timed beside it.
        def e2e_run(blocking):
            def step(k):
                frame()
                if rank == 0:
                    if blocking:
                        chk(lib.b200_read_frame(h, host_frames[k & 1].data_ptr()))
                    else:
                        chk(lib.b200_read_frame_async(h, host_frames[k & 1].data_ptr()))
            step(0)
            chk(lib.b200_wait_frame(h))
            ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            ea.record(stream)
            for k in range(e2e_steps):
                step(k)
            chk(lib.b200_wait_frame(h))
            eb.record(stream)
            barrier()
            return ea.elapsed_time(eb) / e2e_steps
        e2e_blocking_ms = e2e_run(True)
        e2e_ms = e2e_run(False)
        if rank == 0:
            for hf in host_frames:
                assert np.isfinite(hf.numpy()).all() and float(hf[..., :3].mean()) > 0.01
            assert torch.equal(host_frames[0], host_frames[1])  # (the same view, the same samples: restarted frames)

    # ---- reduce over ranks: slowest rank's time, all ranks' rays ------------------------------------------------
    vec = torch.tensor([total_ms, e2e_ms or 0.0, render_ms, gather_ms, gather_kernel_ms, wall_ms], device=dev, dtype=torch.float64)
    vmin = vec.clone()
    cnt = torch.tensor([rays_mine, launches], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(vec, op=dist.ReduceOp.MAX)
        dist.all_reduce(vmin, op=dist.ReduceOp.MIN)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    vec, vmin, cnt = vec.tolist(), vmin.tolist(), cnt.tolist()

    if rank == 0:
        steps = args.steps
        ms_per_step = vec[0] / steps
        rays_per_frame = cnt[0] / steps
        value = rays_per_frame / (ms_per_step * 1e-3) / 1e6
        peak, peak_src = peaks()
        line = {
            "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world,
            "steps": steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args), "isect": "reference-exact" if args.isect == 0 else "watertight",
                       "sharding": ("scene replicated, frame sharded by interleaved 32x8 tiles over %d ranks; per frame one gather of the float4 accumulator over NVLink (%s)"
                                    % (world, "fused peer-memory pull + finalize kernel" if p2p else "NCCL all-reduce")) if world > 1 else "1 GPU, no sharding",
                       "l2": "no flush: the path buffers of one pass (%.1f GB of rays, hits, path state and contribution slots per rank) exceed the 126 MB L2 many times over; "
                             "the scene (%.0f MB nodes + %.0f MB triangles) is meant to stay L2-resident"
                             % (W * H * spp / world / max(1, st["framePasses"]) * 272 / 1e9, st["wideNodeCount"] * 80 / 1e6, st["triangleCount"] * 48 / 1e6),
                       "rays_per_frame": rays_per_frame, "passes_per_frame": st["framePasses"], "wave_chains": st["frameChains"],
                       "frame_graph_replays": st["frameGraphLaunches"], "wide_nodes": st["wideNodeCount"], "bvh_build_s": round(build_s, 2),
                       "timed_region_wall_s": vec[5] * 1e-3},
            "frame_check": frame_check["result"] if frame_check else "skipped",
            "frame": {"check": frame_check, "render_ms_slowest_rank": vec[2] / steps, "render_ms_fastest_rank": vmin[2] / steps,
                      "gather": (sr.gather if world > 1 else "none"),
                      "gather_ms_barrier_to_barrier": vec[3] / steps if world > 1 else 0.0,
                      "gather_note": "barrier_to_barrier = from a rank's own arrival to the end of the second arrival: waiting for the slowest rank + pull + finalize (CUDA "
                                     "events inside the frame, the rank that waited longest); kernel = the pull + finalize kernel alone, after everyone has arrived",
                      "rank0_stats": {k: st[k] for k in ("primaryRayCount", "traceTime0", "bounce1RayCount", "traceTime1", "deepRayCount", "traceTimeX",
                                                         "totalShadowRays", "shadowTraceTime", "shadeTime", "renderTime")}},
            "gpu_launches": int(cnt[1]),
            "clocks": clock_info,
        }
        if p2p:
            gk = vec[4] / steps
            pulled = frame_bytes * (world - 1) / world
            line["frame"]["gather_kernel_ms"] = gk
            line["frame"]["gather_nvlink_gbs_per_gpu"] = pulled / (gk * 1e-3) / 1e9 if gk > 0 else None
            line["frame"]["gather_bytes_pulled_per_gpu"] = pulled
        if e2e_ms:
            line["e2e"] = {"value": rays_per_frame / (vec[1] * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 68 + 96 * st["framePasses"],
                           "d2h_bytes_per_step": frame_bytes, "ms_per_step": vec[1],
                           "blocking_read": {"ms_per_step": e2e_blocking_ms, "value": rays_per_frame / (e2e_blocking_ms * 1e-3) / 1e6,
                                             "note": "b200_read_frame after every b200_render (rank 0's clock): the frame crosses the bus before the next one starts"},
                           "api": "b200_render (ViewPyramid from host memory; ends with the peer-memory gather when N > 1) + b200_read_frame_async into two alternating pinned host buffers "
                                  "(rank 0; frame k travels while frame k + 1 renders, b200_wait_frame closes the timed region): what the RenderSystem "
                                  "sees through CoreAPI::Render; the C++ vtable face forwards to the same calls (tests/test_plugin_gpu.py)",
                           "note": "inputs of this path are a 68-byte view pyramid plus 96 bytes of pass parameters per pass; the output is the finished RGBA32F frame"}
        if waves_out:
            ext = waves_out["extend"]
            achieved = ext["rays"] * EXTEND_BYTES_PER_RAY / (ext["ms"] * 1e-3) / 1e9
            line["rays"] = waves_out["rays"]
            line["waves"] = waves_out["block"]
            if "e2e_waves" in waves_out:
                line["e2e_waves"] = waves_out["e2e_waves"]
            line["roofline"] = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": measured_traffic(),
                                "kernel": "traceKernel<closest> (extend wave)", "algorithmic_bytes_per_ray": EXTEND_BYTES_PER_RAY,
                                "algorithmic_bytes_per_launch": ext["rays"] * EXTEND_BYTES_PER_RAY / ext["launches"], "avg_launch_ms": ext["ms"] / ext["launches"],
                                "launches_timed": ext["launches"], "peak_source": peak_src,
                                "measured_on": "the extend launches of the `waves` block (each timed alone with CUDA events on its launch stream, rank 0)",
                                "per_class_mrays_s": {k: v["mrays_s_per_gpu"] for k, v in waves_out["rays"].items()},
                                "note": "stream roofline of a latency-bound traversal kernel (dependent chain of a pass at 6 warps per scheduler); see DESIGN.md",
                                "ncu": measured_traffic("extend_ncu")}
        if not args.no_refit:
            # the north_star's refit kernel on the same mesh: new positions already on the device, one launch that climbs the tree
            # (refit.cu::refitWalkKernel); bytes by SURVEY.md 8(d)'s formula: 48 B per triangle (3 x float4 vertex read) + one read and one write per node
            v4 = np.ones((verts.shape[0] * 3, 4), np.float32)
            v4[:, :3] = verts.reshape(-1, 3)
            dv = torch.from_numpy(v4).to(dev)
            ms = []
            for _ in range(12):
                rc.RefitDevice(0, dv.data_ptr(), v4.shape[0])
                ms.append(rc.GetCoreStats()["refitTime"])
            ms = float(np.mean(ms[2:]))
            nodes = st["wideNodeCount"]
            b = 48 * verts.shape[0] + 2 * 80 * nodes
            line["refit"] = {"workload": "GPU refit of the C2 mesh (%d triangles, %d wide nodes), vertices resident on the device" % (verts.shape[0], nodes), "ms": ms,
                             "launches": 1, "bytes_survey_formula": b, "gbs": b / (ms * 1e-3) / 1e9, "frac_of_hbm_peak": b / (ms * 1e-3) / 1e9 / peak,
                             "bytes_with_writes": 144 * verts.shape[0] + 224 * nodes,
                             "note": "survey formula = 48 T + 2 x 80 x nodes; with the record / shading-normal writes and the per-node float boxes the kernel moves 144 T + 224 nodes"}
            # the build itself: SetGeometry wall time of a fresh core, vertex upload, tree, refit fill and read-back of the
            # root box included; the default builder (build.cu, binned SAH on the device) and the host builder (bvh_builder.cpp) on this box's threads
            def build_ms(setting):
                best = 1e9
                for _ in range(3):
                    rb = core.RenderCoreB200(local)
                    rb.Setting("gpuBuild", setting)
                    t0 = time.perf_counter()
                    rb.SetGeometry(0, v4)
                    best = min(best, (time.perf_counter() - t0) * 1e3)
                    n = rb.GetCoreStats()["wideNodeCount"]
                    rb.Shutdown()
                return best, n
            dms, dn = build_ms(4)
            hms, hn = build_ms(0)
            line["build"] = {"workload": "SetGeometry of the C2 mesh (%d triangles), wall time, best of 3" % verts.shape[0], "device_sah_ms": dms, "device_wide_nodes": dn,
                             "host_builder_ms": hms, "host_wide_nodes": hn, "host_threads": os.cpu_count(),
                             "note": "the timed frames run on the default (device-built) tree; trace rate against the host-built tree: profiles/r02_experiments.md"}
        if world == 1 and not args.no_filter:
            # filter mode (Setting filterCore: the reference's RenderCore_Optix7Filter -- one diffuse bounce, per-pixel features, SVGF-style
            # reconstruction + TAA, filter.cu) on the same scene at the wave leg's size, 1 spp: device time of the filter and of the frame
            rc.SetAccumulatorBuffer(0)  # (the one-rank renderer above rendered into a torch buffer: the filter keeps its own two planes)
            rc.SetTarget(args.wave_width, args.wave_height, 1)
            rc.Setting("filterCore", 1)
            fview = scenes.view_pyramid(width=args.wave_width, height=args.wave_height, **scenes.C2_CAMERA)
            fms, rms = [], []
            for _ in range(8):
                rc.Render(fview, converge=0)
                fst = rc.GetCoreStats()
                fms.append(fst["filterTime"]), rms.append(fst["renderTime"])
            rc.Setting("filterCore", 0)
            px = args.wave_width * args.wave_height
            line["filter"] = {"workload": "C2 scene %dx%d x 1 spp in filter mode (TAA on)" % (args.wave_width, args.wave_height), "filter_ms": float(np.median(fms[2:])),
                              "frame_ms": float(np.median(rms[2:])), "kernels": 6, "mpixels_s": px / (float(np.median(fms[2:])) * 1e-3) / 1e6,
                              "note": "prepare is HBM-bound (78 % of the peak), the three reconstruction passes are issue-bound (profiles/r02_ncu_filter.txt)"}
        if world == 1 and not args.no_cpu_baseline:
            r = cpu_pass(args, verts, args.cpu_sample, steps=2, warmup=1)
            line["cpu_baseline"] = {"value": r["value"], "unit": "Mrays/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"],
                                    "per_class": r["per_class"], "bvh_build_s": r["bvh_build_s"]}
        if args.stats and waves:
            rc.Setting("countSteps", 1)
            line["steps_per_ray"] = {}
            for w in waves:
                if w["kind"] in ("closest", "any"):  # (shadow rays counted as closest-hit rays over their segment: the walk without the early exit)
                    rc.TraceClosest(w["O4"], w["D4"])
                    nn, nt = rc.StepCounters(w["n"])
                    line["steps_per_ray"][f"{w['cls']}{w['bounce']}"] = {"nodes": float(nn.mean()), "tris": float(nt.mean()), "nodes_max": int(nn.max()),
                                                                             "nodes_pct": [float(x) for x in np.percentile(nn, [50, 90, 99, 99.9, 99.99])]}
            rc.Setting("countSteps", 0)
        if args.check and waves:
            line["check"] = check_against_cpu(rc, core, verts, waves)
        print(json.dumps(line), flush=True)
    sr.Shutdown() if world > 1 else rc.Shutdown()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def check_against_cpu(rc, core, verts, waves, n=20000):
    """Optional: a sample of every wave against the CPU checker (test infrastructure; not in any timed region)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle import Oracle
    o = Oracle("port")
    o.SetGeometry(0, verts); o.SetInstance(0, 0); o.UpdateToplevel()
    out = {}
    rng = np.random.default_rng(7)
    for w in waves:
        idx = np.sort(rng.choice(w["n"], min(n, w["n"]), replace=False))
        O4, D4 = w["O4"][idx], w["D4"][idx]
        t, prim, _ = o.trace(O4[:, :3], D4[:, :3], nthreads=os.cpu_count() or 1)
        key = f"{w['cls']}{w['bounce']}"
        if w["kind"] == "closest":
            h = core.unpack_hits(rc.TraceClosest(O4, D4))
            out[key] = int(((h["prim"] != prim) | (h["t"].view(np.uint32) != t.view(np.uint32))).sum())
        else:
            out[key] = int((rc.TraceAny(O4, D4) != (t < D4[:, 3])).sum())
    return out


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
