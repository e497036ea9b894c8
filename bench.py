#!/usr/bin/env python
"""bench.py -- Mrays/s of the wavefront extend/shadow path (BASELINE.json metric) on B200.

One STEP = one pass of the hot path over one batch of synthetic rays on the C2 scene (BASELINE.json configs[1]:
tessellated-sphere field, 1 000 000 triangles, 1920x1080): the primary extend wave, the shadow (any-hit) wave
of one ray per lit primary hit, and the four incoherent diffuse-bounce extend waves (the headline class).
The ray sets are generated once, outside the timed region (lighthouse2_b200/workload.py), and are resident
in HBM when the timed region starts.  `value` = all rays of the step / device time (CUDA events, max over
ranks); `rays` breaks it down per class the way the reference UI does (count / traceTime per wave class,
apps/imguiapp/main.cpp:230-233).  `e2e` = the same pass through the host-buffer C-ABI calls
(b200_trace_closest / b200_trace_any: pinned host rays in, host hits out, copies inside the timed region).

Multi-GPU (torchrun, one process per GPU): the scene is replicated, every rank traces its own sample of the
frame (rank r = sub-pixel sample r: rays shard by spp) -- weak scaling, no data-path collective.

`--impl reference` times the reference's own CPU traversal (oracle/_ref = unmodified lib/RenderCore_Bart/bvh.cpp
compiled in place; else the plain-C port) on the box's host cores over a bounded pixel sample of the same pass.
"""
from __future__ import annotations

import argparse
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


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--tris", type=int, default=1_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--bounces", type=int, default=4)
    ap.add_argument("--isect", type=int, default=0, help="0 = reference arithmetic (bit-exact parity), 1 = watertight")
    ap.add_argument("--cpu-sample", type=int, default=65536, help="pixels in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-detail", action="store_true", help="per-wave wall time of the host-buffer calls for several chunk schedules (stderr)")
    ap.add_argument("--no-frame", action="store_true")
    ap.add_argument("--setting", action="append", default=[], help="name=value, applied to the core before the scene is built")
    ap.add_argument("--stats", action="store_true", help="also report nodes visited / triangles tested per ray (untimed counting run)")
    ap.add_argument("--check", action="store_true", help="compare a sample of the GPU hits with the CPU checker")
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
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 8:
                continue
            try:
                sm.append(float(c[0])); mx.append(float(c[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def frame_rays(args, sample_index=0):
    view = scenes.view_pyramid(width=args.width, height=args.height, **scenes.C2_CAMERA)
    # sub-pixel sample position: rank/sample r uses a different (irrational) offset so pixel grids never coincide
    ox = (0.5 + 0.6180339887 * sample_index) % 1.0
    oy = (0.5 + 0.7548776662 * sample_index) % 1.0
    return scenes.primary_rays(view, args.width, args.height, offset=(ox, oy))


def workload_name(args):
    return (f"C2 sphere field {args.tris} tris, {args.width}x{args.height} x 1 spp per GPU: primary extend + 1 shadow ray per lit hit + "
            f"{args.bounces} diffuse-bounce extend waves")


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
               sample=f"{len(pix)} random pixels of the frame, full pass ({total} rays/step: {counts}), {steps} steps, "
                      f"{cores} host threads; reference BVH build {o.blas_build_seconds:.1f} s not included",
               per_class={k: counts[k] * steps / per_class[k] / 1e6 for k in counts if per_class[k] > 0},
               ms_per_step=sec * 1e3, bvh_build_s=o.blas_build_seconds)
    o.close()
    return res


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    verts = scenes.sphere_field(args.tris)
    steps, warmup = max(1, min(args.steps, 5)), max(1, min(args.warmup, 1))
    r = cpu_pass(args, verts, args.cpu_sample, steps, warmup)
    line = {
        "impl": "reference", "metric": "Mrays/s (primary+shadow+diffuse waves, C2 1M-tri scene)", "value": r["value"], "unit": "Mrays/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
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
def run_b200(args):
    import torch
    import torch.distributed as dist
    from lighthouse2_b200 import core

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- RenderCore_B200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

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

    # ---- synthetic input: this rank's pass (its own sub-pixel sample), traced by the CUDA core itself -----------
    O, D = frame_rays(args, sample_index=rank)

    def gpu_tracer(o_, d_):
        h = core.unpack_hits(rc.TraceClosest(*core.pack_rays(o_, d_, tmin=workload.EPS)))
        return h["t"], h["prim"]

    waves = workload.make_pass(gpu_tracer, verts, O, D, scenes.C2_LIGHT, args.bounces, seed=1 + rank)
    counts = workload.pass_ray_counts(waves)
    total_rays = sum(counts.values())

    stream = torch.cuda.current_stream()
    rc.SetStream(stream.cuda_stream)
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

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident pass -------------------------------------------------------------------------------
    for _ in range(args.warmup):
        for w in waves:
            launch(w)
    ev = [[(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in waves] for _ in range(args.steps)]
    e_begin, e_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = rc.GetCoreStats()["kernelLaunches"]
    barrier()
    clocks = ClockSampler(local) if rank == 0 else None
    e_begin.record(stream)
    for s in range(args.steps):
        for i, w in enumerate(waves):
            ev[s][i][0].record(stream)
            launch(w)
            ev[s][i][1].record(stream)
    e_end.record(stream)
    barrier()
    total_ms = e_begin.elapsed_time(e_end)
    clock_info = clocks.stop() if clocks else None
    launches = rc.GetCoreStats()["kernelLaunches"] - launches0
    cls_ms, cls_rays = {}, {}
    ext_ms, ext_rays = 0.0, 0
    for s in range(args.steps):
        for i, w in enumerate(waves):
            ms = ev[s][i][0].elapsed_time(ev[s][i][1])
            cls_ms[w["cls"]] = cls_ms.get(w["cls"], 0.0) + ms
            cls_rays[w["cls"]] = cls_rays.get(w["cls"], 0) + w["n"]
            if w["kind"] == "closest":
                ext_ms += ms
                ext_rays += w["n"]

    # ---- e2e: host buffers through the C ABI (H2D + wave + D2H + sync inside each call) ------------------------
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
        e2e_steps = max(3, min(args.steps, 10))
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
        if rank == 0 and args.check:
            # the queued results against the device-resident pass
            for w in waves:
                launch(w)
            torch.cuda.synchronize()
            for w in waves:
                assert torch.equal(w["hOut"].view(torch.int32), w["dOut"].cpu().view(torch.int32)), "queued host-buffer wave differs from the device wave"
        # context for the e2e number: what a plain pinned-memory copy achieves on this box (the e2e floor is h2d_bytes / this)
        big = max(waves, key=lambda w_: w_["n"])
        torch.cuda.synchronize()
        ea.record(stream)
        for _ in range(5):
            big["dO"].copy_(big["hO"], non_blocking=True)
        eb.record(stream)
        torch.cuda.synchronize()
        pcie_h2d = 5 * big["n"] * 16 / (ea.elapsed_time(eb) * 1e-3) / 1e9

        if args.e2e_detail and rank == 0:
            def copy_ms(w):  # the wave's own upload alone, and its kernel alone
                torch.cuda.synchronize(); t0 = time.perf_counter()
                for _ in range(5):
                    w["dO"].copy_(w["hO"], non_blocking=True); w["dD"].copy_(w["hD"], non_blocking=True)
                torch.cuda.synchronize(); up = (time.perf_counter() - t0) / 5 * 1e3
                t0 = time.perf_counter()
                for _ in range(5):
                    launch(w)
                torch.cuda.synchronize(); k = (time.perf_counter() - t0) / 5 * 1e3
                t0 = time.perf_counter()
                for _ in range(5):
                    w["hOut"].copy_(w["dOut"], non_blocking=True)
                torch.cuda.synchronize(); dn = (time.perf_counter() - t0) / 5 * 1e3
                return up, k, dn
            print("wave            n      up_ms  kern_ms  down_ms", file=sys.stderr)
            for w in waves:
                print("%-9s %9d  %7.3f  %7.3f  %7.3f" % ((w["cls"] + str(w["bounce"]), w["n"]) + copy_ms(w)), file=sys.stderr)
            for zc, two, parts in ((0, 0, 6), (0, 0, 8), (0, 1, 4), (0, 1, 6), (0, 1, 8), (0, 1, 12), (1, 0, 4), (1, 1, 2), (1, 1, 4), (1, 1, 8)):
                rc.Setting("zeroCopy", zc); rc.Setting("twoStreams", two); rc.Setting("zeroCopyBelow", 1 << 30)
                rc.Setting("pinnedChunks" if zc else "hostChunks", parts)
                row = []
                for w in waves:
                    fn = lib.b200_trace_closest if w["kind"] == "closest" else lib.b200_trace_any
                    best = 1e9
                    for _ in range(6):
                        t0 = time.perf_counter()
                        fn(h, w["hO"].data_ptr(), w["hD"].data_ptr(), w["n"], w["hOut"].data_ptr())
                        best = min(best, (time.perf_counter() - t0) * 1e3)
                    row.append(best)
                print("zc %d two %d parts %2d: " % (zc, two, parts) + " ".join("%6.3f" % x for x in row) + "  sum %.3f" % sum(row), file=sys.stderr)
            rc.Setting("zeroCopyBelow", 400000); rc.Setting("zeroCopy", 1); rc.Setting("twoStreams", 1); rc.Setting("pinnedChunks", 1); rc.Setting("hostChunks", 4)

    # ---- the same scene through Render(): full wavefront frame (ray generation, shade, fused connect), tile-sharded over the
    #      ranks, accumulator gathered by the fused peer-memory pull + finalize kernel (NCCL all-reduce if the peers cannot be mapped).  Reported beside the wave numbers, not as `value`.
    frame = None
    if not args.no_frame:
        from lighthouse2_b200 import multigpu
        sr = multigpu.ShardedRenderer(rank, world, local, mode="tiles", existing_core=rc)
        sr.SetTarget(args.width, args.height, 1)
        rc.SetPointLight(scenes.C2_LIGHT, (4000.0, 4000.0, 3600.0))
        rc.SetSky((0.4, 0.5, 0.7))
        rc.Setting("maxPathLength", args.bounces + 1)
        rc.Setting("russianRoulette", 0)  # fixed-depth pass: the frame traces the same waves as the wave benchmark above
        view = scenes.view_pyramid(width=args.width, height=args.height, **scenes.C2_CAMERA)
        for _ in range(2):
            sr.Render(view, 1)
        fsteps = max(3, min(args.steps, 10))
        barrier()
        t0 = time.perf_counter()
        r_ms = g_ms = 0.0
        rays_f = 0
        for _ in range(fsteps):
            a_, b_ = sr.Render(view, 1)
            r_ms += a_; g_ms += b_
            rays_f += rc.GetCoreStats()["totalRays"]
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3 / fsteps
        st = rc.GetCoreStats()
        fv = torch.tensor([wall_ms, r_ms / fsteps, g_ms / fsteps], device=dev, dtype=torch.float64)
        fr = torch.tensor([rays_f / fsteps], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(fv, op=dist.ReduceOp.MAX)
            dist.all_reduce(fr, op=dist.ReduceOp.SUM)
        fv = fv.tolist()
        frame = {"ms_per_frame_wall": fv[0], "render_ms": fv[1], "gather_ms": fv[2], "rays_per_frame": fr.item(),
                 "mrays_s": fr.item() / (fv[0] * 1e-3) / 1e6, "gather_bytes": args.width * args.height * 16,
                 "rank0_stats": {k: st[k] for k in ("primaryRayCount", "traceTime0", "bounce1RayCount", "traceTime1", "deepRayCount", "traceTimeX",
                                                    "totalShadowRays", "shadowTraceTime", "shadeTime", "renderTime")},
                 "gather": sr.gather if world > 1 else "none",
                 "note": "Render() + gather of the float4 accumulator (p2p = one kernel pulls the foreign tiles over NVLink peer memory and finalizes the frame; "
                         "nccl = all-reduce + finalize), 1 spp, tile-sharded; wall clock incl. host sync per frame"}

    # ---- reduce over ranks: slowest rank's time, all ranks' rays ------------------------------------------------
    vec = torch.tensor([total_ms, e2e_ms or 0.0, ext_ms] + [cls_ms.get(c, 0.0) for c in ("primary", "shadow", "diffuse")], device=dev, dtype=torch.float64)
    cnt = torch.tensor([total_rays, ext_rays] + [cls_rays.get(c, 0) for c in ("primary", "shadow", "diffuse")], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(vec, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    vec, cnt = vec.tolist(), cnt.tolist()
    total_ms, e2e_ms_max, ext_ms = vec[0], vec[1], vec[2]
    all_rays_step = cnt[0]

    if rank == 0:
        ms_per_step = total_ms / args.steps
        value = all_rays_step / (ms_per_step * 1e-3) / 1e6
        peak, peak_src = peaks()
        # dominant kernel = the extend (closest-hit) wave kernel; algorithmic bytes = 48 B/ray (this rank's launches)
        my_ext_ms = sum(ev[s][i][0].elapsed_time(ev[s][i][1]) for s in range(args.steps) for i, w in enumerate(waves) if w["kind"] == "closest")
        my_ext_rays = sum(w["n"] for w in waves if w["kind"] == "closest") * args.steps
        achieved = my_ext_rays * EXTEND_BYTES_PER_RAY / (my_ext_ms * 1e-3) / 1e9
        rays = {}
        for k, c in enumerate(("primary", "shadow", "diffuse")):
            if cnt[2 + k] > 0 and vec[3 + k] > 0:
                rays[c] = {"rays_per_step": cnt[2 + k] / args.steps, "mrays_s": cnt[2 + k] / (vec[3 + k] * 1e-3) / 1e6,
                           "mrays_s_per_gpu": cnt[2 + k] / world / (vec[3 + k] * 1e-3) / 1e6}
        st = rc.GetCoreStats()
        line = {
            "metric": "Mrays/s (primary+shadow+diffuse waves, C2 1M-tri scene)", "value": value, "unit": "Mrays/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args), "isect": "reference-exact" if args.isect == 0 else "watertight",
                       "sharding": "scene replicated, rank r traces sub-pixel sample r (spp-sharded), no data-path collective",
                       "l2": "no flush: ray+hit streams of one step (%.0f MB) exceed the 126 MB L2; the scene (%.0f MB nodes + %.0f MB triangles) "
                             "is meant to stay L2-resident" % (all_rays_step / world * 48 / 1e6, st["wideNodeCount"] * 80 / 1e6, st["triangleCount"] * 48 / 1e6),
                       "wide_nodes": st["wideNodeCount"], "bvh_build_s": round(build_s, 2)},
            "rays": rays,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": measured_traffic(),
                         "kernel": "traceKernel<closest> (extend wave)", "algorithmic_bytes_per_ray": EXTEND_BYTES_PER_RAY, "peak_source": peak_src,
                         "note": "stream roofline of a latency-bound traversal kernel (dependent chain of a pass at 6 warps per scheduler); see DESIGN.md", "ncu": measured_traffic("extend_ncu")},
            "gpu_launches": int(launches),
            "clocks": clock_info,
        }
        if e2e_ms_max:
            h2d = sum(w["n"] * 32 for w in waves)
            d2h = sum(w["n"] * 16 if w["kind"] == "closest" else ((w["n"] + 31) // 32) * 4 for w in waves)
            line["e2e"] = {"value": all_rays_step / (e2e_ms_max * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": h2d,
                           "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms_max, "pcie_h2d_gbs_measured": pcie_h2d,
                           "pcie_floor_ms": h2d / (pcie_h2d * 1e9) * 1e3,
                           "api": "b200_trace_closest_async / b200_trace_any_async + one b200_wait per step, pinned host buffers (the six waves of a step queued back to back)",
                           "sync_calls": {"ms_per_step": e2e_sync_ms, "value": (all_rays_step / world) / (e2e_sync_ms * 1e-3) / 1e6,
                                          "note": "rank 0, b200_trace_closest / b200_trace_any: every wave synchronises before the next one is submitted"}}
        if world == 1 and not args.no_cpu_baseline:
            r = cpu_pass(args, verts, args.cpu_sample, steps=2, warmup=1)
            line["cpu_baseline"] = {"value": r["value"], "unit": "Mrays/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"],
                                    "per_class": r["per_class"], "bvh_build_s": r["bvh_build_s"]}
        if frame:
            line["frame"] = frame
        if args.stats:
            rc.Setting("countSteps", 1)
            line["steps_per_ray"] = {}
            for w in waves:
                if w["kind"] == "closest":
                    rc.TraceClosest(w["O4"], w["D4"])
                    nn, nt = rc.StepCounters(w["n"])
                    line["steps_per_ray"][f"{w['cls']}{w['bounce']}"] = {"nodes": float(nn.mean()), "tris": float(nt.mean()), "nodes_max": int(nn.max()),
                                                                             "nodes_pct": [float(x) for x in np.percentile(nn, [50, 90, 99, 99.9, 99.99])]}
            rc.Setting("countSteps", 0)
        if args.check:
            line["check"] = check_against_cpu(rc, core, verts, waves)
        print(json.dumps(line), flush=True)
    rc.Shutdown()
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
