#!/bin/bash
# tools/ab_frame.sh v1 v2 ... : frame bench (Render through the C ABI) for each variant under variants/ ("base" = in-tree lib)
for v in "$@"; do
  if [ "$v" = base ]; then unset B200_LIB; else export B200_LIB=$PWD/variants/$v/libRenderCore_B200.so; fi
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/abf_$v.json 2> gpurun_out/abf_$v.err || tail -3 gpurun_out/abf_$v.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/abf_$v.json").read().strip().splitlines()[-1]); f=d["frame"]
    print("$v", "frame render_ms %.3f"%f["render_ms"], "shade %.3f"%f["rank0_stats"]["shadeTime"])
except Exception as e: print("$v", "FAILED", e)
PY
done
