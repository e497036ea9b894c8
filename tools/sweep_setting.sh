#!/bin/bash
# tools/sweep_setting.sh name v1 v2 ... : wave bench (device-resident) for each value of one Setting; prints class Mrays/s
name=$1; shift
for v in "$@"; do
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-frame --setting $name=$v > gpurun_out/sw_${name}_$v.json 2> gpurun_out/sw_${name}_$v.err || tail -3 gpurun_out/sw_${name}_$v.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/sw_${name}_$v.json").read().strip().splitlines()[-1])
    print("$name=$v", "total %.0f"%d["value"], {k:round(x["mrays_s"]) for k,x in d["rays"].items()})
except Exception as e: print("$name=$v", "FAILED", e)
PY
done
