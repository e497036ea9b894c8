#!/bin/bash
# tools/ab.sh v1 v2 ... : bench each variant under variants/ (and "base" = the in-tree lib) on the GPU box; prints class Mrays/s
for v in "$@"; do
  if [ "$v" = base ]; then unset B200_LIB; else export B200_LIB=$PWD/variants/$v/libRenderCore_B200.so; fi
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --check > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err || tail -3 gpurun_out/ab_$v.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/ab_$v.json").read().strip().splitlines()[-1])
    print("$v", "total %.0f"%d["value"], {k:round(x["mrays_s"]) for k,x in d["rays"].items()}, "build_s", d["config"]["bvh_build_s"], "nodes", d["config"]["wide_nodes"], "mismatches", sum((d.get("check") or {}).values()))
except Exception as e: print("$v", "FAILED", e)
PY
done
