#!/bin/bash
# tools/sweep.sh "name=v1" "name=v2" ... : bench with one core Setting each (GPU box), prints class Mrays/s and per-ray step counts
for kv in "$@"; do
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-frame --stats --setting "$kv" > gpurun_out/sw.json 2> gpurun_out/sw.err || tail -3 gpurun_out/sw.err
  python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/sw.json") if l.startswith("{")][-1])
print("$kv", "total %.0f"%d["value"], {k:round(x["mrays_s"]) for k,x in d["rays"].items()}, "nodes", d["config"]["wide_nodes"], {k:(round(v["nodes"],1),round(v["tris"],1)) for k,v in d["steps_per_ray"].items()})
PY
done
