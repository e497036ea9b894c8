#!/bin/bash
# tools/sweep_e2e.sh "a=1,b=2" "c=3" ... : host-buffer (e2e) bench for each comma-separated group of core Settings; prints queued / synchronous ms per step
for g in "$@"; do
  args=""
  if [ "$g" != base ]; then for kv in ${g//,/ }; do args="$args --setting $kv"; done; fi
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-frame $args > gpurun_out/e2e_sw.json 2> gpurun_out/e2e_sw.err || tail -3 gpurun_out/e2e_sw.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/e2e_sw.json").read().strip().splitlines()[-1]); e=d["e2e"]
    print("$g", "queued %.3f ms (%.0f Mrays/s)  sync %.3f ms  floor %.3f ms" % (e["ms_per_step"], e["value"], e["sync_calls"]["ms_per_step"], e["pcie_floor_ms"]))
except Exception as ex: print("$g", "FAILED", ex)
PY
done
