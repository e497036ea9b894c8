#!/bin/bash
# SAH inside the radix tree's cells (Setting gpuBuild = 3): how many levels are rebuilt (Setting gpuBuildLevels; 1 = the bottoms only,
# the default), against the host tree (gpuBuild = 0) and the plain radix tree (1).  Subtree size and bin count: build variants with
# tools/build_variant.sh NAME "-DSAHB_CLUSTER=512" / "-DSAHB_BINS_N=32" and run with B200_LIB (128 .. 2048 items and 32 bins: within 0.3 %).
run() { # label, settings...
  label=$1; shift
  timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-check --no-refit --no-filter --steps 5 --warmup 2 --wave-steps 30 $(for s in "$@"; do echo --setting $s; done) 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); w=d['waves']['rays']
print('C2 $label: frame %.2f ms (%.0f Mrays/s); waves primary %.0f shadow %.0f diffuse %.0f Mrays/s'%(d['ms_per_step'], d['value'], w['primary']['mrays_s'], w['shadow']['mrays_s'], w['diffuse']['mrays_s']))"
}
run "host tree" gpuBuild=0
run "radix tree" gpuBuild=1
run "SAH bottoms (1 level)" gpuBuild=3
run "SAH, 2 levels" gpuBuild=3 gpuBuildLevels=2
run "SAH, all levels" gpuBuild=3 gpuBuildLevels=64
run "binned SAH from the top" gpuBuild=4
