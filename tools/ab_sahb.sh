#!/bin/bash
# SAH bottoms under the radix tree (Setting gpuBuild = 3): subtree size and bin count, against the host tree (gpuBuild = 0).
# variants/sahb_* are built by tools/build_variant.sh with -DSAHB_CLUSTER=... / -DSAHB_BINS_N=...
run() { # label, lib, gpuBuild
  B200_LIB=$2 timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-check --no-refit --no-filter --steps 5 --warmup 2 --wave-steps 30 --setting gpuBuild=$3 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); w=d['waves']['rays']
print('C2 $1: frame %.2f ms (%.0f Mrays/s); waves primary %.0f shadow %.0f diffuse %.0f Mrays/s'%(d['ms_per_step'], d['value'], w['primary']['mrays_s'], w['shadow']['mrays_s'], w['diffuse']['mrays_s']))"
}
run "host tree" "" 0
run "sahb 256 x 16 bins" "" 3
for v in variants/sahb_*; do run "$(basename $v)" $v/libRenderCore_B200.so 3; done
