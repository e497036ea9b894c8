#!/bin/bash
# builder 4 (binned SAH from the top): bins of the top phase, size of the subtrees finished in shared memory, bins there.
# variants/sah_* are built by tools/build_variant.sh with -DTOP_BINS_N=.. / -DSAHB_CLUSTER=.. / -DSAHB_BINS_N=..
run() { # label, lib
  B200_LIB=$2 timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-check --no-refit --no-filter --steps 5 --warmup 2 --wave-steps 30 --setting gpuBuild=4 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); w=d['waves']['rays']
print('C2 $1: frame %.2f ms (%.0f Mrays/s); waves primary %.0f shadow %.0f diffuse %.0f Mrays/s'%(d['ms_per_step'], d['value'], w['primary']['mrays_s'], w['shadow']['mrays_s'], w['diffuse']['mrays_s']))"
}
run "32 top bins, 256, 16 bins" ""
for v in variants/sah_*; do run "$(basename $v)" $v/libRenderCore_B200.so; done
