#!/bin/bash
# End-of-wave prefetch of the next node (traverse.cu, B200_DRAIN_PREFETCH = n: warps out of rays with fewer than n active lanes): the 1080p waves of
# bench.py (where the per-wave constant shows) for n = 0 (off), 12 and 33 (every exhausted warp).  Builds: make -C lighthouse2_b200/csrc OUT=variants/pfN DEFS=-DB200_DRAIN_PREFETCH=N
for v in pf0 pf12 pf33 pf0 pf12 pf33; do
  B200_LIB=$PWD/variants/$v/libRenderCore_B200.so python bench.py --no-cpu-baseline --no-e2e --no-check --steps 3 --warmup 1 --wave-steps 30 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); w=d['waves']['rays']
print('$v: frame %.2f ms; waves primary %.0f shadow %.0f diffuse %.0f Mrays/s'%(d['ms_per_step'], w['primary']['mrays_s'], w['shadow']['mrays_s'], w['diffuse']['mrays_s']))"
done
