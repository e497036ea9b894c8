#!/bin/bash
# Work donation at the end of a wave (traverse.cu, B200_STEAL): the 1080p waves of bench.py, the 4K frame, and the C2 1080p frame, with (default
# build) and without (make -C lighthouse2_b200/csrc OUT=$PWD/variants/nosteal DEFS=-DB200_STEAL=0).
for v in steal nosteal steal nosteal; do
  if [ $v = nosteal ]; then export B200_LIB=$PWD/variants/nosteal/libRenderCore_B200.so; else unset B200_LIB; fi
  python bench.py --no-cpu-baseline --no-e2e --no-check --no-refit --steps 3 --warmup 1 --wave-steps 30 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); w=d['waves']['rays']
print('$v: 4K frame %.2f ms; waves primary %.0f shadow %.0f diffuse %.0f Mrays/s'%(d['ms_per_step'], w['primary']['mrays_s'], w['shadow']['mrays_s'], w['diffuse']['mrays_s']))"
  python tools/frame_sweep.py --chains 0,1,2 --frames 30 2>/dev/null | grep "\"frameGraph\": 1, \"waveTimers\": 0" | cut -c1-220
done
