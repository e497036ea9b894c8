#!/bin/bash
for v in base pn pt pnt; do
  if [ "$v" = base ]; then unset B200_LIB; else export B200_LIB=$PWD/variants/$v/libRenderCore_B200.so; fi
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-frame 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); r=d['rays']
print('$v C2 value %.0f primary %.0f shadow %.0f diffuse %.0f'%(d['value'], r['primary']['mrays_s'], r['shadow']['mrays_s'], r['diffuse']['mrays_s']))"
  timeout 400 python tools/bench_configs.py c3 --frames 5 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); m=d['mrays_s']
print('$v C3 frame %.3f ms primary %.0f b1 %.0f deep %.0f shadow %.0f'%(d['render_ms'], m['primary'], m['bounce1'], m['deep'], m['shadow']))"
done
