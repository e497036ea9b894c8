#!/bin/bash
# Trace rates by the tree they run on: host SAH builder (Setting gpuBuild = 0), Morton radix tree (1), agglomerative clustering (2);
# tools/ab_sahb.sh has the SAH builders (3, 4).
# C2 scene: the 1080p waves and the 4K frame of bench.py; C3 soup (10 M triangles): the 1080p frame of tools/bench_configs.py.
for g in 0 1 2 0 2; do
  python bench.py --no-cpu-baseline --no-e2e --no-check --no-refit --steps 5 --warmup 2 --wave-steps 30 --setting gpuBuild=$g 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); w=d['waves']['rays']
print('C2 gpuBuild=$g: frame %.2f ms (%.0f Mrays/s); waves primary %.0f shadow %.0f diffuse %.0f Mrays/s'%(d['ms_per_step'], d['value'], w['primary']['mrays_s'], w['shadow']['mrays_s'], w['diffuse']['mrays_s']))"
done
for g in 0 1 2; do
  python tools/bench_configs.py c3 --frames 6 gpuBuild=$g 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); m=d['mrays_s']
print('C3 gpuBuild=$g: frame %.3f ms; primary %.0f bounce1 %.0f deep %.0f shadow %.0f Mrays/s; nodes %s'%(d['render_ms'], m['primary'], m['bounce1'], m['deep'], m['shadow'], d.get('wide_nodes')))"
done
