#!/bin/bash
# C3 (10 M-triangle soup, 581 MB scene: not L2-resident): ray / hit streams with the evict-first policy (B200_STREAM_RAYS, the default) vs plain
# loads (variants/nostream), and the access-policy window over the node pool (Setting l2Persist).  Needs B200_BVH_CACHE to avoid four 5 s builds.
export B200_BVH_CACHE=$PWD/gpurun_out/bvhcache; mkdir -p $B200_BVH_CACHE
run() { python tools/bench_configs.py c3 --frames 8 $2 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); m=d['mrays_s']
print('$1 $2: C3 frame %.3f ms (waves serial %.3f) primary %.0f b1 %.0f deep %.0f shadow %.0f'%(d['render_ms'], d['render_ms_waves_serial'], m['primary'], m['bounce1'], m['deep'], m['shadow']))"; }
unset B200_LIB; run stream ""
export B200_LIB=$PWD/variants/nostream/libRenderCore_B200.so; run nostream ""
unset B200_LIB; run stream l2Persist=0.5; run stream l2Persist=1; run stream ""
