for v in base b8 b7 b5 base; do
  if [ "$v" = base ]; then unset B200_LIB; else export B200_LIB=$PWD/variants/$v/libRenderCore_B200.so; fi
  python tools/bench_configs.py c3 --frames 5 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); m=d['mrays_s']
print('$v C3 frame %.3f ms primary %.0f b1 %.0f deep %.0f shadow %.0f'%(d['render_ms'], m['primary'], m['bounce1'], m['deep'], m['shadow']))"
done
