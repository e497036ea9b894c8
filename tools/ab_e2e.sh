for v in base ws6 ws2 base ws6; do
  if [ "$v" = base ]; then unset B200_LIB; else export B200_LIB=$PWD/variants/$v/libRenderCore_B200.so; fi
  bash tools/sweep_e2e.sh base | sed "s/^base/$v/"
done
