#!/bin/bash
# tools/build_variant.sh NAME "-DB200_REFILL_BELOW=16 ..."  -> build/variants/NAME/libRenderCore_B200.so  (run with B200_LIB=<that path>)
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/variants/$1
mkdir -p $OUT
make -s -j8 -C $ROOT/lighthouse2_b200/csrc OUT=$OUT DEFS="$2"
rm -f $OUT/*.o
echo $OUT/libRenderCore_B200.so
