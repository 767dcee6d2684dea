#!/bin/bash
# developer A/B: build a variant of libvr_b200.so with extra nvcc flags into exp_libs/libvr_<name>.so (out of tree build)
# usage: tools/mkexp.sh name "-DVR_SOMETHING=1 ..."
set -e
name=$1; flags=$2
root=$(cd "$(dirname "$0")/.." && pwd)
tmp=$(mktemp -d /tmp/vrexp_XXXX)
mkdir -p $tmp/volume_renderer_b200 $tmp/include $root/exp_libs
cp -r $root/volume_renderer_b200/csrc $tmp/volume_renderer_b200/
cp $root/include/*.h $tmp/include/
make -C $tmp/volume_renderer_b200/csrc clean > /dev/null
make -C $tmp/volume_renderer_b200/csrc -j4 EXTRA="$flags" > $tmp/build.log 2>&1 || (tail -20 $tmp/build.log; false)
cp $tmp/volume_renderer_b200/csrc/libvr_b200.so $root/exp_libs/libvr_$name.so
grep -h "k_raymarch.*Li1ELi2ENS_7ElemF32ELi0E" -A2 $tmp/volume_renderer_b200/csrc/raymarch_lookup.ptxas.log | grep -h "spill\|Used" | head -2
rm -rf $tmp
