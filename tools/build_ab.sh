#!/bin/bash
# tools/build_ab.sh NAME "EXTRA nvcc flags"  ->  volume_renderer_b200/csrc/ab/libvr_b200_NAME.so  (A/B builds; use with VR_B200_LIB=...)
set -e
name=$1; extra=$2
root=$(cd "$(dirname "$0")/.." && pwd)
tmp=/tmp/ab_build_$name
rm -rf $tmp && mkdir -p $tmp/volume_renderer_b200 $tmp/include
cp -r $root/volume_renderer_b200/csrc $tmp/volume_renderer_b200/
cp $root/include/vr_abi.h $tmp/include/
rm -f $tmp/volume_renderer_b200/csrc/*.o $tmp/volume_renderer_b200/csrc/*.so
make -C $tmp/volume_renderer_b200/csrc -j8 EXTRA="$extra" > $tmp/build.log 2>&1 || (tail -20 $tmp/build.log; exit 1)
mkdir -p $root/volume_renderer_b200/csrc/ab
cp $tmp/volume_renderer_b200/csrc/libvr_b200.so $root/volume_renderer_b200/csrc/ab/libvr_b200_$name.so
grep -h -A2 "k_raymarchILi1ELi2ENS_7ElemF32ELi0ELi2E" $tmp/volume_renderer_b200/csrc/raymarch_lookup_a2.ptxas.log | grep -E "Used|spill" | head -2
echo "built ab/libvr_b200_$name.so"
