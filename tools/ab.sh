#!/bin/bash
# developer A/B: time the default bench (lookup gradient, 2 lights) and the computed-gradient one with each library in exp_libs/
# usage: tools/ab.sh tag lib1 lib2 ...   (results in gpurun_out/<tag>_<lib>_{l,c}.json)
tag=$1; shift
for lib in "$@"; do
  VR_B200_LIB=exp_libs/libvr_$lib.so python bench.py --no-cpu-baseline --no-upload-probe --steps 12 > gpurun_out/${tag}_${lib}_l.json 2>/dev/null
  VR_B200_LIB=exp_libs/libvr_$lib.so python bench.py --no-cpu-baseline --no-upload-probe --steps 12 --gradient compute --lights 1 > gpurun_out/${tag}_${lib}_c.json 2>/dev/null
done
