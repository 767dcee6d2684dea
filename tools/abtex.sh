#!/bin/bash
# developer A/B of the texture-unit kernels: tools/abtex.sh tag lib1 lib2 ...  (results in gpurun_out/<tag>_<lib>_{l,c}.json)
tag=$1; shift
for lib in "$@"; do
  VR_B200_LIB=exp_libs/libvr_$lib.so python bench.py --texture-unit 1 --no-cpu-baseline --no-upload-probe --no-gather-probe --steps 16 > gpurun_out/${tag}_${lib}_l.json 2>/dev/null
  VR_B200_LIB=exp_libs/libvr_$lib.so python bench.py --texture-unit 1 --no-cpu-baseline --no-upload-probe --no-gather-probe --steps 16 --gradient compute --lights 1 > gpurun_out/${tag}_${lib}_c.json 2>/dev/null
done
