#!/bin/bash
# usage: scratch/mkvariants.sh name1:"-DX=1" name2:"-DX=0 -DY=2" ...   -> scratch/variants/libvslam_<name>.so
set -e
cd /root/repo
python visual-slam_b200/build.py >/dev/null
mkdir -p scratch/variants
B=visual-slam_b200/build
for spec in "$@"; do
  name="${spec%%:*}"; flags="${spec#*:}"
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -fmad=false $flags -c visual-slam_b200/csrc/orb_level.cu -o scratch/variants/orb_level_$name.o
  objs=$(ls $B/*.o | grep -v orb_level.o)
  /usr/local/cuda/bin/nvcc -shared -o scratch/variants/libvslam_$name.so $objs scratch/variants/orb_level_$name.o -gencode arch=compute_100a,code=sm_100a
  echo "$name: $(cuobjdump -res-usage scratch/variants/libvslam_$name.so 2>/dev/null | grep -A1 vs_level_kernel | grep -o 'REG:[0-9]*\|STACK:[0-9]*' | tr '\n' ' ')"
done
