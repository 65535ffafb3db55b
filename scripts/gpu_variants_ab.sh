#!/bin/bash
# every prebuilt variants/*.so x every line of the settings file $1 (see gpu_ab.sh)
mkdir -p gpurun_out
cp salz_b200/libsalz_cuda.so /tmp/libsalz_cuda.orig.so
for v in variants/*.so; do
  cp $v salz_b200/libsalz_cuda.so
  echo "##### $v"
  bash scripts/gpu_ab.sh "$@" 
  cp gpurun_out/ab.txt gpurun_out/ab_$(basename $v .so).txt
done 2>&1 | tee gpurun_out/variants_ab.txt
cp /tmp/libsalz_cuda.orig.so salz_b200/libsalz_cuda.so
