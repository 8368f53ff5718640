#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/ring_ab.txt
for v in t128 t256 t1024 t1024s3 main; do
  for g in 0 1; do
    if [ $v = main ]; then unset NB_LIB_PATH; else export NB_LIB_PATH=$PWD/tools/ubench/variants/$v.so; fi
    echo "== $v generic=$g" >> gpurun_out/ring_ab.txt
    NB_FORCE_GENERIC=$g timeout 300 python tools/tune.py 1048576 2 4,64,4 >> gpurun_out/ring_ab.txt 2>&1
  done
done
cat gpurun_out/ring_ab.txt
