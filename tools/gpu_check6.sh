#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --legs c5 --no-cpu-baseline --steps 6"
( timeout 600 $B ) > gpurun_out/w_default.json 2> gpurun_out/w_default.err
for v in hashold red64 agg1 agg2; do ( TJB_LIB=$PWD/tatajuba_b200/lib/libtjb_gpu_$v.so timeout 600 $B ) > gpurun_out/w_$v.json 2> gpurun_out/w_$v.err; done
ls gpurun_out/w_*.json
