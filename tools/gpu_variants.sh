#!/bin/bash
# kernel-variant comparison on one box: the product library and every tatajuba_b200/lib/libtjb_gpu_<name>.so
# (make -C tatajuba_b200/csrc variant NAME=... EXTRA_NVFLAGS=...), per-kernel CUDA-event times at BASELINE configs[1]
# (default) or with the prof_step.py arguments given, e.g.  bash tools/gpu_variants.sh -k 32 -m 3 --coverage 300
mkdir -p gpurun_out
: > gpurun_out/variants.log
for lib in tatajuba_b200/lib/libtjb_gpu.so tatajuba_b200/lib/libtjb_gpu_*.so; do
  [ -f "$lib" ] || continue
  echo "== $lib $*" >> gpurun_out/variants.log
  for rep in 1 2; do
    ( TJB_LIB=$PWD/$lib timeout 300 python tools/prof_step.py --steps 6 --stats "$@" ) 2>&1 | grep -E "tjb_scan|aggregate|pack_kernel|n_raw" >> gpurun_out/variants.log
  done
done
cat gpurun_out/variants.log
