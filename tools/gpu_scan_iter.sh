#!/bin/bash
# one scan-kernel iteration on a GPU box: parity tests of the product library, per-kernel times of every variant
# library, then one `ncu --set full` capture of the scan (source page for tools/ncu_lines.py)
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --durations=3 ) > gpurun_out/iter_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/iter_pytest.log
tail -n 5 gpurun_out/iter_pytest.log
bash tools/gpu_variants.sh > /dev/null 2>&1
cat gpurun_out/variants.log
if [ "$1" = ncu ]; then
  ( timeout 600 ncu --set full --clock-control none --import-source on -k regex:tjb_scan_kernel --launch-skip 2 --launch-count 1 \
      -o gpurun_out/iter_scan -f python tools/prof_step.py --steps 3 ) > gpurun_out/iter_ncu.log 2>&1
  ( timeout 600 ncu --set full --clock-control none --import-source on -k regex:aggregate_kernel --launch-skip 2 --launch-count 1 \
      -o gpurun_out/iter_agg -f python tools/prof_step.py --steps 3 ) >> gpurun_out/iter_ncu.log 2>&1
  ls -la gpurun_out/*.ncu-rep
fi
if [ "$2" = pack ]; then ( timeout 300 python tools/pack_bench.py 1 2 4 8 12 16 24 31 ) > gpurun_out/pack_bench.log 2>&1; nproc >> gpurun_out/pack_bench.log; lscpu | grep -E "Model name|^CPU\(s\)|Thread|Socket|NUMA" >> gpurun_out/pack_bench.log; cat gpurun_out/pack_bench.log; fi
