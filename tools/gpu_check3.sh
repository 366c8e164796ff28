#!/bin/bash
mkdir -p gpurun_out
( timeout 2400 python -m pytest tests -m gpu -x -q --durations=8 ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
( timeout 1200 python bench.py ) > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "rc=$?" >> gpurun_out/bench_full.err
for bs in 5 6 8; do ( TJB_BUCKET_SHIFT=$bs timeout 600 python bench.py --legs none --no-cpu-baseline --steps 5 ) > gpurun_out/bench_bs$bs.json 2> gpurun_out/bench_bs$bs.err; done
( timeout 600 python bench.py --legs none --no-cpu-baseline --contexts 1 ) > gpurun_out/bench_ctx1.json 2> gpurun_out/bench_ctx1.err
( timeout 600 python bench.py --legs none --no-cpu-baseline --contexts 3 ) > gpurun_out/bench_ctx3.json 2> gpurun_out/bench_ctx3.err
tail -n 4 gpurun_out/pytest_gpu.log; tail -n 3 gpurun_out/bench_full.err
