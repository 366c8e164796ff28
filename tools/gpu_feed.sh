#!/bin/bash
# file feed on the GPU box: decoder rates and tjb_count_sample_files on plain / .gz paired-end samples, then the files leg
mkdir -p gpurun_out
( timeout 900 python tools/feed_bench.py ) > gpurun_out/feed_bench.log 2>&1
( timeout 900 python bench.py --legs files --steps 5 --warmup 3 ) > gpurun_out/bench_files.json 2> gpurun_out/bench_files.err
cat gpurun_out/feed_bench.log; python tools/show_bench.py gpurun_out/bench_files.json | tail -3
