#!/bin/bash
# multi-GPU look (gpurun --gpus N -- bash tools/gpu_multi.sh N): the GPU tests (the 2-GPU sharded test and the
# multi-device command-line test need >= 2 GPUs), then both bench arms under torchrun like the driver launches them
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/gpus_$N.txt 2>&1; nproc >> gpurun_out/gpus_$N.txt
if [ "$2" != "notests" ]; then
( timeout 2400 python -m pytest tests -m gpu -x -q --durations=5 ) > gpurun_out/pytest_gpu_$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$N.log
fi
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
( timeout 1500 $TR bench.py --gpus $N --steps 10 --warmup 3 ) > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "rc=$?" >> gpurun_out/bench_${N}gpu.err
( timeout 1500 $TR bench.py --impl reference --gpus $N --steps 3 --warmup 1 ) > gpurun_out/bench_${N}gpu_ref.json 2> gpurun_out/bench_${N}gpu_ref.err
tail -n 3 gpurun_out/pytest_gpu_$N.log 2>/dev/null; tail -n 3 gpurun_out/bench_${N}gpu.err
