#!/bin/bash
# is the per-step time of the headline a fixed cost per timed region?  K = 10 and K = 100 on N GPUs, host wall per step
N=${1:-1}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
[ "$N" = 1 ] && TR=python
for K in 10 100; do
  ( BENCH_DEBUG=1 timeout 900 $TR bench.py --gpus $N --steps $K --warmup 3 --legs none --no-cpu-baseline ) > gpurun_out/steps_${N}_$K.json 2> gpurun_out/steps_${N}_$K.err
  grep "\[bench\]" gpurun_out/steps_${N}_$K.err | cut -c1-400
  python tools/show_bench.py gpurun_out/steps_${N}_$K.json | head -1
done
