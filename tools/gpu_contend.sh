#!/bin/bash
# what N processes, one per GPU, do to each other's steps (no torch.distributed involved): alone, then together
N=${1:-2}
mkdir -p gpurun_out
{
echo "== alone"; python tools/step_loop.py 0 300
echo "== $N together"
START_AT=$(python -c "import time; print(time.time() + 25)")
for g in $(seq 0 $((N-1))); do START_AT=$START_AT python tools/step_loop.py $g 300 & done; wait
echo "== $N together, TJB_TRACE on gpu 0 (last lines)"
START_AT=$(python -c "import time; print(time.time() + 25)")
for g in $(seq 1 $((N-1))); do START_AT=$START_AT python tools/step_loop.py $g 300 & done
START_AT=$START_AT TJB_TRACE=1 python tools/step_loop.py 0 300 2>&1 | tail -4; wait
} > gpurun_out/contend_$N.log 2>&1
cat gpurun_out/contend_$N.log
