#!/bin/bash
# second look: GPU tests on the new build, bench with the two candidate-expansion variants, host-side stage times
mkdir -p gpurun_out
( timeout 1800 python -m pytest tests -m gpu -x -q --durations=8 ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
( timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline ) > gpurun_out/bench_direct.json 2> gpurun_out/bench_direct.err
( TJB_LIB=$PWD/tatajuba_b200/lib/libtjb_gpu_queue.so timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline ) > gpurun_out/bench_queue.json 2> gpurun_out/bench_queue.err
( TJB_RESULT_PATH=dma timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline ) > gpurun_out/bench_dma.json 2> gpurun_out/bench_dma.err
for pe in 0 2 4; do ( TJB_PLAIN_EVERY=$pe timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline ) > gpurun_out/bench_plain$pe.json 2> gpurun_out/bench_plain$pe.err; done
( TJB_TRACE=1 timeout 600 python - <<'PY'
import torch, sys
sys.path.insert(0, '.')
from tatajuba_b200 import Context, TatajubaOptions, synth
import time
buf = synth.sample_buffer(5_000_000, 100, seed=1)
d = torch.from_numpy(buf).cuda()
ctx = Context(0); opt = TatajubaOptions()
for mode in (False, True):
    ctx.set_result_mode(mode)
    for i in range(4):
        t0 = time.perf_counter(); hc = ctx.count_device(d.data_ptr(), d.numel(), opt); dt = time.perf_counter() - t0
        sys.stderr.write(f"[py] arena={mode} call {i}: {dt*1e6:.0f} us wall, n_elem={hc.n_elem}\n")
PY
) > gpurun_out/trace.log 2>&1
tail -n 5 gpurun_out/pytest_gpu.log
