#!/bin/bash
# One-GPU round check (gpurun -- bash tools/gpu_round.sh [tests] [bench] [launches] [ncu] [sanitize]): the GPU tests, both bench arms,
# the ncu launch list of the bench command and one `ncu --set full` capture of every kernel of the step.  Each ncu pass
# runs only after its own command has exited 0 without ncu; nothing printed under ncu is a bench value.
what=${*:-tests bench launches ncu}
mkdir -p gpurun_out
has () { case " $what " in *" $1 "*) return 0;; esac; return 1; }
if has tests; then
  ( timeout 2400 python -m pytest tests -m gpu -x -q --durations=5 ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
  tail -n 4 gpurun_out/pytest_gpu.log
fi
if has bench; then
  ( timeout 1200 python bench.py ) > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "rc=$?" >> gpurun_out/bench_full.err
  ( timeout 900 python bench.py --impl reference --steps 3 --warmup 1 ) > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
  tail -n 3 gpurun_out/bench_full.err
fi
if has launches; then
  if timeout 600 python bench.py --steps 2 --warmup 3 --legs none --no-cpu-baseline > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; then
    ( timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
        python bench.py --steps 2 --warmup 3 --legs none --no-cpu-baseline ) > gpurun_out/ncu_launches.log 2>&1
  fi
fi
if has ncu; then
  cat > /tmp/one_sample.py <<'PY'
import sys, torch
sys.path.insert(0, '.')
from tatajuba_b200 import Context, TatajubaOptions, synth
buf = synth.sample_buffer(5_000_000, 100, seed=1)   # BASELINE configs[1]: 5 Mbp x 100x, 150 bp reads
d = torch.from_numpy(buf).cuda()
ctx = Context(0); opt = TatajubaOptions()
ctx.set_result_mode(True)
for i in range(3):
    hc = ctx.count_device(d.data_ptr(), d.numel(), opt)
print(hc.n_raw, hc.n_elem)
PY
  NCU="ncu --set full --clock-control none --import-source on"
  if timeout 300 python /tmp/one_sample.py > gpurun_out/one_sample.log 2>&1; then
    # third sample of the context: sizes settled.  One capture per kernel of the step
    ( timeout 900 $NCU -k regex:tjb_scan_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/r02b_scan -f python /tmp/one_sample.py ) > gpurun_out/ncu_scan.log 2>&1
    ( timeout 900 $NCU -k regex:"aggregate_kernel|wpartition_kernel|wbucket_reduce_kernel|pack_kernel|plan_kernel|totals_kernel" --launch-skip 12 --launch-count 6 -o gpurun_out/r02b_reduce -f python /tmp/one_sample.py ) > gpurun_out/ncu_reduce.log 2>&1
  fi
  ls -la gpurun_out/*.ncu-rep
fi
if has sanitize; then
  # (compute-sanitizer is closed on this pool: the step is not part of the default list)
  # memcheck and racecheck over a small sample (every kernel of the step, the warp-private shared-memory protocols of the scan)
  for tool in memcheck racecheck; do
    ( timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/prof_step.py --genome 300000 --coverage 30 --steps 2 ) > gpurun_out/sanitizer_$tool.log 2>&1
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|n_raw" gpurun_out/sanitizer_$tool.log
  done
fi
