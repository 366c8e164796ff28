#!/bin/bash
mkdir -p gpurun_out
cat > /tmp/one_sample.py <<'PY'
import sys, torch
sys.path.insert(0, '.')
from tatajuba_b200 import Context, TatajubaOptions, synth
buf = synth.sample_buffer(5_000_000, 100, seed=1)
d = torch.from_numpy(buf).cuda()
ctx = Context(0); opt = TatajubaOptions()
ctx.set_result_mode(True)
for i in range(3):
    hc = ctx.count_device(d.data_ptr(), d.numel(), opt)
print(hc.n_raw, hc.n_elem)
PY
NCU="ncu --set full --clock-control none --import-source on"
( TJB_TRACE=1 timeout 300 python /tmp/one_sample.py ) > gpurun_out/trace2.log 2>&1
( timeout 900 $NCU -k regex:"aggregate_kernel|wbucket_reduce_kernel" --launch-skip 4 --launch-count 2 -o gpurun_out/r02b_agg -f python /tmp/one_sample.py ) > gpurun_out/ncu_agg.log 2>&1
( timeout 900 $NCU -k regex:tjb_scan_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/r02b_scan -f python /tmp/one_sample.py ) > gpurun_out/ncu_scan.log 2>&1
ls -la gpurun_out/*.ncu-rep
