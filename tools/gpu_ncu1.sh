#!/bin/bash
# ncu captures of the scan (both candidate-expansion variants), bucket and pack kernels on the config-2 sample
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
NCU="ncu --set full --clock-control none --import-source on --launch-skip-before-match 0"
( timeout 900 $NCU -k regex:tjb_scan_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/r02a_scan_direct -f python /tmp/one_sample.py ) > gpurun_out/ncu_scan_direct.log 2>&1
( TJB_LIB=$PWD/tatajuba_b200/lib/libtjb_gpu_queue.so timeout 900 $NCU -k regex:tjb_scan_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/r02a_scan_queue -f python /tmp/one_sample.py ) > gpurun_out/ncu_scan_queue.log 2>&1
( timeout 900 $NCU -k regex:"wbucket_reduce_kernel|pack_kernel|wpartition_kernel" --launch-skip 6 --launch-count 3 -o gpurun_out/r02a_reduce -f python /tmp/one_sample.py ) > gpurun_out/ncu_reduce.log 2>&1
ls -la gpurun_out/*.ncu-rep
