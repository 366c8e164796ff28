#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --legs none --no-cpu-baseline --steps 10"
( timeout 600 $B ) > gpurun_out/v_default.json 2> gpurun_out/v_default.err
( TJB_LIB=$PWD/tatajuba_b200/lib/libtjb_gpu_chunk4.so timeout 600 $B ) > gpurun_out/v_chunk4.json 2> gpurun_out/v_chunk4.err
( TJB_LIB=$PWD/tatajuba_b200/lib/libtjb_gpu_chunk8.so timeout 600 $B ) > gpurun_out/v_chunk8.json 2> gpurun_out/v_chunk8.err
( TJB_TABLE_FACTOR=8 timeout 600 $B ) > gpurun_out/v_tab8.json 2> gpurun_out/v_tab8.err
( TJB_TABLE_FACTOR=16 timeout 600 $B ) > gpurun_out/v_tab16.json 2> gpurun_out/v_tab16.err
( timeout 900 python bench.py --legs c4 --no-cpu-baseline --steps 3 ) > gpurun_out/v_c4.json 2> gpurun_out/v_c4.err
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
( timeout 900 ncu --set full --clock-control none --import-source on -k regex:"aggregate_kernel" --launch-skip 2 --launch-count 1 -o gpurun_out/r02c_agg -f python /tmp/one_sample.py ) > gpurun_out/ncu_agg2.log 2>&1
ls gpurun_out/*.json | head -30
