mkdir -p gpurun_out
for v in "" _raw3; do
  lib=$PWD/tatajuba_b200/lib/libtjb_gpu$v.so
  TJB_LIB=$lib timeout 600 ncu --set full --clock-control none -k regex:tjb_scan_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/c5_scan$v -f python tools/prof_step.py --steps 3 -k 32 -m 3 --coverage 300 > gpurun_out/ncu_c5$v.log 2>&1
done
ls -la gpurun_out/c5_scan*.ncu-rep
