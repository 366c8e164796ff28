#!/usr/bin/env python
"""steps of the hot path on one GPU, host wall clock per step (device-resident reads, arena results): to see what
several processes on one box do to each other.  usage: python tools/step_loop.py <device> [steps] [trace]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

dev = int(sys.argv[1])
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 300
torch.cuda.set_device(dev)
from tatajuba_b200 import Context, TatajubaOptions, synth  # noqa: E402

buf = synth.sample_buffer(5_000_000, 100, seed=1)
d = torch.from_numpy(buf).cuda()
ctx = Context(dev)
opt = TatajubaOptions()
ctx.set_result_mode(True)
for _ in range(5):
    ctx.count_device(d.data_ptr(), d.numel(), opt)
start_at = float(os.environ.get("START_AT", "0"))
while time.time() < start_at:  # all processes start their loops together
    pass
ts = []
for i in range(steps):
    t0 = time.perf_counter()
    hc = ctx.count_device(d.data_ptr(), d.numel(), opt)
    ts.append(time.perf_counter() - t0)
ts.sort()
print(f"gpu {dev}: median {ts[len(ts) // 2] * 1e3:.3f} ms  p10 {ts[len(ts) // 10] * 1e3:.3f}  p90 {ts[len(ts) * 9 // 10] * 1e3:.3f}  mean {sum(ts) / len(ts) * 1e3:.3f}", flush=True)
if len(sys.argv) > 3:
    os.environ["TJB_TRACE"] = "1"
