#!/usr/bin/env python
"""One workload, a few passes of the hot path on cuda:0 — the command the ncu captures under profiles/ are taken from.
usage: python tools/prof_step.py [--steps 2] [--genome 5000000] [--coverage 100] [-k 25] [-m 4] [--error 0.0]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from tatajuba_b200 import Context, TatajubaOptions, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--genome", type=int, default=5_000_000)
ap.add_argument("--coverage", type=int, default=100)
ap.add_argument("-k", type=int, default=25)
ap.add_argument("-m", type=int, default=4)
ap.add_argument("--error", type=float, default=0.0)
ap.add_argument("--stats", action="store_true", help="per-kernel CUDA-event times of the last step")
a = ap.parse_args()
buf = synth.sample_buffer(a.genome, a.coverage, seed=1, error_rate=a.error)
d = torch.from_numpy(buf).cuda()
ctx = Context(0)
opt = TatajubaOptions(a.k, a.m, 5, True)
for i in range(a.steps):
    if a.stats and i == a.steps - 1:
        ctx.set_profiling(2)
        ctx.reset_kernel_stats()
    hc = ctx.count_device(d.data_ptr(), d.numel(), opt)
if a.stats:
    for s in ctx.kernel_stats():
        print(f"  {s['name']:24s} launches={s['launches']:4d} ms={s['ms_total']:8.3f}")
print(f"n_raw={hc.n_raw} n_distinct={hc.n_distinct} n_elem={hc.n_elem} n_idx={hc.n_idx} coverage={hc.coverage} ms_device={hc.ms_device:.3f}")
