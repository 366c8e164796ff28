#!/usr/bin/env python
"""End-to-end from FASTQ files (host decode + H2D + kernels): wall time of tjb_count_sample_files on a paired-end sample
written plain and .gz.  usage: python tools/feed_bench.py [--genome 2000000] [--coverage 60]"""
import argparse
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tatajuba_b200 import Context, TatajubaOptions, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--genome", type=int, default=2_000_000)
ap.add_argument("--coverage", type=int, default=60)
a = ap.parse_args()
rows = synth.paired_reads(synth.inject_tract_polymorphism(synth.random_genome(a.genome, 5)), a.coverage, seed=8)
ctx = Context(0)
opt = TatajubaOptions(paired_end=True)
with tempfile.TemporaryDirectory() as d:
    for ext in (".fastq", ".fastq.gz"):
        r1, r2 = os.path.join(d, "s_R1" + ext), os.path.join(d, "s_R2" + ext)
        synth.write_fastq(r1, rows[0::2]); synth.write_fastq(r2, rows[1::2])
        for rep in range(2):
            t0 = time.perf_counter()
            hc = ctx.count_files([r1, r2], opt)
            dt = time.perf_counter() - t0
        print(f"{ext:10s} {hc.n_bases / 1e6:8.1f} Mbases  wall {dt * 1e3:8.1f} ms  = {hc.n_bases / dt / 1e9:6.3f} Gbases/s   "
              f"decode (sum over the {2} decoder threads) {hc.ms_decode:8.1f} ms  n_elem={hc.n_elem}")
