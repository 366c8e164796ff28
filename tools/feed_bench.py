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

# the decoder alone (no GPU work): one file through tjb_reader_fill with our inflate and with zlib, and the same reads
# with realistic (random) quality strings, which deflate codes as literals instead of long matches
import ctypes as C  # noqa: E402
import gzip  # noqa: E402

import numpy as np  # noqa: E402

from tatajuba_b200 import _lib  # noqa: E402

lib = _lib.load()


def reader_rate(path):
    r = lib.tjb_reader_open(os.fsencode(path))
    cap = 16 << 20
    buf = np.empty(cap, np.uint8)
    nr, nb = C.c_int64(0), C.c_int64(0)
    t0 = time.perf_counter()
    while lib.tjb_reader_fill(r, buf.ctypes.data, cap, C.byref(nr), C.byref(nb)) > 0:
        pass
    dt = time.perf_counter() - t0
    lib.tjb_reader_close(r)
    return nb.value / dt / 1e9


with tempfile.TemporaryDirectory() as d:
    half = rows[0::2][:400_000]
    const_q = os.path.join(d, "const.fastq.gz")
    synth.write_fastq(const_q, half)
    rng = np.random.default_rng(1)
    rand_q = os.path.join(d, "randq.fastq.gz")
    L = half.shape[1] - 1
    qual = rng.choice(np.frombuffer(b"FFFFFFFF:,#", np.uint8), size=(half.shape[0], L))
    with gzip.open(rand_q, "wb", compresslevel=1) as f:
        parts = []
        for i in range(half.shape[0]):
            parts.append(b"@r%d\n" % i + half[i, :L].tobytes() + b"\n+\n" + qual[i].tobytes() + b"\n")
        f.write(b"".join(parts))
    for name, p in (("constant quality", const_q), ("mixed quality", rand_q)):
        out = {}
        for z in ("0", "1"):
            os.environ["TJB_ZLIB"] = z
            out[z] = max(reader_rate(p) for _ in range(3))
        print(f"decoder alone, one .gz file, {name:17s}: own inflate {out['0']:6.3f} Gbases/s   zlib {out['1']:6.3f} Gbases/s")
    os.environ.pop("TJB_ZLIB", None)
