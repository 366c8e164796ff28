#!/usr/bin/env python
"""tjb_merge_tracts alone on the shape of BASELINE configs[3] (96 samples x ~19.4 K tracts of 21 K sites), by thread
count (TJB_MERGE_THREADS).  No GPU work.  usage: python tools/merge_bench.py [threads ...]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from tatajuba_b200 import genome_set  # noqa: E402

rng = np.random.default_rng(0)
tables = []
for n in [19400] * 96:
    loc = (np.sort(rng.choice(21000, size=n, replace=False)) * 8).astype(np.int32)
    tables.append({"location": loc, "base": np.zeros(n, np.int32), "mode_context_length": np.full(n, 4, np.int32),
                   "name": np.zeros((n, 1), np.uint8)})
ref = None
for th in [int(x) for x in sys.argv[1:]] or [1, 2, 4, 8, 16]:
    os.environ["TJB_MERGE_THREADS"] = str(th)
    best = 1e9
    for _ in range(5):
        t0 = time.perf_counter()
        r = genome_set.merge_tracts(tables, 2)
        best = min(best, time.perf_counter() - t0)
    if ref is None:
        ref = r
    assert all(np.array_equal(a, b) for a, b in zip(r, ref))
    print(f"{th:3d} threads: {best * 1e3:7.1f} ms for {len(r[0])} rows, {int(r[2][-1]) + 1} tracts")
