#!/usr/bin/env python
"""Host packer alone: GB/s of 2-bit packing (csrc/tjb_hostpack.cpp) by thread count, on the bytes of BASELINE configs[1]
(no GPU work).  usage: python tools/pack_bench.py [threads ...]"""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from tatajuba_b200 import _lib, synth  # noqa: E402

lib = _lib.load()
lib.tjb_hostpack_bench.restype = C.c_double
lib.tjb_hostpack_bench.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_size_t]
buf = synth.sample_buffer(5_000_000, 100, seed=1)
t0 = time.perf_counter()
s = int(buf[: buf.size // 8 * 8].view(np.uint64).sum())
print(f"numpy sum over the bytes (one thread): {buf.size / (time.perf_counter() - t0) / 1e9:.1f} GB/s")
threads = [int(x) for x in sys.argv[1:]] or [1, 2, 4, 8, 12, 16, 22, 30]
for t in threads:
    sec = lib.tjb_hostpack_bench(buf.ctypes.data, buf.size, t, 5, 32 << 20)
    print(f"{t:3d} workers + caller: {buf.size / sec / 1e9:7.1f} GB/s   ({sec * 1e3:.2f} ms per sample)")
