"""Host-side 2-bit packing of sequence chunks (csrc/tjb_hostpack.cpp): packed codes + exceptions must rebuild the input
byte for byte, for the scalar restatement, the AVX2 loop and any number of packer threads."""
import ctypes as C

import numpy as np
import pytest

from tatajuba_b200 import _lib

LETTERS = np.frombuffer(b"ACTG", np.uint8)  # by gray code (byte >> 1) & 3


def unpack(packed, pos, val, n):
    codes = (packed[:, None] >> (2 * np.arange(4, dtype=np.uint8))) & 3
    out = LETTERS[codes.reshape(-1)][:n].copy()
    out[pos] = val
    return out


def run(src, threads, cap=None):
    lib = _lib.load()
    fn = lib.tjb_hostpack_selftest
    fn.restype = C.c_long
    fn.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
    n = len(src)
    cap = n + 8 if cap is None else cap
    packed = np.zeros((n + 3) // 4 + 8, np.uint8)
    pos = np.zeros(cap + 8, np.uint32)
    val = np.zeros(cap + 8, np.uint8)
    ne = fn(src.ctypes.data, n, threads, packed.ctypes.data, pos.ctypes.data, val.ctypes.data, cap)
    return ne, packed[: (n + 3) // 4], pos, val


def reads(rng, n_reads, read_len=150, p_other=0.0):
    rows = rng.integers(0, 4, size=(n_reads, read_len + 1))
    b = np.frombuffer(b"ACGT", np.uint8)[rows]
    b[:, -1] = ord("\n")
    if p_other:
        m = rng.random(b.shape) < p_other
        b[m] = rng.integers(0, 256, size=int(m.sum()), dtype=np.uint8)
    return b.reshape(-1).copy()


@pytest.mark.parametrize("threads", [0, 1, 3, 8])
@pytest.mark.parametrize("n_reads,p_other", [(1, 0.0), (7, 0.0), (2000, 0.0), (2000, 0.02), (333, 0.5), (9000, 0.001)])
def test_pack_round_trip(threads, n_reads, p_other):
    rng = np.random.default_rng(n_reads * 31 + threads)
    src = reads(rng, n_reads, p_other=p_other)
    for cut in (0, 1, 2, 3, 33):  # every length residue, and sub-ranges that do not end on 32 bytes
        s = src[: len(src) - cut] if cut < len(src) else src
        ne, packed, pos, val = run(s, threads)
        assert ne == int((~np.isin(s, np.frombuffer(b"ACGT", np.uint8))).sum())
        assert np.array_equal(unpack(packed, pos[:ne], val[:ne], len(s)), s)
        assert len(np.unique(pos[:ne])) == ne  # every exception once (blocks are appended in any order)


def test_pack_reports_exception_overflow():
    src = np.full(4096, ord("N"), np.uint8)
    for threads in (0, 4):
        ne, *_ = run(src, threads, cap=100)
        assert ne == -1


def test_pack_empty():
    ne, packed, _, _ = run(np.zeros(0, np.uint8), 2)
    assert ne == 0 and len(packed) == 0


@pytest.mark.parametrize("n_rows", [0, 1, 2, 3, 8, 1001])
@pytest.mark.parametrize("shift", [0, 1])  # destination starting on an even / odd element (16-byte aligned or not)
def test_expand_rows(n_rows, shift):
    """24 varying bytes per result element -> the reference's 40-byte hopo_element with its constant tail of -1s"""
    lib = _lib.load()
    fn = lib.tjb_hostpack_expand_hook
    fn.restype = None
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    rng = np.random.default_rng(n_rows)
    src = rng.integers(0, 2**63, size=(n_rows + shift, 3), dtype=np.int64)
    dst = np.full((n_rows + shift + 1, 5), 7, np.int64)
    assert dst.ctypes.data % 16 == 0
    fn(dst[shift:].ctypes.data, src[shift:].ctypes.data, n_rows)
    assert np.array_equal(dst[shift:shift + n_rows, :3], src[shift:])
    assert np.all(dst[shift:shift + n_rows, 3:] == -1)
    assert np.all(dst[:shift] == 7) and np.all(dst[shift + n_rows:] == 7)  # nothing outside the range is written
