"""tjb_tract_rows_noalloc (include/tjb_gpu.h): the scan + count of a small batch of reads with all scratch in one
static device array — no cudaMalloc inside — against the existing entry points tjb_sample_begin / _scan_device /
_export on the same reads.  Integer work: bit-exact.  Two calls in a row on different batches (the shared scratch keeps
the last call's contents and must be cleared by the call itself), then the first batch again.
Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

from tatajuba_b200 import _lib, synth

TJB_ERR_NOALLOC_LIMIT = -6


def table_of(rows):
    """rows {key words..., count word} in any order, keys possibly repeated -> sorted unique keys with summed counts"""
    rows = np.ascontiguousarray(rows)
    keys, inv = np.unique(rows[:, :-1], axis=0, return_inverse=True)
    cnt = np.zeros(len(keys), np.uint64)
    np.add.at(cnt, inv.reshape(-1), rows[:, -1])
    return keys, cnt


def existing_rows(ctx, opt, d):
    import torch
    ctx.sample_begin(opt, d.numel())
    ctx.sample_scan_device(d.data_ptr(), d.numel())
    n, words = ctx.sample_export()
    out = torch.empty((max(n, 1), words), dtype=torch.int64, device=d.device)
    ctx.sample_export(out.data_ptr(), n)  # (the next tjb_sample_begin of the context starts a new sample)
    return out[:n].cpu().numpy().view(np.uint64), words


def noalloc_rows(lib, opt, d, words, max_rows=1 << 18):
    import torch
    rows = torch.full((max_rows, words), -1, dtype=torch.int64, device=d.device)
    status = torch.full((3,), -7, dtype=torch.int64, device=d.device)
    o = opt.c_struct()
    rc = lib.tjb_tract_rows_noalloc(d.device.index, C.c_void_p(d.data_ptr()), d.numel(), C.byref(o), C.c_void_p(rows.data_ptr()), max_rows,
                                    C.c_void_p(status.data_ptr()), C.c_void_p(torch.cuda.current_stream(d.device).cuda_stream))
    assert rc == 0, lib.tjb_strerror(rc)
    st = status.cpu().numpy()  # the copy waits for the stream the call was enqueued on
    return rows[: max(int(st[0]), 0)].cpu().numpy().view(np.uint64), st


@pytest.mark.gpu
@pytest.mark.parametrize("k,m", [(25, 4), (32, 3)])  # two and three key words per row
def test_noalloc_rows_equal_the_existing_entry_points(k, m):
    import torch
    from tatajuba_b200 import Context, TatajubaOptions
    lib = _lib.load()
    opt = TatajubaOptions(k, m, 2, True)
    ctx = Context(0)
    batches = [synth.sample_buffer(60_000, 20, seed=3, error_rate=0.002), synth.sample_buffer(9_000, 15, seed=4),
               np.frombuffer(b"\n".join(synth.nrun_reads(k, seed=5, n_reads=300)) + b"\n", np.uint8).copy()]
    order = [0, 1, 2, 0]  # two different calls in a row, a batch with N runs, and the first one again
    for i in order:
        d = torch.from_numpy(np.ascontiguousarray(batches[i])).cuda()
        want, words = existing_rows(ctx, opt, d)
        got, st = noalloc_rows(lib, opt, d, words)
        assert st[0] == len(got) and st[0] > 0, st
        wk, wc = table_of(want)
        gk, gc = table_of(got)
        assert int(st[1]) == int((wc & np.uint64(0xffffffff)).sum() + (wc >> np.uint64(32)).sum())  # every observation counted once
        assert np.array_equal(gk, wk), f"batch {i}: keys differ"
        assert np.array_equal(gc, wc), f"batch {i}: counts differ"


@pytest.mark.gpu
def test_noalloc_limits_are_reported():
    import torch
    from tatajuba_b200 import TatajubaOptions
    lib = _lib.load()
    opt = TatajubaOptions(25, 4, 2, True)
    o = opt.c_struct()
    limit = int(lib.tjb_noalloc_max_bytes())
    assert limit == 4 << 20
    d = torch.zeros(limit + 1, dtype=torch.uint8, device="cuda")
    rows = torch.zeros((16, 3), dtype=torch.int64, device="cuda")
    status = torch.zeros(3, dtype=torch.int64, device="cuda")
    args = (C.byref(o), C.c_void_p(rows.data_ptr()), 16, C.c_void_p(status.data_ptr()), None)
    rc = lib.tjb_tract_rows_noalloc(0, C.c_void_p(d.data_ptr()), limit + 1, *args)
    assert rc == TJB_ERR_NOALLOC_LIMIT and b"4 MiB" in lib.tjb_strerror(rc)
    # a batch that fits the limit but not the caller's 16 rows: status -1, and nothing written past the 16 rows
    buf = synth.sample_buffer(30_000, 20, seed=8)
    d = torch.from_numpy(buf).cuda()
    guard = torch.full((32, 3), 5, dtype=torch.int64, device="cuda")
    rc = lib.tjb_tract_rows_noalloc(0, C.c_void_p(d.data_ptr()), d.numel(), C.byref(o), C.c_void_p(guard.data_ptr()), 16, C.c_void_p(status.data_ptr()), None)
    assert rc == 0
    torch.cuda.synchronize()
    st = status.cpu().numpy()
    assert st[0] == -1 and st[1] > 16
    assert bool((guard[16:] == 5).all())
    # an empty batch: zero rows
    rc = lib.tjb_tract_rows_noalloc(0, None, 0, *args)
    assert rc == 0
    torch.cuda.synchronize()
    assert status.cpu().numpy().tolist() == [0, 0, 0]
