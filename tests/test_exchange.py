"""Cross-sample exchange (SURVEY.md section 8 e): sample -> rank dealing, query rows, allgather of the per-sample
tables (gloo, world_size 2, CPU) and - on the GPU - the union of distinct mapper queries vs the numpy restatement."""
import glob
import os

import numpy as np
import pytest

from oracle import pyoracle
from tatajuba_b200 import genome_set

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def fake_table(sample, n_shared=200, n_own=50, k=25):
    """seeded rows: a block shared by every sample (same population) plus sample-specific rows"""
    def block(rng, n):
        t = np.empty((n, 3), np.uint64)
        mask = np.uint64((1 << (2 * k)) - 1) if k < 32 else np.uint64(0xFFFFFFFFFFFFFFFF)
        t[:, 0] = rng.integers(0, 1 << 63, n, dtype=np.uint64) * np.uint64(2) & mask
        t[:, 1] = rng.integers(0, 1 << 63, n, dtype=np.uint64) * np.uint64(2) + np.uint64(1) & mask
        t[:, 2] = rng.integers(0, 2, n).astype(np.uint64) | (rng.integers(2, 40, n).astype(np.uint64) << np.uint64(8))
        return t
    shared = block(np.random.default_rng(7), n_shared)
    own = block(np.random.default_rng(1000 + sample), n_own + sample)
    return np.concatenate([shared[sample % 3::2], own])


def test_samples_of_rank_partition():
    for n, w in ((96, 8), (5, 2), (3, 4), (1, 1)):
        seen = sorted(s for r in range(w) for s in genome_set.samples_of_rank(n, r, w))
        assert seen == list(range(n))


def test_query_rows_match_oracle(built):
    from tatajuba_b200 import HopoCounter, TatajubaOptions
    files = sorted(glob.glob(os.path.join(GOLD, "*.npz")))
    assert files
    for f in files:
        z = np.load(f)
        if "status" not in z.files or int(z["status"]) != 0:
            continue
        hc = HopoCounter(TatajubaOptions())
        hc.status, hc.elem = 0, np.ascontiguousarray(z["elem"])
        hc.idx_initial = np.ascontiguousarray(z["idx_initial"], np.int32)
        hc.n_elem, hc.n_idx = len(hc.elem), len(hc.idx_initial)
        want = pyoracle.query_rows_oracle(hc.elem, hc.idx_initial)
        got = genome_set.query_rows(hc)
        assert np.array_equal(got, want), f


def _worker(rank, world, n_samples, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        local = {s: fake_table(s) for s in genome_set.samples_of_rank(n_samples, rank, world)}
        tables = genome_set.allgather_tables(local, n_samples)
        ok = all(np.array_equal(tables[s], fake_table(s)) for s in range(n_samples))
        union, index = pyoracle.union_oracle(tables)  # every rank holds the same tables -> the same union
        q.put((rank, ok, len(union), int(sum(int(i.sum()) for i in index))))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_samples", [5, 2, 1])
def test_allgather_tables_gloo_world2(n_samples):
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, n_samples, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(r[1] for r in res), res
    assert res[0][2:] == res[1][2:]  # identical union on both ranks
    want_union, _ = pyoracle.union_oracle([fake_table(s) for s in range(n_samples)])
    assert res[0][2] == len(want_union)


def test_union_oracle_properties():
    tables = [fake_table(s) for s in range(4)]
    union, index = pyoracle.union_oracle(tables)
    assert len(np.unique(union, axis=0)) == len(union)
    for t, ix in zip(tables, index):
        assert np.array_equal(union[ix], t)


@pytest.mark.gpu
@pytest.mark.parametrize("k", [25, 32, 8])
def test_union_queries_device_vs_oracle(built, k):
    from tatajuba_b200 import Context
    ctx = Context(0)
    tables = [fake_table(s, 3000, 700, k) for s in range(6)] + [np.zeros((0, 3), np.uint64)]
    # negative length fields (the reference's signed 10-bit field wraps at 512) and duplicated rows inside a sample
    tables[1][:5, 2] = np.uint64(1) | (np.uint64(0x3f0) << np.uint64(8))
    tables[2] = np.concatenate([tables[2], tables[2][:17]])
    want_union, want_index = pyoracle.union_oracle(tables)
    union, index = genome_set.union_queries(ctx, tables, k)
    assert np.array_equal(union, want_union)
    for a, b in zip(index, want_index):
        assert np.array_equal(a, b)
    assert ctx.launch_count() > 0
    ctx.close()


@pytest.mark.gpu
def test_union_of_counted_samples(built, restatement):
    """samples of one population counted on the GPU -> query rows -> union; checked against the restatement's
    boundary state pushed through the numpy union"""
    from tatajuba_b200 import Context, TatajubaOptions, synth
    ctx = Context(0)
    opt = TatajubaOptions()
    tabs, want_tabs = [], []
    for s in range(3):
        buf = synth.sample_buffer(150_000, 25, seed=5, sample=s)
        hc = ctx.count_host(buf, opt)
        tabs.append(genome_set.query_rows(hc))
        w = restatement.run_buffer(buf, opt.kmer_size, opt.min_tract_size, opt.min_coverage, opt.remove_biased)
        want_tabs.append(pyoracle.query_rows_oracle(w.elem, w.idx_initial))
        assert np.array_equal(tabs[-1], want_tabs[-1])
    union, index = genome_set.union_queries(ctx, tabs, opt.kmer_size)
    want_union, want_index = pyoracle.union_oracle(want_tabs)
    assert np.array_equal(union, want_union)
    assert len(union) < sum(len(t) for t in tabs)  # the population shares contexts
    for a, b in zip(index, want_index):
        assert np.array_equal(a, b)
    ctx.close()


@pytest.mark.parametrize("n_samples,world", [(96, 2), (96, 8), (5, 4), (3, 8), (7, 1)])
def test_gather_plan_restores_the_global_sample_order(n_samples, world):
    """the index arithmetic of allgather_tables_device (N > 1 ranks), on a simulated receive buffer"""
    import torch
    rng = np.random.default_rng(n_samples * 31 + world)
    tabs = {s: torch.from_numpy(rng.integers(0, 1 << 40, (int(rng.integers(0, 50)), 3))) for s in range(n_samples)}
    per_rank = (n_samples + world - 1) // world
    all_counts = np.zeros((world, per_rank), dtype=np.int64)
    for r in range(world):
        for i, s in enumerate(genome_set.samples_of_rank(n_samples, r, world)):
            all_counts[r, i] = tabs[s].shape[0]
    pad_rows = max(int(all_counts.sum(axis=1).max()), 1)
    recv = torch.full((world, pad_rows, 3), -1, dtype=torch.int64)
    for r in range(world):
        ids = genome_set.samples_of_rank(n_samples, r, world)
        if ids and all_counts[r].sum():
            recv[r, :int(all_counts[r].sum())] = torch.cat([tabs[s] for s in ids])
    shift, sizes = genome_set.gather_plan(all_counts, n_samples, world, pad_rows)
    total = int(sizes.sum())
    index = torch.repeat_interleave(torch.from_numpy(shift), torch.from_numpy(sizes), output_size=total) + torch.arange(total)
    got = recv.view(-1, 3).index_select(0, index)
    assert torch.equal(got, torch.cat([tabs[s] for s in range(n_samples)]))
    assert [int(x) for x in sizes] == [tabs[s].shape[0] for s in range(n_samples)]
