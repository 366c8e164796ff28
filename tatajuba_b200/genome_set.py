"""Per-sample fan-out and cross-sample exchange: host mirror of the sample loop of new_genome_set_from_files()
(reference src/genome_set.c:46-112) for one process per GPU.

The reference deals samples to OpenMP threads (`schedule(dynamic)`, genome_set.c:67, 82); here they are dealt to
ranks (sample s -> rank s mod world, samples of a rank processed in order), each rank counts its samples on its own
B200 through the C ABI, and the compact per-sample query tables (one row per indexed context group, see
include/tjb_gpu.h) are exchanged with ONE allgather (NCCL over NVLink; gloo in the CPU tests).  Every rank then
holds every sample's table in global sample order and computes the union of distinct mapper queries on its GPU
(tjb_union_queries_device), the input of the host-side mapper that follows the path (hopo_counter.c:523).

torch.distributed is plumbing only; no compute happens in this file.
"""
import ctypes as C

import numpy as np

from . import _lib
from .hopo_counter import TJB_OK

ROW_WORDS = 3  # {context[0], context[1], base | length_field << 8}


def samples_of_rank(n_samples, rank, world):
    """global sample ids this rank counts (round-robin, the static equivalent of the reference's dynamic schedule)"""
    return list(range(rank, n_samples, world))


def query_rows(hc):
    """(n_idx, 3) uint64 query rows of one counted sample: the first = longest record of every indexed context
    group (src/hopo_counter.c:513-520).  Dropped samples (status != 0) give an empty table (genome_set.c:136)."""
    if hc.status != TJB_OK or hc.n_idx == 0:
        return np.zeros((0, ROW_WORDS), np.uint64)
    lib = _lib.load()
    out = _lib.TjbCounterOut()
    out.n_elem, out.n_idx = hc.n_elem, hc.n_idx
    out.elem = hc.elem.ctypes.data
    out.idx_initial = hc.idx_initial.ctypes.data_as(C.POINTER(C.c_int32))
    rows = np.empty((hc.n_idx, ROW_WORDS), np.uint64)
    rc = lib.tjb_query_rows(C.byref(out), C.c_void_p(rows.ctypes.data))
    if rc != hc.n_idx:
        raise RuntimeError(f"tjb_query_rows: {rc}")
    return rows


def allgather_tables(local_tables, n_samples, group=None, device=None):
    """local_tables: {global sample id: (n, 3) uint64 array} of this rank's samples.  Returns the list of all
    n_samples tables in global sample order, identical on every rank.  Two collectives: the row counts, then one
    allgather of the rank's concatenated rows padded to the largest rank payload."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        assert sorted(local_tables) == list(range(n_samples)), "single process: every sample must be local"
        return [np.ascontiguousarray(local_tables[s], np.uint64).reshape(-1, ROW_WORDS) for s in range(n_samples)]
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    mine = samples_of_rank(n_samples, rank, world)
    assert sorted(local_tables) == mine, f"rank {rank} must hold samples {mine}"
    per_rank = (n_samples + world - 1) // world
    counts = torch.zeros(per_rank, dtype=torch.int64)
    for i, s in enumerate(mine):
        counts[i] = len(local_tables[s])
    counts = counts.to(device)
    all_counts = torch.empty(world * per_rank, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(all_counts, counts, group=group)
    all_counts = all_counts.cpu().numpy().reshape(world, per_rank)
    pad_rows = int(all_counts.sum(axis=1).max())
    # uint64 rows travel as int64 (same bits): NCCL and gloo both move 8-byte integers
    payload = np.zeros((pad_rows, ROW_WORDS), np.uint64)
    off = 0
    for s in mine:
        t = np.ascontiguousarray(local_tables[s], np.uint64).reshape(-1, ROW_WORDS)
        payload[off:off + len(t)] = t
        off += len(t)
    send = torch.from_numpy(payload.view(np.int64)).to(device)
    recv = torch.empty((world, pad_rows, ROW_WORDS), dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group)
    recv = recv.cpu().numpy().view(np.uint64)
    tables = [None] * n_samples
    for r in range(world):
        off = 0
        for i, s in enumerate(samples_of_rank(n_samples, r, world)):
            n = int(all_counts[r, i])
            tables[s] = recv[r, off:off + n].copy()
            off += n
    return tables


def union_queries(ctx, tables, kmer_size):
    """distinct query rows over all samples, in the reference's descending (base, ctx0, ctx1, length) order, and per
    sample the index of each of its rows in that union.  Runs on ctx's GPU (no CPU path)."""
    import torch
    lib = _lib.load()
    sizes = [len(t) for t in tables]
    n = int(sum(sizes))
    if n == 0:
        return np.zeros((0, ROW_WORDS), np.uint64), [np.zeros(0, np.int32) for _ in tables]
    rows = np.concatenate([np.ascontiguousarray(t, np.uint64).reshape(-1, ROW_WORDS) for t in tables])
    dev = torch.device("cuda", ctx.device)
    d_rows = torch.from_numpy(rows.view(np.int64)).to(dev)
    d_union = torch.empty_like(d_rows)
    d_index = torch.empty(n, dtype=torch.int32, device=dev)
    n_union = C.c_int64(0)
    torch.cuda.synchronize(dev)
    rc = lib.tjb_union_queries_device(ctx.h, C.c_void_p(d_rows.data_ptr()), n, int(kmer_size),
                                      C.c_void_p(d_union.data_ptr()), C.byref(n_union), C.c_void_p(d_index.data_ptr()))
    ctx._check(rc)
    union = d_union[: n_union.value].cpu().numpy().view(np.uint64)
    index = d_index.cpu().numpy()
    per_sample, off = [], 0
    for m in sizes:
        per_sample.append(index[off:off + m])
        off += m
    return union, per_sample


# ---- device-resident exchange: nothing of the tables crosses PCIe between the count and the union ---------------------
def query_rows_device(ctx, hc):
    """(n_idx, 3) int64 CUDA tensor of the query rows of the sample `ctx` counted LAST (must be `hc`), written by the
    device from the device-resident result (tjb_query_rows_device)"""
    import torch
    n = hc.n_idx if hc.status == TJB_OK else 0
    t = torch.empty((max(n, 1), ROW_WORDS), dtype=torch.int64, device=torch.device("cuda", ctx.device))
    got = ctx.query_rows_device(t.data_ptr(), n) if n else 0
    return t[:got]


def allgather_tables_device(local_tables, n_samples, group=None):
    """like allgather_tables, for {sample id: (n, 3) int64 CUDA tensor}; the rows stay on the devices (NCCL moves them
    GPU to GPU).  Returns (rows of all samples concatenated in global sample order, rows per sample) — only the
    per-sample row counts visit the host."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        assert sorted(local_tables) == list(range(n_samples)), "single process: every sample must be local"
        ts = [local_tables[s].reshape(-1, ROW_WORDS) for s in range(n_samples)]
        return (torch.cat(ts) if ts else torch.empty((0, ROW_WORDS), dtype=torch.int64)), [int(t.shape[0]) for t in ts]
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    mine = samples_of_rank(n_samples, rank, world)
    assert sorted(local_tables) == mine, f"rank {rank} must hold samples {mine}"
    dev = local_tables[mine[0]].device if mine else torch.device("cuda", torch.cuda.current_device())
    per_rank = (n_samples + world - 1) // world
    counts = torch.zeros(per_rank, dtype=torch.int64)
    for i, s in enumerate(mine):
        counts[i] = local_tables[s].shape[0]
    all_counts = torch.empty(world * per_rank, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(all_counts, counts.to(dev), group=group)
    all_counts = all_counts.cpu().numpy().reshape(world, per_rank)
    pad_rows = max(int(all_counts.sum(axis=1).max()), 1)
    send = torch.empty((pad_rows, ROW_WORDS), dtype=torch.int64, device=dev)  # the padding is never read
    n_mine = int(all_counts[rank].sum())
    if n_mine:
        torch.cat([local_tables[s].reshape(-1, ROW_WORDS) for s in mine], out=send[:n_mine])
    recv = torch.empty((world, pad_rows, ROW_WORDS), dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group)
    shift, sizes = gather_plan(all_counts, n_samples, world, pad_rows)
    total = int(sizes.sum())
    if total == 0:
        return torch.empty((0, ROW_WORDS), dtype=torch.int64, device=dev), [0] * n_samples
    meta = torch.from_numpy(np.stack([shift, sizes])).to(dev)
    index = torch.repeat_interleave(meta[0], meta[1], output_size=total) + torch.arange(total, dtype=torch.int64, device=dev)
    return recv.view(-1, ROW_WORDS).index_select(0, index), [int(x) for x in sizes]


def gather_plan(all_counts, n_samples, world, pad_rows):
    """Global sample order out of the rank-major receive buffer of the allgather with ONE gather.  all_counts[r, i] =
    rows of the i-th sample of rank r; every rank's rows are contiguous from r * pad_rows.  Returns (shift, sizes), one
    entry per sample in global order: output row p of sample s comes from receive row p + shift[s]."""
    sizes, starts = np.zeros(n_samples, dtype=np.int64), np.zeros(n_samples, dtype=np.int64)
    for r in range(world):
        ids = samples_of_rank(n_samples, r, world)
        if not ids:
            continue
        c = np.asarray(all_counts[r, :len(ids)], dtype=np.int64)
        sizes[ids] = c
        starts[ids] = r * pad_rows + np.concatenate(([0], np.cumsum(c)[:-1]))
    out_first = np.concatenate(([0], np.cumsum(sizes)[:-1])) if n_samples else np.zeros(0, dtype=np.int64)
    return starts - out_first, sizes


def union_queries_device(ctx, d_rows, kmer_size):
    """d_rows: (n, 3) int64 CUDA tensor.  -> (union rows (n_union, 3) int64 CUDA tensor in the reference's descending
    order, index of every input row in the union (n,) int32 CUDA tensor)"""
    import torch
    lib = _lib.load()
    n = int(d_rows.shape[0])
    d_rows = d_rows.contiguous()
    d_union = torch.empty_like(d_rows)
    d_index = torch.empty(max(n, 1), dtype=torch.int32, device=d_rows.device)
    n_union = C.c_int64(0)
    torch.cuda.current_stream(d_rows.device).synchronize()  # the rows were produced on torch's stream
    rc = lib.tjb_union_queries_device(ctx.h, C.c_void_p(d_rows.data_ptr()), n, int(kmer_size),
                                      C.c_void_p(d_union.data_ptr()), C.byref(n_union), C.c_void_p(d_index.data_ptr()))
    ctx._check(rc)
    return d_union[: n_union.value], d_index[:n]


# ---- one sample in pieces: read chunks on several GPUs (SURVEY.md 8e "by read chunk within a sample") -----------------
def count_sample_sharded(ctx, d_ptr, n_bytes, opt, group=None, total_bytes_hint=0):
    """Every rank scans ITS chunk of the sample's reads (d_ptr / n_bytes: device-resident, whole reads) into its own
    aggregation table; the tables are exchanged as rows with one allgather (NCCL over NVLink; gloo cannot move device
    rows) and every rank adds the other ranks' rows to its table (counts per strand add up to what one scan of all the
    reads would have counted), so that QC, context groups and coverage (hopo_counter.c:367-438) run on the complete
    table.  Every rank returns the same HopoCounter, bit-identical to counting the whole sample on one GPU."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    dev = torch.device("cuda", ctx.device)
    for _attempt in range(4):
        ctx.sample_begin(opt, total_bytes_hint or n_bytes * world)
        ctx.sample_scan_device(d_ptr, n_bytes)
        if world == 1:
            return ctx.sample_finalise()
        try:
            n_rows, words = ctx.sample_export()
            failed = 0
        except Exception:  # table too small for this chunk: every rank starts again (the context remembers the size)
            n_rows, words, failed = 0, tjb_rec_words(opt.kmer_size) + 1, 1
        info = torch.tensor([n_rows, failed, ctx.table_capacity()], dtype=torch.int64, device=dev)
        all_info = torch.empty((world, 3), dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(all_info.view(-1), info, group=group)
        all_info = all_info.cpu().numpy()
        total = int(all_info[:, 0].sum())  # upper bound of the distinct observations of the whole sample
        if all_info[:, 1].any() or (all_info[:, 2] < total).any():
            # some rank's table cannot take the combined rows: every rank knows it and starts again, sized for the sum
            ctx.reserve_entries(max(total, 2 * int(all_info[:, 0].max())))
            continue
        pad = max(int(all_info[:, 0].max()), 1)
        send = torch.zeros((pad, words), dtype=torch.int64, device=dev)
        if n_rows:
            ctx.sample_export(send.data_ptr(), pad)
        recv = torch.empty((world, pad, words), dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group)
        torch.cuda.current_stream(dev).synchronize()
        for r in range(world):
            if r != rank and all_info[r, 0]:
                ctx.sample_import(recv[r].data_ptr(), int(all_info[r, 0]))
        return ctx.sample_finalise()
    raise RuntimeError("count_sample_sharded: aggregation table overflow persisted")


def tjb_rec_words(k):
    return 2 if k <= 26 else 3


# ---- many samples per GPU: the reference's `#pragma omp parallel for schedule(dynamic)` over samples ------------------
def count_samples(contexts, n_samples, count_one, on_done=None):
    """Deals samples 0..n_samples-1 dynamically to `contexts` (tjb contexts of one or several devices, e.g. two per
    GPU so that sample i's copy-out and host work overlap sample i+1's copy-in and scan), one host thread per context,
    like the reference's OpenMP loop (src/genome_set.c:66-68).  count_one(ctx, sample) -> HopoCounter runs on the
    context's thread (the library call releases the GIL); on_done(ctx, sample, hc), if given, runs on the same thread
    right after (e.g. to take the device-resident query rows before the context counts its next sample).
    Returns the list of results in sample order."""
    import threading
    results, nxt, lock, errors = [None] * n_samples, [0], threading.Lock(), []

    def worker(ctx):
        try:
            while True:
                with lock:
                    s = nxt[0]
                    nxt[0] += 1
                if s >= n_samples:
                    return
                hc = count_one(ctx, s)
                results[s] = on_done(ctx, s, hc) if on_done else hc
        except Exception as e:  # noqa: BLE001
            errors.append(e)
    threads = [threading.Thread(target=worker, args=(c,)) for c in contexts[1:]]
    for t in threads:
        t.start()
    worker(contexts[0])
    for t in threads:
        t.join()
    if errors:
        raise errors[0]
    return results


def context_histograms(elem40, kmer_size=25, min_tract_size=4, max_distance_per_flank=1, levenshtein_distance=2):
    """new_genomic_context_list() after finalise_hopo_counter() (src/context_histogram.c:237-305, 364-421): the mapped
    elements of one sample (HOPO_DTYPE array, hc.elem[ref_start:] in the order the location sort left them in) ->
    dict of flat arrays, one row per context histogram (tjb_context_histograms)"""
    lib = _lib.load()
    e = np.ascontiguousarray(elem40)
    assert e.dtype.itemsize == 40
    opt = _lib.TjbHistOpts(int(kmer_size), int(min_tract_size), int(max_distance_per_flank), int(levenshtein_distance))
    t = _lib.TjbContextTable()
    rc = lib.tjb_context_histograms(C.byref(opt), C.c_void_p(e.ctypes.data if len(e) else 0), len(e), C.byref(t))
    if rc < 0:
        raise RuntimeError(f"tjb_context_histograms: {rc}")
    out = t.to_numpy()
    lib.tjb_context_table_free(C.byref(t))
    return out


def surviving_samples(alive):
    """remove_empty_genomes + index labelling (src/genome_set.c:97-100, 132-142): samples without any mapped HT are
    dropped BEFORE sample ids are assigned, so ids depend on which samples survive.  alive: one flag per sample.
    Returns {original sample -> compacted id}."""
    ids, j = {}, 0
    for s, a in enumerate(alive):
        if a:
            ids[s] = j
            j += 1
    if not ids:
        raise RuntimeError("No sample has any tract mapped to reference genome")  # genome_set.c:137
    return ids


def merge_tracts(tables, levenshtein_distance=2):
    """g_tract_vector_concatenate_tracts + tract ids (src/genome_set.c:195-289) over the per-sample post-mapping tables
    of the surviving samples.  tables: list of dicts with int32 arrays `location` (ascending), `base`,
    `mode_context_length` and a bytes array `name` of shape (n, stride).  Returns (sample, index, tract_id)."""
    lib = _lib.load()
    n_hist = np.array([len(t["location"]) for t in tables], np.int32)
    total = int(n_hist.sum())
    cat = lambda k, dt: np.ascontiguousarray(np.concatenate([np.asarray(t[k], dt) for t in tables]) if total else np.zeros(0, dt))  # noqa: E731
    loc, base, mlen = cat("location", np.int32), cat("base", np.int32), cat("mode_context_length", np.int32)
    widths = {t["name"].shape[1] for t in tables if len(t["location"])}
    stride = max(widths or [1])
    if len(widths) == 1:  # the usual case (one kmer size): one concatenation instead of a zeroed array filled piece by piece
        names = np.ascontiguousarray(np.concatenate([np.asarray(t["name"]).view(np.uint8).reshape(len(t["location"]), -1)
                                                     for t in tables if len(t["location"])]))
    else:
        names = np.zeros((total, stride), np.uint8)
        off = 0
        for t in tables:
            n = len(t["location"])
            if n:
                names[off:off + n, : t["name"].shape[1]] = np.asarray(t["name"]).view(np.uint8).reshape(n, -1)
            off += n
    out = [np.empty(total, np.int32) for _ in range(3)]
    p = lambda a: C.c_void_p(a.ctypes.data)  # noqa: E731
    rc = lib.tjb_merge_tracts(len(tables), p(n_hist), p(loc), p(base), p(mlen), p(names), stride, int(levenshtein_distance),
                              p(out[0]), p(out[1]), p(out[2]))
    if rc != total:
        raise RuntimeError(f"tjb_merge_tracts: {rc}")
    return tuple(out)
