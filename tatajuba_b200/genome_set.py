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
    stride = max([t["name"].shape[1] for t in tables if len(t["location"])] or [1])
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
