"""Seeded per-sample post-mapping context tables for the cross-sample merge tests (shared by the golden generator)."""
import numpy as np


def make_tables(seed, n_samples, n_sites, k=8, dup=True):
    rng = np.random.default_rng(seed)
    sites = np.sort(rng.choice(np.arange(50, 50 + 40 * n_sites), n_sites, replace=False)).astype(np.int32)
    site_base = rng.integers(0, 2, n_sites)
    site_len = rng.integers(3, 12, n_sites)
    alphabet = np.frombuffer(b"ACGT", np.uint8)
    site_name = alphabet[rng.integers(0, 4, (n_sites, 2 * k + 3))]
    site_name[:, k] = site_name[:, k + 2] = ord("-")
    tables = []
    for s in range(n_samples):
        keep = rng.random(n_sites) < 0.8
        if s == 2:
            keep[:] = False if n_samples > 3 else keep  # an empty table in the middle of the list
        loc = sites[keep] + rng.choice([0, 0, 0, 1, -1, 2, 3, 7], keep.sum()).astype(np.int32)
        base = site_base[keep].copy()
        base[rng.random(keep.sum()) < 0.03] ^= 1
        mlen = np.maximum(2, site_len[keep] + rng.integers(-2, 3, keep.sum()))
        name = site_name[keep].copy()
        for r in np.flatnonzero(rng.random(keep.sum()) < 0.4):  # substitutions and a one-base shift (an indel)
            if rng.random() < 0.5:
                name[r, rng.integers(0, name.shape[1], rng.integers(1, 5))] = alphabet[rng.integers(0, 4)]
            else:
                name[r, 1:] = name[r, :-1].copy()
        if dup and keep.sum() > 4:  # two histograms of one sample at the same location
            j = rng.integers(0, keep.sum() - 1)
            loc[j + 1] = loc[j]
        order = np.argsort(loc, kind="stable")
        tables.append({"location": loc[order].astype(np.int32), "base": base[order].astype(np.int32),
                       "mode_context_length": mlen[order].astype(np.int32), "name": name[order]})
    return tables


CASES = [(1, 2, 40), (2, 5, 300), (3, 8, 1500), (4, 1, 25), (5, 3, 10)]
