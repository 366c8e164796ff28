"""Cross-sample merge (reference src/genome_set.c:195-289): tjb_merge_tracts vs golden vectors made by the compiled
reference, and vs the compiled reference itself when oracle/_ref is available.  Host-only (no GPU needed)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(__file__))
from merge_cases import CASES, make_tables  # noqa: E402
from oracle import pyoracle  # noqa: E402
from tatajuba_b200 import genome_set  # noqa: E402

GOLD = os.path.join(os.path.dirname(__file__), "golden", "merge", "merge_tracts.npz")


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("lev", [2, 0])
def test_merge_matches_golden(built, case, lev):
    seed, ns, nsites = case
    z = np.load(GOLD)
    s, i, t = genome_set.merge_tracts(make_tables(seed, ns, nsites), lev)
    assert np.array_equal(s, z[f"c{seed}_l{lev}_sample"])
    assert np.array_equal(i, z[f"c{seed}_l{lev}_index"])
    assert np.array_equal(t, z[f"c{seed}_l{lev}_tract"])


@pytest.mark.parametrize("seed", range(20, 30))
def test_merge_matches_compiled_reference(reference, seed):
    rng = np.random.default_rng(seed)
    tables = make_tables(seed, int(rng.integers(1, 9)), int(rng.integers(5, 400)), k=int(rng.integers(2, 33)))
    lev = int(rng.integers(0, 4))
    got = genome_set.merge_tracts(tables, lev)
    want = pyoracle.reference_merge_tracts(tables, lev)
    for a, b in zip(got, want):
        assert np.array_equal(a, b)


def test_merge_properties(built):
    tables = make_tables(9, 6, 500)
    s, i, t = genome_set.merge_tracts(tables)
    loc = np.array([tables[a]["location"][b] for a, b in zip(s, i)])
    assert np.all(np.diff(loc) >= 0)                      # ordered by location
    assert np.all(np.diff(t) >= 0) and t[0] == 0 and np.all(np.diff(t) <= 1)
    seen = sorted(zip(s.tolist(), i.tolist()))
    assert seen == [(a, b) for a in range(len(tables)) for b in range(len(tables[a]["location"]))]  # a permutation


def test_surviving_samples():
    assert genome_set.surviving_samples([True, False, True, True]) == {0: 0, 2: 1, 3: 2}
    with pytest.raises(RuntimeError):
        genome_set.surviving_samples([False, False])


@pytest.mark.parametrize("seed", range(40, 46))
def test_merge_many_ties_matches_compiled_reference(reference, seed):
    """few distinct locations, many samples, several rows of one sample at the same location: the order the sequential
    two-way merges leave equal locations in (genome_set.c:274-281) is reproduced group by group"""
    rng = np.random.default_rng(seed)
    n_samples, k = int(rng.integers(2, 40)), 6
    tables = []
    for _ in range(n_samples):
        n = int(rng.integers(0, 60))
        loc = np.sort(rng.integers(0, 25, size=n)).astype(np.int32)  # duplicates inside a sample are likely
        name = np.frombuffer(b"ACGTAC.A.GTCAGT\0", np.uint8)
        tables.append({"location": loc, "base": rng.integers(0, 2, size=n).astype(np.int32),
                       "mode_context_length": rng.integers(2, 9, size=n).astype(np.int32), "name": np.tile(name, (n, 1))})
    if not any(len(t["location"]) for t in tables):
        tables[0] = {"location": np.array([3], np.int32), "base": np.zeros(1, np.int32), "mode_context_length": np.full(1, 4, np.int32),
                     "name": np.tile(np.frombuffer(b"ACGTAC.A.GTCAGT\0", np.uint8), (1, 1))}
    got = genome_set.merge_tracts(tables, 2)
    want = pyoracle.reference_merge_tracts(tables, 2)
    for a, b in zip(got, want):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("seed", range(40, 44))
def test_merge_sparse_locations_matches_compiled_reference(reference, seed):
    """locations spread over far more positions than there are rows (few samples of a large genome): the library sorts
    them by radix passes instead of one counting sort; same result as the reference's sequential merges"""
    rng = np.random.default_rng(seed)
    tables = make_tables(seed, int(rng.integers(2, 7)), int(rng.integers(50, 300)))
    for t in tables:
        t["location"] = (t["location"].astype(np.int64) * 4001 % (1 << 30)).astype(np.int32)
        order = np.argsort(t["location"], kind="stable")
        for key in t:
            t[key] = np.ascontiguousarray(t[key][order])
    got = genome_set.merge_tracts(tables, 2)
    want = pyoracle.reference_merge_tracts(tables, 2)
    for a, b in zip(got, want):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("seed,threads", [(50, 16), (51, 5), (52, 32)])
def test_merge_in_slices_matches_compiled_reference(reference, seed, threads, monkeypatch):
    """tables large enough for the sliced, multi-threaded path (many samples of one genome): same result as the
    reference's sequential merges and as one thread, whether or not the tract-id chain can be cut at a slice boundary
    (sites ~40 apart, mode lengths up to 13 / up to 60: chains are cut / carried over)"""
    rng = np.random.default_rng(seed)
    tables = make_tables(seed, 30, 4000)
    if seed == 51:  # long tracts: neighbouring sites overlap, no boundary is a safe cut
        for t in tables:
            t["mode_context_length"] = (t["mode_context_length"] + rng.integers(20, 50, len(t["location"]))).astype(np.int32)
    assert sum(len(t["location"]) for t in tables) > 16 * 4096
    want = pyoracle.reference_merge_tracts(tables, 2)
    monkeypatch.setenv("TJB_MERGE_THREADS", str(threads))
    got = genome_set.merge_tracts(tables, 2)
    monkeypatch.setenv("TJB_MERGE_THREADS", "1")
    one = genome_set.merge_tracts(tables, 2)
    for a, b, c in zip(got, want, one):
        assert np.array_equal(a, b)
        assert np.array_equal(c, b)
