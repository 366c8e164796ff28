"""Rows 11-13 of the hot-path table: tjb_context_histograms against the reference's own
new_genomic_context_list() / finalise_genomic_context_hist() / genomic_context_merge_histograms_at_same_location()
(src/context_histogram.c:224-305, 364-421), run unmodified with the deterministic stand-in mapper (oracle/shim).
The shim hands back both ends: the mapped elements in location order (our input) and the histograms (our target).
CPU only; live against oracle/_ref, plus one committed golden for boxes without /root/reference."""
import os

import numpy as np
import pytest

from oracle import pyoracle
from tatajuba_b200 import genome_set, synth

GOLD = os.path.join(os.path.dirname(__file__), "golden", "hist", "context_histograms.npz")
KEYS = ["location", "loc2d", "base", "neg_strand", "mismatches", "multi", "indel", "integral", "n_context",
        "mode_context_count", "mode_context_length", "mode_context_id", "h_min", "h_max", "names", "ctx_first", "contexts",
        "h_first", "h_length", "h_freq"]


def dataset(tmp, genome_len, coverage, seed, fraction, error_rate):
    base = synth.random_genome(genome_len, seed)
    fa, gff = os.path.join(tmp, "ref.fasta"), os.path.join(tmp, "ref.gff3")
    synth.write_reference(fa, gff, base)
    g = synth.inject_tract_polymorphism(base, fraction, seed=100 + seed)
    return synth.paired_reads(g, coverage, seed=1000 + seed, error_rate=error_rate).reshape(-1), fa, gff


def assert_same_table(got, want, what):
    assert got["n_hist"] == want["n_hist"], f"{what}: n_hist {got['n_hist']} != {want['n_hist']}"
    for k in KEYS:
        if want["n_hist"] > 0:
            assert np.array_equal(got[k], want[k]), f"{what}: {k} differs"


needs_ref = pytest.mark.skipif(not os.path.exists(pyoracle.PIPELINE_SO), reason="oracle/_ref/libref_pipeline.so not built")


@needs_ref
@pytest.mark.parametrize("seed", range(12))
def test_context_histograms_match_compiled_reference(tmp_path, built, seed):
    """12 random samples: options drawn per seed (k 7..32, m 2..5, -d 0..3, -l 0..4, both QC modes), substitution
    errors up to 2 % so that near-identical contexts (Hamming and indel merges, :256-266) and neighbouring histograms
    (:364-385) occur"""
    rng = np.random.default_rng(700 + seed)
    k = int(rng.choice([7, 12, 21, 25, 26, 27, 32]))
    m, d, lev = int(rng.integers(2, 6)), int(rng.integers(0, 4)), int(rng.integers(0, 5))
    d = min(d, k // 2)
    rb = bool(rng.integers(0, 2))
    buf, fa, gff = dataset(str(tmp_path), int(rng.integers(30_000, 120_000)), 40, seed, 0.1, float(rng.choice([0.0, 0.005, 0.02])))
    elem, want = pyoracle.reference_pipeline(buf, fa, gff, k, m, 3, rb, d, lev)
    assert elem is not None and len(elem) > 50
    got = genome_set.context_histograms(elem, k, m, d, lev)
    assert_same_table(got, want, f"seed {seed} k={k} m={m} d={d} l={lev}")
    assert want["n_hist"] > 10
    if seed == 0:  # the last histogram of the sample is dropped (:379-382): fewer rows than a plain merge would give
        assert int(want["integral"].sum()) < int(pyoracle.unpack_bits(elem["bits"])["count"].sum())


def test_context_histograms_match_golden(built):
    z = np.load(GOLD)
    got = genome_set.context_histograms(z["elem"], int(z["k"]), int(z["m"]), int(z["d"]), int(z["lev"]))
    assert_same_table(got, {k: z[k] for k in KEYS + ["n_hist"]} | {"n_hist": int(z["n_hist"])}, "golden")


def test_context_histograms_edge_cases(built):
    assert genome_set.context_histograms(np.zeros(0, pyoracle.HOPO_DTYPE))["n_hist"] == 0
    z = np.load(GOLD)
    one = genome_set.context_histograms(z["elem"][:1], int(z["k"]), int(z["m"]), int(z["d"]), int(z["lev"]))
    assert one["n_hist"] == 0  # a single histogram: n_hist = j = 0 (:379-382)
