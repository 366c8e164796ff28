"""Pins the plain-C restatement (oracle/tjb_oracle.c) against the reference's own code compiled here
(oracle/_ref/libref_hotpath.so) — the reference ships no golden vectors for this path, so the
compiled reference IS the pin.  CPU only."""
import os

import numpy as np
import pytest

from oracle import pyoracle
from tatajuba_b200 import synth

FIXTURE = os.path.join(os.path.dirname(__file__), "golden", "err1750956.reads.txt.gz")
REF_FIXTURE = "/root/reference/tests/files/err1750956.fastq.gz"

# values recorded in SURVEY.md Appendix B for the reference's own fixture
KNOWN = {(25, 4): (8456, 334, 1, 6), (32, 3): (17744, 718, 2, 6), (2, 1): (58965, 1073, 286, 18976),
         (13, 2): (53531, 2573, 25, 10)}


def _fixture_buffer():
    import gzip
    with gzip.open(FIXTURE, "rb") as f:
        return f.read()


@pytest.mark.parametrize("km", sorted(KNOWN))
def test_restatement_known_answers_on_reference_fixture(restatement, km):
    k, m = km
    got = restatement.run_buffer(_fixture_buffer(), k, m, 5, True)
    assert (got.n_raw, got.n_elem, got.n_idx, got.coverage) == KNOWN[km]


@pytest.mark.skipif(not os.path.exists(REF_FIXTURE), reason="/root/reference not present")
@pytest.mark.parametrize("km", sorted(KNOWN))
def test_reference_known_answers_from_file(reference, km):
    k, m = km
    got = reference.run_files([REF_FIXTURE], k, m, 5, True)
    assert (got.n_raw, got.n_elem, got.n_idx, got.coverage) == KNOWN[km]


@pytest.mark.parametrize("k,m,cov,rb", [(25, 4, 5, True), (32, 3, 5, True), (2, 1, 5, True), (13, 2, 3, False),
                                        (26, 4, 1, False), (27, 2, 2, True), (5, 32, 0, False)])
def test_restatement_equals_reference_on_fixture(restatement, reference, k, m, cov, rb):
    buf = _fixture_buffer()
    a = restatement.run_buffer(buf, k, m, cov, rb)
    b = reference.run_buffer(buf, k, m, cov, rb)
    pyoracle.assert_same_boundary(a, b, f"fixture k={k} m={m}")


@pytest.mark.parametrize("k,m,cov,rb,err", [(25, 4, 5, True, 0.0), (25, 4, 5, False, 0.002), (32, 3, 5, True, 0.001),
                                            (8, 2, 10, True, 0.0), (16, 5, 2, False, 0.01)])
def test_restatement_equals_reference_on_synthetic(restatement, reference, k, m, cov, rb, err):
    buf = synth.sample_buffer(150_000, 25, seed=5, sample=1, error_rate=err).tobytes()
    a = restatement.run_buffer(buf, k, m, cov, rb)
    b = reference.run_buffer(buf, k, m, cov, rb)
    assert a.n_raw > 0 and a.status == 0
    pyoracle.assert_same_boundary(a, b, f"synthetic k={k} m={m}")


@pytest.mark.parametrize("k", [2, 7, 25, 26, 27, 32])
def test_restatement_equals_reference_on_adversarial(restatement, reference, k):
    reads = synth.adversarial_reads(k)
    buf = b"\n".join(reads * 3) + b"\n"  # x3 so that something survives the QC filters
    for m, cov, rb in [(1, 1, False), (4, 2, False), (3, 0, True)]:
        a = restatement.run_buffer(buf, k, m, cov, rb)
        b = reference.run_buffer(buf, k, m, cov, rb)
        assert a.n_skipped_nonacgt == 0  # the adversarial set holds no qualifying non-ACGTU run (reference UB)
        pyoracle.assert_same_boundary(a, b, f"adversarial k={k} m={m}")


def test_drop_paths_match_reference(restatement, reference):
    cases = {1: b"ACGTACGTAC\nACGT\n\n",                                # no tract at all
             2: (b"GATC" * 8 + b"AAAAA" + b"CTAG" * 8 + b"\n"),            # one observation: fails both QC modes
             3: (b"GATC" * 8 + b"AAAAA" + b"CTAG" * 8 + b"\n") * 2}        # survives -b, below min coverage
    for want, buf in cases.items():
        for rb in (True, False):
            a = restatement.run_buffer(buf, 25, 4, 50, rb)
            b = reference.run_buffer(buf, 25, 4, 50, rb)
            assert a.status == b.status, (want, rb)
            assert a.n_raw == b.n_raw
    assert restatement.run_buffer(cases[1], 25, 4, 5, True).status == 1
    assert restatement.run_buffer(cases[2], 25, 4, 5, False).status == 2
    assert restatement.run_buffer(cases[3], 25, 4, 50, False).status == 3


def test_nonacgt_runs_are_counted_and_skipped(restatement):
    fl = b"GATC" * 8
    buf = fl + b"NNNNN" + fl + b"\n" + fl + b"AAAA" + fl + b"\n"
    recs, skipped = restatement.scan_buffer(buf, 25, 4)
    assert skipped == 1 and len(recs) == 1 and recs[0]["length"] == 4


@pytest.mark.parametrize("k,m,rb", [(25, 4, False), (32, 3, True), (2, 1, False), (7, 2, False), (13, 5, False), (27, 6, False)])
def test_nonacgt_runs_after_a_tract_match_reference(restatement, reference, k, m, rb):
    """the DEFINED half of src/hopo_counter.c:233-248: a qualifying run of a byte that is none of ACGTU which follows a
    tract of the same read re-emits that tract's context, base and flag with the run's length"""
    for seed in range(4):
        buf = b"\n".join(synth.nrun_reads(k, seed)) + b"\n"
        a = restatement.run_buffer(buf, k, m, 1, rb)
        b = reference.run_buffer(buf, k, m, 1, rb)
        assert a.n_skipped_nonacgt == 0
        pyoracle.assert_same_boundary(a, b, f"N runs k={k} m={m} seed={seed}")
    fl = b"GATC" * 8
    buf = fl + b"AAAA" + fl + b"NNNNN" + fl + b"\n" + fl + b"TTTT" + fl + b"NNNN" + fl + b"nnnnnn" + fl + b"CC" + fl + b"---" + fl + b"\n"
    recs, skipped = restatement.scan_buffer(buf, 25, 4)
    assert skipped == 0 and [int(x) for x in recs["length"]] == [4, 5, 4, 4, 6]
    assert recs[1]["ctx0"] == recs[0]["ctx0"] and recs[3]["flag"] == 2 and recs[4]["ctx1"] == recs[2]["ctx1"]
    pyoracle.assert_same_boundary(restatement.run_buffer(buf, 25, 4, 1, False), reference.run_buffer(buf, 25, 4, 1, False), "hand-made")


@pytest.mark.parametrize("seed", range(8))
def test_restatement_equals_reference_on_arbitrary_bytes(restatement, reference, seed):
    """flanks and run neighbourhoods with every class of byte (lower case, U, IUPAC, punctuation, control bytes, bytes
    below 0x80); qualifying runs of non-ACGTU bytes are excluded (undefined behaviour in the reference, fenced), and so
    are bytes >= 0x80: the reference indexes dna_in_2_bits[] with a signed char (src/hopo_counter.c:236-245), i.e. out
    of bounds; the restatement and the GPU path code them as "not ACGT" (0)"""
    rng = np.random.default_rng(500 + seed)
    odd = np.array(list(b"acgtuUNnRYKMSWBDHV*-.09@[`{\r\t ") + [0x08, 0x0E, 0x1D, 0x0B, 0x02, 0x4A, 0x4B, 0x7F, 0x00], np.uint8)
    acgtu = set(b"ACGTUacgtu")
    reads = []
    for _ in range(int(rng.integers(300, 1500))):
        L = int(rng.integers(0, 400))
        r = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=L)
        for _ in range(int(rng.integers(0, 5))):  # homopolymer runs of real bases (either case, U)
            if L < 4:
                break
            s = int(rng.integers(0, L - 1))
            r[s:s + int(rng.integers(2, 30))] = rng.choice(np.frombuffer(b"ACGTacgtUu", np.uint8))
        hit = rng.random(L) < [0.0, 0.01, 0.05, 0.2][seed % 4]
        r[hit] = rng.choice(odd, size=int(hit.sum()))
        for i in range(1, L):  # no run of two or more equal non-ACGTU bytes
            if r[i] == r[i - 1] and int(r[i]) not in acgtu:
                r[i] = ord("A") if r[i - 1] != ord("A") else ord("C")
        reads.append(r.tobytes())
    buf = b"\n".join(reads) + b"\n"
    k, m = int(rng.integers(2, 33)), int(rng.integers(1, 6))
    a = restatement.run_buffer(buf, k, m, 1, False)
    b = reference.run_buffer(buf, k, m, 1, False)
    assert a.n_skipped_nonacgt == 0
    pyoracle.assert_same_boundary(a, b, f"arbitrary bytes seed={seed} k={k} m={m}")
