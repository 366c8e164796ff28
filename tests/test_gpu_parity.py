"""GPU parity tests proper (-m gpu): the CUDA path through the C ABI vs the oracle, bit-exact.
Nothing here reads /root/reference (absent on the GPU box): the checkers are the plain-C restatement, the committed
golden vectors (made from the compiled reference) and the prebuilt oracle/_ref library when it travelled."""
import glob
import os

import numpy as np
import pytest

from oracle import pyoracle
from tatajuba_b200 import Context, TatajubaOptions, synth

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def ctx(built):
    import torch
    assert torch.cuda.is_available()
    c = Context(0)
    yield c
    c.close()


def as_boundary(hc):
    return pyoracle.Boundary(hc.status, hc.n_raw, hc.elem, hc.idx_initial, hc.idx_final, hc.coverage, hc.n_skipped_nonacgt)


def check(ctx, restatement, buf, k, m, cov, rb, what, device_too=True):
    import torch
    opt = TatajubaOptions(k, m, cov, rb)
    want = restatement.run_buffer(buf, k, m, cov, rb)
    got = as_boundary(ctx.count_host(buf, opt))
    pyoracle.assert_same_boundary(got, want, what + " [host]")
    assert got.n_skipped_nonacgt == want.n_skipped_nonacgt
    if device_too and len(buf):
        d = torch.frombuffer(bytearray(buf), dtype=torch.uint8).cuda()
        got = as_boundary(ctx.count_device(d.data_ptr(), d.numel(), opt))
        pyoracle.assert_same_boundary(got, want, what + " [device]")
    return want


from test_golden_and_abi import GOLDENS, golden_input, load_golden  # noqa: E402


@pytest.mark.parametrize("name", GOLDENS)
def test_gpu_matches_reference_goldens(ctx, name):
    want, k, m, cov, rb = load_golden(os.path.join(GOLD, name))
    got = as_boundary(ctx.count_host(golden_input(name), TatajubaOptions(k, m, cov, rb)))
    pyoracle.assert_same_boundary(got, want, name)


@pytest.mark.parametrize("k,m,cov,rb,err", [(25, 4, 5, True, 0.0), (25, 4, 5, False, 0.003), (32, 3, 5, True, 0.001),
                                            (26, 2, 3, True, 0.0), (27, 4, 5, True, 0.0), (2, 1, 5, True, 0.0),
                                            (13, 2, 10, False, 0.01), (8, 32, 0, False, 0.0)])
def test_gpu_matches_oracle_on_synthetic(ctx, restatement, k, m, cov, rb, err):
    buf = synth.sample_buffer(400_000, 30, seed=9, sample=2, error_rate=err).tobytes()
    want = check(ctx, restatement, buf, k, m, cov, rb, f"synthetic k={k} m={m}")
    if m <= 4:
        assert want.n_raw > 1000


@pytest.mark.parametrize("k", [2, 3, 7, 16, 25, 26, 27, 31, 32])
def test_gpu_matches_oracle_on_adversarial(ctx, restatement, k):
    reads = synth.adversarial_reads(k, wrap_lengths=True)
    buf = b"\n".join(reads * 3) + b"\n"
    for m, cov, rb in [(1, 1, False), (4, 2, False), (3, 0, True), (2, 1, True)]:
        check(ctx, restatement, buf, k, m, cov, rb, f"adversarial k={k} m={m}")


def test_gpu_nonacgt_runs_and_separators(ctx, restatement):
    fl = b"GATC" * 8
    reads = [fl + b"NNNNN" + fl, fl + b"AAAA" + fl, fl + b"nnnn" + fl + b"TTTTT" + fl, b"", b"", b"", fl + b"AAAA" + fl,
             b"N" * 200, fl + b"----" + fl, fl + b"AAAA" + fl[:-1] + b"N"]
    buf = b"\n".join(reads * 4) + b"\n"
    want = check(ctx, restatement, buf, 25, 4, 1, False, "non-ACGT")
    assert want.n_skipped_nonacgt == 4 * 3
    check(ctx, restatement, buf[:-1], 25, 4, 1, False, "no trailing newline")


@pytest.mark.parametrize("k,m", [(25, 4), (32, 3), (2, 1), (7, 2), (26, 5)])
def test_gpu_nonacgt_runs_after_a_tract(ctx, restatement, k, m):
    """defined half of src/hopo_counter.c:233-248 (re-emission of the previous tract's context with the run's length)
    and, with lead_tract=False, the fenced half (first tract of a read) mixed in"""
    for seed, lead in [(0, True), (1, True), (2, False), (3, False)]:
        buf = b"\n".join(synth.nrun_reads(k, seed, lead_tract=lead)) + b"\n"
        want = check(ctx, restatement, buf, k, m, 1, False, f"N runs k={k} seed={seed}")
        assert lead is False or want.n_skipped_nonacgt == 0
    # a long read: the walk back to the previous tract crosses scan windows
    rng = np.random.default_rng(5)
    fl = (b"GATC" * 8)[:max(k, 2)]
    body = b"".join(bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), size=int(rng.integers(10, 3000))).tobytes()) + b"N" * int(rng.integers(2, 40))
                    for _ in range(30))
    check(ctx, restatement, fl + b"CCCCCC" + fl + body + fl + b"\n", k, m, 1, False, "long read with N runs")


def test_gpu_drop_paths(ctx, restatement):
    one = b"GATC" * 8 + b"AAAAA" + b"CTAG" * 8 + b"\n"
    for buf, rb, cov, status in [(b"ACGTACGTAC\nACGT\n\n", True, 5, 1), (b"", True, 5, 1), (one, False, 5, 2),
                                 (one, True, 5, 2), (one * 2, False, 50, 3)]:
        want = check(ctx, restatement, buf, 25, 4, cov, rb, f"drop {status}")
        assert want.status == status


SW_TILE, SW_HALO_L, SW_HALO_R = 4000, 32, 64  # scan window geometry (csrc/tjb_scan.cuh): owned bytes, staged halos


def test_gpu_tile_boundaries(ctx, restatement):
    """tracts whose start sits on the last / first owned byte of a scan window (t * 4000 - 1, t * 4000, t * 4000 + 3999),
    with runs of 31 (fast path) and 32 bases (slow path) at k = 32, which needs the whole 31 + 32 + 1 = 64-byte right
    halo; flanks reaching into the left halo; an odd byte inside a flank at the first and the last interior window"""
    rng = np.random.default_rng(4)

    def rnd(n):  # random bases without runs of 2+ (so that only the planted tracts are found near the boundaries)
        out = bytearray()
        while len(out) < n:
            c = b"ACGT"[rng.integers(0, 4)]
            if not out or out[-1] != c:
                out.append(c)
        return bytes(out)
    n_windows = 5
    for k, m in [(32, 2), (25, 4)]:
        for t in (1, 2, n_windows - 1):
            for start in (t * SW_TILE - 1, t * SW_TILE, t * SW_TILE + SW_TILE - 1):
                for run, base in [(31, b"A"), (32, b"T"), (5, b"C"), (31, b"G")]:
                    for odd_at in (None, start - 3, start + run + k - 2):
                        body = bytearray(rnd(n_windows * SW_TILE + 300))
                        prev = bytes([body[start - 1]])
                        base_ = base if base != prev and base != bytes([body[start + run]]) else (b"C" if prev != b"C" and body[start + run] != ord("C") else b"G")
                        body[start:start + run] = base_ * run
                        if odd_at is not None:
                            body[odd_at] = ord("n")
                        buf = bytes(body) + b"\n"
                        check(ctx, restatement, buf, k, m, 1, False, f"k={k} window {t} start {start} run {run} odd {odd_at}", device_too=False)
    fl = b"GATC" * 8
    long_run = fl + b"G" * 20000 + fl + b"\n"
    check(ctx, restatement, long_run * 2, 25, 4, 1, False, "20k run")
    # reads that end / begin exactly at a window boundary
    for cut in (SW_TILE - 1, SW_TILE, SW_TILE + 1, 2 * SW_TILE - SW_HALO_L, 2 * SW_TILE + SW_HALO_R):
        a, b = rnd(cut - 40) + fl + b"AAAAA" + fl[:3], fl + b"TTTTTT" + fl + rnd(500)
        check(ctx, restatement, a + b"\n" + b + b"\n", 25, 4, 1, False, f"read boundary at {len(a)}", device_too=False)


def test_gpu_files_path_matches_oracle(ctx, restatement, tmp_path):
    rows = synth.paired_reads(synth.inject_tract_polymorphism(synth.random_genome(300_000, 5)), 30, seed=8)
    r1, r2 = str(tmp_path / "s_R1.fastq.gz"), str(tmp_path / "s_R2.fastq")
    synth.write_fastq(r1, rows[0::2]); synth.write_fastq(r2, rows[1::2])
    opt = TatajubaOptions(paired_end=True)
    hc = ctx.count_files([r1, r2], opt)
    want = restatement.run_buffer(rows[0::2].tobytes() + rows[1::2].tobytes(), 25, 4, 5, True)
    pyoracle.assert_same_boundary(as_boundary(hc), want, "files")
    assert hc.name == r1 and hc.n_reads == rows.shape[0] and hc.n_bases == rows.shape[0] * 150


def test_gpu_read_longer_than_a_chunk(ctx, restatement, tmp_path):
    """a 40 Mbp sequence (an assembly given as a sample) between ordinary reads: longer than the 32 MiB chunks of the
    host-buffer and file paths, which scan it on its own; the reference accepts reads of any length (src/kseq.h:172-212)"""
    import torch
    big = synth.inject_tract_polymorphism(synth.random_genome(40_000_000, 21), 0.02, seed=3)
    rows = synth.paired_reads(synth.random_genome(100_000, 22), 20, seed=4)
    half = rows.shape[0] // 2
    nl = np.frombuffer(b"\n", np.uint8)
    buf = np.concatenate([rows[:half].reshape(-1), big, nl, rows[half:].reshape(-1), big, nl])  # twice: its tracts pass -b
    opt = TatajubaOptions(25, 4, 2, False)
    want = restatement.run_buffer(buf.tobytes(), 25, 4, 2, False)
    assert want.n_raw > 400_000
    d = torch.from_numpy(buf).cuda()
    pyoracle.assert_same_boundary(as_boundary(ctx.count_device(d.data_ptr(), d.numel(), opt)), want, "long read, device buffer")
    pyoracle.assert_same_boundary(as_boundary(ctx.count_host(buf, opt)), want, "long read, host buffer")
    fa = str(tmp_path / "assembly.fasta")
    with open(fa, "wb") as f:
        for r in rows[:half]:
            f.write(b">r\n" + r.tobytes())
        f.write(b">contig\n" + b"\n".join(big[j:j + 80].tobytes() for j in range(0, big.size, 80)) + b"\n")
        for r in rows[half:]:
            f.write(b">r\n" + r.tobytes())
        f.write(b">contig_again\n" + big.tobytes() + b"\n")  # one line this time
    hc = ctx.count_files([fa], TatajubaOptions(25, 4, 2, False, paired_end=False))
    pyoracle.assert_same_boundary(as_boundary(hc), want, "long read, FASTA file")
    assert hc.n_reads == rows.shape[0] + 2 and hc.n_bases == rows.shape[0] * 150 + 2 * big.size
    assert hc.n_elem > 100_000  # the contig's tracts, seen twice


@pytest.mark.parametrize("coverage,k,m", [(150, 32, 3), (120, 27, 2)])
def test_gpu_mid_size_three_word_records(ctx, restatement, coverage, k, m):
    """120-150 MB of reads with k >= 27: eight or more windows per warp, so the scan reserves its record positions in
    blocks of 256 and fills the unused rest of every warp's last block with empty records (tjb_scan.cuh); m = 2 makes
    the records many (14 M)"""
    import torch
    buf = synth.sample_buffer(1_000_000, coverage, seed=7, error_rate=0.002)
    opt = TatajubaOptions(k, m, 5, True)
    want = restatement.run_buffer(buf, k, m, 5, True)
    d = torch.from_numpy(buf).cuda()
    for rep in range(2):  # the second sample of the context: sizes from the first
        pyoracle.assert_same_boundary(as_boundary(ctx.count_device(d.data_ptr(), d.numel(), opt)), want, f"k={k} m={m} sample {rep}")


def test_gpu_matches_compiled_reference_when_it_travelled(ctx):
    if not pyoracle.Reference.available():
        pytest.skip("oracle/_ref not present")
    ref = pyoracle.Reference()
    buf = synth.sample_buffer(300_000, 40, seed=13, error_rate=0.002).tobytes()
    for k, m, rb in [(25, 4, True), (32, 3, True), (25, 4, False)]:
        want = ref.run_buffer(buf, k, m, 5, rb)
        got = as_boundary(ctx.count_host(buf, TatajubaOptions(k, m, 5, rb)))
        pyoracle.assert_same_boundary(got, want, f"reference k={k}")


def test_gpu_full_size_properties(ctx):
    """BASELINE config 2 size (5 Mbp, 100x): oracle too slow to be the checker at every run, so size-independent
    properties: order strictly descending, counts sum to n_raw minus filtered, groups partition correctly, and the
    result is invariant under shuffling the reads."""
    import torch
    buf = synth.sample_buffer(5_000_000, 100, seed=1)
    opt = TatajubaOptions()
    d = torch.from_numpy(buf).cuda()
    a = ctx.count_device(d.data_ptr(), d.numel(), opt)
    assert a.status == 0 and a.n_raw > 3_000_000
    e = a.elem
    f = pyoracle.unpack_bits(e["bits"])
    key = np.stack([f["base"].astype(np.uint64), e["ctx0"], e["ctx1"], (f["length"] + 512).astype(np.uint64)], 1)
    order = np.lexsort(key.T[::-1])
    assert np.array_equal(order[::-1], np.arange(len(e)))  # strictly descending = sorted ascending reversed, no ties
    assert np.all(f["canon_flag"] == 3) and np.all(f["count"] >= 2)
    assert np.all(a.idx_final > a.idx_initial) and np.all(a.idx_initial[1:] >= a.idx_final[:-1])
    rows = buf.reshape(-1, 151)
    perm = np.random.default_rng(0).permutation(rows.shape[0])
    d2 = torch.from_numpy(np.ascontiguousarray(rows[perm]).reshape(-1)).cuda()
    b = ctx.count_device(d2.data_ptr(), d2.numel(), opt)
    pyoracle.assert_same_boundary(as_boundary(a), as_boundary(b), "shuffle invariance")


@pytest.mark.parametrize("k,m,depth", [(25, 2, 150), (32, 2, 150)])
def test_gpu_many_bins_and_record_overflow(ctx, restatement, k, m, depth):
    """m = 2 on ~105 MB of reads: ~19 % of the positions start a tract, so the first record-array estimate overflows
    (the sample is redone with the exact size) and the partition plan needs more than 2^16 bins (plan_big_kernel)"""
    buf = synth.sample_buffer(700_000, depth, seed=21)
    check(ctx, restatement, buf, k, m, 5, True, f"many bins k={k} m={m}", device_too=False)


def test_gpu_deep_long_flank_properties(ctx):
    """BASELINE config 5 shape at reduced genome size (k = 32, m = 3, 300x): descending order, QC invariants, shuffle
    invariance; the full-size sample (1.5 Gbases) takes ~9 ms on the device"""
    import torch
    buf = synth.sample_buffer(1_000_000, 300, seed=3)
    opt = TatajubaOptions(32, 3, 5, True)
    d = torch.from_numpy(buf).cuda()
    a = ctx.count_device(d.data_ptr(), d.numel(), opt)
    assert a.status == 0 and a.n_raw > 5_000_000
    e = a.elem
    f = pyoracle.unpack_bits(e["bits"])
    key = np.stack([f["base"].astype(np.uint64), e["ctx0"], e["ctx1"], (f["length"] + 512).astype(np.uint64)], 1)
    assert np.array_equal(np.lexsort(key.T[::-1])[::-1], np.arange(len(e)))
    assert np.all(f["canon_flag"] == 3)
    rows = buf.reshape(-1, 151)
    perm = np.random.default_rng(1).permutation(rows.shape[0])
    d2 = torch.from_numpy(np.ascontiguousarray(rows[perm]).reshape(-1)).cuda()
    b = ctx.count_device(d2.data_ptr(), d2.numel(), opt)
    pyoracle.assert_same_boundary(as_boundary(a), as_boundary(b), "shuffle invariance (deep, k=32)")


def _fuzz_buffer(rng, n_reads, alphabet, weights, max_len):
    reads = []
    for _ in range(n_reads):
        L = int(rng.integers(0, max_len))
        r = rng.choice(alphabet, size=L, p=weights)
        # homopolymer runs of every kind, also of odd bytes, at random places
        for _ in range(int(rng.integers(0, 4))):
            if L < 4:
                break
            s = int(rng.integers(0, L - 1))
            r[s:s + int(rng.integers(2, 40))] = rng.choice(alphabet, p=weights)
        reads.append(r.tobytes())
    return b"\n".join(reads) + (b"\n" if rng.random() < 0.7 else b"")


@pytest.mark.parametrize("seed", range(12))
def test_gpu_fuzz_arbitrary_bytes(ctx, restatement, seed):
    """random reads over an alphabet with every class of byte the scan distinguishes: ACGT, lower case, U/u, N and
    IUPAC letters, '*', '-', '.', digits, control bytes that look like the separator to the pass-1 look-up
    (0x08, 0x0E, 0x1D), '\\r', bytes >= 0x80; random read lengths (also empty and shorter than k)"""
    rng = np.random.default_rng(1000 + seed)
    odd = list(b"acgtuUNnRYKMSWBDHV*-.09@[`{\r\t ") + [0x08, 0x0E, 0x1D, 0x0B, 0x02, 0x4A, 0x4B, 0x80, 0xC1, 0xFF, 0x7F, 0x00]
    alphabet = np.array(list(b"ACGT") + odd, np.uint8)
    p_odd = [0.0, 0.002, 0.02, 0.3][seed % 4]
    weights = np.array([(1 - p_odd) / 4] * 4 + [p_odd / len(odd)] * len(odd))
    weights /= weights.sum()
    buf = _fuzz_buffer(rng, int(rng.integers(200, 3000)), alphabet, weights, int(rng.choice([30, 160, 700])))
    k, m = int(rng.integers(2, 33)), int(rng.integers(1, 7))
    check(ctx, restatement, buf, k, m, int(rng.integers(0, 4)), bool(rng.integers(0, 2)), f"fuzz seed={seed} k={k} m={m}")


def test_gpu_packed_host_transfer_is_transparent(ctx, restatement, monkeypatch):
    """Host buffers of 8 MiB and more cross PCIe 2-bit packed + exceptions and are rebuilt on the device
    (csrc/tjb_hostpack.cpp): three chunks of reads with N, lower case and stray bytes sprinkled in must give the result
    of the unpacked transfer, of the device-resident path and, on a prefix the oracle finishes quickly, of the oracle."""
    import torch
    buf = synth.sample_buffer(700_000, 100, seed=5).copy()  # 70 MB: three 32 MiB chunks
    rng = np.random.default_rng(9)
    pos = rng.integers(0, len(buf), size=20_000)
    keep = buf[pos] != ord("\n")
    buf[pos[keep]] = rng.choice(np.frombuffer(b"NnacgtRY.\x00\xff", np.uint8), size=int(keep.sum()))
    opt = TatajubaOptions(25, 3, 5, True)
    monkeypatch.setenv("TJB_HOST_PACK", "0")
    plain = as_boundary(ctx.count_host(buf, opt))
    monkeypatch.delenv("TJB_HOST_PACK")
    launches0 = dict((s["name"], s["launches"]) for s in ctx.kernel_stats()).get("unpack_kernels", 0)
    packed = as_boundary(ctx.count_host(buf, opt))
    launches1 = dict((s["name"], s["launches"]) for s in ctx.kernel_stats()).get("unpack_kernels", 0)
    assert launches1 >= launches0 + 6, "the packed transfer did not run (3 chunks x unpack + exceptions)"
    pyoracle.assert_same_boundary(packed, plain, "packed vs plain transfer")
    d = torch.from_numpy(buf).cuda()
    pyoracle.assert_same_boundary(packed, as_boundary(ctx.count_device(d.data_ptr(), d.numel(), opt)), "packed vs device")
    other = Context(0)  # the packer pool is shared by the contexts of a process
    try:
        pyoracle.assert_same_boundary(as_boundary(other.count_host(buf, opt)), plain, "packed (second context) vs plain")
        # two host threads feeding at once (the reference's OpenMP loop): sessions overlap, chunks take turns in the pool
        import threading
        res = [None, None]

        def feed(i, c):
            res[i] = as_boundary(c.count_host(buf, opt))
        th = [threading.Thread(target=feed, args=(i, c)) for i, c in enumerate((ctx, other))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        for i in range(2):
            pyoracle.assert_same_boundary(res[i], plain, f"packed, concurrent feeder {i} vs plain")
    finally:
        other.close()
    small = buf[: 151 * 60_000]  # 9 MB: still packed, small enough for the oracle
    check(ctx, restatement, small, 25, 3, 5, True, "packed transfer vs oracle", device_too=False)


def test_gpu_result_size_changes_between_samples(ctx, restatement):
    """a context sizes the result arrays and the last kernels of a sample from the previous sample's counts (they are
    allocated / launched before the counts are known): alternate small and 20x larger samples, and an empty one"""
    small = synth.sample_buffer(20_000, 40, seed=21)
    big = synth.sample_buffer(400_000, 60, seed=22)
    for i, buf in enumerate((small, big, small, big, big, small)):
        check(ctx, restatement, buf, 25, 4, 5, True, f"alternating sample {i}")
    check(ctx, restatement, small[:0], 25, 4, 5, True, "empty after large", device_too=False)
    check(ctx, restatement, big, 25, 4, 5, True, "large after empty")


# ---- full-size parity against the compiled reference (oracle/_ref travels to the GPU box) ----------------------------
def _reference_or_skip():
    if not pyoracle.Reference.available():
        pytest.skip("oracle/_ref not present")
    return pyoracle.Reference()


@pytest.mark.parametrize("err", [0.0, 0.001, 0.01])
def test_gpu_full_size_config2_matches_reference(ctx, err):
    """BASELINE configs[1] at its own size (5 Mbp, 100x, 2x150 PE = 500 Mbases), without and with substitution errors
    (0.1 % / 1 %: errors create singleton contexts, the opposite of the 64:1 de-duplication of the clean sample):
    the compiled reference itself is the checker (~3-6 s per run)"""
    import torch
    ref = _reference_or_skip()
    buf = synth.sample_buffer(5_000_000, 100, seed=1, error_rate=err)
    want = ref.run_buffer(buf.tobytes(), 25, 4, 5, True)
    opt = TatajubaOptions(25, 4, 5, True)
    d = torch.from_numpy(buf).cuda()
    pyoracle.assert_same_boundary(as_boundary(ctx.count_device(d.data_ptr(), d.numel(), opt)), want, f"config 2 err={err} [device]")
    del d
    pyoracle.assert_same_boundary(as_boundary(ctx.count_host(buf, opt)), want, f"config 2 err={err} [host]")


def test_gpu_full_size_config5_matches_reference(ctx):
    """BASELINE configs[4] at its own size: k = 32 (full 64-bit keys, 3-word records), m = 3, 300x of a 5 Mbp genome =
    1.5 Gbases, ~39 M tract observations; compiled reference ~40 s"""
    import torch
    ref = _reference_or_skip()
    buf = synth.sample_buffer(5_000_000, 300, seed=3)
    want = ref.run_buffer(buf.tobytes(), 32, 3, 5, True)
    opt = TatajubaOptions(32, 3, 5, True)
    d = torch.from_numpy(buf).cuda()
    got = ctx.count_device(d.data_ptr(), d.numel(), opt)
    assert got.n_raw > 30_000_000
    pyoracle.assert_same_boundary(as_boundary(got), want, "config 5 [device]")


# ---- the aggregation table and what hangs off it -------------------------------------------------------------------
def test_gpu_table_too_small_is_redone(ctx, restatement):
    """a sample with more distinct observations than its table was sized for is detected at the synchronisation that
    ends it and counted again with a larger table"""
    buf = synth.sample_buffer(400_000, 40, seed=31, error_rate=0.02)
    ctx.reserve_entries(16)  # the smallest table: 65536 slots
    want = check(ctx, restatement, buf, 25, 3, 2, False, "table overflow", device_too=False)
    assert want.n_distinct > 40_000
    ctx.reserve_entries(16)
    check(ctx, restatement, buf, 32, 2, 2, True, "table overflow k=32")


@pytest.mark.parametrize("k", [25, 32])
def test_gpu_oversize_bucket_goes_through_the_global_path(ctx, restatement, k):
    """thousands of distinct contexts that agree on the 10 bases next to the tract fall into ONE bin of the partition
    plan (it looks at the top <= 20 bits of the context): a bucket that does not fit in shared memory"""
    rng = np.random.default_rng(77)
    fixed = b"GTCAGTCAGT"
    reads = []
    for i in range(6000):
        left = bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), size=k - 10).tobytes()) + fixed
        right = b"C" + bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), size=k + 3).tobytes())
        r = left + b"A" * int(rng.integers(4, 9)) + right
        reads += [r, r, bytes(synth.revcomp(np.frombuffer(r, np.uint8)).tobytes())]
    buf = b"\n".join(reads) + b"\n"
    want = check(ctx, restatement, buf, k, 4, 2, True, f"oversize bucket k={k}")
    assert want.n_elem > 5000
    check(ctx, restatement, buf, k, 4, 2, False, f"oversize bucket k={k} -b")


def test_gpu_result_arena_mode(ctx, restatement):
    """TJB_RESULT_ARENA: the arrays are views of the context's pinned arena (written by the device), same contents"""
    buf = synth.sample_buffer(300_000, 30, seed=41)
    opt = TatajubaOptions()
    want = restatement.run_buffer(buf, 25, 4, 5, True)
    ctx.set_result_mode(True)
    try:
        for _ in range(3):
            got = ctx.count_host(buf, opt)
            pyoracle.assert_same_boundary(as_boundary(got), want, "arena mode")
        small = ctx.count_host(buf[: 151 * 2000], opt)  # a dropped / tiny sample after a large one
        assert small.status in (0, 2, 3)
    finally:
        ctx.set_result_mode(False)
    pyoracle.assert_same_boundary(as_boundary(ctx.count_host(buf, opt)), want, "back to malloc mode")


def _again(fn, tries=4):
    """the explicit begin / scan / finalise entry points report a table that was sized too small (the context then
    remembers the size): the caller begins the sample again"""
    from tatajuba_b200.hopo_counter import TjbError
    for _ in range(tries):
        try:
            return fn()
        except TjbError as e:
            if "begin the sample again" not in str(e):
                raise
    raise AssertionError("table overflow persisted")


def test_gpu_sample_in_pieces(ctx, restatement):
    """begin / scan chunks / finalise, and the export -> import hand-over between two contexts (what ranks do over
    NCCL when one sample is split by read chunk, genome_set.count_sample_sharded): identical to counting it whole"""
    import torch
    for k, m, rb, err in [(25, 4, True, 0.0), (32, 3, False, 0.005)]:
        buf = synth.sample_buffer(300_000, 40, seed=51, error_rate=err)
        opt = TatajubaOptions(k, m, 5, rb)
        want = restatement.run_buffer(buf, k, m, 5, rb)
        cut = (buf.size // 3) // 151 * 151
        a, b = torch.from_numpy(buf[:cut].copy()).cuda(), torch.from_numpy(buf[cut:].copy()).cuda()

        def two_chunks():
            ctx.sample_begin(opt, buf.size)
            ctx.sample_scan_device(a.data_ptr(), a.numel())
            ctx.sample_scan_device(b.data_ptr(), b.numel())
            return ctx.sample_finalise()
        pyoracle.assert_same_boundary(as_boundary(_again(two_chunks)), want, f"two chunks, one table k={k}")
        other = Context(0)
        try:
            def hand_over():
                ctx.sample_begin(opt, buf.size)
                ctx.sample_scan_device(a.data_ptr(), a.numel())
                other.sample_begin(opt, buf.size)
                other.sample_scan_device(b.data_ptr(), b.numel())
                n_rows, words = other.sample_export()
                assert words == (3 if k <= 26 else 4) and n_rows > 1000
                rows = torch.empty((n_rows, words), dtype=torch.int64, device="cuda")
                other.sample_export(rows.data_ptr(), n_rows)
                ctx.reserve_entries(0)
                if ctx.table_capacity() < 2 * n_rows:  # what the ranks agree on before they import
                    ctx.reserve_entries(4 * n_rows)
                    ctx.sample_begin(opt, buf.size)
                    ctx.sample_scan_device(a.data_ptr(), a.numel())
                ctx.sample_import(rows.data_ptr(), n_rows)
                res = ctx.sample_finalise()
                other.sample_finalise()
                return res
            pyoracle.assert_same_boundary(as_boundary(_again(hand_over)), want, f"export -> import k={k}")
        finally:
            other.close()


def test_gpu_query_rows_device_and_union(ctx, restatement):
    """query rows written on the device = tjb_query_rows on the host result; device-resident union = numpy restatement"""
    import torch
    from tatajuba_b200 import genome_set
    tables, dev_tables = [], {}
    for s in range(3):
        buf = synth.sample_buffer(200_000, 30, seed=61, sample=s, fraction=0.05)
        hc = ctx.count_host(buf, TatajubaOptions())
        dev_tables[s] = genome_set.query_rows_device(ctx, hc).clone()
        tables.append(genome_set.query_rows(hc))
        assert np.array_equal(dev_tables[s].cpu().numpy().view(np.uint64), tables[s])
    rows, sizes = genome_set.allgather_tables_device(dev_tables, 3)
    assert sizes == [len(t) for t in tables]
    d_union, d_index = genome_set.union_queries_device(ctx, rows, 25)
    want_union, want_index = pyoracle.union_oracle(tables)
    assert np.array_equal(d_union.cpu().numpy().view(np.uint64), want_union)
    assert np.array_equal(d_index.cpu().numpy(), np.concatenate(want_index))


def test_gpu_many_samples_two_contexts(ctx, restatement):
    """the reference's dynamic sample loop (genome_set.c:66-68) with two contexts on one device: every sample's result
    is what one context gives on its own"""
    from tatajuba_b200 import genome_set
    bufs = [synth.sample_buffer(150_000, 30, seed=71, sample=s, fraction=0.05) for s in range(7)]
    opt = TatajubaOptions()
    other = Context(0)
    try:
        res = genome_set.count_samples([ctx, other], len(bufs), lambda c, s: c.count_host(bufs[s], opt))
    finally:
        other.close()
    for s, hc in enumerate(res):
        pyoracle.assert_same_boundary(as_boundary(hc), restatement.run_buffer(bufs[s], 25, 4, 5, True), f"sample {s}")


def test_gpu_count_field_wraps(ctx, restatement):
    """more than 2^19 copies of one (context, length): the reference's signed 20-bit count field wraps (hopo_counter.h:40),
    the wrapped value goes through the QC filter, the group sums and the coverage (the coverage then needs the scan of
    the whole table instead of the running maximum the pack kernel keeps)"""
    fwd, rev = b"CGAAAATC", b"GATTTTCG"
    other = b"TCAAAAAGT"
    buf = (fwd + b"\n") * 300_100 + (rev + b"\n") * 300_000 + (other + b"\n") * 7 + (bytes(synth.revcomp(np.frombuffer(other, np.uint8)).tobytes()) + b"\n") * 5
    for rb in (True, False):
        want = check(ctx, restatement, buf, 2, 4, 1, rb, f"count wrap rb={rb}")
        cnt = pyoracle.unpack_bits(want.elem["bits"])["count"]
        assert want.n_raw == 600_112 and (cnt.min() < 0) == rb  # "-b" keeps count > 1 only: the wrapped one is dropped
    if pyoracle.Reference.available():
        pyoracle.assert_same_boundary(restatement.run_buffer(buf, 2, 4, 1, True), pyoracle.Reference().run_buffer(buf, 2, 4, 1, True), "count wrap vs reference")
