"""The drop-in proof: the reference's own command line, built twice from /root/reference by oracle/Makefile —
unmodified (`oracle/_ref/tatajuba_ref`) and with integration/tatajuba-libtjb_gpu.patch applied and linked against
libtjb_gpu.so (`oracle/_ref/tatajuba_gpu`) — both with the same deterministic stand-in mapper behind wrapper_bwa.h
(oracle/shim/standin_mapper.c; the BWA fork is missing from the reference checkout).  Every file they write must be
identical: the seven TSV / BED outputs (src/genome_set.c:379-433, 590-735) and the per-sample VCFs
(src/analyse_variable_tracts.c:13).  The binaries are prebuilt here and travel to the GPU box."""
import gzip
import os
import subprocess

import numpy as np
import pytest

from tatajuba_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_CLI = os.path.join(ROOT, "oracle", "_ref", "tatajuba_ref")
GPU_CLI = os.path.join(ROOT, "oracle", "_ref", "tatajuba_gpu")
OUTPUTS = ["per_sample_average_length.tsv", "per_sample_modal_frequency.tsv", "per_sample_proportional_coverage.tsv",
           "selected_tracts_unknown.tsv", "selected_tracts_annotated.tsv", "tract_list.tsv", "variable_tracts.bed"]


def make_dataset(d, genome_len, coverage, n_samples, paired=True, fraction=0.02, seed=1, gz=True, error_rate=0.0):
    base = synth.random_genome(genome_len, seed)
    synth.write_reference(os.path.join(d, "ref.fasta"), os.path.join(d, "ref.gff3"), base)
    files = []
    ext = ".fastq.gz" if gz else ".fastq"
    for s in range(n_samples):
        g = synth.inject_tract_polymorphism(base, fraction, seed=100 + s)
        rows = synth.paired_reads(g, coverage, seed=1000 + s, error_rate=error_rate)
        if paired:
            r1, r2 = os.path.join(d, f"smp{s}_R1{ext}"), os.path.join(d, f"smp{s}_R2{ext}")
            synth.write_fastq(r1, rows[0::2]); synth.write_fastq(r2, rows[1::2])
            files += [r1, r2]
        else:
            r1 = os.path.join(d, f"smp{s}{ext}")
            synth.write_fastq(r1, rows)
            files.append(r1)
    return files


def run_cli(binary, d, outdir, files, extra=(), threads=None):
    env = dict(os.environ)
    if threads:
        env["OMP_NUM_THREADS"] = str(threads)
    cmd = [binary, "-g", os.path.join(d, "ref.gff3"), "-f", os.path.join(d, "ref.fasta"), "-o", os.path.join(d, outdir), "-V"]
    cmd += list(extra) + files
    p = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=900)
    assert p.returncode == 0, f"{os.path.basename(binary)} failed:\n{p.stderr[-2000:]}"
    return p


def read_outputs(d, outdir):
    out = {}
    for name in sorted(os.listdir(os.path.join(d, outdir))):
        path = os.path.join(d, outdir, name)
        with (gzip.open(path, "rb") if name.endswith(".gz") else open(path, "rb")) as f:
            out[name] = f.read()
    return out


def warnings_of(p):
    """what the reference reports on stderr about its samples (contexts found / not found by the mapper, excluded
    samples), order-independent: samples finish in any order, and a message is printed in several pieces that threads
    interleave, so only the pieces printed by one call are compared"""
    import re
    found = sorted(re.findall(r"(\d+) found and\s+(\d+) context\+tracts were not found", p.stderr))
    excluded = sorted(re.findall(r"(No HTs[^\n]*?excluded\.|From the HTs[^\n]*?excluded\.|doesn't contain any HT mapped[^\n]*)", p.stderr))
    return found, excluded


def assert_same_outputs(a, b, what):
    assert sorted(a) == sorted(b), f"{what}: different sets of files {sorted(a)} vs {sorted(b)}"
    for name in a:
        assert a[name] == b[name], f"{what}: {name} differs"
    for name in OUTPUTS:
        assert name in a, f"{what}: {name} was not written"


@pytest.mark.skipif(not os.path.exists(REF_CLI), reason="oracle/_ref/tatajuba_ref not built (needs /root/reference)")
def test_stock_cli_with_standin_mapper_is_deterministic(tmp_path):
    """CPU only: the unmodified reference with the stand-in mapper writes all its files, finds nearly every context in
    the reference it was sampled from, and gives the same bytes with 1 and 2 OpenMP threads"""
    d = str(tmp_path)
    files = make_dataset(d, 200_000, 20, 2)
    p1 = run_cli(REF_CLI, d, "out1", files, ["-p"], threads=1)
    p2 = run_cli(REF_CLI, d, "out2", files, ["-p"], threads=2)
    assert_same_outputs(read_outputs(d, "out1"), read_outputs(d, "out2"), "1 vs 2 threads")
    found = warnings_of(p1)[0]
    assert len(found) == 2
    for n_found, n_missing in found:
        assert int(n_found) > 20 * int(n_missing)
    assert warnings_of(p1) == warnings_of(p2)


needs_both = pytest.mark.skipif(not (os.path.exists(REF_CLI) and os.path.exists(GPU_CLI)), reason="command-line builds not present")


@pytest.mark.gpu
@needs_both
@pytest.mark.parametrize("case", ["c1_paired", "single_end_b", "k32_m3", "errors_and_a_dropped_sample"])
def test_patched_cli_writes_the_same_files(tmp_path, case):
    """BASELINE configs[0] (2 paired-end samples, 1 Mbp, 30x, fasta + GFF3, default -k / -m) and variations: every
    output file of the GPU build equals the stock build's, byte for byte; so do the warnings"""
    d = str(tmp_path)
    if case == "c1_paired":
        files, extra = make_dataset(d, 1_000_000, 30, 2), ["-p"]
    elif case == "single_end_b":
        files, extra = make_dataset(d, 300_000, 30, 3, paired=False, gz=False), ["-b", "-i", "3"]
    elif case == "k32_m3":
        files, extra = make_dataset(d, 300_000, 40, 2, fraction=0.05), ["-p", "-k", "32", "-m", "3", "-d", "2"]
    else:
        files, extra = make_dataset(d, 300_000, 30, 3, error_rate=0.003), ["-p"]
        tiny = os.path.join(d, "smp9_R1.fastq")  # a sample without any tract: excluded with a warning by both builds
        with open(tiny, "w") as f:
            f.write("@r0\nACGTACGTACGTAGCATCGATCGATCAGCTAGCATCGACTAGCATCAGCATCG\n+\n" + "I" * 52 + "\n")
        with open(os.path.join(d, "smp9_R2.fastq"), "w") as f:
            f.write("@r0\nTGCATGCATCGATCAGCTAGCTAGCATCGATCGATCGACTAGCTAGCATCGA\n+\n" + "I" * 52 + "\n")
        files = files[:2] + [tiny, os.path.join(d, "smp9_R2.fastq")] + files[2:]
    pr = run_cli(REF_CLI, d, "out_ref", files, extra, threads=2)
    pg = run_cli(GPU_CLI, d, "out_gpu", files, extra, threads=2)
    assert_same_outputs(read_outputs(d, "out_ref"), read_outputs(d, "out_gpu"), case)
    assert warnings_of(pr) == warnings_of(pg), f"{case}: warnings differ"


@pytest.mark.gpu
@needs_both
def test_patched_cli_many_threads_many_devices(tmp_path):
    """the reference's execution model: ONE process, `#pragma omp parallel for schedule(dynamic)` over 8 samples with 4
    threads (src/genome_set.c:66-68), every thread calling tjb_count_sample() on device (thread number mod visible
    GPUs): one context per (thread, device), no locks.  Same files as the stock build on 1 thread."""
    d = str(tmp_path)
    files = make_dataset(d, 200_000, 25, 8, fraction=0.05)
    run_cli(REF_CLI, d, "out_ref", files, ["-p"], threads=1)
    run_cli(GPU_CLI, d, "out_gpu", files, ["-p", "-t", "4"], threads=4)
    assert_same_outputs(read_outputs(d, "out_ref"), read_outputs(d, "out_gpu"), "8 samples, 4 threads")


@pytest.mark.gpu
def test_binding_state_at_the_mapper_call(tmp_path):
    """new_hopo_counter_from_files_gpu() + finalise_hopo_counter() of the PATCHED sources, called from C: the
    hopo_counter the mapper receives (elements, group indices, coverage) equals the one of the unmodified reference"""
    from oracle import pyoracle
    if not (pyoracle.Reference.available() and pyoracle.PatchedReference.available()):
        pytest.skip("oracle/_ref libraries not present")
    stock, patched = pyoracle.Reference(), pyoracle.PatchedReference()
    d = str(tmp_path)
    files = make_dataset(d, 300_000, 30, 2, error_rate=0.002)
    for fl, k, m, cov, rb in [(files[0:2], 25, 4, 5, True), (files[2:4], 32, 3, 5, True), (files[0:1], 25, 4, 3, False)]:
        want = stock.run_files(fl, k, m, cov, rb)
        got = patched.run_files_gpu(fl, k, m, cov, rb)
        got.n_raw = want.n_raw  # the binding does not keep the pre-finalise element count
        pyoracle.assert_same_boundary(got, want, f"binding k={k} m={m}")
    empty = os.path.join(d, "empty.fastq")
    open(empty, "w").close()
    assert patched.run_files_gpu([empty]).status == 1 and stock.run_files([empty]).status == 1
