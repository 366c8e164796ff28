"""CPU-only: the host FASTQ/FASTA(.gz) decoder (csrc/tjb_fastq.c) against the reference's parser semantics
(src/kseq.h:172-212) — checked through the compiled reference where available and through hand-made cases."""
import ctypes as C
import gzip
import os

import numpy as np
import pytest

from tatajuba_b200 import _lib, synth


def decode(path, cap=1 << 20):
    lib = _lib.load()
    r = lib.tjb_reader_open(os.fsencode(path))
    assert r
    buf = np.empty(cap, dtype=np.uint8)
    out, nr, nb = [], C.c_int64(0), C.c_int64(0)
    while True:
        got = lib.tjb_reader_fill(r, buf.ctypes.data, cap, C.byref(nr), C.byref(nb))
        assert got >= 0
        if got == 0:
            break
        out.append(buf[:got].tobytes())
    lib.tjb_reader_close(r)
    return b"".join(out), nr.value, nb.value


def test_plain_and_gz_fastq_roundtrip(tmp_path, built):
    rows = synth.paired_reads(synth.random_genome(20000, 3), 10, seed=2)
    for name in ("a.fastq", "a.fastq.gz"):
        p = str(tmp_path / name)
        synth.write_fastq(p, rows)
        got, nr, nb = decode(p, cap=100_000)  # small capacity: many fills, reads straddling chunk ends
        assert got == rows.tobytes()
        assert nr == rows.shape[0] and nb == rows.shape[0] * (rows.shape[1] - 1)


def test_record_rules(tmp_path, built):
    cases = {
        b"@r1\nACGT\n+\nIIII\n@r2\nGG\nTT\n+r2\nII\nII\n": b"ACGT\nGGTT\n",        # multi-line sequence and quality
        b">f1 comment\nACGT\nAC\n\n>f2\nTT\n": b"ACGTAC\nTT\n",                       # FASTA, empty line skipped
        b"@r1\r\nACGT\r\n+\r\nIIII\r\n@r2\r\nGG\r\n+\r\nII\r\n": b"ACGT\nGG\n",      # CRLF
        b"@r1\nACGT\n+\nIIII\n@r2\nACGTA\n+\nIII\n@r3\nAC\n+\nII\n": b"ACGT\n",      # short quality: stop there
        b"@r1\nACGT\n+\nIIII\n@r2\nACG\n+\n": b"ACGT\n",                               # no quality at EOF
        b"junk\n@r1\nAC\n+\n@I\n@r2\nGG\n+\nII": b"AC\nGG\n",                          # '@' as first quality char
        b"": b"",
        b"@r1\n": b"\n",                                                               # header only: empty sequence
    }
    for i, (text, want) in enumerate(cases.items()):
        p = str(tmp_path / f"c{i}.fq")
        open(p, "wb").write(text)
        got, _, _ = decode(p)
        assert got == want, (text, got)


@pytest.mark.skipif(not os.path.exists("/root/reference/tests/files/err1750956.fastq.gz"), reason="no /root/reference")
def test_reference_fixture_decodes_to_golden(built):
    got, nr, nb = decode("/root/reference/tests/files/err1750956.fastq.gz")
    with gzip.open(os.path.join(os.path.dirname(__file__), "golden", "err1750956.reads.txt.gz"), "rb") as f:
        assert got == f.read()
    assert (nr, nb) == (1253, 300461)


def test_record_rules_match_compiled_reference(tmp_path, reference, restatement, built):
    """same odd files through the reference's own reader+scan vs our reader + restatement scan"""
    fl = b"GATC" * 8
    rec = fl + b"AAAAA" + fl
    texts = [b"@a\n" + rec + b"\n+\n" + b"I" * len(rec) + b"\n" + b"@b\n" + rec[:40] + b"\n" + rec[40:] + b"\n+\n" + b"I" * len(rec) + b"\n",
             b">a\n" + rec[:30] + b"\n" + rec[30:] + b"\n>b\n" + rec + b"\n",
             b"@a\r\n" + rec + b"\r\n+\r\n" + b"I" * len(rec) + b"\r\n",
             b"@a\n" + rec + b"\n+\n" + b"I" * (len(rec) - 1) + b"\n@b\n" + rec + b"\n+\n" + b"I" * len(rec) + b"\n"]
    for i, t in enumerate(texts):
        p = str(tmp_path / f"x{i}.fq")
        open(p, "wb").write(t)
        ref = reference.run_files([p], 25, 4, 1, False)
        ours, _, _ = decode(p)
        assert restatement.run_buffer(ours, 25, 4, 1, False).n_raw == ref.n_raw, i


def decode_with_long_reads(path, cap):
    """like decode(), for files with reads longer than the chunk: tjb_reader_fill() hands out what the chunk holds,
    then TJB_ERR_TOO_LARGE (-5), and tjb_reader_take_long() the read itself"""
    lib = _lib.load()
    r = lib.tjb_reader_open(os.fsencode(path))
    assert r
    buf = np.empty(cap, dtype=np.uint8)
    out, nr, nb, n_long = [], C.c_int64(0), C.c_int64(0), 0
    while True:
        got = lib.tjb_reader_fill(r, buf.ctypes.data, cap, C.byref(nr), C.byref(nb))
        if got == -5:
            q = C.c_void_p()
            n = lib.tjb_reader_take_long(r, C.byref(q), C.byref(nr), C.byref(nb))
            assert n + 1 > cap
            out.append(C.string_at(q, n) + b"\n")
            n_long += 1
            continue
        assert got >= 0
        if got == 0:
            break
        out.append(buf[:got].tobytes())
    assert lib.tjb_reader_take_long(r, C.byref(C.c_void_p()), None, None) == 0  # nothing waiting
    lib.tjb_reader_close(r)
    return b"".join(out), nr.value, nb.value, n_long


@pytest.mark.parametrize("gz", [False, True])
def test_reads_longer_than_a_chunk(tmp_path, built, gz):
    """the reference's parser accepts reads of any length (src/kseq.h:172-212); here a read that does not fit into a
    chunk is handed out on its own, in order, between the chunks of the reads around it"""
    rng = np.random.default_rng(5)
    seqs = [bytes(rng.choice(list(b"ACGT"), size=n).astype(np.uint8)) for n in (50, 3000, 80, 999, 1000, 5000, 10, 7)]
    for fasta in (False, True):
        p = str(tmp_path / (("a.fasta" if fasta else "a.fastq") + (".gz" if gz else "")))
        with (gzip.open(p, "wb") if gz else open(p, "wb")) as f:
            for i, q in enumerate(seqs):
                if fasta:  # multi-line FASTA: the long reads span many lines
                    f.write(b">s%d\n" % i + b"\n".join(q[j:j + 60] for j in range(0, len(q), 60)) + b"\n")
                else:
                    f.write(b"@s%d\n" % i + q + b"\n+\n" + b"I" * len(q) + b"\n")
        got, nr, nb, n_long = decode_with_long_reads(p, cap=1000)  # reads of >= 1000 bases do not fit (length + 1 > cap)
        assert got == b"".join(q + b"\n" for q in seqs)
        assert nr == len(seqs) and nb == sum(len(q) for q in seqs) and n_long == 3
