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
