"""CPU-only: the gzip stream of the file feed (csrc/tjb_inflate.c, our own DEFLATE decoder) against zlib on every
block type, on concatenated members, on header variants and on damaged files.  The reference reads its samples with
gzopen / gzread (src/hopo_counter.c:141-153); whatever gzread delivers for a valid file, tjb_gz_next must deliver."""
import ctypes as C
import gzip
import os
import struct
import zlib

import numpy as np
import pytest

from tatajuba_b200 import _lib, synth

TJB_ERR_IO = -3


def inflate(path):
    """-> (bytes delivered, final return code of tjb_gz_next: 0 = clean end, TJB_ERR_IO = damaged)"""
    lib = _lib.load()
    g = lib.tjb_gz_open(os.fsencode(path))
    assert g, "not taken as a gzip file"
    out, p = [], C.c_void_p()
    while True:
        n = lib.tjb_gz_next(g, C.byref(p))
        if n <= 0:
            break
        out.append(C.string_at(p, n))
    lib.tjb_gz_close(g)
    return b"".join(out), n


def fastq_text(genome=40_000, coverage=30, seed=5):
    rows = synth.paired_reads(synth.random_genome(genome, seed), coverage, seed=seed + 1, error_rate=0.01)
    rng = np.random.default_rng(seed)
    recs = []
    for i, r in enumerate(rows[:4000]):  # realistic qualities: mostly literals, unlike the constant "IIII" lines
        q = rng.integers(33, 74, size=len(r) - 1, dtype=np.uint8).tobytes()
        recs.append(b"@read%d/1\n" % i + r[:-1].tobytes() + b"\n+\n" + q + b"\n")
    return b"".join(recs)


def deflate(data, level=6, strategy=zlib.Z_DEFAULT_STRATEGY, memlevel=8):
    c = zlib.compressobj(level, zlib.DEFLATED, 31, memlevel, strategy)
    return c.compress(data) + c.flush()


@pytest.fixture(scope="module")
def text():
    return fastq_text()


@pytest.mark.parametrize("kind", ["stored", "level1", "level6", "level9", "fixed", "huffman_only", "rle", "memlevel1", "python_gzip"])
def test_every_block_type_matches_zlib(tmp_path, built, text, kind):
    blob = {"stored": lambda: deflate(text, 0), "level1": lambda: deflate(text, 1), "level6": lambda: deflate(text, 6),
            "level9": lambda: deflate(text, 9), "fixed": lambda: deflate(text, 6, zlib.Z_FIXED),
            "huffman_only": lambda: deflate(text, 6, zlib.Z_HUFFMAN_ONLY), "rle": lambda: deflate(text, 6, zlib.Z_RLE),
            "memlevel1": lambda: deflate(text, 6, memlevel=1),  # small blocks: many dynamic headers
            "python_gzip": lambda: gzip.compress(text, 1)}[kind]()
    p = str(tmp_path / "x.gz")
    open(p, "wb").write(blob)
    got, rc = inflate(p)
    assert rc == 0
    assert got == text


def test_long_codes_binary_runs_and_far_matches(tmp_path, built):
    """skewed byte frequencies give 13-15 bit codes (sub-tables); long runs give distance-1 matches of length 258;
    a repeat 32 KiB back uses the largest distances; random bytes end up in stored blocks"""
    rng = np.random.default_rng(11)
    skew = rng.choice(256, size=400_000, p=np.r_[0.55, 0.25, np.full(254, 0.2 / 254)]).astype(np.uint8).tobytes()
    runs = b"".join(bytes([int(b)]) * int(n) for b, n in zip(rng.integers(0, 256, 300), rng.integers(1, 3000, 300)))
    block = rng.integers(0, 256, 32768 - 7, dtype=np.uint8).tobytes()
    far = block + b"tatajuba" * 3 + block + block[:1000]
    short = b"".join(bytes(rng.integers(65, 69, int(n), dtype=np.uint8)) * 40 for n in rng.integers(2, 8, 200))  # distances 2..7
    for i, data in enumerate([skew, runs, far, short, skew + runs + far + short]):
        for level in (1, 9):
            p = str(tmp_path / f"b{i}_{level}.gz")
            open(p, "wb").write(deflate(data, level))
            got, rc = inflate(p)
            assert rc == 0 and got == data, (i, level)


def test_members_headers_and_what_follows(tmp_path, built, text):
    parts = [text[i:i + 65_280] for i in range(0, 600_000, 65_280)]  # bgzip-sized members
    members = b"".join(deflate(x, 6) for x in parts) + deflate(b"")   # ... and an empty one, as bgzip appends
    # a member with every optional header field (RFC 1952 2.3.1): FEXTRA, FNAME, FCOMMENT, FHCRC
    body = zlib.compress(b"ACGT" * 1000, 6)[2:-4]
    hdr = b"\x1f\x8b\x08\x1e" + b"\0" * 4 + b"\0\x03" + struct.pack("<H", 6) + b"BC\x02\0\x10\0" + b"name.fq\0" + b"a comment\0"
    hdr += struct.pack("<H", zlib.crc32(hdr) & 0xffff)
    full = hdr + body + struct.pack("<II", zlib.crc32(b"ACGT" * 1000), 4000)
    cases = {"members": (members, b"".join(parts)), "header_fields": (full, b"ACGT" * 1000),
             "garbage_after": (deflate(text[:50_000]) + b"\0" * 100 + b"not gzip", text[:50_000]),  # ignored like gzread does
             "empty": (deflate(b""), b"")}
    for name, (blob, want) in cases.items():
        p = str(tmp_path / f"{name}.gz")
        open(p, "wb").write(blob)
        got, rc = inflate(p)
        assert rc == 0 and got == want, name


def test_buffer_boundaries(tmp_path, built):
    """more than one 4 MiB piece: matches reach back across pieces (the 32 KiB history travels with them), a member ends
    exactly on a piece boundary, a stored block spans two pieces"""
    rng = np.random.default_rng(3)
    unit = rng.integers(65, 85, 30_000, dtype=np.uint8).tobytes()
    data = (unit * 300)[: (4 << 20) * 2 + 12345]
    exact = (unit * 150)[: 4 << 20]
    for i, blob_want in enumerate([(deflate(data, 6), data), (deflate(exact, 6) + deflate(unit, 1), exact + unit), (deflate(data, 0), data)]):
        p = str(tmp_path / f"big{i}.gz")
        open(p, "wb").write(blob_want[0])
        got, rc = inflate(p)
        assert rc == 0 and got == blob_want[1], i


def test_damaged_files_are_reported(tmp_path, built, text):
    good = deflate(text, 6)
    flipped = bytearray(good)
    flipped[len(good) // 2] ^= 0x10
    bad_crc = bytearray(good)
    bad_crc[-6] ^= 1
    bad_size = bytearray(good)
    bad_size[-1] ^= 1
    over = bytearray(deflate(b"A" * 100, 6, zlib.Z_FIXED))
    cases = {"truncated": good[: len(good) * 2 // 3], "truncated_in_trailer": good[:-3], "flipped_bit": bytes(flipped),
             "bad_crc": bytes(bad_crc), "bad_size": bytes(bad_size), "reserved_block_type": good[:10] + b"\x07" + good[11:]}
    for name, blob in cases.items():
        p = str(tmp_path / f"{name}.gz")
        open(p, "wb").write(blob)
        got, rc = inflate(p)
        assert rc == TJB_ERR_IO, name
        if name.startswith("truncated") or name.startswith("bad"):
            assert text.startswith(got) or got == text, name  # everything decodable was delivered first, nothing invented
    del over


def test_not_gzip_is_left_to_zlib(tmp_path, built):
    lib = _lib.load()
    p = str(tmp_path / "plain.fq")
    open(p, "wb").write(b"@r\nACGT\n+\nIIII\n" * 10)
    assert not lib.tjb_gz_open(os.fsencode(p))
    assert not lib.tjb_gz_open(os.fsencode(str(tmp_path / "missing.gz")))
    assert not lib.tjb_gz_open(os.fsencode(str(tmp_path)))


def test_reader_gives_the_same_reads_through_both_inflaters(tmp_path, built, text, monkeypatch):
    from test_fastq_reader import decode
    p = str(tmp_path / "r.fastq.gz")
    open(p, "wb").write(b"".join(deflate(text[i:i + 300_000], 1) for i in range(0, len(text), 300_000)))
    ours = decode(p)
    monkeypatch.setenv("TJB_ZLIB", "1")
    assert decode(p) == ours
    assert ours[1] == 4000
