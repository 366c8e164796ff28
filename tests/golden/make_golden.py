"""Generates the committed golden vectors from the reference's own code compiled in this container
(oracle/_ref/libref_hotpath.so, built by oracle/Makefile from /root/reference).  Run from the repo root:

    python tests/golden/make_golden.py

Outputs (small, committed):
  err1750956.reads.txt.gz   the reference fixture tests/files/err1750956.fastq.gz decoded to "SEQ\\n" lines
  fixture_k{K}_m{M}.npz     BWA-boundary state of the compiled reference on that fixture
  synth_k{K}_m{M}.npz       same on a seeded synthetic sample (tatajuba_b200.synth.sample_buffer(60000, 20, seed=21))
  adversarial_k{K}.npz      same on synth.adversarial_reads(K) x3
  nrun_k{K}_m{M}.npz        same on synth.nrun_reads(K, seed=17): runs of N / n / - / R ... that FOLLOW a tract of the same read
                            (the defined half of src/hopo_counter.c:233-248: the previous tract's context is re-emitted)
"""
import gzip
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from tatajuba_b200 import synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/tests/files/err1750956.fastq.gz"


def decode_fastq(path):
    out = []
    with gzip.open(path, "rb") as f:
        lines = f.read().split(b"\n")
    for i in range(0, len(lines) - 3, 4):
        assert lines[i].startswith(b"@") and lines[i + 2].startswith(b"+")
        out.append(lines[i + 1])
    return b"\n".join(out) + b"\n"


def save(name, b, **extra):
    np.savez_compressed(os.path.join(HERE, name), status=b.status, n_raw=b.n_raw, coverage=b.coverage, elem=b.elem,
                        idx_initial=b.idx_initial, idx_final=b.idx_final, **extra)


def main():
    ref = pyoracle.Reference()
    buf = decode_fastq(SRC)
    with gzip.GzipFile(os.path.join(HERE, "err1750956.reads.txt.gz"), "wb", mtime=0) as f:
        f.write(buf)
    for k, m in [(25, 4), (32, 3), (2, 1), (13, 2)]:
        from_file = ref.run_files([SRC], k, m, 5, True)
        from_buf = ref.run_buffer(buf, k, m, 5, True)
        pyoracle.assert_same_boundary(from_file, from_buf, "file vs decoded buffer")
        save(f"fixture_k{k}_m{m}.npz", from_file, k=k, m=m, min_coverage=5, remove_biased=1)
    sbuf = synth.sample_buffer(60_000, 20, seed=21).tobytes()
    for k, m, rb in [(25, 4, 1), (32, 3, 1), (25, 4, 0)]:
        save(f"synth_k{k}_m{m}_rb{rb}.npz", ref.run_buffer(sbuf, k, m, 5, bool(rb)), k=k, m=m, min_coverage=5, remove_biased=rb)
    for k in (25, 32):
        abuf = b"\n".join(synth.adversarial_reads(k) * 3) + b"\n"
        save(f"adversarial_k{k}.npz", ref.run_buffer(abuf, k, 2, 2, False), k=k, m=2, min_coverage=2, remove_biased=0)
    for k, m in [(25, 4), (32, 3), (7, 2)]:
        nbuf = b"\n".join(synth.nrun_reads(k, seed=17)) + b"\n"
        save(f"nrun_k{k}_m{m}.npz", ref.run_buffer(nbuf, k, m, 1, False), k=k, m=m, min_coverage=1, remove_biased=0)


if __name__ == "__main__":
    main()
