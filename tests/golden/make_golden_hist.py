"""Golden vector for tests/test_context_histograms.py: the compiled reference (oracle/_ref/libref_pipeline.so: unmodified
sources + stand-in mapper) run end to end on a seeded sample.  Run from the repo root: python tests/golden/make_golden_hist.py"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from tatajuba_b200 import synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    k, m, d, lev = 25, 4, 1, 2
    with tempfile.TemporaryDirectory() as tmp:
        base = synth.random_genome(60_000, 5)
        fa, gff = os.path.join(tmp, "ref.fasta"), os.path.join(tmp, "ref.gff3")
        synth.write_reference(fa, gff, base)
        g = synth.inject_tract_polymorphism(base, 0.1, seed=105)
        buf = synth.paired_reads(g, 40, seed=1005, error_rate=0.01).reshape(-1)
        elem, table = pyoracle.reference_pipeline(buf, fa, gff, k, m, 3, True, d, lev)
    np.savez_compressed(os.path.join(HERE, "hist", "context_histograms.npz"), elem=elem, k=k, m=m, d=d, lev=lev, **table)
    print(len(elem), "mapped elements ->", table["n_hist"], "histograms")


if __name__ == "__main__":
    main()
