"""Golden vectors of the cross-sample merge, made by the COMPILED REFERENCE (oracle/_ref: the reference's own
g_tract_vector_concatenate_tracts + context_histograms_overlap).  Run in the build container:
    python tests/golden/make_golden_merge.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
from merge_cases import CASES, make_tables  # noqa: E402
from oracle import pyoracle  # noqa: E402

out = {}
for seed, ns, nsites in CASES:
    for lev in (2, 0):
        s, i, t = pyoracle.reference_merge_tracts(make_tables(seed, ns, nsites), lev)
        out[f"c{seed}_l{lev}_sample"], out[f"c{seed}_l{lev}_index"], out[f"c{seed}_l{lev}_tract"] = s, i, t
np.savez_compressed(os.path.join(HERE, "merge", "merge_tracts.npz"), **out)
print("wrote", len(out), "arrays")
