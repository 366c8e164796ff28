"""CPU-only: the restatement against the committed golden vectors (generated from the compiled reference by
tests/golden/make_golden.py), and the C-ABI library: loads, exports every declared symbol, fails loudly without a GPU."""
import ctypes as C
import glob
import gzip
import os
import re

import numpy as np
import pytest

from oracle import pyoracle
from tatajuba_b200 import _lib, synth

GOLD = os.path.join(os.path.dirname(__file__), "golden")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def golden_input(name):
    if name.startswith("fixture"):
        with gzip.open(os.path.join(GOLD, "err1750956.reads.txt.gz"), "rb") as f:
            return f.read()
    if name.startswith("synth"):
        return synth.sample_buffer(60_000, 20, seed=21).tobytes()
    k = int(re.search(r"_k(\d+)", name).group(1))
    if name.startswith("nrun"):
        return b"\n".join(synth.nrun_reads(k, seed=17)) + b"\n"
    return b"\n".join(synth.adversarial_reads(k) * 3) + b"\n"


def load_golden(path):
    z = np.load(path)
    b = pyoracle.Boundary(int(z["status"]), int(z["n_raw"]), z["elem"], z["idx_initial"], z["idx_final"], int(z["coverage"]))
    return b, int(z["k"]), int(z["m"]), int(z["min_coverage"]), bool(int(z["remove_biased"]))


GOLDENS = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLD, "*.npz")))


def test_goldens_exist():
    assert len(GOLDENS) >= 12


@pytest.mark.parametrize("name", GOLDENS)
def test_restatement_matches_golden(restatement, name):
    want, k, m, cov, rb = load_golden(os.path.join(GOLD, name))
    got = restatement.run_buffer(golden_input(name), k, m, cov, rb)
    pyoracle.assert_same_boundary(got, want, name)


def test_abi_exports_every_declared_symbol(built):
    hdr = open(os.path.join(ROOT, "include", "tjb_gpu.h")).read()
    declared = set(re.findall(r"\b(tjb_[a-z_]+)\s*\(", hdr))
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    lib = _lib.load()
    for sym in declared:
        assert hasattr(lib, sym), sym
    assert b"sm_100a" in lib.tjb_version()
    assert lib.tjb_strerror(-1).startswith(b"CUDA error")


def test_struct_sizes_match_header(built):
    assert C.sizeof(_lib.TjbOpts) == 20
    assert C.sizeof(_lib.TjbCounterOut) == 16 + 4 * 8 + 5 * 8 + 2 * 8 + 8
    assert C.sizeof(_lib.TjbKernelStat) == 48 + 3 * 8


def test_no_cpu_fallback_without_gpu(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _lib.load()
    h = C.c_void_p()
    assert lib.tjb_ctx_create(0, C.byref(h)) == -1  # TJB_ERR_CUDA: the product path refuses to run on the CPU
    out = _lib.TjbCounterOut()
    opt = _lib.TjbOpts(25, 4, 5, 1, 0)
    files = (C.c_char_p * 1)(b"/nonexistent.fastq")
    assert lib.tjb_count_sample(files, 1, C.byref(opt), 0, C.byref(out)) == -1


def test_product_never_imports_oracle():
    for p in glob.glob(os.path.join(ROOT, "tatajuba_b200", "**", "*"), recursive=True):
        if os.path.isfile(p) and p.endswith((".py", ".cu", ".cuh", ".c", ".h")):
            src = open(p, errors="ignore").read()
            assert "pyoracle" not in src and "tjb_oracle" not in src and "ref_hotpath" not in src, p
