"""ctypes bindings of the two CPU checkers (TEST INFRASTRUCTURE ONLY).

* ``Restatement``  -> oracle/_build/libtjb_oracle.so (our plain-C restatement, oracle/tjb_oracle.c)
* ``Reference``    -> oracle/_ref/libref_hotpath.so  (the unmodified reference sources, oracle/Makefile)

Only tests/, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may import
this module; nothing under ``tatajuba_b200/`` does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "_build", "libtjb_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libref_hotpath.so")

# numpy view of the reference's 40-byte hopo_element (src/hopo_counter.h:34-49)
HOPO_DTYPE = np.dtype([("ctx0", "<u8"), ("ctx1", "<u8"), ("bits", "<u8"), ("read_offset", "<i4"),
                       ("loc_ref_id", "<i4"), ("loc_pos", "<i4"), ("loc_last", "<i4")])
assert HOPO_DTYPE.itemsize == 40
BITS_MASK = np.uint64((1 << 52) - 1)  # bits 52..63 of the bit-field word are padding (uninitialised in the reference)


def build(ref: bool = True) -> None:
    """make the restatement, and the compiled reference when /root/reference is present"""
    subprocess.run(["make", "-s", "-C", HERE, "oracle"] + (["ref"] if ref else []), check=True)


def unpack_bits(bits):
    """bit-field word -> dict of signed field values (base, length, count, mismatches, multi, neg_strand, canon_flag)"""
    bits = np.asarray(bits, dtype=np.uint64)

    def field(lo, width, signed=True):
        v = ((bits >> np.uint64(lo)) & np.uint64((1 << width) - 1)).astype(np.int64)
        if signed:
            v = (v ^ (1 << (width - 1))) - (1 << (width - 1))
        return v
    return dict(base=field(0, 2), length=field(2, 10), count=field(12, 20), mismatches=field(32, 12),
                multi=field(44, 3), neg_strand=field(47, 2), canon_flag=field(49, 3))


class _RefDump(C.Structure):
    _fields_ = [("status", C.c_int), ("n_raw", C.c_int), ("n_elem", C.c_int), ("n_idx", C.c_int),
                ("coverage", C.c_int), ("elem", C.c_void_p), ("idx_initial", C.POINTER(C.c_int)),
                ("idx_final", C.POINTER(C.c_int)), ("secs_scan", C.c_double), ("secs_finalise", C.c_double)]


class _Rec(C.Structure):
    _fields_ = [("ctx", C.c_uint64 * 2), ("base", C.c_int32), ("length", C.c_int32), ("count", C.c_int32),
                ("flag", C.c_int32)]


REC_DTYPE = np.dtype([("ctx0", "<u8"), ("ctx1", "<u8"), ("base", "<i4"), ("length", "<i4"), ("count", "<i4"),
                      ("flag", "<i4")])
assert REC_DTYPE.itemsize == C.sizeof(_Rec)


class _Result(C.Structure):
    _fields_ = [("status", C.c_int), ("n_raw", C.c_int64), ("n_skipped_nonacgt", C.c_int64),
                ("n_distinct", C.c_int64), ("n_elem", C.c_int32), ("n_idx", C.c_int32), ("coverage", C.c_int32),
                ("elem", C.POINTER(_Rec)), ("idx_initial", C.POINTER(C.c_int32)),
                ("idx_final", C.POINTER(C.c_int32))]


class Boundary:
    """State at the BWA boundary, in comparable form."""

    def __init__(self, status, n_raw, elem40, idx_initial, idx_final, coverage, n_skipped=0, secs=(0.0, 0.0)):
        self.status = int(status)
        self.n_raw = int(n_raw)
        self.elem = elem40  # structured array, HOPO_DTYPE
        self.idx_initial = idx_initial
        self.idx_final = idx_final
        self.coverage = int(coverage)
        self.n_skipped_nonacgt = int(n_skipped)
        self.secs_scan, self.secs_finalise = secs

    @property
    def n_elem(self):
        return len(self.elem)

    @property
    def n_idx(self):
        return len(self.idx_initial)

    def canonical(self):
        """tuple of plain arrays with padding bits masked (what "bit-exact" means for this struct)"""
        e = self.elem
        return (self.status, self.n_raw, self.coverage, e["ctx0"].tolist(), e["ctx1"].tolist(),
                (e["bits"] & BITS_MASK).tolist(), e["read_offset"].tolist(), e["loc_ref_id"].tolist(),
                e["loc_pos"].tolist(), e["loc_last"].tolist(), self.idx_initial.tolist(), self.idx_final.tolist())


def assert_same_boundary(a: "Boundary", b: "Boundary", what=""):
    assert a.status == b.status, f"{what}: status {a.status} != {b.status}"
    assert a.n_raw == b.n_raw, f"{what}: n_raw {a.n_raw} != {b.n_raw}"
    if a.status != 0:
        return
    assert a.n_elem == b.n_elem, f"{what}: n_elem {a.n_elem} != {b.n_elem}"
    for f in ("ctx0", "ctx1", "read_offset", "loc_ref_id", "loc_pos", "loc_last"):
        assert np.array_equal(a.elem[f], b.elem[f]), f"{what}: field {f} differs"
    assert np.array_equal(a.elem["bits"] & BITS_MASK, b.elem["bits"] & BITS_MASK), f"{what}: bit-fields differ"
    assert np.array_equal(a.idx_initial, b.idx_initial), f"{what}: idx_initial differs"
    assert np.array_equal(a.idx_final, b.idx_final), f"{what}: idx_final differs"
    assert a.coverage == b.coverage, f"{what}: coverage {a.coverage} != {b.coverage}"


class Reference:
    """the compiled, unmodified reference"""

    def __init__(self):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO)
        self.lib = C.CDLL(REF_SO)
        self.lib.ref_sizeof_hopo_element.restype = C.c_size_t
        assert self.lib.ref_sizeof_hopo_element() == 40

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    def _collect(self, d):
        n, g = (d.n_elem, d.n_idx) if d.status == 0 else (0, 0)
        elem = np.zeros(n, dtype=HOPO_DTYPE)
        if n:
            C.memmove(elem.ctypes.data, d.elem, 40 * n)
        ii = np.array(d.idx_initial[:g], dtype=np.int32) if g else np.zeros(0, np.int32)
        jj = np.array(d.idx_final[:g], dtype=np.int32) if g else np.zeros(0, np.int32)
        out = Boundary(d.status, d.n_raw, elem, ii, jj, d.coverage if d.status == 0 else 0,
                       secs=(d.secs_scan, d.secs_finalise))
        self.lib.ref_dump_free(C.byref(d))
        return out

    def run_files(self, files, k=25, m=4, min_coverage=5, remove_biased=True):
        arr = (C.c_char_p * len(files))(*[os.fsencode(f) for f in files])
        d = _RefDump()
        self.lib.ref_hotpath_run_files(arr, len(files), k, m, min_coverage, int(remove_biased), C.byref(d))
        return self._collect(d)

    def run_buffer(self, buf, k=25, m=4, min_coverage=5, remove_biased=True):
        """buf: bytes, or a contiguous numpy uint8 array (passed without a copy)"""
        d = _RefDump()
        if isinstance(buf, np.ndarray):
            assert buf.dtype == np.uint8 and buf.flags["C_CONTIGUOUS"]
            b, n = C.c_void_p(buf.ctypes.data if buf.size else 0), buf.size
        else:
            b, n = ((C.c_char * len(buf)).from_buffer_copy(buf) if len(buf) else (C.c_char * 1)()), len(buf)
        self.lib.ref_hotpath_run_buffer(b, C.c_size_t(n), k, m, min_coverage, int(remove_biased), C.byref(d))
        return self._collect(d)


class PatchedReference(Reference):
    """the reference sources with integration/tatajuba-libtjb_gpu.patch applied, behind the same boundary-capturing shim
    (oracle/_ref/libref_hotpath_gpu.so): what the mapper sees when the binding counts on the GPU"""
    SO = os.path.join(HERE, "_ref", "libref_hotpath_gpu.so")

    def __init__(self):
        if not os.path.exists(self.SO):
            raise FileNotFoundError(self.SO)
        self.lib = C.CDLL(self.SO)
        self.lib.ref_sizeof_hopo_element.restype = C.c_size_t
        assert self.lib.ref_sizeof_hopo_element() == 40

    @staticmethod
    def available():
        return os.path.exists(PatchedReference.SO)

    def run_files_gpu(self, files, k=25, m=4, min_coverage=5, remove_biased=True, device=0):
        arr = (C.c_char_p * len(files))(*[os.fsencode(f) for f in files])
        d = _RefDump()
        self.lib.ref_hotpath_run_files_gpu(arr, len(files), k, m, min_coverage, int(remove_biased), int(device), C.byref(d))
        return self._collect(d)


class Restatement:
    """our plain-C restatement"""

    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        self.lib = C.CDLL(ORACLE_SO)
        self.lib.tjbo_scan_buffer.restype = C.c_int64

    def run_buffer(self, buf, k=25, m=4, min_coverage=5, remove_biased=True):
        buf = np.frombuffer(buf, dtype=np.uint8) if not isinstance(buf, np.ndarray) else buf
        r = _Result()
        self.lib.tjbo_count_buffer(C.c_void_p(buf.ctypes.data), C.c_size_t(buf.size), k, m, min_coverage,
                                   int(remove_biased), C.byref(r))
        n, g = (r.n_elem, r.n_idx) if r.status == 0 else (0, 0)
        elem = np.zeros(n, dtype=HOPO_DTYPE)
        if n:
            self.lib.tjbo_pack_hopo_elements(r.elem, C.c_int64(n), C.c_void_p(elem.ctypes.data))
        ii = np.array(r.idx_initial[:g], dtype=np.int32) if g else np.zeros(0, np.int32)
        jj = np.array(r.idx_final[:g], dtype=np.int32) if g else np.zeros(0, np.int32)
        out = Boundary(r.status, r.n_raw, elem, ii, jj, r.coverage if r.status == 0 else 0, r.n_skipped_nonacgt)
        out.n_distinct = int(r.n_distinct)
        self.lib.tjbo_free_result(C.byref(r))
        return out

    def scan_buffer(self, buf, k=25, m=4):
        """raw scan records in emission order -> (structured array REC_DTYPE, n_skipped)"""
        buf = np.frombuffer(buf, dtype=np.uint8) if not isinstance(buf, np.ndarray) else buf
        p = C.POINTER(_Rec)()
        skipped = C.c_int64(0)
        n = self.lib.tjbo_scan_buffer(C.c_void_p(buf.ctypes.data), C.c_size_t(buf.size), k, m, C.byref(p),
                                      C.byref(skipped))
        out = np.zeros(n, dtype=REC_DTYPE)
        if n:
            C.memmove(out.ctypes.data, p, n * REC_DTYPE.itemsize)
        self.lib.tjbo_free(p)
        return out, int(skipped.value)


# ---- cross-sample exchange (SURVEY.md section 8 e): numpy restatement, test infrastructure only -----------------
def query_rows_oracle(elem40, idx_initial):
    """one row per indexed group from its first = longest record (reference src/hopo_counter.c:513-520):
    {context[0], context[1], base | length_field << 8}"""
    e = np.asarray(elem40).view(np.uint64).reshape(-1, 5)[np.asarray(idx_initial, dtype=np.int64)]
    rows = np.empty((len(e), 3), np.uint64)
    rows[:, 0], rows[:, 1] = e[:, 0], e[:, 1]
    rows[:, 2] = (e[:, 2] & np.uint64(1)) | (((e[:, 2] >> np.uint64(2)) & np.uint64(0x3ff)) << np.uint64(8))
    return rows


def union_oracle(tables):
    """distinct rows over all tables in the reference's descending (base, ctx0, ctx1, length) order
    (compare_hopo_element_decreasing, src/hopo_counter.c:28-38; length is a signed 10-bit field) and, per table,
    the index of each row in the union"""
    rows = np.concatenate([np.asarray(t, np.uint64).reshape(-1, 3) for t in tables]) if tables else np.zeros((0, 3), np.uint64)
    if len(rows) == 0:
        return rows, [np.zeros(0, np.int32) for _ in tables]
    base = (rows[:, 2] & np.uint64(1)).astype(np.int64)
    lf = ((rows[:, 2] >> np.uint64(8)) & np.uint64(0x3ff)).astype(np.int64)
    length = np.where(lf >= 512, lf - 1024, lf)
    order = np.lexsort((length, rows[:, 1], rows[:, 0], base))[::-1]  # last key is the primary one
    srt = rows[order]
    new = np.ones(len(srt), bool)
    new[1:] = np.any(srt[1:] != srt[:-1], axis=1)
    seg = np.cumsum(new) - 1
    union = srt[new]
    index = np.empty(len(rows), np.int32)
    index[order] = seg
    out, off = [], 0
    for t in tables:
        out.append(index[off:off + len(t)])
        off += len(t)
    return union, out


def reference_merge_tracts(tables, levenshtein_distance=2):
    """the compiled reference's own cross-sample merge (oracle/shim ref_merge_tracts); same inputs / outputs as
    tatajuba_b200.genome_set.merge_tracts"""
    lib = C.CDLL(REF_SO)
    n_hist = np.array([len(t["location"]) for t in tables], np.int32)
    total = int(n_hist.sum())
    cat = lambda k: np.ascontiguousarray(np.concatenate([np.asarray(t[k], np.int32) for t in tables]))  # noqa: E731
    loc, base, mlen = cat("location"), cat("base"), cat("mode_context_length")
    stride = max(t["name"].shape[1] for t in tables if len(t["location"]))
    names = np.zeros((total, stride), np.uint8)
    off = 0
    for t in tables:
        n = len(t["location"])
        if n:
            names[off:off + n, : t["name"].shape[1]] = np.asarray(t["name"]).view(np.uint8).reshape(n, -1)
        off += n
    out = [np.empty(total, np.int32) for _ in range(3)]
    p = lambda a: C.c_void_p(a.ctypes.data)  # noqa: E731
    lib.ref_merge_tracts.restype = C.c_int
    rc = lib.ref_merge_tracts(C.c_int(len(tables)), p(n_hist), p(loc), p(base), p(mlen), p(names), C.c_int(stride),
                              C.c_int(levenshtein_distance), p(out[0]), p(out[1]), p(out[2]))
    assert rc == total
    return tuple(out)


# ---- per-sample context histograms (rows 11-13): the unmodified reference end to end with the stand-in mapper ---------
PIPELINE_SO = os.path.join(HERE, "_ref", "libref_pipeline.so")


def reference_pipeline(buf, fasta, gff, k=25, m=4, min_coverage=5, remove_biased=True, max_distance_per_flank=1,
                       levenshtein_distance=2):
    """reads -> new_genomic_context_list() of the compiled reference (oracle/shim/ref_pipeline.c).  Returns
    (mapped elements in location order as a HOPO_DTYPE array, dict of the histogram table) or (None, None) when the
    reference excludes the sample"""
    from tatajuba_b200._lib import TjbContextTable  # layout of the table only

    class _Dump(C.Structure):
        _fields_ = [("n_mapped", C.c_int), ("elem", C.c_void_p), ("hist", TjbContextTable)]
    lib = C.CDLL(PIPELINE_SO)
    buf = np.frombuffer(buf, dtype=np.uint8) if not isinstance(buf, np.ndarray) else buf
    d = _Dump()
    rc = lib.ref_pipeline_run(C.c_void_p(buf.ctypes.data), C.c_size_t(buf.size), os.fsencode(fasta), os.fsencode(gff), k, m,
                              min_coverage, int(remove_biased), max_distance_per_flank, levenshtein_distance, C.byref(d))
    if rc < 0:
        return None, None
    elem = np.zeros(d.n_mapped, dtype=HOPO_DTYPE)
    if d.n_mapped:
        C.memmove(elem.ctypes.data, d.elem, 40 * d.n_mapped)
    table = d.hist.to_numpy()
    lib.ref_pipeline_free(C.byref(d))
    return elem, table
