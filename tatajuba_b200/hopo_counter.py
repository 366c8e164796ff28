"""Host-side mirror of the reference's hopo_counter interface (src/hopo_counter.h:51-81) over libtjb_gpu.

    hc = new_or_append_hopo_counter_from_file(None, "s1_R1.fastq.gz", opt)   # src/hopo_counter.c:135
    hc = new_or_append_hopo_counter_from_file(hc,   "s1_R2.fastq.gz", opt)   # paired end: same counter
    finalise_hopo_counter(hc)                                                # src/hopo_counter.c:339 (up to :415)
    hc.elem, hc.n_elem, hc.idx_initial, hc.idx_final, hc.n_idx, hc.coverage  # state at the BWA boundary (:523)

Everything is computed by the CUDA library; nothing here touches oracle/ and there is no CPU fallback.
"""
import ctypes as C
import os
import warnings

import numpy as np

from . import _lib
from ._lib import TjbCounterOut, TjbKernelStat, TjbOpts

# numpy view of the reference's 40-byte hopo_element (src/hopo_counter.h:34-49)
HOPO_DTYPE = np.dtype([("ctx0", "<u8"), ("ctx1", "<u8"), ("bits", "<u8"), ("read_offset", "<i4"),
                       ("loc_ref_id", "<i4"), ("loc_pos", "<i4"), ("loc_last", "<i4")])

TJB_OK, TJB_NO_HT, TJB_NO_HT_AFTER_QC, TJB_NO_CTX_ABOVE_COVERAGE = 0, 1, 2, 3


class TatajubaOptions:
    """the fields of tatajuba_options_t the path reads (src/hopo_counter.h:20-32); defaults as src/main.c:56-60"""

    def __init__(self, kmer_size=25, min_tract_size=4, min_coverage=5, remove_biased=True, paired_end=False):
        self.kmer_size, self.min_tract_size, self.min_coverage = kmer_size, min_tract_size, min_coverage
        self.remove_biased, self.paired_end = remove_biased, paired_end

    def c_struct(self):
        return TjbOpts(self.kmer_size, self.min_tract_size, self.min_coverage, int(self.remove_biased),
                       int(self.paired_end))


class TjbError(RuntimeError):
    pass


class Context:
    """one tjb_ctx: a (host thread, device) pair with its stream, pinned double buffers and device workspace"""

    def __init__(self, device=0):
        self.lib = _lib.load()
        # the host pool (2-bit packing of host buffers, copy-out of the results) spins while a sample is processed, so
        # ranks that share one host (torchrun: LOCAL_WORLD_SIZE) split the cores between them; with fewer than 9
        # threads the library sends host buffers as they are and uses the pool for the copy-out only
        if "TJB_HOST_PACK" not in os.environ and int(os.environ.get("LOCAL_WORLD_SIZE", "1")) > 1:
            cpus = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
            share = cpus // int(os.environ["LOCAL_WORLD_SIZE"]) - 1
            os.environ["TJB_HOST_PACK"] = str(max(share, 0))
        h = C.c_void_p()
        rc = self.lib.tjb_ctx_create(int(device), C.byref(h))
        if rc < 0:
            raise TjbError(f"tjb_ctx_create(device={device}) failed: {self.lib.tjb_strerror(rc).decode()}")
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.lib.tjb_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc < 0:
            raise TjbError(f"{self.lib.tjb_strerror(rc).decode()}: {self.lib.tjb_ctx_last_error(self.h).decode()}")
        return rc

    def set_profiling(self, level):
        self.lib.tjb_ctx_set_profiling(self.h, int(level))

    def set_result_mode(self, arena):
        """arena=True: result arrays are views of the context's pinned result arena (written by the device, no host
        copy), valid until this context counts its next sample; False (default): malloc()ed copies like the reference's"""
        self._check(self.lib.tjb_ctx_set_result_mode(self.h, 1 if arena else 0))

    def reset_kernel_stats(self):
        self.lib.tjb_ctx_reset_kernel_stats(self.h)

    def launch_count(self):
        return int(self.lib.tjb_ctx_launch_count(self.h))

    def h2d_bytes(self):
        """bytes count_host() has sent over PCIe since the last reset_kernel_stats() (packed transfers send fewer)"""
        return int(self.lib.tjb_ctx_htod_bytes(self.h))

    def stream(self):
        return self.lib.tjb_ctx_stream(self.h)

    def kernel_stats(self):
        arr = (TjbKernelStat * 16)()
        n = self.lib.tjb_ctx_kernel_stats(self.h, arr, 16)
        return [dict(name=arr[i].name.decode(), launches=arr[i].launches, ms_total=arr[i].ms_total,
                     bytes_algorithmic=arr[i].bytes_algorithmic) for i in range(min(n, 16))]

    # the three entry points -------------------------------------------------------------------------
    def count_files(self, files, opt):
        arr = (C.c_char_p * len(files))(*[os.fsencode(f) for f in files])
        out = TjbCounterOut()
        o = opt.c_struct()
        self._check(self.lib.tjb_count_sample_files(self.h, arr, len(files), C.byref(o), C.byref(out)))
        return HopoCounter._from_out(self.lib, out, opt)

    def count_host(self, seqs, opt):
        """seqs: bytes / numpy uint8 / pinned torch uint8 tensor of reads separated by '\\n'"""
        ptr, n = _host_ptr(seqs)
        out = TjbCounterOut()
        o = opt.c_struct()
        self._check(self.lib.tjb_count_sample_host(self.h, ptr, n, C.byref(o), C.byref(out)))
        return HopoCounter._from_out(self.lib, out, opt)

    def count_device(self, d_ptr, n_bytes, opt):
        """d_ptr: integer device address (e.g. tensor.data_ptr()), 16-byte aligned"""
        out = TjbCounterOut()
        o = opt.c_struct()
        self._check(self.lib.tjb_count_sample_device(self.h, C.c_void_p(int(d_ptr)), int(n_bytes), C.byref(o),
                                                     C.byref(out)))
        return HopoCounter._from_out(self.lib, out, opt)


    # one sample in pieces (read chunks on several GPUs): begin -> scan... -> [export -> exchange -> import] -> finalise
    def sample_begin(self, opt, n_bytes_hint=0):
        o = opt.c_struct()
        self._opt = opt
        self._check(self.lib.tjb_sample_begin(self.h, C.byref(o), int(n_bytes_hint)))

    def table_capacity(self):
        return int(self.lib.tjb_ctx_table_capacity(self.h))

    def reserve_entries(self, n):
        self._check(self.lib.tjb_ctx_reserve_entries(self.h, int(n)))

    def sample_scan_device(self, d_ptr, n_bytes):
        self._check(self.lib.tjb_sample_scan_device(self.h, C.c_void_p(int(d_ptr)), int(n_bytes)))

    def sample_export(self, d_ptr=None, max_rows=0):
        """table of the sample so far as rows; without a destination only (n_rows, uint64 words per row) are returned,
        with one (device address, capacity in rows) the rows are copied there as well.  Synchronous."""
        n, w = C.c_int64(0), C.c_int32(0)
        self._check(self.lib.tjb_sample_export(self.h, C.c_void_p(int(d_ptr)) if d_ptr else None, int(max_rows),
                                               C.byref(n), C.byref(w)))
        return int(n.value), int(w.value)

    def sample_import(self, d_ptr, n_rows):
        self._check(self.lib.tjb_sample_import(self.h, C.c_void_p(int(d_ptr)), int(n_rows)))

    def sample_finalise(self):
        out = TjbCounterOut()
        self._check(self.lib.tjb_sample_finalise(self.h, C.byref(out)))
        return HopoCounter._from_out(self.lib, out, self._opt)

    def query_rows_device(self, d_ptr, max_rows):
        """query rows of the sample counted last, written on the device; returns the number of rows"""
        return self._check(self.lib.tjb_query_rows_device(self.h, C.c_void_p(int(d_ptr)), int(max_rows)))


def _host_ptr(seqs):
    if isinstance(seqs, (bytes, bytearray)):
        a = np.frombuffer(seqs, dtype=np.uint8)
        _host_ptr.keep = a
        return C.c_void_p(a.ctypes.data), a.size
    if isinstance(seqs, np.ndarray):
        assert seqs.dtype == np.uint8 and seqs.flags["C_CONTIGUOUS"]
        return C.c_void_p(seqs.ctypes.data), seqs.size
    # torch tensor (CPU, possibly pinned)
    assert not seqs.is_cuda and seqs.is_contiguous()
    return C.c_void_p(seqs.data_ptr()), seqs.numel()


class _OutOwner:
    """keeps a tjb_counter_out alive while numpy views of its arrays exist; free()s it afterwards"""

    def __init__(self, lib, out):
        self.lib, self.out = lib, out

    def __del__(self):
        try:
            self.lib.tjb_counter_out_free(C.byref(self.out))
        except Exception:
            pass


class HopoCounter:
    """struct hopo_counter_struct at the BWA boundary (src/hopo_counter.h:51-59)"""

    def __init__(self, opt):
        self.opt = opt
        self.files = []
        self.name = None
        self.status = None
        self.elem = np.zeros(0, dtype=HOPO_DTYPE)
        self.idx_initial = np.zeros(0, np.int32)
        self.idx_final = np.zeros(0, np.int32)
        self.n_elem = self.n_idx = self.coverage = 0
        self.ref_start = 0
        self.finalised = False

    @classmethod
    def _from_out(cls, lib, out, opt):
        hc = cls(opt)
        hc.status = out.status
        hc.n_elem, hc.n_idx, hc.coverage = out.n_elem, out.n_idx, out.coverage
        # zero-copy views of the malloc()'ed arrays; the owner frees them when the last view is gone
        owner = _OutOwner(lib, out)
        if out.n_elem:
            raw = (C.c_uint8 * (40 * out.n_elem)).from_address(out.elem)
            raw._owner = owner
            hc.elem = np.ctypeslib.as_array(raw).view(HOPO_DTYPE)
        if out.n_idx:
            for name in ("idx_initial", "idx_final"):
                arr = (C.c_int32 * out.n_idx).from_address(C.addressof(getattr(out, name).contents))
                arr._owner = owner
                setattr(hc, name, np.ctypeslib.as_array(arr))
        hc.name = out.name.decode() if out.name else None
        for f in ("n_reads", "n_bases", "n_raw", "n_distinct", "n_skipped_nonacgt", "ms_decode", "ms_device"):
            setattr(hc, f, getattr(out, f))
        hc.finalised = True
        return hc


_default_ctx = {}


def _ctx(device=0):
    if device not in _default_ctx:
        _default_ctx[device] = Context(device)
    return _default_ctx[device]


def new_or_append_hopo_counter_from_file(hc, filename, opt):
    """src/hopo_counter.c:135-157: the first file names the sample; appending after finalise is an error (:152)"""
    if hc is None:
        hc = HopoCounter(opt)
        hc.name = filename
    if hc.finalised:
        raise TjbError("This counter has been compared to another; cannot add more reads to it")
    if len(hc.files) >= 2:
        raise TjbError("at most two files (R1, R2) per sample")
    hc.files.append(filename)
    return hc


def finalise_hopo_counter(hc, device=0):
    """src/hopo_counter.c:339-415 (everything before the mapper call): fills hc in place and warns like the
    reference on the three "sample will be excluded" paths"""
    res = _ctx(device).count_files(hc.files, hc.opt)
    name, files = hc.name, hc.files
    hc.__dict__.update(res.__dict__)
    hc.name, hc.files = name, files
    pair = " together with its pair." if hc.opt.paired_end else "."
    if hc.status == TJB_NO_HT:
        warnings.warn(f"No HTs were found in sample {name}{pair[:-1]}, not even before QC. This sample will be excluded.")
    elif hc.status == TJB_NO_HT_AFTER_QC:
        warnings.warn(f"No HTs found in sample {name} after QC. This sample will be excluded.")
    elif hc.status == TJB_NO_CTX_ABOVE_COVERAGE:
        warnings.warn(f"From the HTs found in sample {name}, not a single one has coverage higher than "
                      f"{hc.opt.min_coverage}. This sample will be excluded.")
    return hc
