"""ctypes binding of tatajuba_b200/lib/libtjb_gpu.so (C ABI: include/tjb_gpu.h).

The library is the product; this module only loads it.  There is no Python or CPU fallback: if the shared
library is missing, loading raises, and every compute entry point fails loudly without a B200-class GPU.
"""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TJB_LIB") or os.path.join(HERE, "lib", "libtjb_gpu.so")  # TJB_LIB: kernel-variant experiments
CSRC = os.path.join(HERE, "csrc")

# every symbol include/tjb_gpu.h declares (checked by tests/test_abi.py)
EXPORTS = [
    "tjb_version", "tjb_strerror", "tjb_ctx_last_error", "tjb_ctx_create", "tjb_ctx_destroy", "tjb_ctx_stream",
    "tjb_ctx_set_profiling", "tjb_ctx_kernel_stats", "tjb_ctx_reset_kernel_stats", "tjb_ctx_launch_count", "tjb_ctx_htod_bytes",
    "tjb_count_sample_files", "tjb_count_sample_host", "tjb_count_sample_device", "tjb_count_sample",
    "tjb_counter_out_free", "tjb_ctx_set_result_mode", "tjb_device_count", "tjb_sample_begin", "tjb_sample_scan_device",
    "tjb_ctx_table_capacity", "tjb_ctx_reserve_entries",
    "tjb_sample_export", "tjb_sample_import", "tjb_sample_finalise", "tjb_query_rows_device", "tjb_query_rows", "tjb_union_queries_device", "tjb_merge_tracts", "tjb_context_histograms", "tjb_context_table_free", "tjb_reader_open", "tjb_reader_fill", "tjb_reader_take_long", "tjb_reader_close",
    "tjb_gz_open", "tjb_gz_next", "tjb_gz_close", "tjb_tract_rows_noalloc", "tjb_noalloc_max_bytes",
]


class TjbOpts(C.Structure):
    _fields_ = [("kmer_size", C.c_int32), ("min_tract_size", C.c_int32), ("min_coverage", C.c_int32),
                ("remove_biased", C.c_int32), ("paired_end", C.c_int32)]


class TjbCounterOut(C.Structure):
    _fields_ = [("status", C.c_int32), ("n_elem", C.c_int32), ("n_idx", C.c_int32), ("coverage", C.c_int32),
                ("elem", C.c_void_p), ("idx_initial", C.POINTER(C.c_int32)), ("idx_final", C.POINTER(C.c_int32)),
                ("name", C.c_char_p), ("n_reads", C.c_int64), ("n_bases", C.c_int64), ("n_raw", C.c_int64),
                ("n_distinct", C.c_int64), ("n_skipped_nonacgt", C.c_int64), ("ms_decode", C.c_double),
                ("ms_device", C.c_double), ("arena", C.c_int32), ("reserved", C.c_int32)]


class TjbHistOpts(C.Structure):
    _fields_ = [("kmer_size", C.c_int32), ("min_tract_size", C.c_int32), ("max_distance_per_flank", C.c_int32),
                ("levenshtein_distance", C.c_int32)]


_I32P, _U64P = C.POINTER(C.c_int32), C.POINTER(C.c_uint64)


class TjbContextTable(C.Structure):
    INT_ARRAYS = ["location", "loc2d", "base", "neg_strand", "mismatches", "multi", "indel", "integral", "n_context",
                  "mode_context_count", "mode_context_length", "mode_context_id", "h_min", "h_max"]
    _fields_ = ([("n_hist", C.c_int32), ("name_stride", C.c_int32)] + [(n, _I32P) for n in INT_ARRAYS] +
                [("names", C.POINTER(C.c_char)), ("ctx_first", _I32P), ("contexts", _U64P), ("h_first", _I32P),
                 ("h_length", _I32P), ("h_freq", _I32P)])

    def to_numpy(self):
        """copies of the arrays as a dict (names as an (n, stride) uint8 array)"""
        import numpy as np
        n = max(self.n_hist, 0)
        out = {"n_hist": self.n_hist}
        if n == 0:
            return out
        arr = lambda p, cnt, dt: np.ctypeslib.as_array(p, shape=(cnt,)).astype(dt).copy() if cnt else np.zeros(0, dt)  # noqa: E731
        for name in self.INT_ARRAYS:
            out[name] = arr(getattr(self, name), 3 * n if name == "loc2d" else n, np.int32)
        out["names"] = np.ctypeslib.as_array(C.cast(self.names, C.POINTER(C.c_uint8)), shape=(n * self.name_stride,)).reshape(n, self.name_stride).copy()
        out["ctx_first"], out["h_first"] = arr(self.ctx_first, n + 1, np.int32), arr(self.h_first, n + 1, np.int32)
        out["contexts"] = arr(self.contexts, 2 * int(out["ctx_first"][-1]), np.uint64)
        out["h_length"], out["h_freq"] = arr(self.h_length, int(out["h_first"][-1]), np.int32), arr(self.h_freq, int(out["h_first"][-1]), np.int32)
        return out


class TjbKernelStat(C.Structure):
    _fields_ = [("name", C.c_char * 48), ("launches", C.c_int64), ("ms_total", C.c_double),
                ("bytes_algorithmic", C.c_int64)]


def build(verbose: bool = False) -> str:
    """compile the CUDA extension in-tree for sm_100a (nvcc cross-compiles without a GPU)"""
    subprocess.run(["make", "-C", CSRC] + ([] if verbose else ["-s"]), check=True)
    return LIB_PATH


_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback for the tatajuba_b200 hot path)")
    lib = C.CDLL(LIB_PATH)
    vp, cp, i32, i64, sz = C.c_void_p, C.c_char_p, C.c_int, C.c_int64, C.c_size_t
    lib.tjb_version.restype = cp
    lib.tjb_strerror.restype = cp
    lib.tjb_strerror.argtypes = [i32]
    lib.tjb_ctx_last_error.restype = cp
    lib.tjb_ctx_last_error.argtypes = [vp]
    lib.tjb_ctx_create.argtypes = [i32, C.POINTER(vp)]
    lib.tjb_ctx_destroy.argtypes = [vp]
    lib.tjb_ctx_destroy.restype = None
    lib.tjb_ctx_stream.restype = vp
    lib.tjb_ctx_stream.argtypes = [vp]
    lib.tjb_ctx_set_profiling.argtypes = [vp, i32]
    lib.tjb_ctx_kernel_stats.argtypes = [vp, C.POINTER(TjbKernelStat), i32]
    lib.tjb_ctx_reset_kernel_stats.argtypes = [vp]
    lib.tjb_ctx_reset_kernel_stats.restype = None
    lib.tjb_ctx_launch_count.restype = i64
    lib.tjb_ctx_launch_count.argtypes = [vp]
    lib.tjb_ctx_htod_bytes.restype = i64
    lib.tjb_ctx_htod_bytes.argtypes = [vp]
    po, pout = C.POINTER(TjbOpts), C.POINTER(TjbCounterOut)
    lib.tjb_count_sample_files.argtypes = [vp, C.POINTER(cp), i32, po, pout]
    lib.tjb_count_sample_host.argtypes = [vp, vp, sz, po, pout]
    lib.tjb_count_sample_device.argtypes = [vp, vp, sz, po, pout]
    lib.tjb_count_sample.argtypes = [C.POINTER(cp), i32, po, i32, pout]
    lib.tjb_counter_out_free.argtypes = [pout]
    lib.tjb_counter_out_free.restype = None
    lib.tjb_ctx_set_result_mode.argtypes = [vp, i32]
    lib.tjb_sample_begin.argtypes = [vp, po, sz]
    lib.tjb_ctx_table_capacity.restype = i64
    lib.tjb_ctx_table_capacity.argtypes = [vp]
    lib.tjb_ctx_reserve_entries.argtypes = [vp, i64]
    lib.tjb_sample_scan_device.argtypes = [vp, vp, sz]
    lib.tjb_sample_export.argtypes = [vp, vp, i64, C.POINTER(i64), C.POINTER(C.c_int32)]
    lib.tjb_sample_import.argtypes = [vp, vp, i64]
    lib.tjb_sample_finalise.argtypes = [vp, pout]
    lib.tjb_query_rows_device.argtypes = [vp, vp, i64]
    lib.tjb_query_rows.argtypes = [pout, vp]
    lib.tjb_union_queries_device.argtypes = [vp, vp, i64, i32, vp, C.POINTER(i64), vp]
    lib.tjb_merge_tracts.argtypes = [i32, vp, vp, vp, vp, vp, i32, i32, vp, vp, vp]
    lib.tjb_context_histograms.argtypes = [C.POINTER(TjbHistOpts), vp, i32, C.POINTER(TjbContextTable)]
    lib.tjb_context_table_free.argtypes = [C.POINTER(TjbContextTable)]
    lib.tjb_context_table_free.restype = None
    lib.tjb_reader_open.restype = vp
    lib.tjb_reader_open.argtypes = [cp]
    lib.tjb_reader_fill.restype = i64
    lib.tjb_reader_fill.argtypes = [vp, vp, sz, C.POINTER(i64), C.POINTER(i64)]
    lib.tjb_reader_take_long.restype = i64
    lib.tjb_reader_take_long.argtypes = [vp, C.POINTER(vp), C.POINTER(i64), C.POINTER(i64)]
    lib.tjb_reader_close.argtypes = [vp]
    lib.tjb_reader_close.restype = None
    lib.tjb_gz_open.restype = vp
    lib.tjb_gz_open.argtypes = [cp]
    lib.tjb_gz_next.restype = i64
    lib.tjb_gz_next.argtypes = [vp, C.POINTER(vp)]
    lib.tjb_gz_close.argtypes = [vp]
    lib.tjb_gz_close.restype = None
    lib.tjb_tract_rows_noalloc.argtypes = [i32, vp, sz, po, vp, i64, vp, vp]
    lib.tjb_noalloc_max_bytes.restype = sz
    _lib = lib
    return lib
