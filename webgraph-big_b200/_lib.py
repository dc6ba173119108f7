"""ctypes binding of libwgbv.so (include/wgbv.h).  Fails loudly when the CUDA library is missing: there is no
CPU fallback anywhere in this package."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("WGBV_LIBRARY") or os.path.join(_HERE, "lib", "libwgbv.so")  # WGBV_LIBRARY: another build of the same library

OK, EINDEX, ESTATE, EIO, EUNSUPPORTED, ECUDA, EARG = 0, -1, -2, -3, -4, -5, -6


class Props(C.Structure):
    """wgbv_props"""
    _fields_ = [("nodes", C.c_int64), ("arcs", C.c_int64), ("window_size", C.c_int32), ("max_ref_count", C.c_int32),
                ("min_interval_length", C.c_int32), ("zeta_k", C.c_int32), ("outdegree_coding", C.c_int32),
                ("block_coding", C.c_int32), ("residual_coding", C.c_int32), ("reference_coding", C.c_int32),
                ("block_count_coding", C.c_int32), ("offset_coding", C.c_int32)]


# the reference's exceptions, one per status code (include/wgbv.h)
class IllegalArgumentException(ValueError):
    pass


class IllegalStateException(RuntimeError):
    pass


class IOException(IOError):
    pass


class UnsupportedOperationException(NotImplementedError):
    pass


class NoSuchElementException(LookupError):
    pass


class CudaError(RuntimeError):
    pass


_EXC = {EINDEX: IllegalArgumentException, ESTATE: IllegalStateException, EIO: IOException,
        EUNSUPPORTED: UnsupportedOperationException, ECUDA: CudaError, EARG: IllegalArgumentException}

# every symbol include/wgbv.h declares: (name, restype, argtypes)
_P = C.c_void_p
SYMBOLS = [
    ("wgbv_last_error", C.c_char_p, []),
    ("wgbv_parse_properties", C.c_int, [C.c_char_p, C.POINTER(Props)]),
    ("wgbv_open", C.c_int, [C.c_char_p, C.c_int, C.c_int, C.c_int, C.POINTER(_P)]),
    ("wgbv_open_mem", C.c_int, [C.POINTER(Props), _P, C.c_uint64, _P, C.c_uint64, C.c_int, C.c_int, C.c_int, C.POINTER(_P)]),
    ("wgbv_plan_shards", C.c_int, [C.POINTER(Props), _P, C.c_uint64, C.c_int, _P, C.POINTER(C.c_int64)]),
    ("wgbv_open_slice", C.c_int, [C.POINTER(Props), C.c_int64, C.c_int64, C.c_int64, _P, C.c_uint64, _P, C.c_uint64, C.c_int, C.POINTER(_P)]),
    ("wgbv_upload_stream", C.c_int, [_P, _P, C.c_uint64]),
    ("wgbv_close", None, [_P]),
    ("wgbv_num_nodes", C.c_int64, [_P]),
    ("wgbv_num_arcs", C.c_int64, [_P]),
    ("wgbv_get_props", C.c_int, [_P, C.POINTER(Props)]),
    ("wgbv_shard_range", C.c_int, [_P, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    ("wgbv_device_bytes", C.c_uint64, [_P]),
    ("wgbv_offsets_bytes", C.c_uint64, [_P]),
    ("wgbv_stream_create", C.c_int, [_P, _P, C.POINTER(_P)]),
    ("wgbv_stream_destroy", None, [_P]),
    ("wgbv_stream_sync", C.c_int, [_P]),
    ("wgbv_decode_degrees", C.c_int, [_P, C.c_int64, C.c_int64, _P, C.POINTER(C.c_int64)]),
    ("wgbv_decode_csr", C.c_int, [_P, C.c_int64, C.c_int64, _P, _P, C.c_uint64]),
    ("wgbv_decode_csr_host", C.c_int, [_P, C.c_int64, C.c_int64, _P, _P, C.c_uint64, C.POINTER(C.c_int64)]),
    ("wgbv_scan_begin", C.c_int, [_P, C.c_int64, C.c_int64, C.c_int64]),
    ("wgbv_scan_begin_upload", C.c_int, [_P, _P, C.c_uint64, C.c_int64, C.c_int64, C.c_int64]),
    ("wgbv_scan_next", C.c_int, [_P, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(_P), C.POINTER(_P), C.POINTER(C.c_int64)]),
    ("wgbv_scan_end", None, [_P]),
    ("wgbv_scan_set_narrow", C.c_int, [_P, C.c_int]),
    ("wgbv_transpose_csr", C.c_int, [_P, C.c_int64, _P, _P, _P, _P]),
    ("wgbv_labels_open_mem", C.c_int, [C.c_int, C.c_int64, C.c_int, C.c_int, _P, C.c_uint64, _P, C.c_uint64, C.POINTER(_P)]),
    ("wgbv_labels_close", None, [_P]),
    ("wgbv_labels_decode", C.c_int, [_P, _P, C.c_int64, C.c_int64, _P, _P, C.c_uint64]),
    ("wgbv_labels_decode_host", C.c_int, [_P, _P, C.c_int64, C.c_int64, _P, _P, C.c_uint64]),
    ("wgbv_outdegrees", C.c_int, [_P, _P, C.c_uint64, _P]),
    ("wgbv_successors_batch_degrees", C.c_int, [_P, _P, C.c_uint64, _P, C.POINTER(C.c_int64)]),
    ("wgbv_successors_batch", C.c_int, [_P, _P, C.c_uint64, _P, _P, C.c_uint64]),
    ("wgbv_successors_batch_host", C.c_int, [_P, _P, C.c_uint64, _P, _P, C.c_uint64, C.POINTER(C.c_int64)]),
    ("wgbv_csr_checksum", C.c_int, [_P, C.c_int64, C.c_int64, _P, _P, C.POINTER(C.c_uint64)]),
    ("wgbv_scan_degrees_static", C.c_int, [_P, C.c_int64, _P, _P]),
    ("wgbv_last_stage_ms", C.c_int, [_P, C.POINTER(C.c_float * 5)]),
    ("wgbv_kernel_launches", C.c_uint64, [_P]),
    ("wgbv_fused_stats", C.c_int, [_P, C.POINTER(C.c_uint64 * 4)]),
    ("wgbv_labels_decode_lists", C.c_int, [_P, _P, C.c_int64, C.c_int64, _P, _P, _P, C.c_uint64, C.POINTER(C.c_int64)]),
    ("wgbv_labels_decode_lists_host", C.c_int, [_P, _P, C.c_int64, C.c_int64, _P, _P, _P, C.c_uint64, C.POINTER(C.c_int64)]),
    ("wgbv_write_offsets", C.c_int, [_P, _P, C.c_uint64, _P, C.c_uint64, C.POINTER(C.c_uint64)]),
    ("wgbv_union_csr", C.c_int, [_P, C.c_int64, _P, _P, _P, _P, _P, _P, C.c_uint64, C.POINTER(C.c_int64)]),
    ("wgbv_symmetrize_csr", C.c_int, [_P, C.c_int64, _P, _P, _P, _P, C.c_uint64, C.POINTER(C.c_int64)]),
    ("wgbv_encode_csr", C.c_int, [C.c_int, _P, C.c_int64, C.c_int64, _P, _P, C.c_int64, _P, C.c_uint64, C.POINTER(C.c_uint64), _P]),
    ("wgbv_algorithmic_bytes", C.c_uint64, [_P, C.c_int64, C.c_int64, C.c_int64]),
]

_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, res, args in SYMBOLS:
            fn = getattr(L, name)  # AttributeError here means the library does not export what wgbv.h declares
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


def check(rc):
    if rc != OK:
        msg = lib().wgbv_last_error().decode("utf-8", "replace")
        raise _EXC.get(rc, RuntimeError)(msg or f"libwgbv error {rc}")
