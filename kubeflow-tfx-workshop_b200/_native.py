"""ctypes binding of libtaxi_xf.so (include/taxi_xf.h).

There is deliberately no fallback: if the CUDA library is missing or fails to load, importing
this module raises, and every compute entry point of the package fails with it.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_int32, c_int64, c_size_t, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtaxi_xf.so")

TXF_MAX_COLS = 16
TXF_MAX_OPS = 32
TXF_POOL_SLACK = 64
TXF_QMAX_TARGETS = 64
TXF_F32, TXF_I64 = 0, 1
(OP_ZSCORE, OP_VOCAB, OP_BUCKETIZE, OP_FILL_I64, OP_FILL_F32, OP_FILL_STR, OP_RATIO_LABEL, OP_VOCAB_CODED) = range(8)


class NativeError(RuntimeError):
    """A libtaxi_xf entry point returned a non-zero status."""


class TxfOp(Structure):
    """struct txf_op (include/taxi_xf.h)."""

    _fields_ = [
        ("kind", c_int32), ("in_dtype", c_int32),
        ("in0", c_void_p), ("valid0", c_void_p), ("in1", c_void_p), ("valid1", c_void_p),
        ("pool", c_void_p), ("out", c_void_p), ("p0", c_void_p), ("p1", c_void_p),
        ("a0", c_int64), ("a1", c_int64), ("a2", c_int64), ("a3", c_int64),
        ("f0", c_float), ("reserved", c_int32),
    ]


class TxfOperand(Structure):
    """struct txf_operand (include/taxi_xf.h)."""

    _fields_ = [("col", c_void_p), ("valid", c_void_p), ("scalar", c_void_p), ("dtype", c_int32), ("reserved", c_int32)]


class TxfCsvColumn(Structure):
    """struct txf_csv_column (include/taxi_xf.h)."""

    _fields_ = [("kind", c_int32), ("reserved", c_int32), ("values", c_void_p), ("lens", c_void_p), ("valid", c_void_p)]


class TxfExampleFeature(Structure):
    """struct txf_example_feature (include/taxi_xf.h)."""

    _fields_ = [("kind", c_int32), ("name_off", c_int32), ("name_len", c_int32), ("reserved", c_int32),
                ("values", c_void_p), ("pool", c_void_p), ("valid", c_void_p)]


class TxfExampleColumn(Structure):
    """struct txf_example_column (include/taxi_xf.h)."""

    _fields_ = [("kind", c_int32), ("name_off", c_int32), ("name_len", c_int32), ("reserved", c_int32),
                ("values", c_void_p), ("lens", c_void_p), ("valid", c_void_p)]


EX_ERRORS = {1: "a record's length field does not match its extent", 2: "malformed protobuf wire data",
             4: "a feature holds a list of another kind than the schema says", 8: "more than one value in a cell",
             16: "data CRC mismatch"}
TXF_EX_BYTES = 2
TXF_EX_MAX_FEATURES = 64
ARITH_ADD, ARITH_SUB, ARITH_MUL, ARITH_DIV = range(4)
CSV_SKIP, CSV_I64, CSV_F32, CSV_BYTES = range(4)
CSV_SEEN_VALUE, CSV_SEEN_INT, CSV_SEEN_DECIMAL, CSV_SEEN_TEXT = 1, 2, 4, 8
CSV_ERRORS = {1: "a line has more or fewer fields than the header", 2: "unterminated quote or text after a closing quote",
              4: "a field of a numeric column is not a number", 8: "integer outside int64",
              16: "decimal literal too long to convert exactly (more than 19 significant digits or exponent beyond +-22)"}
TXF_CSV_MAX_COLS = 64

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python __graft_entry__.py` (or "
        "`python kubeflow-tfx-workshop_b200/build.py`). This package has no CPU fallback.")

_lib = ctypes.CDLL(LIB_PATH)

_PP = POINTER(c_void_p)
_SIGNATURES = {
    "txf_last_error": (c_char_p, []),
    "txf_abi_version": (c_int, []),
    "txf_launch_count": (c_int64, []),
    "txf_moments_workspace_bytes": (c_size_t, [c_int]),
    "txf_moments_update": (c_int, [_PP, _PP, POINTER(c_int32), c_int, c_int64, c_void_p, c_void_p, c_void_p]),
    "txf_moments_update_noalloc": (c_int, [_PP, _PP, POINTER(c_int32), c_int, c_int64, c_void_p, c_void_p]),
    "txf_moments_merge": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "txf_moments_finalize": (c_int, [c_void_p, c_int, c_void_p, c_void_p]),
    "txf_msel_state_bytes": (c_size_t, [c_int, c_int]),
    "txf_msel_hist_bytes": (c_size_t, [c_int, c_int]),
    "txf_msel_begin": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "txf_msel_hist": (c_int, [c_void_p, c_int, c_int, c_int, _PP, _PP, POINTER(c_int32), c_int, c_int64, c_void_p, c_void_p]),
    "txf_msel_resolve": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p]),
    "txf_msel_result": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "txf_qsample": (c_int, [_PP, _PP, POINTER(c_int32), c_int, c_int64, c_int64, c_uint64, c_void_p, c_void_p]),
    "txf_qcount": (c_int, [_PP, _PP, POINTER(c_int32), c_int, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "txf_qresolve": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "txf_bsel_hist_bytes": (c_size_t, [c_int, c_int, c_int]),
    "txf_bsel_hist": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "txf_bsel_resolve": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "txf_qfinal": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "txf_qsplit_edges": (c_int, [c_void_p, c_int, c_int64, c_int, c_void_p, c_void_p]),
    "txf_qsplit_scratch_bytes": (c_size_t, [c_int, c_int]),
    "txf_qsplit_flag_words": (c_size_t, [c_int]),
    "txf_qsplit_count": (c_int, [_PP, _PP, POINTER(c_int32), c_int, c_int64, c_int, c_void_p, c_void_p, c_void_p]),
    "txf_qsplit_resolve": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "txf_qsplit_collect": (c_int, [_PP, _PP, POINTER(c_int32), c_int, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "txf_vocab_table_bytes": (c_size_t, [c_int64]),
    "txf_vocab_lookup_bytes": (c_size_t, [c_int64]),
    "txf_vocab_sort_scratch_bytes": (c_size_t, [c_int64]),
    "txf_vocab_local_codes": (c_int64, []),
    "txf_debug_set_identity_bits": (c_int, [c_int]),
    "txf_vocab_count": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "txf_vocab_final_codes": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p]),
    "txf_vocab_compact": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "txf_vocab_export_meta": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "txf_vocab_export_bytes": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p]),
    "txf_vocab_export_blob": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "txf_vocab_insert_blob": (c_int, [c_void_p, c_int64, c_int64, c_int, c_int64, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p]),
    "txf_vocab_insert": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p]),
    "txf_vocab_filter_ge": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "txf_vocab_sort": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "txf_vocab_build_lookup": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p]),
    "txf_vocab_lookup_from_terms": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p]),
    "txf_transform": (c_int, [POINTER(TxfOp), c_int, c_int64, c_void_p]),
    "txf_copy_string_bytes": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "txf_transform_stream_workspace_bytes": (c_size_t, []),
    "txf_transform_stream": (c_int, [POINTER(TxfOp), c_int, c_int64, c_void_p, c_void_p]),
    "txf_transform_launch_info": (c_int, [POINTER(c_int), POINTER(c_int), POINTER(c_int)]),
    "txf_minmax_workspace_bytes": (c_size_t, [c_int]),
    "txf_minmax_begin": (c_int, [c_void_p, c_int, c_void_p]),
    "txf_minmax_update": (c_int, [_PP, _PP, POINTER(c_int32), c_int, c_int64, c_void_p, c_void_p, c_void_p]),
    "txf_minmax_merge": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "txf_map_binary": (c_int, [c_int, POINTER(TxfOperand), POINTER(TxfOperand), c_void_p, c_int64, c_void_p]),
    "txf_map_cast": (c_int, [POINTER(TxfOperand), c_int, c_void_p, c_int64, c_void_p]),
    "txf_i64_format_lengths": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "txf_i64_format_write": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p]),
    "txf_map_scale_minmax": (c_int, [POINTER(TxfOperand), c_void_p, c_float, c_float, c_void_p, c_int64, c_void_p]),
    "txf_csv_block_bytes": (c_int64, []),
    "txf_csv_count_lines": (c_int, [c_void_p, c_int64, c_void_p, c_void_p]),
    "txf_csv_line_starts": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p]),
    "txf_csv_classify": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int, c_void_p, c_void_p, c_void_p]),
    "txf_csv_parse": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int, POINTER(TxfCsvColumn), c_void_p, c_void_p]),
    "txf_csv_copy_strings": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "txf_example_sizes": (c_int, [POINTER(TxfExampleFeature), c_int, c_int64, c_void_p, c_void_p]),
    "txf_example_encode": (c_int, [POINTER(TxfExampleFeature), c_int, c_void_p, c_int64, c_void_p, c_void_p, c_void_p]),
    "txf_tfrecord_block_bytes": (c_int64, []),
    "txf_tfrecord_count_candidates": (c_int, [c_void_p, c_int64, c_void_p, c_void_p]),
    "txf_tfrecord_candidates": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p]),
    "txf_tfrecord_link": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "txf_example_decode": (c_int, [POINTER(TxfExampleColumn), c_int, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p]),
    "txf_gather_bytes": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "txf_fingerprint64_batch": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
}

EXPORTED = sorted(_SIGNATURES)

for _name, (_res, _args) in _SIGNATURES.items():
    _fn = getattr(_lib, _name)  # AttributeError here == the library does not match the header
    _fn.restype = _res
    _fn.argtypes = _args


def last_error() -> str:
    return (_lib.txf_last_error() or b"").decode("utf-8", "replace")


def call(name: str, *args):
    """Call a status-returning entry point; raise NativeError with the library's message."""
    rc = getattr(_lib, name)(*args)
    if rc != 0:
        raise NativeError(f"{name} failed ({rc}): {last_error()}")


def query(name: str, *args):
    """Call a size/version query (no status code)."""
    return getattr(_lib, name)(*args)


def ptr_array(ptrs):
    """Host array of device pointers (None -> NULL)."""
    arr = (c_void_p * len(ptrs))()
    for i, p in enumerate(ptrs):
        arr[i] = p if p else None
    return arr


def i32_array(vals):
    return (c_int32 * len(vals))(*vals)


def launch_count() -> int:
    return int(_lib.txf_launch_count())
