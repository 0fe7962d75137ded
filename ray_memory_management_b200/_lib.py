"""ctypes binding of libds_shuffle.so (the C ABI declared in include/ds_shuffle.h).

There is no CPU fallback: if the CUDA library has not been built (run
``python -c "import __graft_entry__ as g; g.build()"`` or ``make -C
ray_memory_management_b200/csrc``) importing this module raises.
"""
import ctypes
import os
from ctypes import c_char_p, c_int32, c_size_t, c_uint32, c_uint64, c_void_p, POINTER

from . import _nvtx

_nested = [False]
_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libds_shuffle.so")

DS_OK = 0
DS_ERR_INVALID_ARG, DS_ERR_UNSUPPORTED, DS_ERR_WORKSPACE, DS_ERR_CUDA, DS_ERR_TOO_LARGE = (
    -1, -2, -3, -4, -5,
)
DS_MAX_ROWS = (1 << 30) - 1
DS_MAX_KEY_BYTES = 10
DS_MAX_PARTITIONS = 256
DS_KEY_I64, DS_KEY_U64, DS_KEY_F64, DS_KEY_FSB = 0, 1, 2, 3
DS_SPLIT_TILE_ROWS = 256
DS_MAX_SPLIT_PARTS = 32
DS_SPLIT_MAX_ROW_BYTES = 200
DS_SORT_NO_PASS_SKIP = 1
DS_SORT_NO_PREFIX_REPAIR = 2
DS_SORT_NO_HYBRID = 4
DS_OP_SUM, DS_OP_MIN, DS_OP_MAX, DS_OP_ABSMAX = 0, 1, 2, 3
DS_SEGMENT_HEADS_STATIC_MAX_ROWS = 1 << 27


class DeviceLibraryError(RuntimeError):
    """A ds_* entry point returned a negative status."""

    def __init__(self, status, message):
        super().__init__(f"ds_shuffle error {status}: {message}")
        self.status = status


if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: the sm_100a CUDA library has not been built. "
        "Build it with __graft_entry__.build(); there is no CPU fallback."
    )

lib = ctypes.CDLL(LIB_PATH)

_P = c_void_p
_SZP = POINTER(c_size_t)

# name -> argtypes; every function returns ds_status except the two below
_PROTOS = {
    "ds_sort_indices_i64": [_P, c_uint64, c_int32, _P, _P, c_int32, _P, _SZP, _P],
    "ds_sort_indices_u64": [_P, c_uint64, c_int32, _P, _P, c_int32, _P, _SZP, _P],
    "ds_sort_indices_f64": [_P, c_uint64, c_int32, _P, _P, c_int32, _P, _SZP, _P],
    "ds_sort_indices_fsb": [_P, c_uint32, c_uint64, c_int32, _P, _P, c_int32, _P, _SZP, _P],
    "ds_gather_rows": [_P, c_uint32, c_uint64, _P, c_uint64, _P, _P],
    "ds_gather_rows_multi": [_P, _P, c_uint32, c_uint32, _P, c_uint64, _P, _P],
    "ds_scatter_rows_to_parts": [_P, c_uint32, c_uint64, _P, c_uint64, c_uint32, _P, _P, _P],
    "ds_split_plan": [c_int32, _P, c_uint32, _P, c_uint64, _P, c_uint32, c_int32, _P, _P, _P, _P],
    "ds_split_rows": [_P, c_uint32, c_uint64, _P, _P, c_uint32, _P, c_uint32, c_uint32, _P],
    "ds_searchsorted_i64": [_P, c_uint64, _P, c_uint32, c_int32, _P, _P],
    "ds_searchsorted_f64": [_P, c_uint64, _P, c_uint32, c_int32, _P, _P],
    "ds_searchsorted_fsb": [_P, c_uint32, c_uint64, _P, c_uint32, c_int32, _P, _P],
    "ds_partition_hist_scatter_i64": [_P, c_uint64, _P, c_uint32, c_int32, _P, _P, _P, _P, _SZP, _P],
    "ds_partition_hist_scatter_fsb": [_P, c_uint32, c_uint64, _P, c_uint32, c_int32, _P, _P, _P, _P, _SZP, _P],
    "ds_split_records": [_P, c_uint32, c_uint32, c_uint64, _P, _P, _P],
    "ds_join_records": [_P, _P, c_uint32, c_uint32, c_uint64, _P, _P],
    "ds_validate_records": [_P, _P, c_uint32, c_uint32, c_uint64, _P, _P],
    "ds_sample_rows": [_P, c_uint32, c_uint64, c_uint32, c_uint64, _P, _P],
    "ds_random_permutation": [c_uint64, c_uint64, c_uint64, _P, _P, _SZP, _P],
    "ds_segment_heads_i64": [_P, c_uint64, _P, _P, _P, _SZP, _P],
    "ds_segment_heads_i64_static": [_P, c_uint64, _P, _P, _P],
    "ds_segmented_reduce_f64": [_P, _P, _P, _P, c_uint32, c_uint64, c_int32, _P, _P, _SZP, _P],
    "ds_segmented_reduce_i64": [_P, _P, _P, _P, c_uint32, c_uint64, c_int32, _P, _P, _SZP, _P],
    "ds_segmented_count_valid": [_P, _P, _P, c_uint32, c_uint64, _P, _P, _SZP, _P],
    "ds_segment_counts": [_P, c_uint32, _P, _P],
    "ds_finalize_mean": [_P, c_int32, _P, c_uint64, _P, _P],
    "ds_segmented_sumsqdiff_f64": [_P, _P, _P, _P, _P, _P, c_uint32, c_uint64, _P, _P, _P, _SZP, _P],
    "ds_segmented_sumsqdiff_i64": [_P, _P, _P, _P, c_uint32, c_uint64, _P, _P, _P, _SZP, _P],
    "ds_scale_f64_by_i64": [_P, _P, c_uint64, _P, _P],
    "ds_finalize_std": [_P, _P, c_uint64, c_int32, _P, _P],
    "ds_finalize_validity": [_P, _P, c_uint64, c_int32, _P, _P],
    "ds_varlen_take_offsets": [_P, c_uint64, _P, c_uint64, _P, _P, _SZP, _P],
    "ds_varlen_take_bytes": [_P, _P, _P, c_uint64, _P, _P, _P],
    "ds_varlen_key_lengths": [_P, _P, c_uint64, _P, _P, _P],
    "ds_varlen_key_chunk": [_P, _P, _P, c_uint64, c_uint64, _P, _P],
    "ds_searchsorted_varlen": [_P, _P, c_uint64, _P, _P, c_uint32, c_int32, _P, _P],
    "ds_group_ids_fixed": [_P, c_uint32, c_int32, c_uint64, _P, _P, _SZP, _P],
    "ds_group_ids_varlen": [_P, _P, c_uint64, _P, _P, _SZP, _P],
    "ds_widen_keys": [_P, c_int32, c_uint64, _P, _P],
}
EXPORTED = ["ds_version", "ds_last_error", "ds_split_table_words"] + sorted(_PROTOS)

lib.ds_version.restype = c_int32
lib.ds_version.argtypes = []
lib.ds_split_table_words.restype = c_uint64
lib.ds_split_table_words.argtypes = [c_uint64, c_uint32]
lib.ds_last_error.restype = c_char_p
lib.ds_last_error.argtypes = []
for _name, _args in _PROTOS.items():
    _fn = getattr(lib, _name)
    _fn.restype = c_int32
    _fn.argtypes = _args


def check(status: int) -> None:
    if status != DS_OK:
        raise DeviceLibraryError(status, lib.ds_last_error().decode())


def call(name: str, *args) -> None:
    """Call a ds_* entry point and raise on a negative status."""
    if _nvtx._on:
        with _nvtx.range(name):
            check(getattr(lib, name)(*args))
        return
    check(getattr(lib, name)(*args))


def call_ws(name: str, alloc, *args_before_ws, stream):
    """Two-phase workspace call: args are (..., ws, &ws_bytes, stream) at the tail.

    ``alloc(nbytes)`` must return an object with ``data_ptr()`` that lives until the
    stream work is done (the caller keeps it).
    """
    if _nvtx._on and not _nested[0]:
        _nested[0] = True
        try:
            with _nvtx.range(name):
                return call_ws(name, alloc, *args_before_ws, stream=stream)
        finally:
            _nested[0] = False
    fn = getattr(lib, name)
    need = c_size_t(0)
    check(fn(*args_before_ws, None, ctypes.byref(need), stream))
    buf = alloc(max(int(need.value), 256))
    have = c_size_t(max(int(need.value), 256))
    check(fn(*args_before_ws, buf.data_ptr(), ctypes.byref(have), stream))
    return buf
