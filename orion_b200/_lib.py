"""ctypes binding of ``libob2.so`` (the C ABI declared in ``include/ob2.h``).

There is no CPU fallback: if the shared library is missing or does not export
the ABI, importing any compute path raises immediately.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libob2.so')

OB2_OK = 0
KIND_REAL, KIND_INTEGER, KIND_CATEGORICAL = 0, 1, 2
DIST_UNIFORM, DIST_RECIPROCAL, DIST_NORM, DIST_CONST, DIST_NONE = 0, 1, 2, 3, 4
OP_IDENTITY, OP_QUANTIZE, OP_TO_REAL, OP_ENUMERATE, OP_ONEHOT = 0, 1, 2, 3, 4
TFLAG_FLOOR = 1
SLOT_DIM_LAST = 1
BASE_NORMAL_VARIATE = 1          # ob2_sample_from_base: norm slots receive standard normal variates
SLOT_CAT_UNIFORM = 2
SET_SLOT_BYTES = 32
SET_HEADER_BYTES = 256
STATIC_DEDUP_MAX = 1 << 20
HOST_SAMPLE_MAX = 4096


class Ob2Error(RuntimeError):
    """An ob2_* entry point returned a non-zero code."""


class SlotDesc(ctypes.Structure):
    """``ob2_slot_desc`` (include/ob2.h), 128 bytes."""
    _fields_ = [
        ('kind', ctypes.c_int32), ('dist', ctypes.c_int32),
        ('dim_first', ctypes.c_int32), ('flags', ctypes.c_int32),
        ('n_cats', ctypes.c_int32), ('cdf_off', ctypes.c_int32),
        ('col_units', ctypes.c_int32), ('t_op', ctypes.c_int32),
        ('t_units', ctypes.c_int32), ('t_flags', ctypes.c_int32),
        ('t_width', ctypes.c_int32), ('reserved', ctypes.c_int32),
        ('p0', ctypes.c_double), ('p1', ctypes.c_double),
        ('loc', ctypes.c_double), ('scale', ctypes.c_double),
        ('low', ctypes.c_double), ('high', ctypes.c_double),
        ('hk', ctypes.c_uint64 * 4),
    ]


assert ctypes.sizeof(SlotDesc) == 128

_vp = ctypes.c_void_p
_i64 = ctypes.c_int64
_u64 = ctypes.c_uint64
_i32 = ctypes.c_int32
_u32 = ctypes.c_uint32

# name -> (restype, argtypes); every symbol include/ob2.h declares
SIGNATURES = {
    'ob2_last_error': (ctypes.c_char_p, []),
    'ob2_version': (ctypes.c_int, []),
    'ob2_space_create': (ctypes.c_int, [ctypes.POINTER(SlotDesc), _i32, ctypes.POINTER(ctypes.c_double), _i32,
                                        ctypes.POINTER(_vp)]),
    'ob2_space_destroy': (None, [_vp]),
    'ob2_space_set_philox_rounds': (ctypes.c_int, [_vp, _i32]),
    'ob2_space_row_bytes': (_i64, [_vp]),
    'ob2_space_trow_bytes': (_i64, [_vp]),
    'ob2_sample': (ctypes.c_int, [_vp, _i64, _i64, _u64, _u64, _vp, _vp, _vp, _vp, _vp, _vp]),
    'ob2_sample_transform': (ctypes.c_int, [_vp, _i64, _i64, _u64, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    'ob2_suggest': (ctypes.c_int, [_vp, _i64, _i64, _u64, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    'ob2_sample_host': (ctypes.c_int, [_vp, _i64, _u64, _u64, _vp, _vp, _vp]),
    'ob2_sample_dedup_host': (ctypes.c_int, [_vp, _vp, _i64, _u64, _u64, _vp, _vp, _vp, _vp]),
    'ob2_contains_host': (ctypes.c_int, [_vp, _i64, _vp, _i32, _vp, _vp]),
    'ob2_sample_from_uniform': (ctypes.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, _vp]),
    'ob2_sample_from_base': (ctypes.c_int, [_vp, _i64, _i64, _vp, ctypes.c_uint32, _vp, _vp, _vp, _vp]),
    'ob2_sample_from_base_host': (ctypes.c_int, [_vp, _i64, _vp, ctypes.c_uint32, _vp, _vp, _vp]),
    'ob2_contains': (ctypes.c_int, [_vp, _i64, _i64, _vp, _i32, _vp, _vp]),
    'ob2_transform': (ctypes.c_int, [_vp, _i64, _i64, _vp, _vp, _vp]),
    'ob2_reverse': (ctypes.c_int, [_vp, _i64, _i64, _vp, _vp, _vp]),
    'ob2_reverse_fingerprint': (ctypes.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp]),
    'ob2_fingerprint': (ctypes.c_int, [_vp, _i64, _i64, _vp, _vp, _vp]),
    'ob2_set_create': (ctypes.c_int, [_vp, _u64, _vp, ctypes.POINTER(_vp)]),
    'ob2_set_destroy': (None, [_vp]),
    'ob2_set_insert': (ctypes.c_int, [_vp, _i64, _vp, _vp]),
    'ob2_set_dedup': (ctypes.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp]),
    'ob2_set_set_dedup_grid': (ctypes.c_int, [_vp, ctypes.c_int32]),
    'ob2_set_dedup_static': (ctypes.c_int, [_vp, _i64, _vp, _vp, _vp, _vp]),
    'ob2_set_probe': (ctypes.c_int, [_vp, _i64, _vp, _vp, _vp]),
    'ob2_set_rehash': (ctypes.c_int, [_vp, _vp, _vp]),
    'ob2_set_export': (ctypes.c_int, [_vp, _u64, _vp, _vp, _vp]),
    'ob2_set_count': (ctypes.c_int, [_vp, ctypes.POINTER(_u64), _vp]),
    'ob2_compact_scratch_bytes': (_i64, [_i64]),
    'ob2_space_wire_row_bytes': (_i64, [_vp]),
    'ob2_space_wire_layout': (ctypes.c_int, [_vp, ctypes.POINTER(_i32), ctypes.POINTER(_i32)]),
    'ob2_compact_wire': (ctypes.c_int, [_vp, _i64, _i64, _vp, _vp, _i64, _vp, _vp, _vp, _vp]),
    'ob2_compact': (ctypes.c_int, [_vp, _i64, _i64, _vp, _vp, _i32, _i64, _vp, _vp, _vp, _vp]),
    'ob2_columns_to_host': (ctypes.c_int, [_vp, _i32, _i64, _i64, _vp, _i64, _vp, _vp]),
    'ob2_bucket_by_owner': (ctypes.c_int, [_i64, _vp, _u32, _i32, _i64, _vp, _vp, _vp, _vp, _vp]),
    'ob2_scatter_verdict': (ctypes.c_int, [_i64, _vp, _vp, _vp, _vp, _vp]),
    'ob2_xchg_create': (ctypes.c_int, [_i32, _i32, _i64, ctypes.POINTER(_vp)]),
    'ob2_xchg_destroy': (None, [_vp]),
    'ob2_xchg_ipc_handle': (ctypes.c_int, [_vp, _vp]),
    'ob2_xchg_open_peers': (ctypes.c_int, [_vp, _vp]),
    'ob2_xchg_set_peers': (ctypes.c_int, [_vp, ctypes.POINTER(_vp)]),
    'ob2_xchg_region': (_vp, [_vp]),
    'ob2_xchg_send': (ctypes.c_int, [_vp, _i64, _vp, _u32, _vp]),
    'ob2_xchg_dedup': (ctypes.c_int, [_vp, _vp, _i32, _vp]),
    'ob2_xchg_collect': (ctypes.c_int, [_vp, _i64, _vp, _vp]),
    'ob2_xchg_send_counts': (ctypes.c_int, [_vp, ctypes.POINTER(_u32), _vp]),
    'ob2_xchg_send_copy': (ctypes.c_int, [_vp, _i64, _vp, _u32, _vp]),
    'ob2_xchg_collect_copy': (ctypes.c_int, [_vp, _i64, _vp, _vp]),
    'ob2_xchg_cap': (_i64, [_vp]),
    'ob2_xchg_flag_overflow': (ctypes.c_int, [_vp, _vp, _vp]),
}

_lib = None


def load():
    """Load libob2.so once; raise loudly if it is absent or incomplete."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "orion_b200: %s not found. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C orion_b200/csrc`). There is no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(rc):
    if rc != OB2_OK:
        msg = load().ob2_last_error()
        raise Ob2Error("libob2 error %d: %s" % (rc, msg.decode('utf-8', 'replace') if msg else '?'))
