"""The C-ABI library loads and exports every symbol include/ob2.h declares
(no compute: runs without a GPU)."""
import ctypes
import os
import re

from orion_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, 'include', 'ob2.h')).read()
    return sorted(set(re.findall(r'\b(ob2_[a-z0-9_]+)\s*\(', text)))


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 18
    for name in names:
        assert hasattr(lib, name), name


def test_python_binding_covers_the_header():
    assert sorted(_lib.SIGNATURES) == _declared()
    lib = _lib.load()
    assert lib.ob2_version() == 1


def test_slot_desc_layout_matches_header():
    assert ctypes.sizeof(_lib.SlotDesc) == 128
    assert _lib.SlotDesc.p0.offset == 48 and _lib.SlotDesc.hk.offset == 96


def test_argument_errors_without_a_gpu():
    lib = _lib.load()
    out = ctypes.c_void_p()
    assert lib.ob2_space_create(None, 0, None, 0, ctypes.byref(out)) == -1
    assert b'bad arguments' in lib.ob2_last_error()
    assert lib.ob2_set_create(None, 100, None, ctypes.byref(out)) == -1
