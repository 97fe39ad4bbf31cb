"""The C-ABI library loads and exports every symbol include/ohcu.h declares (CPU only,
no compute calls)."""
import os
import re

from hevc_b200 import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "ohcu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(r"\b(ohcu_[a-z0-9_]+)\s*\(", text))


def test_header_and_ctypes_table_agree():
    assert _declared() == set(abi.SYMBOLS)


def test_library_exports_every_symbol():
    lib = abi.load()
    for name in _declared():
        assert getattr(lib, name) is not None
    assert lib.ohcu_abi_version() == abi.OHCU_ABI_VERSION


def test_create_fails_loudly_without_device_or_with_bad_params():
    lib = abi.load()
    import ctypes as C
    h = C.c_void_p()
    sp = abi.SeqParams(width=64, height=64, bit_depth=11, chroma_format_idc=1, log2_ctb_size=6,
                       log2_min_cb_size=3, log2_min_pu_size=2, dpb_slots=2, jobs=1)
    assert lib.ohcu_create(C.byref(h), 0, C.byref(sp)) == -1          # EINVAL: bit depth 11
    assert b"bit depth" in lib.ohcu_last_error()
    sp.bit_depth = 8
    if lib.ohcu_device_count() == 0:
        assert lib.ohcu_create(C.byref(h), 0, C.byref(sp)) == -5       # ENODEV, never a CPU fallback
