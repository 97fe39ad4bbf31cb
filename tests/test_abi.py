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


def test_struct_layout_matches_the_c_header(tmp_path):
    """sizeof / offsetof of every struct of include/ohcu.h as the C compiler sees them (tests/c/abi_layout.c, plain
    C99) against the ctypes mirror in hevc_b200/abi.py"""
    import ctypes as C
    import json
    import subprocess
    exe = str(tmp_path / "abi_layout")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "c", "abi_layout.c"), "-o", exe])
    layout = json.loads(subprocess.check_output([exe]))
    text = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "ohcu.h")).read(), flags=re.S)
    declared = set(re.findall(r"}\s*(ohcu_[a-z0-9_]+)\s*;", text))
    assert declared == set(layout) == set(abi.STRUCTS), "a struct of ohcu.h is missing from abi_layout.c or abi.STRUCTS"
    for name, cls in abi.STRUCTS.items():
        assert C.sizeof(cls) == layout[name]["sizeof"], name
        offs = {f[0]: getattr(cls, f[0]).offset for f in cls._fields_}
        assert offs == layout[name]["fields"], f"{name}: {offs} != {layout[name]['fields']}"


def test_cmake_builds_the_same_library(tmp_path):
    """north_star: "the build (CMake with nvcc for sm_100a)".  CMakeLists.txt must configure, build for sm_100a only and
    export exactly the symbols of the in-tree build (nvcc cross-compiles without a GPU)."""
    import shutil
    import subprocess
    if not shutil.which("cmake") or not shutil.which("nvcc"):
        import pytest
        pytest.skip("cmake / nvcc not on PATH")
    b = str(tmp_path / "b")
    subprocess.check_call(["cmake", "-S", ROOT, "-B", b, "-DCMAKE_CUDA_COMPILER=" + shutil.which("nvcc")], stdout=subprocess.DEVNULL)
    subprocess.check_call(["cmake", "--build", b, "-j", "8"], stdout=subprocess.DEVNULL)
    so = os.path.join(b, "libohcu.so")
    exported = {l.split()[-1] for l in subprocess.check_output(["nm", "-D", "--defined-only", so], text=True).splitlines()
                if " T ohcu_" in l}
    assert exported == _declared()
    elfs = subprocess.check_output(["cuobjdump", "-lelf", so], text=True)
    assert "sm_100a" in elfs and not re.search(r"sm_(?!100a)\d+", elfs), elfs
    assert os.path.exists(os.path.join(b, "libohcu_drv.so"))
