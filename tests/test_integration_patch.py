"""The decoder integration as code (integration/): ohcu_openhevc.patch must apply to the reference's libavcodec and the
four patched files must compile, with the backend switched on (CONFIG_OHCU=1, against ohcu_glue.h / ohcu.h) and off
(the stock decoder unchanged).  CPU only; needs /root/reference (skipped on the GPU box)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not os.path.isdir("/root/reference/libavcodec"), reason="/root/reference absent")
def test_patch_applies_and_patched_reference_compiles():
    d = os.path.join(ROOT, "integration")
    subprocess.check_call(["make", "-s", "-C", d, "clean"])
    subprocess.check_call(["make", "-s", "-C", d, "compile"])
    for f in ("hevcdec", "hevc_cabac", "hevc_filter", "hevc_refs"):
        assert os.path.exists(os.path.join(d, "_build", f + ".o")) and os.path.exists(os.path.join(d, "_build", f + ".stock.o"))
    # the switched-on objects call the glue, the stock ones do not
    def undefined(o):
        return subprocess.check_output(["nm", "-u", os.path.join(d, "_build", o)], text=True)
    assert "ff_ohcu_emit_pu" in undefined("hevcdec.o") and "ff_ohcu_emit_tu" in undefined("hevc_cabac.o")
    assert "ff_ohcu_fetch_frame" in undefined("hevc_refs.o")
    for f in ("hevcdec", "hevc_cabac", "hevc_filter", "hevc_refs"):
        assert "ff_ohcu" not in undefined(f + ".stock.o")
    if os.path.exists(os.path.join(ROOT, "hevc_b200", "libohcu.so")):
        subprocess.check_call(["make", "-s", "-C", d, "harness"])
        assert os.path.exists(os.path.join(d, "_build", "libohglue.so"))
