"""The real boundary test: the PATCHED reference decoder code (integration/ohcu_openhevc.patch on a scratch copy of
libavcodec, CONFIG_OHCU=1) and the decoder-side glue emit the work lists into libohcu.so — coded blocks through the
reference's own ff_hevc_hls_transform{,_c}, PUs / slice tables / loop-filter arrays from the reference's own structures —
and the picture that comes back must be the one the UNPATCHED reference reconstructs (digests of tests/golden/frames.json,
written by oracle/_ref).  First C caller of the ABI on the decoder side (integration/glue_harness.c)."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import frame_cases
import oracle_lib

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GLUE_SO = os.path.join(ROOT, "integration", "_build", "libohglue.so")
GOLDEN = os.path.join(ROOT, "tests", "golden", "frames.json")
# dense_coeffs_*: coefficients beyond col_limit, which no parser produces (the glue derives col_limit from the block)
CASES = [n for n in frame_cases.FRAME_CASES if not n.startswith("dense_coeffs")]


def _replay(lib, job, refs, out_init):
    sp = job.sp
    nplanes = 3 if sp.chroma_format_idc else 1
    out = [np.ascontiguousarray(p.copy()) for p in out_init]
    dpb = (oracle_lib.Picture * max(1, sp.dpb_slots))()
    keep = []
    for i in range(sp.dpb_slots):
        if i < len(refs) and refs[i] is not None:
            for c in range(nplanes):
                a = np.ascontiguousarray(refs[i][c]); keep.append(a)
                dpb[i].data[c] = a.ctypes.data; dpb[i].stride[c] = a.strides[0]
    fj = oracle_lib.FrameJobC()
    fj.sp = sp
    fj.n_slices = len(job.slices); fj.slices = job.slices.ctypes.data
    fj.n_pu = len(job.pus); fj.pus = job.pus.ctypes.data
    fj.n_tu = len(job.tus); fj.tus = job.tus.ctypes.data
    fj.coeffs = job.coeffs.ctypes.data; fj.ncoeff = len(job.coeffs)
    db = job.deblock_in(); so = job.sao_in()
    fj.deblock = C.pointer(db); fj.sao = C.pointer(so)
    fj.dpb = dpb
    for c in range(nplanes):
        fj.out.data[c] = out[c].ctypes.data; fj.out.stride[c] = out[c].strides[0]
    fj.stages = 0
    if len(job.ius):
        fj.n_iu = len(job.ius); fj.ius = job.ius.ctypes.data
    rc = lib.ohglue_replay_frame(C.byref(fj), 0)
    assert rc == 0, f"ohglue_replay_frame -> {rc}"
    return out


@pytest.mark.skipif(not os.path.exists(GLUE_SO), reason="integration/_build/libohglue.so not built (make -C integration harness)")
@pytest.mark.parametrize("name", CASES)
def test_patched_reference_emits_the_same_picture(name):
    from hevc_b200 import abi
    abi.load()
    lib = C.CDLL(GLUE_SO)
    lib.ohglue_replay_frame.restype = C.c_int
    lib.ohglue_replay_frame.argtypes = [C.POINTER(oracle_lib.FrameJobC), C.c_int]
    gold = json.load(open(GOLDEN))["cases"][name]
    job, refs = frame_cases.build_case(name)
    got = _replay(lib, job, refs, refs[-1])
    assert frame_cases.picture_digest(got) == gold["all"], f"{name}: the patched reference + glue + device differ from the unpatched reference"
