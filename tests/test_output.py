"""Cropped copy-out: ohcu_slot_output_cropped against the oracle's restatement of
the reference's window offset (hevc_refs.c:335-341) and row memcpy (openhevc.c:722-735), and
against plain array slicing."""
import numpy as np
import pytest

import oracle_lib
from hevc_b200 import abi

# (width, height, bit depth, chroma_format_idc, [(left, top, width, height), ...])
CASES = [
    (416, 240, 8, 1, [(0, 0, 416, 240), (0, 0, 416, 232), (16, 8, 384, 224), (2, 2, 410, 236), (6, 4, 34, 2)]),
    (416, 240, 10, 1, [(0, 0, 416, 240), (8, 0, 400, 240), (2, 6, 402, 230), (414, 238, 2, 2)]),
    (200, 120, 10, 2, [(0, 0, 200, 120), (2, 1, 194, 117), (16, 3, 160, 3)]),
    (200, 120, 8, 3, [(0, 0, 200, 120), (1, 1, 197, 117), (3, 0, 16, 120), (5, 7, 1, 1)]),
    (200, 120, 8, 0, [(0, 0, 200, 120), (7, 3, 33, 5)]),
    (1920, 1088, 10, 1, [(0, 0, 1920, 1080)]),
]


def _picture(w, h, bd, cf, seed):
    rng = np.random.default_rng(seed)
    hs = 1 if cf in (1, 2) else 0
    vs = 1 if cf == 1 else 0
    dt = np.uint8 if bd == 8 else np.uint16
    shapes = [(h, w)] + ([(h >> vs, w >> hs)] * 2 if cf else [])
    return [rng.integers(0, 1 << bd, s).astype(dt) for s in shapes], hs, vs


def seq_params(w, h, bd, cf):
    return abi.SeqParams(width=w, height=h, bit_depth=bd, chroma_format_idc=cf, log2_ctb_size=6, log2_min_cb_size=3,
                         log2_min_pu_size=2, dpb_slots=2, jobs=1)


def _slices(planes, crop, hs, vs):
    l, t, w, h = crop
    out = []
    for i, p in enumerate(planes):
        a, b = (hs, vs) if i else (0, 0)
        out.append(p[t >> b:(t + h) >> b, l >> a:(l + w) >> a])
    return out


@pytest.mark.parametrize("case", CASES, ids=lambda c: f"{c[0]}x{c[1]}_{c[2]}b_cf{c[3]}")
def test_oracle_crop_is_slicing(case):
    w, h, bd, cf, crops = case
    planes, hs, vs = _picture(w, h, bd, cf, 7)
    orc = oracle_lib.Lib("oracle")
    for crop in crops:
        got = orc.output_cropped(planes, crop, hs, vs)
        for g, want in zip(got, _slices(planes, crop, hs, vs)):
            assert g.shape == want.shape and np.array_equal(g, want)


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=lambda c: f"{c[0]}x{c[1]}_{c[2]}b_cf{c[3]}")
def test_device_crop(case):
    from hevc_b200.backend import Context
    w, h, bd, cf, crops = case
    planes, hs, vs = _picture(w, h, bd, cf, 11)
    orc = oracle_lib.Lib("oracle")
    with Context(seq_params(w, h, bd, cf), dpb_slots=2, jobs=1) as ctx:
        ctx.upload(0, planes)
        for crop in crops:
            want = orc.output_cropped(planes, crop, hs, vs)
            got = ctx.output_cropped(0, crop)
            for g, x in zip(got, want):
                assert g.shape == x.shape and np.array_equal(g, x), crop
            # planes that follow each other in pinned memory leave as one copy
            pin = ctx.pinned_cropped(crop)
            ctx.output_cropped(0, crop, out=pin, wait=False)
            ctx.slot_sync(0)
            for g, x in zip(pin, want):
                assert np.array_equal(g, x), crop


@pytest.mark.gpu
def test_device_crop_errors():
    from hevc_b200.backend import Context
    with Context(seq_params(416, 240, 8, 1), dpb_slots=2, jobs=1) as ctx:
        for crop in [(0, 0, 418, 240), (-2, 0, 16, 16), (1, 0, 16, 16), (0, 0, 16, 15), (0, 0, 0, 16)]:
            with pytest.raises(abi.OhcuError):
                ctx.output_cropped(0, crop)


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,bd,cf", [(416, 240, 8, 1), (200, 72, 10, 1), (192, 128, 10, 3), (320, 184, 10, 2), (64, 64, 8, 0), (8, 8, 8, 1)])
def test_picture_md5_on_the_device(w, h, bd, cf):
    """ohcu_slot_md5 = verify_md5 (hevcdec.c:4035-4098): per plane, MD5 over the rows; the checker is hashlib (RFC 1321)"""
    import ctypes as C
    import hashlib
    from hevc_b200 import abi, backend, synth
    job = synth.make_job(w, h, bd, cf, seed=3)
    pic = synth.make_picture(job.sp, 555)
    with backend.Context(job.sp) as ctx:
        ctx.upload(0, pic)
        out = (C.c_uint8 * 48)()
        abi.check(ctx.lib.ohcu_slot_md5(ctx.handle, 0, out))
    got = bytes(out)
    for c, p in enumerate(pic):
        assert got[16 * c:16 * c + 16] == hashlib.md5(np.ascontiguousarray(p).tobytes()).digest(), f"plane {c}"
