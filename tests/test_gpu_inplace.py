"""The zero-copy host path (ohcu_frame_arenas_get / ohcu_frame_commit_lists): lists written straight into the
job's pinned arenas, validated and indexed by prepare_kernel on the device.  Same pictures as the copying
path (reference digests), device-side rejection of malformed lists, per-picture PPS parameters, PCM sample
placement and slot release."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import frame_cases
import levels_cases
from hevc_b200 import abi, backend, synth

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _diff(a, b):
    for c, (x, y) in enumerate(zip(a, b)):
        if not np.array_equal(x, y):
            d = np.argwhere(x != y)
            return f"plane {c}: {len(d)} samples differ, first {d[:6].tolist()}"
    return None


@pytest.mark.parametrize("name", list(frame_cases.FRAME_CASES))
def test_inplace_lists_parity(name):
    gold = json.load(open(os.path.join(GOLDEN, "frames.json")))["cases"][name]
    job, refs = frame_cases.build_case(name)
    init = refs[-1]
    with backend.Context(job.sp) as ctx:
        got = ctx.reconstruct(job, refs, out_init=init, stages=15, sparse=True, inplace=True)
        assert frame_cases.picture_digest(got) == gold["all"], f"{name} (in place, sparse) differs from the reference fixture"
        got = ctx.reconstruct(job, refs, out_init=init, stages=3, sparse=False, inplace=True)
        assert frame_cases.picture_digest(got) == gold["mc_res"], f"{name} (in place, dense) differs from the reference fixture"


@pytest.mark.parametrize("name", ["cfg3_3840x2160_10b_420_weighted", "cfg4b_1920x1080_10b_444"])
def test_inplace_fullsize(name):
    gold = json.load(open(os.path.join(GOLDEN, "frames_fullsize.json")))["cases"][name]
    job, refs = frame_cases.build_case(name)
    with backend.Context(job.sp) as ctx:
        got = ctx.reconstruct(job, refs, out_init=refs[-1], stages=15, sparse=True, inplace=True)
    assert frame_cases.picture_digest(got) == gold["all"]


@pytest.mark.parametrize("name", ["main10_416x240_420_weighted", "slices_tiles_8b_420", "rext_10b_444"])
def test_inplace_levels_and_derived_bs(name, oracle):
    """dequantisation and the bs derivation on the device, fed in place: the slice / tile CTB flags come from prepare_kernel"""
    job, refs = frame_cases.build_case(name)
    lv, quant, scaling = levels_cases.frame_levels(job, seed=77)
    job.coeffs = oracle.dequant_tus(job.tus, quant, lv, job.sp.bit_depth, scaling)
    job.vbs, job.hbs = oracle.boundary_strengths(job)
    want = oracle.replay(job, refs, out_init=refs[-1], stages=15)
    job.coeffs = lv
    with backend.Context(job.sp) as ctx:
        got = ctx.reconstruct(job, refs, out_init=refs[-1], stages=15, levels=(quant, scaling), derive_bs=True, inplace=True)
    assert _diff(got, want) is None, _diff(got, want)


def _submit(ctx, job, mutate):
    """one frame through the in-place path with `mutate(arenas, counts)` applied before the commit;
    returns (ohcu_frame_sync's code, ohcu_slot_sync's code, error text)"""
    lib, h = ctx.lib, ctx.handle
    out_slot = ctx.sp.dpb_slots - 1
    j = ctx.begin(out_slot)
    a, counts = ctx.append_inplace(j, job, coeffs="sparse", commit=False)
    mutate(a, counts)
    abi.check(lib.ohcu_frame_commit_lists(h, j, C.byref(counts)))
    ctx.stage(j)
    ctx.launch(j, out_slot)
    rc = lib.ohcu_frame_sync(h, j)
    msg = lib.ohcu_last_error().decode() if rc else ""
    rc_slot = lib.ohcu_slot_sync(h, out_slot)
    return rc, rc_slot, msg


def test_device_validation_rejects_malformed_lists(oracle):
    job, refs = frame_cases.build_case("cfg1_416x240_8b_420")
    want = oracle.replay(job, refs)
    PU, TU = abi.PU, abi.TU

    def pu_at(a, i):
        return PU.from_address(a.pus + i * C.sizeof(PU))

    def tu_at(a, i):
        return TU.from_address(a.tus + i * C.sizeof(TU))

    def crossing(a, n):          # a PU that crosses a CTB boundary would leave the CTB's cell map
        pu_at(a, 3).x0 = 60; pu_at(a, 3).w = 8

    def unaligned(a, n):
        pu_at(a, 5).x0 += 2

    def bad_slot(a, n):          # reference index that names no DPB slot
        st = abi.SliceTab.from_address(a.slices)
        for l in range(2):
            for i in range(16):
                st.ref_slot[l][i] = abi.NO_REF

    def self_reference(a, n):
        st = abi.SliceTab.from_address(a.slices)
        for l in range(2):
            for i in range(16):
                st.ref_slot[l][i] = job.sp.dpb_slots - 1

    def bad_index(a, n):
        abi.CtbIndex.from_address(a.ctbs).n_pu = 10 ** 6

    def bad_pairs(a, n):
        abi.CtbIndex.from_address(a.ctbs + C.sizeof(abi.CtbIndex)).n_coef += 1
        n.n_coef += 1

    def bad_tu(a, n):
        tu_at(a, 2).log2_size = 7

    def tu_outside_area(a, n):
        tu_at(a, 1).coeff_off = 12288

    def false_promise(a, n):     # all_inter although one CTB loses a PU
        ix = abi.CtbIndex.from_address(a.ctbs + 2 * C.sizeof(abi.CtbIndex))
        ix.n_pu -= 1
        n.all_inter = 1

    with backend.Context(job.sp, jobs=2) as ctx:
        for i, r in enumerate(refs[:-1]):
            ctx.upload(i, r)
        for mutate, text in ((crossing, "PU"), (unaligned, "PU"), (bad_slot, "PU"), (self_reference, "PU"), (bad_index, "index"),
                             (bad_pairs, "pair"), (bad_tu, "TU"), (tu_outside_area, "TU"), (false_promise, "all_inter")):
            rc, rc_slot, msg = _submit(ctx, job, mutate)
            assert rc == -1 and rc_slot == -1, (mutate.__name__, rc, rc_slot)
            assert text in msg, (mutate.__name__, msg)
            # the failure is the job's, not the context's: the next frame reconstructs
            rc, rc_slot, msg = _submit(ctx, job, lambda a, n: None)
            assert rc == 0 and rc_slot == 0, msg
            assert _diff(ctx.download(job.sp.dpb_slots - 1), want) is None


def test_commit_errors_on_the_host():
    job, refs = frame_cases.build_case("cfg1_416x240_8b_420")
    with backend.Context(job.sp) as ctx:
        lib, h = ctx.lib, ctx.handle
        j = ctx.begin(ctx.sp.dpb_slots - 1)
        n = abi.FrameCounts(n_pu=10 ** 9, n_slices=1)
        assert lib.ohcu_frame_commit_lists(h, j, C.byref(n)) == -6               # EOVERFLOW
        n = abi.FrameCounts(n_pu=4, n_tu=4, n_slices=1, coef_form=0)
        assert lib.ohcu_frame_commit_lists(h, j, C.byref(n)) == -1               # coded blocks need a coefficient form
        n = abi.FrameCounts(n_pu=4, n_slices=0)
        assert lib.ohcu_frame_commit_lists(h, j, C.byref(n)) == -1
        # appended lists and committed lists do not mix
        assert lib.ohcu_append_pu(h, j, job.pus.ctypes.data, 8) == 0
        n = abi.FrameCounts(n_pu=4, n_slices=1)
        assert lib.ohcu_frame_commit_lists(h, j, C.byref(n)) == -4               # ESTATE
        assert lib.ohcu_frame_abort(h, j) == 0


def test_append_rejects_records_that_would_corrupt_memory():
    """blocks that are not 4-aligned or cross a CTB boundary, reference indices that name no slot"""
    job, refs = frame_cases.build_case("cfg1_416x240_8b_420")
    with backend.Context(job.sp) as ctx:
        lib, h = ctx.lib, ctx.handle
        j = ctx.begin(ctx.sp.dpb_slots - 1)
        bad = job.pus[:1].copy(); bad["x0"] = 60; bad["w"] = 8
        assert lib.ohcu_append_pu(h, j, bad.ctypes.data, 1) == -1
        bad = job.pus[:1].copy(); bad["y0"] += 2
        assert lib.ohcu_append_pu(h, j, bad.ctypes.data, 1) == -1
        tu = job.tus[:1].copy(); tu["x0"] += 2
        assert lib.ohcu_append_tu(h, j, tu.ctypes.data, 1, job.coeffs.ctypes.data, len(job.coeffs)) == -1
        tu = job.tus[job.tus["log2_size"] == 3][:1].copy(); tu["x0"] = 60 if tu["c_idx"][0] == 0 else 28
        assert lib.ohcu_append_tu(h, j, tu.ctypes.data, 1, job.coeffs.ctypes.data, len(job.coeffs)) == -1
        # a reference index whose slice-table entry names no DPB slot is caught at stage
        sl = job.slices.copy(); sl["ref_slot"][:] = 0xff
        tab = abi.SliceTab.from_buffer_copy(sl[0].tobytes())
        abi.check(lib.ohcu_set_slice(h, j, 0, C.byref(tab)))
        assert lib.ohcu_append_pu(h, j, job.pus.ctypes.data, len(job.pus)) == 0
        assert lib.ohcu_frame_stage(h, j) == -1
        assert b"DPB slot" in lib.ohcu_last_error()


def test_pic_params_override(oracle):
    """PPS fields are per picture (hevc_filter.c:72, 541-543): a context created with other values reconstructs the
    picture correctly once the job carries the picture's own parameters"""
    job, refs = frame_cases.build_case("slices_10b_420_bypass")
    want = oracle.replay(job, refs, out_init=refs[-1])
    sp = abi.SeqParams.from_buffer_copy(job.sp)
    sp.pcm_or_bypass = 0; sp.cb_qp_offset = 5; sp.cr_qp_offset = -3; sp.no_tile_filter = 0
    pp = abi.PicParams(job.sp.pcm_or_bypass, job.sp.cb_qp_offset, job.sp.cr_qp_offset, job.sp.no_tile_filter)
    with backend.Context(sp) as ctx:
        out_slot = sp.dpb_slots - 1
        for i, r in enumerate(refs[:-1]):
            ctx.upload(i, r)
        for inplace in (False, True):
            ctx.upload(out_slot, refs[-1])
            j = ctx.begin(out_slot)
            abi.check(ctx.lib.ohcu_set_pic_params(ctx.handle, j, C.byref(pp)))
            ctx.sp.pcm_or_bypass = job.sp.pcm_or_bypass        # the harness consults it to decide whether is_pcm goes along
            if inplace:
                ctx.append_inplace(j, job)
            else:
                ctx.append(j, job)
            ctx.sp.pcm_or_bypass = 0
            ctx.flush(j)
            ctx.sync(j)
            got = ctx.download(out_slot)
            assert _diff(got, want) is None, f"inplace={inplace}: {_diff(got, want)}"


def test_pcm_samples(oracle):
    """ohcu_append_pcm: PCM blocks replace prediction + residual of their samples ahead of the loop filters
    (hls_pcm_sample -> put_pcm, hevcdec.c:1686-1721): checked against the oracle replay of the deblock + SAO
    stages on the picture with the PCM samples put in by numpy."""
    job = synth.make_job(256, 128, 10, 1, seed=91, sp_pcm_or_bypass=1)
    refs = [synth.make_picture(job.sp, 910 + s) for s in range(job.sp.dpb_slots)]
    rng = np.random.default_rng(5)
    blocks = [(64, 32, 4), (8, 8, 3), (224, 96, 5), (128, 0, 3)]             # luma x0, y0, log2 (CU-aligned)
    pred = oracle.replay(job, refs, out_init=refs[-1], stages=3)
    pcm = np.zeros((job.sp.height >> job.sp.log2_min_pu_size, job.sp.width >> job.sp.log2_min_pu_size), np.uint8)
    recs = []
    for x0, y0, l2 in blocks:
        n = 1 << l2
        for c in range(3):
            s = 1 if c else 0
            blk = (rng.integers(0, 256, size=(n >> s, n >> s)) << 2).astype(np.uint16)      # pcm_bit_depth 8 -> << (10 - 8)
            pred[c][y0 >> s:(y0 >> s) + (n >> s), x0 >> s:(x0 >> s) + (n >> s)] = blk
            recs.append((x0 >> s, y0 >> s, l2 - s, c, np.ascontiguousarray(blk)))
        pcm[y0 >> 2:(y0 + n) >> 2, x0 >> 2:(x0 + n) >> 2] = 1
    job.is_pcm = pcm.reshape(-1).copy()                                    # pcm_loop_filter_disable: deblock / SAO skip them
    want = oracle.replay(job, refs, out_init=pred, stages=12)
    with backend.Context(job.sp) as ctx:
        out_slot = job.sp.dpb_slots - 1
        for i, r in enumerate(refs[:-1]):
            ctx.upload(i, r)
        ctx.upload(out_slot, refs[-1])
        j = ctx.begin(out_slot)
        ctx.append(j, job)
        for x, y, l2, c, blk in recs:
            abi.check(ctx.lib.ohcu_append_pcm(ctx.handle, j, x, y, l2, c, blk.ctypes.data))
        assert ctx.lib.ohcu_append_pcm(ctx.handle, j, 250, 0, 3, 0, recs[0][4].ctypes.data) == -1      # outside the plane
        ctx.flush(j)
        ctx.sync(j)
        got = ctx.download(out_slot)
    assert _diff(got, want) is None, _diff(got, want)


def test_slot_release():
    job, refs = frame_cases.build_case("tiny_8x8")
    with backend.Context(job.sp) as ctx:
        lib, h = ctx.lib, ctx.handle
        ctx.upload(0, refs[0])
        assert lib.ohcu_slot_release(h, 0) == 0
        out = [np.empty(ctx.plane_shape(c), ctx.dtype) for c in range(3)]
        ptrs, strides = backend._plane_args(out)
        assert lib.ohcu_slot_download(h, 0, C.byref(ptrs), C.byref(strides)) == -4       # ESTATE: nothing to read
        ctx.upload(0, refs[0])                                                            # written again
        assert lib.ohcu_slot_download(h, 0, C.byref(ptrs), C.byref(strides)) == 0
        assert lib.ohcu_slot_release(h, 99) == -1


def test_two_contexts_on_two_devices(oracle):
    """one context per decoder and GPU in ONE process (function attributes are per device)"""
    lib = abi.load()
    if lib.ohcu_device_count() < 2:
        pytest.skip("needs two GPUs")
    job, refs = frame_cases.build_case("rext_10b_444")            # > 48 KB of dynamic shared memory in residual_kernel
    want = oracle.replay(job, refs, out_init=refs[-1])
    with backend.Context(job.sp, device=0) as a, backend.Context(job.sp, device=1) as b:
        for ctx in (b, a, b):
            got = ctx.reconstruct(job, refs, out_init=refs[-1])
            assert _diff(got, want) is None
