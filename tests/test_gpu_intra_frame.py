"""Frame-level intra prediction on the device: ohcu_iu records, residual planes, the CTB wavefront.
The pictures themselves are covered by the intra_* entries of frame_cases.FRAME_CASES / FULLSIZE_CASES in
test_gpu_parity.py, test_gpu_inplace.py and test_bs.py (reference digests); here: the error behaviour of the new
records, an I picture used as a reference by the next picture, and intra blocks together with device-side
dequantisation and derived boundary strengths."""
import ctypes as C

import numpy as np
import pytest

import frame_cases
import levels_cases
from hevc_b200 import abi, backend, synth

pytestmark = pytest.mark.gpu


def _diff(a, b):
    for c, (x, y) in enumerate(zip(a, b)):
        if not np.array_equal(x, y):
            d = np.argwhere(x != y)
            return f"plane {c}: {len(d)} samples differ, first {d[:6].tolist()}"
    return None


def test_context_without_intra_refuses_the_records():
    job, refs = frame_cases.build_case("intra_8b_420")
    sp = abi.SeqParams.from_buffer_copy(job.sp)
    sp.intra = 0
    with backend.Context(sp) as ctx:
        lib, h = ctx.lib, ctx.handle
        j = ctx.begin(sp.dpb_slots - 1)
        assert lib.ohcu_append_iu(h, j, job.ius.ctypes.data, len(job.ius)) == abi.ESTATE
        assert b"intra" in lib.ohcu_last_error()
        # a coded TB that asks for the residual planes is refused as well
        rc = lib.ohcu_append_tu(h, j, job.tus.ctypes.data, len(job.tus), job.coeffs.ctypes.data, len(job.coeffs))
        assert rc == abi.EINVAL
        a = abi.FrameArenas()
        abi.check(lib.ohcu_frame_arenas_get(h, j, C.byref(a)))
        assert not a.ius and not a.iu_ctbs and a.max_ius == 0
        n = abi.FrameCounts()
        n.n_iu = 1
        assert lib.ohcu_frame_commit_lists(h, j, C.byref(n)) == abi.ESTATE
        lib.ohcu_frame_abort(h, j)


def test_intra_at_14_bits_is_refused():
    """ff_hevc_pred_init (hevcpred.c:62-76) instantiates 8, 9, 10 and 12 bits: nothing to be identical to above that"""
    p = synth.default_params(64, 64, 14, 1, sp_intra=1)
    h = C.c_void_p()
    lib = abi.load()
    assert lib.ohcu_create(C.byref(h), 0, C.byref(p.sp)) == abi.EINVAL
    assert b"intra" in lib.ohcu_last_error()


def _bad_records(job):
    """(name, mutated copy of job.ius)"""
    sp = job.sp
    luma = np.flatnonzero(job.ius["c_idx"] == 0)
    out = []

    def mut(name, i, **kw):
        ius = job.ius.copy()
        for k, v in kw.items():
            ius[k][i] = v
        out.append((name, ius))

    i = int(luma[5])
    mut("mode", i, mode=35)
    mut("size", i, log2_size=6)
    mut("plane", i, c_idx=3)
    mut("unaligned", i, x0=job.ius["x0"][i] + 2)
    mut("outside", i, x0=sp.width)
    mut("flags", i, cand=0x40)
    top = int(luma[np.flatnonzero(job.ius["y0"][luma] == 0)[0]])
    mut("up_at_row_0", top, cand=abi.CAND_UP)
    left = int(luma[np.flatnonzero(job.ius["x0"][luma] == 0)[0]])
    mut("left_at_column_0", left, cand=abi.CAND_LEFT)
    # the last row of blocks of a CTB cannot have its bottom-left neighbour yet
    n = 1 << job.ius["log2_size"][luma].astype(int)
    low = luma[((job.ius["y0"][luma].astype(int) + n) % 64 == 0) & (job.ius["x0"][luma] > 0) & (job.ius["y0"][luma].astype(int) + n < sp.height)]
    mut("bottom_left_below_the_ctb", int(low[0]), cand=abi.CAND_BOTTOM_LEFT)
    big = luma[(job.ius["log2_size"][luma] < 5) & ((job.ius["x0"][luma].astype(int) + n) % 64 == 0)]
    j = int(big[0])
    mut("crossing", j, log2_size=job.ius["log2_size"][j] + 1)
    return out


def test_append_iu_rejects_malformed_records():
    job, refs = frame_cases.build_case("intra_8b_420")
    with backend.Context(job.sp) as ctx:
        lib, h = ctx.lib, ctx.handle
        for name, ius in _bad_records(job):
            j = ctx.begin(job.sp.dpb_slots - 1)
            assert lib.ohcu_append_iu(h, j, ius.ctypes.data, len(ius)) == abi.EINVAL, name
            assert b"intra block record" in lib.ohcu_last_error(), name
            lib.ohcu_frame_abort(h, j)
        # the blocks of a CTB have to arrive together
        j = ctx.begin(job.sp.dpb_slots - 1)
        order = np.r_[np.arange(1, len(job.ius)), 0]
        ius = np.ascontiguousarray(job.ius[order])
        assert lib.ohcu_append_iu(h, j, ius.ctypes.data, len(ius)) == abi.EINVAL
        assert b"contiguously" in lib.ohcu_last_error()
        lib.ohcu_frame_abort(h, j)


def test_device_validation_rejects_malformed_intra_blocks(oracle):
    job, refs = frame_cases.build_case("intra_8b_420")
    want = oracle.replay(job, refs, out_init=refs[-1])
    out_slot = job.sp.dpb_slots - 1

    def submit(ctx, ius, index_mutation=None):
        lib, h = ctx.lib, ctx.handle
        ctx.upload(out_slot, refs[-1])
        j = ctx.begin(out_slot)
        a, counts = ctx.append_inplace(j, job, coeffs="sparse", commit=False)
        C.memmove(a.ius, ius.ctypes.data, ius.nbytes)
        if index_mutation:
            index_mutation(a)
        abi.check(lib.ohcu_frame_commit_lists(h, j, C.byref(counts)))
        ctx.stage(j)
        ctx.launch(j, out_slot)
        rc = lib.ohcu_frame_sync(h, j)
        return rc, (lib.ohcu_last_error().decode() if rc else "")

    with backend.Context(job.sp, jobs=2) as ctx:
        for i, r in enumerate(refs[:-1]):
            ctx.upload(i, r)
        for name, ius in _bad_records(job):
            rc, msg = submit(ctx, ius)
            assert rc == abi.EINVAL and "intra block" in msg, (name, rc, msg)
            rc, msg = submit(ctx, job.ius)                       # the failure is the job's, not the context's
            assert rc == 0, msg
            assert _diff(ctx.download(out_slot), want) is None, name

        def bad_index(a):
            abi.IuIndex.from_address(a.iu_ctbs).n_iu = 10 ** 6
        rc, msg = submit(ctx, job.ius, bad_index)
        assert rc == abi.EINVAL and "intra block" in msg, (rc, msg)


def test_intra_picture_is_a_reference_for_the_next(oracle):
    """an all-intra picture (I), then a picture with inter and intra CUs (P/B) that predicts from it: the wavefront's
    output goes through the loop filter and the halo fill like any other picture"""
    w, h = 320, 192
    i_job = synth.make_job(w, h, 10, 1, seed=41, intra_pct=100)
    p_job = synth.make_job(w, h, 10, 1, seed=42, intra_pct=30, n_refs=2, weighted=1)
    ref0 = synth.make_picture(i_job.sp, 410)
    for s in range(len(p_job.slices)):
        p_job.slices["ref_slot"][s, :, :] = np.array([[2 if i % 2 == 0 else 0 for i in range(16)],
                                                      [0 if i % 2 == 0 else 2 for i in range(16)]])
    with backend.Context(i_job.sp, dpb_slots=4, jobs=2) as ctx:
        ctx.upload(0, ref0)
        j = ctx.begin(2); ctx.append_inplace(j, i_job, coeffs="sparse"); ctx.flush(j)
        j = ctx.begin(3); ctx.append(j, p_job, sparse=True); ctx.flush(j)
        ctx.sync()
        got_i, got_p = ctx.download(2), ctx.download(3)
    want_i = oracle.replay(i_job, [ref0, None, None, None])
    assert _diff(got_i, want_i) is None, _diff(got_i, want_i)
    want_p = oracle.replay(p_job, [ref0, None, want_i, None])
    assert _diff(got_p, want_p) is None, _diff(got_p, want_p)


@pytest.mark.parametrize("name", ["intra_8b_420", "intra_slices_tiles_8b", "intra_10b_422_rext"])
def test_intra_with_levels_and_derived_bs(name, oracle):
    job, refs = frame_cases.build_case(name)
    lv, quant, scaling = levels_cases.frame_levels(job, seed=78)
    job.coeffs = oracle.dequant_tus(job.tus, quant, lv, job.sp.bit_depth, scaling)
    job.vbs, job.hbs = oracle.boundary_strengths(job)
    want = oracle.replay(job, refs, out_init=refs[-1], stages=15)
    job.coeffs = lv
    with backend.Context(job.sp) as ctx:
        got = ctx.reconstruct(job, refs, out_init=refs[-1], stages=15, levels=(quant, scaling), derive_bs=True, inplace=True)
    assert _diff(got, want) is None, _diff(got, want)


def test_intra_residual_stage_alone_on_a_given_picture(oracle):
    """stages = 2: no MC; the intra blocks predict from the supplied picture and from each other"""
    job, refs = frame_cases.build_case("intra_only_10b_420_strong")
    with backend.Context(job.sp) as ctx:
        got = ctx.reconstruct(job, refs, out_init=refs[-1], stages=2)
    want = oracle.replay(job, refs, out_init=refs[-1], stages=2)
    assert _diff(got, want) is None, _diff(got, want)


def test_intra_blocks_read_pcm_neighbours(oracle):
    """PCM CUs are placed after the inter reconstruction and BEFORE the intra blocks (hls_pcm_sample runs in decoding order,
    hevcdec.c:1686-1721; its samples are what an intra block next to it predicts from).  Square 2Nx2N inter CUs of a picture
    with intra CUs become PCM; expected picture: oracle replay in the same four steps."""
    import copy
    job = synth.make_job(256, 192, 10, 1, seed=93, intra_pct=50, sp_pcm_or_bypass=1)
    refs = [synth.make_picture(job.sp, 930 + s) for s in range(job.sp.dpb_slots)]
    rng = np.random.default_rng(6)
    sq = np.flatnonzero((job.pus["w"] == job.pus["h"]) & (job.pus["w"] >= 8) & (job.pus["w"] <= 32))
    picks = [int(i) for i in sq[rng.choice(len(sq), size=min(12, len(sq)), replace=False)]]
    # 1. everything that is not intra
    inter = copy.copy(job)
    inter.ius = job.ius[:0]
    pic = oracle.replay(inter, refs, out_init=refs[-1], stages=3)
    # 2. the PCM samples
    pcm = np.zeros((job.sp.height >> 2, job.sp.width >> 2), np.uint8)
    recs = []
    for i in picks:
        x0, y0, n = int(job.pus["x0"][i]), int(job.pus["y0"][i]), int(job.pus["w"][i])
        l2 = n.bit_length() - 1
        for c in range(3):
            s = 1 if c else 0
            blk = (rng.integers(0, 256, size=(n >> s, n >> s)) << 2).astype(np.uint16)
            pic[c][y0 >> s:(y0 >> s) + (n >> s), x0 >> s:(x0 >> s) + (n >> s)] = blk
            recs.append((x0 >> s, y0 >> s, l2 - s, c, np.ascontiguousarray(blk)))
        pcm[y0 >> 2:(y0 + n) >> 2, x0 >> 2:(x0 + n) >> 2] = 1
    # 3. the intra blocks with their residuals, 4. the loop filters
    intra = copy.copy(job)
    intra.tus = np.ascontiguousarray(job.tus[(job.tus["flags"] & abi.TUF_INTRA) != 0])
    pic = oracle.replay(intra, refs, out_init=pic, stages=2)
    job.is_pcm = pcm.reshape(-1).copy()
    want = oracle.replay(job, refs, out_init=pic, stages=12)
    with backend.Context(job.sp) as ctx:
        out_slot = job.sp.dpb_slots - 1
        for i, r in enumerate(refs[:-1]):
            ctx.upload(i, r)
        ctx.upload(out_slot, refs[-1])
        j = ctx.begin(out_slot)
        ctx.append(j, job, sparse=True)
        for x, y, l2, c, blk in recs:
            abi.check(ctx.lib.ohcu_append_pcm(ctx.handle, j, x, y, l2, c, blk.ctypes.data))
        ctx.flush(j)
        ctx.sync(j)
        got = ctx.download(out_slot)
    assert _diff(got, want) is None, _diff(got, want)
