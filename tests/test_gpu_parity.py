"""Parity of the CUDA path (through the C ABI) against the CPU oracle and the committed
golden fixtures.  Bit-exact: all arithmetic on this path is integer."""
import json
import os

import numpy as np
import pytest

import frame_cases
from hevc_b200 import backend, synth

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
STAGE_SETS = {"mc": 1, "mc_res": 3, "mc_res_db": 7, "all": 15}


def _diff(a, b):
    for c, (x, y) in enumerate(zip(a, b)):
        if not np.array_equal(x, y):
            d = np.argwhere(x != y)
            return f"plane {c}: {len(d)} samples differ, first {d[:6].tolist()}"
    return None


@pytest.mark.parametrize("name", list(frame_cases.FRAME_CASES))
def test_frame_parity(name, oracle):
    gold = json.load(open(os.path.join(GOLDEN, "frames.json")))["cases"][name]
    job, refs = frame_cases.build_case(name)
    init = refs[-1]
    with backend.Context(job.sp) as ctx:
        for key, st in STAGE_SETS.items():
            got = ctx.reconstruct(job, refs, out_init=init, stages=st)
            want = oracle.replay(job, refs, out_init=init, stages=st)
            assert _diff(got, want) is None, f"{name}/{key}: {_diff(got, want)}"
            assert frame_cases.picture_digest(got) == gold[key], f"{name}/{key} differs from the reference fixture"


@pytest.mark.parametrize("name", list(frame_cases.FULLSIZE_CASES))
def test_fullsize_parity(name, oracle):
    """BASELINE.json configs[1..4] at their NAMED sizes (1080p 8-bit, 4K Main10 weighted, 1080p 10-bit 4:2:2 and
    4:4:4, 8K Main10), one picture each: plane for plane against the oracle replay, and the digests the compiled
    reference wrote (tests/golden/frames_fullsize.json).  Dense blocks for MC + residual, sparse pairs for the
    full pipeline, so both coefficient forms run at size."""
    gold = json.load(open(os.path.join(GOLDEN, "frames_fullsize.json")))["cases"][name]
    job, refs = frame_cases.build_case(name)
    init = refs[-1]
    with backend.Context(job.sp) as ctx:
        got = ctx.reconstruct(job, refs, out_init=init, stages=3)
        assert frame_cases.picture_digest(got) == gold["mc_res"], f"{name}/mc_res differs from the reference fixture"
        got = ctx.reconstruct(job, refs, out_init=init, stages=15, sparse=True)
        want = oracle.replay(job, refs, out_init=init, stages=15)
        assert _diff(got, want) is None, f"{name}/all: {_diff(got, want)}"
        assert frame_cases.picture_digest(got) == gold["all"], f"{name}/all differs from the reference fixture"


@pytest.mark.parametrize("name", ["main10_416x240_420_weighted", "rext_10b_422", "rext_8b_444_bypass",
                                  "slices_tiles_8b_420", "slices_10b_420_bypass", "mono_8b"])
@pytest.mark.parametrize("stage", [2, 4, 8, 12])
def test_single_stage_on_given_picture(name, stage, oracle):
    """residual / deblock / SAO alone on a supplied picture (no MC)"""
    job, refs = frame_cases.build_case(name)
    init = refs[-1]
    with backend.Context(job.sp) as ctx:
        got = ctx.reconstruct(job, refs, out_init=init, stages=stage)
    want = oracle.replay(job, refs, out_init=init, stages=stage)
    assert _diff(got, want) is None, _diff(got, want)


def test_cfg1_golden_picture():
    z = np.load(os.path.join(GOLDEN, "cfg1_416x240.npz"))
    job, refs = frame_cases.build_case("cfg1_416x240_8b_420")
    with backend.Context(job.sp) as ctx:
        got = ctx.reconstruct(job, refs, out_init=refs[-1])
    for c, p in enumerate(got):
        assert np.array_equal(p, z[f"plane{c}"])


def test_unordered_pu_list_is_sorted(oracle):
    job, refs = frame_cases.build_case("cfg1_416x240_8b_420")
    rng = np.random.default_rng(3)
    job.pus = job.pus[rng.permutation(len(job.pus))].copy()
    with backend.Context(job.sp) as ctx:
        got = ctx.reconstruct(job, refs, stages=1)
    want = oracle.replay(job, refs, stages=1)
    assert _diff(got, want) is None


def test_frames_in_flight_and_reference_chain(oracle):
    """frame k+1 predicts from frame k's output slot: ordering by events on DPB slots"""
    sp_job = synth.make_job(256, 128, 8, 1, seed=11)
    refs = [synth.make_picture(sp_job.sp, 50 + s) for s in range(2)]
    jobs = [synth.make_job(256, 128, 8, 1, seed=20 + k, n_refs=2) for k in range(4)]
    with backend.Context(sp_job.sp, dpb_slots=6, jobs=3) as ctx:
        ctx.upload(0, refs[0]); ctx.upload(1, refs[1])
        chain = [refs[0], refs[1]]
        for k, jb in enumerate(jobs):
            # slots: references are the two most recent pictures
            for s in range(len(jb.slices)):
                jb.slices["ref_slot"][s, :, :] = np.array([[k + 1 if i % 2 == 0 else k for i in range(16)],
                                                           [k if i % 2 == 0 else k + 1 for i in range(16)]])
            j = ctx.begin(k + 2)
            ctx.append(j, jb)
            ctx.flush(j)
        ctx.sync()
        outs = [ctx.download(k + 2) for k in range(4)]
    for k, jb in enumerate(jobs):
        dpb = [None] * 6
        for i, p in enumerate(chain):
            dpb[i] = p
        jb.sp.dpb_slots = 6
        want = oracle.replay(jb, dpb)
        assert _diff(outs[k], want) is None, f"frame {k}: {_diff(outs[k], want)}"
        chain.append(want)


@pytest.mark.parametrize("name", list(frame_cases.FRAME_CASES))
def test_sparse_coefficients_parity(name):
    """The same frames with the coefficients sent as (position, level) pairs
    (ohcu_append_tu_sparse): identical pictures, checked against the reference fixtures."""
    gold = json.load(open(os.path.join(GOLDEN, "frames.json")))["cases"][name]
    job, refs = frame_cases.build_case(name)
    init = refs[-1]
    with backend.Context(job.sp) as ctx:
        for key in ("mc_res", "all"):
            got = ctx.reconstruct(job, refs, out_init=init, stages=STAGE_SETS[key], sparse=True)
            assert frame_cases.picture_digest(got) == gold[key], f"{name}/{key} (sparse) differs from the reference fixture"


def test_sparse_residual_only_and_ungrouped_list(oracle):
    """Residual alone on a given picture; a TU list that is not grouped by CTB is grouped by the
    harness (the library itself refuses it, see test_sparse_append_errors)."""
    job, refs = frame_cases.build_case("rext_10b_444")
    init = refs[-1]
    want = oracle.replay(job, refs, out_init=init, stages=2)
    with backend.Context(job.sp) as ctx:
        got = ctx.reconstruct(job, refs, out_init=init, stages=2, sparse=True)
        assert _diff(got, want) is None, _diff(got, want)
        no_cross = (job.tus["flags"] & 8) == 0
        if no_cross.all():
            rng = np.random.default_rng(5)
            job.tus = job.tus[rng.permutation(len(job.tus))].copy()
            got = ctx.reconstruct(job, refs, out_init=init, stages=2, sparse=True)
            assert _diff(got, want) is None, _diff(got, want)


def test_sparse_append_errors():
    from hevc_b200 import abi
    job, refs = frame_cases.build_case("cfg1_416x240_8b_420")
    with backend.Context(job.sp) as ctx:
        lib, h = ctx.lib, ctx.handle
        recs, pairs = ctx.sparse_lists(job)
        j = ctx.begin(ctx.sp.dpb_slots - 1)
        # pair count that does not match the records
        assert lib.ohcu_append_tu_sparse(h, j, recs.ctypes.data, len(recs), pairs.ctypes.data, len(pairs) - 1) == -1
        # a CTB met twice
        twice = np.concatenate([recs[:1], recs[-1:], recs[:1]])
        n = int(twice["pad"].sum())
        assert lib.ohcu_append_tu_sparse(h, j, twice.ctypes.data, 3, pairs.ctypes.data, n) == -1
        # block outside the CTB's coefficient area
        bad = recs[:1].copy()
        bad["coeff_off"] = 12288
        assert lib.ohcu_append_tu_sparse(h, j, bad.ctypes.data, 1, pairs.ctypes.data, int(bad["pad"][0])) == -1
        # nothing was appended by the failed calls; dense and sparse do not mix
        assert lib.ohcu_append_tu_sparse(h, j, recs.ctypes.data, len(recs), pairs.ctypes.data, len(pairs)) == 0
        assert lib.ohcu_append_tu(h, j, job.tus.ctypes.data, len(job.tus), job.coeffs.ctypes.data, len(job.coeffs)) == -4
        assert lib.ohcu_frame_abort(h, j) == 0
        j = ctx.begin(ctx.sp.dpb_slots - 1)
        assert lib.ohcu_append_tu(h, j, job.tus.ctypes.data, len(job.tus), job.coeffs.ctypes.data, len(job.coeffs)) == 0
        assert lib.ohcu_append_tu_sparse(h, j, recs.ctypes.data, len(recs), pairs.ctypes.data, len(pairs)) == -4
        assert lib.ohcu_frame_abort(h, j) == 0


def test_threaded_submission_ring(oracle):
    """Four host threads submit 96 frames through a ring of 8 jobs / 8 slots with asynchronous
    downloads (the flow of bench.py's e2e leg); every downloaded picture is compared."""
    from concurrent.futures import ThreadPoolExecutor
    jobs = [synth.make_job(256, 128, 10, 1, seed=40 + k) for k in range(4)]
    sp = jobs[0].sp
    refs = [synth.make_picture(sp, 70 + s) for s in range(2)]
    E = 8
    want = []
    for jb in jobs:
        jb.sp.dpb_slots = 2 + E
        want.append(oracle.replay(jb, list(refs) + [None] * E))
    bad = []
    with backend.Context(sp, dpb_slots=2 + E, jobs=E) as ctx:
        ctx.upload(0, refs[0]); ctx.upload(1, refs[1])
        outs = [ctx.pinned_picture() for _ in range(E)]
        sparse = [ctx.sparse_lists(jb) for jb in jobs]
        pending = [None] * E

        def submit(i):
            k = i % E
            if pending[k] is not None:
                ctx.slot_sync(2 + k)
                d = _diff(outs[k], want[pending[k] % len(jobs)])
                if d is not None:
                    bad.append((pending[k], d))
            j = ctx.begin(2 + k)
            ctx.append(j, jobs[i % len(jobs)], sparse=sparse[i % len(jobs)] if i & 1 else False)
            ctx.flush(j)
            ctx.download_async(2 + k, outs[k])
            pending[k] = i

        # frames that share a slot must not be in flight together on the host side: E frames per wave
        with ThreadPoolExecutor(4) as pool:
            for wave in range(12):
                list(pool.map(submit, range(wave * E, (wave + 1) * E)))
        for k in range(E):
            ctx.slot_sync(2 + k)
            d = _diff(outs[k], want[pending[k] % len(jobs)])
            if d is not None:
                bad.append((pending[k], d))
    assert not bad, bad[:3]
