"""Pins the oracle (oracle/hevc_oracle.c + replay.c) against the reference's OWN compiled
code (oracle/_ref/liboh_ref.so built from /root/reference by oracle/Makefile):
every HEVCDSPContext leaf on seeded inputs, and whole frames through the reference's
luma_mc_* / chroma_mc_* drivers and ff_hevc_hls_filters.  CPU only."""
import numpy as np
import pytest

import frame_cases
import leaf_cases
import oracle_lib


def test_ref_table_is_the_reference_struct(reflib):
    # sizeof(HEVCDSPContext) on x86-64 as probed in SURVEY.md section 8(a1)
    assert reflib.lib.ohref_sizeof_dsp_context() == 3568


def test_dct_matrix_rule(oracle):
    # spot values of transform[32][32] (hevcdsp.c:32-97)
    f = oracle.lib.or_dct_coef
    assert [f(0, c) for c in (0, 7, 31)] == [64, 64, 64]
    assert [f(1, c) for c in (0, 1, 15, 16, 31)] == [90, 90, 4, -4, -90]
    assert [f(16, c) for c in (0, 1, 2, 3)] == [64, -64, -64, 64]
    assert [f(8, c) for c in (0, 1, 2, 3)] == [83, 36, -36, -83]
    assert [f(31, c) for c in (0, 1, 31)] == [4, -13, -4]


def test_leaves_match_reference(oracle, reflib):
    bad = []
    cases = leaf_cases.all_cases()
    for name, run in cases:
        if run(oracle) != run(reflib):
            bad.append(name)
    assert not bad, f"{len(bad)}/{len(cases)} leaf cases differ: {bad[:10]}"


@pytest.mark.parametrize("name", list(frame_cases.FRAME_CASES))
def test_frame_matches_reference_drivers(name, oracle, reflib):
    job, refs = frame_cases.build_case(name)
    init = refs[-1]
    for stages in (0, oracle_lib.STAGE_MC, oracle_lib.STAGE_MC | oracle_lib.STAGE_RESIDUAL,
                   oracle_lib.STAGE_DEBLOCK, oracle_lib.STAGE_SAO):
        a = oracle.replay(job, refs, out_init=init, stages=stages)
        b = reflib.replay(job, refs, out_init=init, stages=stages)
        for c, (x, y) in enumerate(zip(a, b)):
            assert np.array_equal(x, y), f"{name} stages={stages} plane {c}: {np.argwhere(x != y)[:4].tolist()}"


def test_filters_do_something(oracle):
    """guards against vacuous parity: each stage must change pixels"""
    job, refs = frame_cases.build_case("main10_416x240_420_weighted")
    prev = None
    for stages in (1, 3, 7, 15):
        cur = oracle.replay(job, refs, stages=stages)
        if prev is not None:
            assert sum(int((x != y).sum()) for x, y in zip(cur, prev)) > 1000
        prev = cur


def test_ctb16_subsampled_chroma_is_rejected(oracle):
    """log2_ctb_size 4 with 4:2:0/4:2:2: the reference's CTB-delayed SAO reads chroma
    whose horizontal edges are not filtered yet (hevc_filter.c:715) — out of contract."""
    from hevc_b200 import synth
    job = synth.make_job(64, 32, 8, 1, seed=1, sp_log2_ctb_size=4, split_pct=(0, 0, 40))
    refs = [synth.make_picture(job.sp, s) for s in range(job.sp.dpb_slots)]
    with pytest.raises(RuntimeError):
        oracle.replay(job, refs)


# ---- the SIMD CPU baseline (oracle/_ref/liboh_ref_simd.so): the reference's own SSE intrinsics wired into its
# table by oracle/ref_simd_init.c.  It is what bench.py times as the CPU arm, so it must reconstruct the SAME pictures.
# Known, documented difference (SURVEY 8c hazard 1): the SSE2 IDCT ignores col_limit (x86/hevc_idct_sse.c:607-720), so
# frames whose blocks carry coefficients BEYOND col_limit (the dense_coeffs_* stress cases, which no parser produces)
# differ from the C table; every other case is bit-identical.
@pytest.fixture(scope="module")
def simdlib():
    if not oracle_lib.have_ref_simd():
        pytest.skip("oracle/_ref/liboh_ref_simd.so not available and /root/reference absent")
    return oracle_lib.Lib("ref_simd")


@pytest.mark.parametrize("name", [n for n in frame_cases.FRAME_CASES if not n.startswith("dense_coeffs")])
def test_simd_reference_matches_c_table(name, simdlib):
    import json
    import os
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "frames.json")))["cases"][name]
    job, refs = frame_cases.build_case(name)
    for key, st in (("mc", 1), ("mc_res", 3), ("mc_res_db", 7), ("all", 15)):
        out = simdlib.replay(job, refs, out_init=refs[-1], stages=st)
        assert frame_cases.picture_digest(out) == gold[key], f"{name}/{key}: SIMD reference differs from the C table"


@pytest.mark.parametrize("name", list(frame_cases.FULLSIZE_CASES))
def test_simd_reference_fullsize(name, simdlib):
    import json
    import os
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "frames_fullsize.json")))["cases"][name]
    job, refs = frame_cases.build_case(name)
    out = simdlib.replay(job, refs, out_init=refs[-1], stages=15)
    assert frame_cases.picture_digest(out) == gold["all"], f"{name}: SIMD reference differs from the C table"


def test_simd_idct_ignores_col_limit(simdlib, reflib):
    """the documented hazard itself: dense coefficients beyond col_limit -> the two reference tables disagree"""
    job, refs = frame_cases.build_case("dense_coeffs_8b")
    a = simdlib.replay(job, refs, out_init=refs[-1], stages=3)
    b = reflib.replay(job, refs, out_init=refs[-1], stages=3)
    assert any(not np.array_equal(x, y) for x, y in zip(a, b))
