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
