"""The levels form: dequantisation (scale_and_clip_coeff, hevc_cabac.c:2030-2055) on the device.

CPU: the oracle's restatement against the reference's own static functions (through
oracle/ref_shim_cabac.c) and against the committed digest of the reference's outputs.
GPU: frames whose coefficients cross the ABI as un-scaled levels, bit-exact against the oracle
replay of the host-dequantised frame and against the scaled-coefficient forms."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import frame_cases
import levels_cases
import oracle_lib
from hevc_b200.sparse import QM_FLAT

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "levels_digest.json")


def _oracle_leaves():
    orc = oracle_lib.Lib("oracle")
    sl, sl_dc, cases = levels_cases.leaf_cases()
    outs = []
    for p, blk in cases:
        qp, mid = levels_cases.ours_params(p)
        b = blk.copy()
        m = None if mid == QM_FLAT else np.ascontiguousarray(sl[p["log2"] - 2, mid])
        dc = int(sl_dc[p["log2"] - 4, mid]) if (m is not None and p["log2"] >= 4) else 16
        orc._dequant_block(b.ctypes.data, p["log2"], p["bd"], qp, m.ctypes.data if m is not None else None, dc)
        outs.append(b)
    return sl, sl_dc, cases, outs


def test_oracle_dequant_matches_committed_reference_digest():
    *_, outs = _oracle_leaves()
    want = json.load(open(GOLDEN))
    assert want["cases"] == len(outs)
    assert levels_cases.digest(outs) == want["sha256"]


@pytest.mark.skipif(not oracle_lib.have_ref(), reason="needs the compiled reference (oracle/_ref)")
def test_oracle_dequant_vs_reference():
    ref = oracle_lib.Lib("ref")
    sl, sl_dc, cases, outs = _oracle_leaves()
    nonzero = saturated = 0
    for (p, blk), got in zip(cases, outs):
        b = blk.copy()
        qp_ref = ref._dequant_block(p["bd"], p["cfi"], p["c_idx"], p["log2"], p["qp_y"], p["qp_offset"], p["intra"],
                                    p["transform_skip"], p["sl_enabled"], sl.ctypes.data, sl_dc.ctypes.data, b.ctypes.data)
        assert qp_ref == levels_cases.ours_params(p)[0], p
        assert np.array_equal(b, got), p
        nonzero += int(np.count_nonzero(b)); saturated += int(np.count_nonzero(np.abs(b.astype(np.int32)) >= 32767))
    assert nonzero > 10000 and saturated > 50          # the cases reach the scaled range and the clip


LEVEL_FRAMES = ["cfg1_416x240_8b_420", "main10_416x240_420_weighted", "rext_10b_444", "rext_8b_444_bypass",
                "dense_coeffs_10b_444", "ctb32_8b", "main12_420"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", LEVEL_FRAMES)
def test_levels_frame_parity(name):
    from hevc_b200 import backend
    job, refs = frame_cases.build_case(name)
    levels, quant, scaling = levels_cases.frame_levels(job, seed=77)
    orc = oracle_lib.Lib("oracle")
    coeffs = orc.dequant_tus(job.tus, quant, levels, job.sp.bit_depth, scaling)
    assert np.count_nonzero(coeffs != levels) > 0
    saved = job.coeffs
    try:
        job.coeffs = coeffs
        want = orc.replay(job, refs)
        with backend.Context(job.sp) as ctx:
            dense = ctx.reconstruct(job, refs)
        job.coeffs = levels
        with backend.Context(job.sp) as ctx:
            got = ctx.reconstruct(job, refs, levels=(quant, scaling))
    finally:
        job.coeffs = saved
    for c, (g, w, d) in enumerate(zip(got, want, dense)):
        assert np.array_equal(g, w), f"{name} plane {c}: levels form differs from the oracle"
        assert np.array_equal(g, d), f"{name} plane {c}: levels form differs from the scaled-coefficient form"


@pytest.mark.gpu
def test_levels_api_errors():
    from hevc_b200 import backend
    job, refs = frame_cases.build_case("cfg1_416x240_8b_420")
    levels, quant, scaling = levels_cases.frame_levels(job, seed=78)
    with backend.Context(job.sp) as ctx:
        lib, h = ctx.lib, ctx.handle
        recs, q, pairs = backend.to_levels(job.tus, quant, levels, *_geometry(job.sp))
        j = ctx.begin(job.sp.dpb_slots - 1)
        bad = q.copy(); bad["qp"][0] = 76
        assert lib.ohcu_append_tu_levels(h, j, recs.ctypes.data, bad.ctypes.data, len(recs), pairs.ctypes.data, len(pairs)) == -1
        bad = q.copy(); bad["matrix_id"][0] = 6
        assert lib.ohcu_append_tu_levels(h, j, recs.ctypes.data, bad.ctypes.data, len(recs), pairs.ctypes.data, len(pairs)) == -1
        assert lib.ohcu_append_tu_levels(h, j, recs.ctypes.data, None, len(recs), pairs.ctypes.data, len(pairs)) == -1
        # levels and scaled pairs do not mix in one frame
        assert lib.ohcu_append_tu_levels(h, j, recs.ctypes.data, q.ctypes.data, len(recs), pairs.ctypes.data, len(pairs)) == 0
        assert lib.ohcu_append_tu_sparse(h, j, recs.ctypes.data, len(recs), pairs.ctypes.data, len(pairs)) == -4
        # a record names a scaling list but none was set: refused at staging
        if (q["matrix_id"] != QM_FLAT).any():
            assert lib.ohcu_frame_stage(h, j) == -1
        else:
            assert lib.ohcu_frame_abort(h, j) == 0


def _geometry(sp):
    hs = 1 if sp.chroma_format_idc in (1, 2) else 0
    vs = 1 if sp.chroma_format_idc == 1 else 0
    return sp.log2_ctb_size, hs, vs, (sp.width + (1 << sp.log2_ctb_size) - 1) >> sp.log2_ctb_size
