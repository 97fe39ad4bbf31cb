"""Host-side conversion of dense coefficient blocks to the sparse (position, level) form."""
import numpy as np

import frame_cases
from hevc_b200 import sparse


def _geometry(sp):
    hs = 1 if sp.chroma_format_idc in (1, 2) else 0
    vs = 1 if sp.chroma_format_idc == 1 else 0
    ctb_w = (sp.width + (1 << sp.log2_ctb_size) - 1) >> sp.log2_ctb_size
    return sp.log2_ctb_size, hs, vs, ctb_w


def _rebuild(recs, pairs, ctb):
    """dense block of every record from the pairs, the way the kernel does it per CTB"""
    out = []
    cursor = 0
    for i, r in enumerate(recs):
        nn = 1 << (2 * int(r["log2_size"]))
        blk = np.zeros(nn, np.int16)
        p = pairs[cursor:cursor + int(r["pad"])]
        cursor += int(r["pad"])
        pos = p["pos"].astype(np.int64) - int(r["coeff_off"])
        assert ((pos >= 0) & (pos < nn)).all()
        blk[pos] = p["level"]
        out.append(blk)
    assert cursor == len(pairs)
    return out


def test_round_trip_and_layout():
    for name in ("cfg1_416x240_8b_420", "rext_10b_444", "ctb16_10b_444", "dense_coeffs_8b"):
        job, _ = frame_cases.build_case(name)
        geo = _geometry(job.sp)
        recs, pairs = sparse.to_sparse(job.tus, job.coeffs, *geo)
        ctb = sparse.ctb_of_tus(recs, *geo)
        assert len(recs) == len(job.tus) and (pairs["level"] != 0).all()
        blocks = _rebuild(recs, pairs, ctb)
        for r, t, blk in zip(recs, job.tus, blocks):
            nn = 1 << (2 * int(t["log2_size"]))
            assert np.array_equal(blk, job.coeffs[int(t["coeff_off"]):int(t["coeff_off"]) + nn])
        # the blocks of a CTB tile its coefficient area in list order, 16-aligned, inside 12288
        first = np.r_[True, ctb[1:] != ctb[:-1]]
        assert (recs["coeff_off"][first] == 0).all()
        nn = 1 << (2 * recs["log2_size"].astype(np.int64))
        nxt = recs["coeff_off"].astype(np.int64) + nn
        assert (recs["coeff_off"][1:][~first[1:]] == nxt[:-1][~first[1:]]).all()
        assert (recs["coeff_off"] % 16 == 0).all() and nxt.max() <= 12288


def test_ungrouped_list_is_grouped():
    job, _ = frame_cases.build_case("cfg1_416x240_8b_420")
    geo = _geometry(job.sp)
    rng = np.random.default_rng(1)
    perm = rng.permutation(len(job.tus))
    recs, pairs = sparse.to_sparse(job.tus[perm], job.coeffs, *geo)
    ctb = sparse.ctb_of_tus(recs, *geo)
    assert len(np.unique(ctb)) == 1 + np.count_nonzero(np.diff(ctb))
    assert int(recs["pad"].sum()) == len(pairs) == np.count_nonzero(
        np.concatenate([job.coeffs[int(t["coeff_off"]):int(t["coeff_off"]) + (1 << (2 * int(t["log2_size"])))] for t in job.tus]))


def test_empty():
    job, _ = frame_cases.build_case("tiny_8x8")
    recs, pairs = sparse.to_sparse(job.tus[:0], job.coeffs, *_geometry(job.sp))
    assert len(recs) == 0 and len(pairs) == 0
