"""Dense coefficient blocks -> the sparse (position, level) form of ohcu_append_tu_sparse.

Host-side helper of the harness (tests, bench.py): a decoder produces the pairs directly
where the residual parser stores a significant coefficient (include/ohcu.h, "Sparse
variant"); synthetic frames are generated dense, so this converts them.
"""
import numpy as np

COEF_DTYPE = np.dtype([("pos", "<u2"), ("level", "<i2")])


def ctb_of_tus(tus, log2_ctb, hs, vs, ctb_w):
    chroma = tus["c_idx"] > 0
    x = tus["x0"].astype(np.int64) << np.where(chroma, hs, 0)
    y = tus["y0"].astype(np.int64) << np.where(chroma, vs, 0)
    return (x >> log2_ctb) + (y >> log2_ctb) * ctb_w


def to_sparse(tus, coeffs, log2_ctb, hs, vs, ctb_w, return_order=False):
    """Return (recs, pairs): recs = the TU records with coeff_off = block base inside the
    CTB's coefficient area and pad = pairs of the block; pairs = COEF_DTYPE array in record
    order.  Lists that are not grouped by CTB are grouped first (stable), cross_tu follows;
    return_order adds the permutation (recs[i] came from tus[order[i]])."""
    tus = np.asarray(tus)
    n_tu = len(tus)
    if n_tu == 0:
        empty = (tus.copy(), np.zeros(0, COEF_DTYPE))
        return empty + (np.zeros(0, np.int64),) if return_order else empty
    ctb = ctb_of_tus(tus, log2_ctb, hs, vs, ctb_w)
    order = np.arange(n_tu)
    if len(np.unique(ctb)) != 1 + np.count_nonzero(np.diff(ctb)):        # a CTB appears in more than one run
        order = np.argsort(ctb, kind="stable")
        new_index = np.empty(n_tu, np.int64)
        new_index[order] = np.arange(n_tu)
        tus = tus[order].copy()
        cross = (tus["flags"] & 8) != 0
        ok = cross & (tus["cross_tu"] < n_tu)
        tus["cross_tu"][ok] = new_index[tus["cross_tu"][ok]]
        ctb = ctb[order]
    nn = (1 << (2 * tus["log2_size"].astype(np.int64)))
    before = np.cumsum(nn) - nn                               # coefficients of the earlier blocks of the list
    first = np.r_[True, ctb[1:] != ctb[:-1]]
    run_start = np.maximum.accumulate(np.where(first, np.arange(n_tu), 0))
    base = before - before[run_start]                         # ... of the earlier blocks of the same CTB
    within = np.arange(int(nn.sum()), dtype=np.int64) - np.repeat(before, nn)
    vals = np.asarray(coeffs)[np.repeat(tus["coeff_off"].astype(np.int64), nn) + within]
    nz = vals != 0
    pairs = np.empty(int(nz.sum()), COEF_DTYPE)
    pairs["pos"] = (np.repeat(base, nn) + within)[nz]
    pairs["level"] = vals[nz]
    recs = tus.copy()
    recs["coeff_off"] = base
    recs["pad"] = np.bincount(np.repeat(np.arange(n_tu), nn)[nz], minlength=n_tu)
    return (recs, pairs, order) if return_order else (recs, pairs)


# ---- the levels form (ohcu_append_tu_levels): un-scaled levels + one quantiser per block
QUANT_DTYPE = np.dtype([("qp", "u1"), ("matrix_id", "u1")])
QM_FLAT = 0xff


def to_levels(tus, quant, levels, log2_ctb, hs, vs, ctb_w):
    """(recs, quant_in_record_order, pairs) of a dense arena of un-scaled levels."""
    recs, pairs, order = to_sparse(tus, levels, log2_ctb, hs, vs, ctb_w, return_order=True)
    return recs, np.ascontiguousarray(np.asarray(quant)[order]), pairs


# ---- the in-place form (ohcu_frame_arenas_get / ohcu_frame_commit_lists): where each CTB's records sit
CTB_INDEX_DTYPE = np.dtype([("first_pu", "<u4"), ("n_pu", "<u4"), ("first_tu", "<u4"), ("n_tu", "<u4"),
                            ("first_coef", "<u4"), ("n_coef", "<u4")])


def _runs(ctb, nctb, what):
    """first record and record count per CTB of a list grouped by CTB"""
    first = np.zeros(nctb, np.uint32)
    count = np.bincount(ctb, minlength=nctb).astype(np.uint32)
    if len(ctb):
        starts = np.flatnonzero(np.r_[True, ctb[1:] != ctb[:-1]])
        if len(starts) != len(np.unique(ctb)):
            raise ValueError(f"the {what} list is not grouped by CTB")
        first[ctb[starts]] = starts
    return first, count


def group_by_ctb(recs, ctb):
    """stable grouping of a record list by CTB (what a parser's CTB walk gives for free)"""
    if len(ctb) and len(np.unique(ctb)) != 1 + np.count_nonzero(np.diff(ctb)):
        order = np.argsort(ctb, kind="stable")
        return recs[order].copy(), ctb[order], order
    return recs, ctb, None


def ctb_index(pus, recs, log2_ctb, hs, vs, ctb_w, nctb, pairs_per_tu=None):
    """The ohcu_ctb_index array a parser fills while it walks the CTBs: PU list, TU list (records as handed to the
    library) and, for the sparse / levels forms, the pairs (recs['pad'] per block)."""
    idx = np.zeros(nctb, CTB_INDEX_DTYPE)
    pc = ((pus["x0"].astype(np.int64) >> log2_ctb) + (pus["y0"].astype(np.int64) >> log2_ctb) * ctb_w) if len(pus) else np.zeros(0, np.int64)
    idx["first_pu"], idx["n_pu"] = _runs(pc, nctb, "PU")
    if len(recs):
        tc = ctb_of_tus(recs, log2_ctb, hs, vs, ctb_w)
        idx["first_tu"], idx["n_tu"] = _runs(tc, nctb, "TU")
        if pairs_per_tu is not None:
            per_ctb = np.bincount(tc, weights=pairs_per_tu, minlength=nctb).astype(np.uint32)
            before = np.cumsum(pairs_per_tu) - pairs_per_tu
            idx["n_coef"] = per_ctb
            has = idx["n_tu"] > 0
            idx["first_coef"][has] = before[idx["first_tu"][has]]
    return idx


IU_INDEX_DTYPE = np.dtype([("first_iu", "<u4"), ("n_iu", "<u4")])


def iu_index(ius, log2_ctb, hs, vs, ctb_w, nctb):
    """The ohcu_iu_index array: where each CTB's intra blocks sit in a list in decoding order."""
    idx = np.zeros(nctb, IU_INDEX_DTYPE)
    if len(ius):
        idx["first_iu"], idx["n_iu"] = _runs(ctb_of_tus(ius, log2_ctb, hs, vs, ctb_w), nctb, "intra block")
    return idx
