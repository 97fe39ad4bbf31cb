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
