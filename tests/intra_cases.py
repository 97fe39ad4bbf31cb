"""Seeded intra-prediction cases shared by the oracle-vs-reference test, the golden digests and the GPU leaf test.

A case = one N x N block of one plane predicted in place from its neighbours: every mode 0..34, every combination of
the five availability flags, block sizes 4..32, luma and chroma of 4:2:0 / 4:2:2 / 4:4:4, positions inside the picture
and against its right / bottom border (clipped top-right / bottom-left runs), noise and smooth content (the strong
smoothing of 32x32 luma needs nearly linear neighbours), smoothing enabled / disabled."""
import hashlib

import numpy as np

DEPTHS = [8, 10, 12]
PLANES = [(1, 0), (1, 1), (2, 1), (3, 1)]          # (chroma_format_idc, c_idx)
PW, PH = 96, 96                                     # plane size in samples


def plane_of(bd, smooth, seed):
    rng = np.random.default_rng(seed)
    dt = np.uint8 if bd == 8 else np.uint16
    if smooth:
        yy, xx = np.mgrid[0:PH, 0:PW]
        a = (xx * 3 + yy * 2) * (1 << (bd - 8)) // 4 + (40 << (bd - 8)) + rng.integers(0, 2, size=(PH, PW))
    else:
        a = rng.integers(0, 1 << bd, size=(PH, PW))
    return np.ascontiguousarray(a.astype(dt))


def positions(n):
    """inside; right border (no top-right samples in the picture); bottom border; partially clipped runs"""
    pos = [(32, 32), (PW - n, 32), (32, PH - n), (PW - n, PH - n)]
    if n >= 8:
        pos += [(PW - n - 4, 32), (32, PH - n - 4)]
    return [(x, y) for x, y in pos if x >= 4 and y >= 4]


def groups():
    """(name, list of call parameter dicts)"""
    out = []
    for bd in DEPTHS:
        for cfi, c_idx in PLANES:
            for log2 in (2, 3, 4, 5):
                n = 1 << log2
                calls = []
                for smooth in (0, 1):
                    for (x, y) in positions(n):
                        for mode in range(35):
                            # all 32 availability patterns inside the picture on noise; the usual ones elsewhere
                            cands = range(32) if (not smooth and (x, y) == (32, 32)) else (31, 14, 10, 6, 0, 30, 15, 2, 8)
                            for cand in cands:
                                calls.append(dict(bd=bd, cfi=cfi, c_idx=c_idx, log2=log2, x=x, y=y, mode=mode, cand=cand,
                                                  smooth=smooth, smoothing_disabled=0, strong=1 if smooth else 0))
                for mode in (0, 2, 10, 18, 26, 34):        # intra_smoothing_disabled_flag
                    calls.append(dict(bd=bd, cfi=cfi, c_idx=c_idx, log2=log2, x=32, y=32, mode=mode, cand=31, smooth=0,
                                      smoothing_disabled=1, strong=0))
                out.append((f"intra_bd{bd}_cfi{cfi}_c{c_idx}_n{n}", calls))
    return out


def run_group(lib, calls, which):
    """all blocks of a group through lib (`which`: 'oracle' | 'ref' | a callable(plane, p) for the device), digest of the outputs"""
    h = hashlib.sha256()
    planes = {}
    for p in calls:
        key = (p["bd"], p["smooth"])
        if key not in planes:
            planes[key] = plane_of(p["bd"], p["smooth"], 77 + p["bd"] + p["smooth"])
        a = planes[key].copy()
        n = 1 << p["log2"]
        if callable(which):
            which(a, p)
        elif which == "oracle":
            filter_ok = int(not p["smoothing_disabled"] and (p["c_idx"] == 0 or p["cfi"] == 3))
            lib.intra_pred(p["bd"], a.ctypes.data, a.strides[0], PW, PH, p["x"], p["y"], p["log2"], p["c_idx"], p["mode"], p["cand"],
                           filter_ok, p["strong"])
        else:
            assert lib.intra_pred_ref(p["bd"], a.ctypes.data, a.strides[0], PW, PH, p["x"], p["y"], p["log2"], p["c_idx"], p["mode"],
                                      p["cand"], p["cfi"], p["smoothing_disabled"], p["strong"]) == 0
        # the block changed, nothing else did
        blk = a[p["y"]:p["y"] + n, p["x"]:p["x"] + n]
        h.update(np.ascontiguousarray(blk).tobytes())
        a[p["y"]:p["y"] + n, p["x"]:p["x"] + n] = planes[key][p["y"]:p["y"] + n, p["x"]:p["x"] + n]
        assert np.array_equal(a, planes[key]), "prediction wrote outside its block"
    return h.hexdigest()[:16]
