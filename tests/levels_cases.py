"""Seeded cases of the levels form (dequantisation on the device, ohcu_append_tu_levels).

Leaf cases drive one N x N block of un-scaled levels through scale_and_clip_coeff with the
parameters of derive_quant_parameters{,_c} (hevc_cabac.c:1472-1557, :2030-2055); frame cases
turn the coefficient arena of a frame case into levels with a quantiser per TU.
"""
import hashlib

import numpy as np

from hevc_b200.sparse import QM_FLAT, QUANT_DTYPE

QP_C = [29, 30, 31, 32, 33, 33, 34, 34, 35, 35, 36, 36, 37, 37]          # hevc_cabac.c:1454


def chroma_qp(qp_y, offset, cfi, bd):
    """quant_ctx->qp of derive_quant_parameters_c (hevc_cabac.c:1511-1527)"""
    qp_bd_offset = 6 * (bd - 8)
    qp_i = min(max(qp_y + offset, -qp_bd_offset), 57)
    if cfi == 1:
        qp = qp_i if qp_i < 30 else (qp_i - 6 if qp_i > 43 else QP_C[qp_i - 30])
    else:
        qp = min(qp_i, 51)
    return qp + qp_bd_offset


def scaling_lists(rng):
    sl = rng.integers(1, 256, size=(4, 6, 64), dtype=np.uint8)
    sl_dc = rng.integers(1, 256, size=(2, 6), dtype=np.uint8)
    return sl, sl_dc


def leaf_cases(seed=20260501, n=600):
    """(params dict, block of levels) per case"""
    rng = np.random.default_rng(seed)
    sl, sl_dc = scaling_lists(rng)
    out = []
    for i in range(n):
        bd = int(rng.choice([8, 9, 10, 12]))
        cfi = int(rng.choice([1, 2, 3]))
        c_idx = int(rng.integers(0, 3))
        log2 = int(rng.integers(2, 6))
        qp_y = int(rng.integers(-6 * (bd - 8), 52))
        off = int(rng.integers(-12, 13)) if c_idx else 0
        p = dict(bd=bd, cfi=cfi, c_idx=c_idx, log2=log2, qp_y=qp_y, qp_offset=off, intra=int(rng.integers(0, 2)),
                 transform_skip=int(rng.random() < 0.15), sl_enabled=int(rng.random() < 0.6))
        nn = 1 << (2 * log2)
        blk = np.zeros(nn, np.int16)
        k = int(rng.integers(1, nn + 1)) if rng.random() < 0.3 else int(rng.integers(1, max(2, nn // 8)))
        pos = rng.choice(nn, size=k, replace=False)
        blk[pos] = np.clip(np.rint(rng.laplace(0, 40 if i % 3 else 4000, size=k)), -32768, 32767).astype(np.int16)
        if i % 7 == 0:
            blk[pos[0]] = rng.choice([-32768, 32767, 1, -1])
        if i % 5 == 0:
            blk[0] = int(rng.integers(-3000, 3000)) or 7                     # DC present: dc_scale of 16x16 / 32x32
        out.append((p, blk))
    return sl, sl_dc, out


def ours_params(p):
    """(qp, matrix_id) of ohcu_tu_quant for a leaf case: the host-side rule of include/ohcu.h"""
    qp = p["qp_y"] + 6 * (p["bd"] - 8) if p["c_idx"] == 0 else chroma_qp(p["qp_y"], p["qp_offset"], p["cfi"], p["bd"])
    use_list = p["sl_enabled"] and not (p["transform_skip"] and p["log2"] > 2)
    return qp, (3 * (0 if p["intra"] else 1) + p["c_idx"]) if use_list else QM_FLAT


def frame_levels(job, seed):
    """A frame case's coefficient arena reinterpreted as un-scaled levels (scaled down so that few
    products saturate, a few extremes kept), one random quantiser per TU, random scaling lists."""
    rng = np.random.default_rng(seed)
    bd = job.sp.bit_depth
    levels = (job.coeffs.astype(np.int32) >> 3).astype(np.int16)
    nzi = np.flatnonzero(job.coeffs)
    if len(nzi):
        pick = rng.choice(nzi, size=max(1, len(nzi) // 200), replace=False)
        levels[pick] = rng.choice(np.array([-32768, 32767, -20000, 20000], np.int16), size=len(pick))
    quant = np.zeros(len(job.tus), QUANT_DTYPE)
    quant["qp"] = rng.integers(0, 52 + 6 * (bd - 8), size=len(job.tus))
    mid = rng.integers(0, 6, size=len(job.tus)).astype(np.uint8)
    flat = rng.random(len(job.tus)) < 0.4
    quant["matrix_id"] = np.where(flat, QM_FLAT, mid)
    return levels, quant, scaling_lists(rng)


def digest(blocks):
    h = hashlib.sha256()
    for b in blocks:
        h.update(np.ascontiguousarray(b).tobytes())
    return h.hexdigest()
