"""Seeded per-entry cases for the HEVCDSPContext leaves, runnable on any checker
library (oracle_lib.Lib('oracle') / Lib('ref')) and, through the same argument
meaning, on the CUDA per-entry shims (hevc_b200.dsp).

Each case is (name, fn) where fn(lib) -> bytes of every output the entry writes.
Inputs come from numpy default_rng(seed) so every consumer sees the same bytes.
"""
import numpy as np

WIDTHS = [2, 4, 6, 8, 12, 16, 24, 32, 48, 64]      # ff_hevc_pel_weight domain, hevcdec.c:49
DEPTHS = [8, 10, 12, 9, 14]      # every depth ff_hevc_dsp_init instantiates (hevcdsp.c:514-545)


def _pix_dtype(bd):
    return np.uint8 if bd == 8 else np.uint16


def _rand_pix(rng, shape, bd, smooth=False):
    maxv = (1 << bd) - 1
    if smooth:
        base = rng.integers(0, maxv + 1)
        a = base + rng.integers(-6 << (bd - 8), (6 << (bd - 8)) + 1, size=shape)
        return np.clip(a, 0, maxv).astype(_pix_dtype(bd))
    return rng.integers(0, maxv + 1, size=shape).astype(_pix_dtype(bd))


def _rand_coeffs(rng, n, sigma=300, outlier=0.02):
    c = np.rint(rng.laplace(0, sigma, size=n)).astype(np.int64)
    m = rng.random(n) < outlier
    c[m] = rng.integers(-32768, 32768, size=int(m.sum()))
    return np.clip(c, -32768, 32767).astype(np.int16)


def _p(a):
    return a.ctypes.data


# ---------------------------------------------------------------- residual
def case_transform_add(seed, bd, log2):
    def run(lib):
        rng = np.random.default_rng(seed)
        n = 1 << log2
        dst = _rand_pix(rng, (n, n + 8), bd)
        co = _rand_coeffs(rng, n * n, sigma=60 << (bd - 8))
        lib.transform_add(bd, _p(dst), _p(co), dst.strides[0], n)
        return dst.tobytes()
    return f"transform_add_bd{bd}_n{1 << log2}_s{seed}", run


def case_rdpcm(seed, log2, mode):
    def run(lib):
        rng = np.random.default_rng(seed)
        co = _rand_coeffs(rng, 1 << (2 * log2), sigma=3000, outlier=0.1)
        lib.transform_rdpcm(8, _p(co), log2, mode)
        return co.tobytes()
    return f"rdpcm_l{log2}_m{mode}_s{seed}", run


def case_skip(seed, bd, log2):
    def run(lib):
        rng = np.random.default_rng(seed)
        co = _rand_coeffs(rng, 1 << (2 * log2), sigma=2000, outlier=0.1)
        lib.transform_skip(bd, _p(co), log2)
        return co.tobytes()
    return f"skip_bd{bd}_l{log2}_s{seed}", run


def case_dst4(seed, bd):
    def run(lib):
        rng = np.random.default_rng(seed)
        co = _rand_coeffs(rng, 16, sigma=800, outlier=0.1)
        lib.idct_4x4_luma(bd, _p(co))
        return co.tobytes()
    return f"dst4_bd{bd}_s{seed}", run


def case_idct(seed, bd, log2, col_limit, dense=True):
    def run(lib):
        rng = np.random.default_rng(seed)
        n = 1 << log2
        co = _rand_coeffs(rng, n * n, sigma=500, outlier=0.05)
        if not dense:      # zero everything past an anti-diagonal, like a real last position
            y, x = np.divmod(np.arange(n * n), n)
            co[(x + y) > max(col_limit - 4, 0)] = 0
        lib.idct(bd, _p(co), log2, col_limit)
        return co.tobytes()
    return f"idct_bd{bd}_l{log2}_c{col_limit}_{'d' if dense else 's'}_s{seed}", run


def case_idct_dc(seed, bd, log2):
    def run(lib):
        rng = np.random.default_rng(seed)
        co = _rand_coeffs(rng, 1 << (2 * log2), sigma=5000, outlier=0.2)
        lib.idct_dc(bd, _p(co), log2)
        return co.tobytes()
    return f"idct_dc_bd{bd}_l{log2}_s{seed}", run


# ---------------------------------------------------------------- SAO
def _sao_offsets(rng, bd, edge):
    m = (1 << (min(bd, 10) - 5)) - 1
    v = rng.integers(0, m + 1, size=4)
    if edge:
        v[2:] = -v[2:]
    else:
        v = v * rng.choice([-1, 1], size=4)
    return np.concatenate([[0], v]).astype(np.int16)


def case_sao_band(seed, bd):
    def run(lib):
        rng = np.random.default_rng(seed)
        w, h = int(rng.integers(1, 65)), int(rng.integers(1, 65))
        src = _rand_pix(rng, (h, w + 3), bd)
        dst = _rand_pix(rng, (h, w + 5), bd)
        off = _sao_offsets(rng, bd, False)
        lib.sao_band_filter(bd, _p(dst), _p(src), dst.strides[0], src.strides[0], _p(off),
                            int(rng.integers(0, 32)), w, h)
        return dst.tobytes()
    return f"sao_band_bd{bd}_s{seed}", run


def case_sao_edge(seed, bd, eo, restore_idx):
    def run(lib):
        rng = np.random.default_rng(seed)
        w, h = int(rng.integers(2, 65)), int(rng.integers(2, 65))
        big = _rand_pix(rng, (h + 2, w + 2), bd, smooth=bool(seed & 1))
        src = big[1:, 1:]                        # 1-sample halo around the block
        dst = _rand_pix(rng, (h, w + 4), bd)
        off = _sao_offsets(rng, bd, True)
        lib.sao_edge_filter(bd, _p(dst), _p(src), dst.strides[0], big.strides[0], _p(off), eo, w, h)
        if restore_idx >= 0:
            borders = rng.integers(0, 2, size=4).astype(np.int32)
            ve = rng.integers(0, 2, size=2).astype(np.uint8)
            he = rng.integers(0, 2, size=2).astype(np.uint8)
            de = rng.integers(0, 2, size=4).astype(np.uint8)
            lib.sao_edge_restore(bd, restore_idx, _p(dst), _p(src), dst.strides[0], big.strides[0],
                                 _p(off), eo, _p(borders), w, h, _p(ve), _p(he), _p(de))
        return dst.tobytes()
    return f"sao_edge_bd{bd}_eo{eo}_r{restore_idx}_s{seed}", run


# ---------------------------------------------------------------- MC
def case_mc(seed, family, bd, luma, width, mx, my):
    """family in put / uni / uni_w / bi / bi_w"""
    def run(lib):
        rng = np.random.default_rng(seed)
        h = int(rng.choice([2, 4, 6, 8, 12, 16, 24, 32, 48, 64])) if not luma else int(rng.choice([4, 8, 12, 16, 32, 64]))
        pad = 8
        big = _rand_pix(rng, (h + 2 * pad, width + 2 * pad), bd)
        src = big[pad:, pad:]
        sstride = big.strides[0]
        dst = _rand_pix(rng, (h, width + 6), bd)
        src2 = rng.integers(-(1 << 13), 1 << 13, size=(64, 64)).astype(np.int16)   # 14-bit intermediates
        denom = int(rng.integers(0, 8))
        wx0, wx1 = (int(x) for x in (1 << denom) + rng.integers(-128, 128, size=2))
        ox0, ox1 = (int(x) for x in rng.integers(-128, 128, size=2))
        if family == "put":
            out = np.zeros((64, 64), dtype=np.int16)
            lib.mc_put(bd, luma, _p(out), _p(src), sstride, h, mx, my, width)
            return out[:h, :width].tobytes()
        if family == "uni":
            lib.mc_uni(bd, luma, _p(dst), dst.strides[0], _p(src), sstride, h, mx, my, width)
        elif family == "uni_w":
            lib.mc_uni_w(bd, luma, _p(dst), dst.strides[0], _p(src), sstride, h, denom, wx0, ox0, mx, my, width)
        elif family == "bi":
            lib.mc_bi(bd, luma, _p(dst), dst.strides[0], _p(src), sstride, _p(src2), h, mx, my, width)
        elif family == "bi_w":
            lib.mc_bi_w(bd, luma, _p(dst), dst.strides[0], _p(src), sstride, _p(src2), h, denom,
                        wx0, wx1, ox0, ox1, mx, my, width)
        else:
            raise ValueError(family)
        return dst.tobytes()
    return f"mc_{family}_{'q' if luma else 'e'}pel_bd{bd}_w{width}_mx{mx}_my{my}_s{seed}", run


# ---------------------------------------------------------------- deblock
def case_deblock(seed, bd, luma, vertical):
    def run(lib):
        rng = np.random.default_rng(seed)
        kind = seed % 3
        # 16x16 patch, edge through the middle; smoothness chosen to hit strong / normal / off
        if kind == 0:
            pix = _rand_pix(rng, (16, 16), bd, smooth=True)
        elif kind == 1:
            pix = _rand_pix(rng, (16, 16), bd, smooth=True)
            step = int(rng.integers(1, 12)) << (bd - 8)
            half = pix.astype(np.int64)
            if vertical:
                half[:, 8:] += step
            else:
                half[8:, :] += step
            pix = np.clip(half, 0, (1 << bd) - 1).astype(_pix_dtype(bd))
        else:
            pix = _rand_pix(rng, (16, 16), bd)
        tc = rng.integers(0, 25, size=2).astype(np.int32)
        if seed % 7 == 0:
            tc[int(rng.integers(0, 2))] = 0
        no_p = (rng.random(2) < 0.15).astype(np.uint8) * rng.integers(1, 3, size=2).astype(np.uint8)
        no_q = (rng.random(2) < 0.15).astype(np.uint8) * rng.integers(1, 3, size=2).astype(np.uint8)
        beta = int(rng.integers(0, 65))
        isz = pix.itemsize
        ptr = _p(pix) + (4 * pix.strides[0] + 8 * isz if vertical else 8 * pix.strides[0] + 4 * isz)
        if luma:
            lib.loop_filter_luma(bd, vertical, ptr, pix.strides[0], beta, _p(tc), _p(no_p), _p(no_q))
        else:
            lib.loop_filter_chroma(bd, vertical, ptr, pix.strides[0], _p(tc), _p(no_p), _p(no_q))
        return pix.tobytes()
    return f"deblock_{'luma' if luma else 'chroma'}_{'v' if vertical else 'h'}_bd{bd}_s{seed}", run


def all_cases(quick=False):
    """The committed case list (order is part of the golden fixture)."""
    cases = []
    depths = DEPTHS if not quick else [8, 10]
    s = 0
    for bd in depths:
        for log2 in (2, 3, 4, 5):
            s += 1; cases.append(case_transform_add(s, bd, log2))
            s += 1; cases.append(case_skip(s, bd, log2))
            s += 1; cases.append(case_idct_dc(s, bd, log2))
            n = 1 << log2
            limits = sorted(set([1, 2, 3, 4, 5, 8, 9, 12, 13, 16, 20, 24, 25, 28, 31, n, n + 1, n + 4]))
            for cl in limits:
                if cl > n + 4:
                    continue
                s += 1; cases.append(case_idct(s, bd, log2, cl, dense=True))
                s += 1; cases.append(case_idct(s, bd, log2, cl, dense=False))
        for k in range(4):
            s += 1; cases.append(case_dst4(s, bd))
            s += 1; cases.append(case_sao_band(s, bd))
        for eo in range(4):
            for r in (-1, 0, 1):
                for k in range(3):
                    s += 1; cases.append(case_sao_edge(s, bd, eo, r))
        for luma in (1, 0):
            for vertical in (1, 0):
                for k in range(24 if not quick else 6):
                    s += 1; cases.append(case_deblock(s, bd, luma, vertical))
    for log2 in (2, 3, 4, 5):
        for mode in (0, 1):
            s += 1; cases.append(case_rdpcm(s, log2, mode))
    # MC: every family x luma/chroma x every width x every phase class; all 16 luma and a
    # sweep of the 64 chroma phases
    for bd in depths:
        for family in ("put", "uni", "uni_w", "bi", "bi_w"):
            for w in WIDTHS:
                if w in (2, 6):
                    phases_l = []
                else:
                    phases_l = [(mx, my) for mx in range(4) for my in range(4)]
                if quick:
                    phases_l = phases_l[::5]
                for (mx, my) in phases_l:
                    s += 1; cases.append(case_mc(s, family, bd, 1, w, mx, my))
                phases_c = [(0, 0), (1, 0), (0, 7), (3, 5), (7, 7), (4, 4), (2, 6), (6, 1), (5, 0), (0, 2)]
                if quick:
                    phases_c = phases_c[::3]
                for (mx, my) in phases_c:
                    s += 1; cases.append(case_mc(s, family, bd, 0, w, mx, my))
    return cases
