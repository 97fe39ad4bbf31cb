"""Synthetic frame jobs (seeded) — Python face of hevc_b200/synth/synth.c.

A ``FrameJob`` is one picture's work lists as numpy arrays whose memory layout is
exactly the C ABI's (include/ohcu.h); tests hand the same arrays to the CUDA
library, to the oracle and to the compiled reference.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from . import abi

HERE = os.path.dirname(os.path.abspath(__file__))
SYNTH_LIB = os.path.join(HERE, "libohsynth.so")

PU_DTYPE = np.dtype([("x0", "<u2"), ("y0", "<u2"), ("w", "u1"), ("h", "u1"), ("pred_flag", "u1"),
                     ("slice", "u1"), ("mv", "<i2", (2, 2)), ("ref_idx", "u1", (2,)), ("pad", "u1", (2,))])
TU_DTYPE = np.dtype([("x0", "<u2"), ("y0", "<u2"), ("c_idx", "u1"), ("log2_size", "u1"), ("mode", "u1"),
                     ("flags", "u1"), ("col_limit", "i1"), ("res_scale_val", "i1"), ("pad", "<u2"),
                     ("coeff_off", "<u4"), ("cross_tu", "<u4")])
SAO_DTYPE = np.dtype([("type_idx", "u1", (3,)), ("eo_class", "u1", (3,)), ("band_position", "u1", (3,)),
                      ("pad", "u1"), ("offset_val", "<i2", (3, 5))])
DB_DTYPE = np.dtype([("beta_offset", "i1"), ("tc_offset", "i1")])
SLICE_DTYPE = np.dtype([("ref_slot", "u1", (2, 16)), ("weight_flag", "u1"), ("luma_log2_weight_denom", "u1"),
                        ("chroma_log2_weight_denom", "u1"), ("pad", "u1"),
                        ("luma_weight", "<i2", (2, 16)), ("luma_offset", "<i2", (2, 16)),
                        ("chroma_weight", "<i2", (2, 16, 2)), ("chroma_offset", "<i2", (2, 16, 2))])
IU_DTYPE = np.dtype([("x0", "<u2"), ("y0", "<u2"), ("c_idx", "u1"), ("log2_size", "u1"), ("mode", "u1"), ("cand", "u1")])
LEAF_DTYPE = np.dtype([("x0", "<u2"), ("y0", "<u2"), ("log2_size", "u1"), ("cbf_luma", "u1"), ("pad", "u1", (2,))])
assert PU_DTYPE.itemsize == 20 and TU_DTYPE.itemsize == 20 and SAO_DTYPE.itemsize == 40
assert SLICE_DTYPE.itemsize == C.sizeof(abi.SliceTab)


class SynthParams(C.Structure):
    _fields_ = [
        ("sp", abi.SeqParams), ("seed", C.c_uint64), ("weighted", C.c_int32), ("n_slices", C.c_int32),
        ("slices_no_lf", C.c_int32), ("tile_cols", C.c_int32), ("tile_rows", C.c_int32),
        ("rext", C.c_int32), ("bypass_pct", C.c_int32), ("far_mv_pct", C.c_int32),
        ("cbf_pct", C.c_int32), ("dc_pct", C.c_int32), ("dst_pct", C.c_int32),
        ("dense_coeffs", C.c_int32), ("bi_pct", C.c_int32), ("split_pct", C.c_int32 * 3),
        ("bs_pct", C.c_int32 * 2), ("sao_pct", C.c_int32 * 2), ("n_refs", C.c_int32),
        ("intra_pct", C.c_int32), ("intra_drop_pct", C.c_int32),
    ]


class _SynthFrame(C.Structure):
    _fields_ = [
        ("n_slices", C.c_int32), ("slices", C.c_void_p),
        ("n_pu", C.c_int32), ("pus", C.c_void_p),
        ("n_tu", C.c_int32), ("tus", C.c_void_p),
        ("ncoeff", C.c_int64), ("coeffs", C.c_void_p),
        ("bs_width", C.c_int32), ("bs_height", C.c_int32), ("vbs", C.c_void_p), ("hbs", C.c_void_p),
        ("min_cb_width", C.c_int32), ("min_cb_height", C.c_int32), ("qp_y_tab", C.c_void_p),
        ("ctb_width", C.c_int32), ("ctb_height", C.c_int32), ("deblock", C.c_void_p),
        ("min_pu_width", C.c_int32), ("min_pu_height", C.c_int32), ("is_pcm", C.c_void_p),
        ("sao", C.c_void_p), ("filter_slice_edges", C.c_void_p),
        ("tab_slice_address", C.c_void_p), ("tile_id", C.c_void_p),
        ("n_leaf", C.c_int32), ("leaves", C.c_void_p),
        ("n_iu", C.c_int32), ("ius", C.c_void_p),
    ]


_lib = None


def build_synth(force=False):
    src = os.path.join(HERE, "synth", "synth.c")
    if force or not os.path.exists(SYNTH_LIB) or os.path.getmtime(SYNTH_LIB) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-std=gnu99", "-fPIC", "-shared", "-o", SYNTH_LIB, src])


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(SYNTH_LIB):
            build_synth()
        lib = C.CDLL(SYNTH_LIB)
        lib.ohsynth_default_params.argtypes = [C.POINTER(SynthParams), C.c_int, C.c_int, C.c_int, C.c_int]
        lib.ohsynth_default_params.restype = None
        lib.ohsynth_generate.argtypes = [C.POINTER(SynthParams)]
        lib.ohsynth_generate.restype = C.POINTER(_SynthFrame)
        lib.ohsynth_free.argtypes = [C.POINTER(_SynthFrame)]
        lib.ohsynth_free.restype = None
        lib.ohsynth_fill_plane.argtypes = [C.c_void_p, C.c_ssize_t, C.c_int, C.c_int, C.c_int, C.c_uint64]
        lib.ohsynth_fill_plane.restype = None
        _lib = lib
    return _lib


def default_params(width, height, bit_depth=8, chroma_format_idc=1, **overrides):
    """ohsynth_default_params + keyword overrides (sp.* fields via ``sp_<name>``)."""
    p = SynthParams()
    _load().ohsynth_default_params(C.byref(p), width, height, bit_depth, chroma_format_idc)
    for k, v in overrides.items():
        if k.startswith("sp_"):
            setattr(p.sp, k[3:], v)
        elif k in ("split_pct", "bs_pct", "sao_pct"):
            for i, x in enumerate(v):
                getattr(p, k)[i] = x
        else:
            if not hasattr(p, k):
                raise AttributeError(k)
            setattr(p, k, v)
    if p.intra_pct:
        p.sp.intra = 1                    # the library sets aside residual planes and an arena for the intra blocks
    return p


def _copy(ptr, dtype, count):
    if not ptr or count == 0:
        return np.zeros(count, dtype=dtype)
    nbytes = int(count) * np.dtype(dtype).itemsize
    buf = (C.c_char * nbytes).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=count).copy()


def chroma_shifts(chroma_format_idc):
    hs = 1 if chroma_format_idc in (1, 2) else 0
    vs = 1 if chroma_format_idc == 1 else 0
    return hs, vs


def plane_dims(sp, c):
    """(width, height) in samples of plane c."""
    if c == 0:
        return sp.width, sp.height
    hs, vs = chroma_shifts(sp.chroma_format_idc)
    return sp.width >> hs, sp.height >> vs


def make_picture(sp, seed):
    """Reference picture: list of 1 or 3 2-D arrays (uint8 / uint16)."""
    lib = _load()
    dt = np.uint8 if sp.bit_depth == 8 else np.uint16
    planes = []
    for c in range(3 if sp.chroma_format_idc else 1):
        w, h = plane_dims(sp, c)
        a = np.zeros((h, w), dtype=dt)
        lib.ohsynth_fill_plane(a.ctypes.data, a.strides[0], w, h, sp.bit_depth, seed * 3 + c)
        planes.append(a)
    return planes


def blank_picture(sp, value=0):
    dt = np.uint8 if sp.bit_depth == 8 else np.uint16
    return [np.full(plane_dims(sp, c)[::-1], value, dtype=dt) for c in range(3 if sp.chroma_format_idc else 1)]


class FrameJob:
    """One picture's work lists (numpy, C-ABI layout) + its sequence parameters."""

    def __init__(self, params):
        lib = _load()
        self.params = params
        self.sp = abi.SeqParams.from_buffer_copy(params.sp)
        fp = lib.ohsynth_generate(C.byref(params))
        f = fp.contents
        try:
            nctb = f.ctb_width * f.ctb_height
            self.slices = _copy(f.slices, SLICE_DTYPE, f.n_slices)
            self.pus = _copy(f.pus, PU_DTYPE, f.n_pu)
            self.tus = _copy(f.tus, TU_DTYPE, f.n_tu)
            self.coeffs = _copy(f.coeffs, np.int16, f.ncoeff)
            self.bs_width, self.bs_height = f.bs_width, f.bs_height
            self.vbs = _copy(f.vbs, np.uint8, f.bs_width * f.bs_height)
            self.hbs = _copy(f.hbs, np.uint8, f.bs_width * f.bs_height)
            self.qp_y_tab = _copy(f.qp_y_tab, np.int8, f.min_cb_width * f.min_cb_height)
            self.deblock = _copy(f.deblock, DB_DTYPE, nctb)
            self.is_pcm = _copy(f.is_pcm, np.uint8, f.min_pu_width * f.min_pu_height)
            self.sao = _copy(f.sao, SAO_DTYPE, nctb)
            self.filter_slice_edges = _copy(f.filter_slice_edges, np.uint8, nctb)
            self.tab_slice_address = _copy(f.tab_slice_address, np.int32, nctb)
            self.tile_id = _copy(f.tile_id, np.int32, nctb)
            self.leaves = _copy(f.leaves, LEAF_DTYPE, f.n_leaf)     # transform-tree leaves, coded or not
            self.ius = _copy(f.ius, IU_DTYPE, f.n_iu)               # intra blocks, decoding order
            self.ctb_width, self.ctb_height = f.ctb_width, f.ctb_height
        finally:
            lib.ohsynth_free(fp)

    # ---- C views -------------------------------------------------------------
    def deblock_in(self):
        d = abi.DeblockIn()
        d.vertical_bs = self.vbs.ctypes.data
        d.horizontal_bs = self.hbs.ctypes.data
        d.qp_y_tab = self.qp_y_tab.ctypes.data
        d.deblock = self.deblock.ctypes.data
        d.is_pcm = self.is_pcm.ctypes.data if self.sp.pcm_or_bypass else None
        return d

    def sao_in(self):
        s = abi.SaoIn()
        s.sao = self.sao.ctypes.data
        s.filter_slice_edges = self.filter_slice_edges.ctypes.data
        s.tab_slice_address = self.tab_slice_address.ctypes.data
        s.tile_id = self.tile_id.ctypes.data
        return s


def make_job(width, height, bit_depth=8, chroma_format_idc=1, seed=1, **overrides):
    return FrameJob(default_params(width, height, bit_depth, chroma_format_idc, seed=seed, **overrides))
