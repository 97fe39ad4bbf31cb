"""ctypes bindings of the CHECKERS: oracle/liboracle.so (our CPU restatement) and
oracle/_ref/liboh_ref.so (the reference's own compiled C code).  Tests only.

``Lib('oracle')`` / ``Lib('ref')`` expose the same leaf entry points with the
prefixes ``or_`` / ``ohref_`` unified under one name, plus ``replay``.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "liboracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "liboh_ref.so")
REF_SIMD_SO = os.path.join(ORACLE_DIR, "_ref", "liboh_ref_simd.so")     # reference C + its SSE intrinsics, -O3
REFERENCE_ROOT = "/root/reference"

import sys
sys.path.insert(0, ROOT)
from hevc_b200 import abi  # noqa: E402


class Picture(C.Structure):
    _fields_ = [("data", C.c_void_p * 3), ("stride", C.c_ssize_t * 3)]


class FrameJobC(C.Structure):
    _fields_ = [
        ("sp", abi.SeqParams),
        ("n_slices", C.c_int32), ("slices", C.c_void_p),
        ("n_pu", C.c_int32), ("pus", C.c_void_p),
        ("n_tu", C.c_int32), ("tus", C.c_void_p),
        ("coeffs", C.c_void_p), ("ncoeff", C.c_int64),
        ("deblock", C.POINTER(abi.DeblockIn)), ("sao", C.POINTER(abi.SaoIn)),
        ("dpb", C.POINTER(Picture)), ("out", Picture), ("stages", C.c_int32),
        ("n_iu", C.c_int32), ("ius", C.c_void_p),
    ]


class BsJobC(C.Structure):
    _fields_ = [
        ("sp", abi.SeqParams),
        ("n_slices", C.c_int32), ("slices", C.c_void_p),
        ("n_pu", C.c_int32), ("pus", C.c_void_p),
        ("n_tu", C.c_int32), ("tus", C.c_void_p),
        ("n_leaf", C.c_int32), ("leaves", C.c_void_p),
        ("filter_slice_edges", C.c_void_p), ("tab_slice_address", C.c_void_p), ("tile_id", C.c_void_p),
        ("deblock_disabled", C.c_void_p), ("vbs", C.c_void_p), ("hbs", C.c_void_p),
        ("n_iu", C.c_int32), ("ius", C.c_void_p),
    ]


STAGE_MC, STAGE_RESIDUAL, STAGE_DEBLOCK, STAGE_SAO = 1, 2, 4, 8


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "oracle"])


def build_ref():
    if not os.path.isdir(REFERENCE_ROOT):
        return False
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "ref", "-j8"])
    return True


def build_ref_simd():
    if not os.path.isdir(REFERENCE_ROOT):
        return False
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "ref_simd", "-j8"])
    return True


def have_ref():
    return os.path.exists(REF_SO) or os.path.isdir(REFERENCE_ROOT)


def have_ref_simd():
    return os.path.exists(REF_SIMD_SO) or os.path.isdir(REFERENCE_ROOT)


_P = C.c_void_p
_I = C.c_int
_S = C.c_ssize_t
_LEAVES = {
    "transform_add": [_I, _P, _P, _S, _I],
    "transform_rdpcm": None,   # differs between the two libs, handled below
    "transform_skip": [_I, _P, _I],
    "idct_4x4_luma": [_I, _P],
    "idct": [_I, _P, _I, _I],
    "idct_dc": [_I, _P, _I],
    "sao_band_filter": [_I, _P, _P, _S, _S, _P, _I, _I, _I],
    "sao_edge_filter": [_I, _P, _P, _S, _S, _P, _I, _I, _I],
    "sao_edge_restore": [_I, _I, _P, _P, _S, _S, _P, _I, _P, _I, _I, _P, _P, _P],
    "mc_put": [_I, _I, _P, _P, _S, _I, _I, _I, _I],
    "mc_uni": [_I, _I, _P, _S, _P, _S, _I, _I, _I, _I],
    "mc_uni_w": [_I, _I, _P, _S, _P, _S, _I, _I, _I, _I, _I, _I, _I],
    "mc_bi": [_I, _I, _P, _S, _P, _S, _P, _I, _I, _I, _I],
    "mc_bi_w": [_I, _I, _P, _S, _P, _S, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I],
    "loop_filter_luma": [_I, _I, _P, _S, _I, _P, _P, _P],
    "loop_filter_chroma": [_I, _I, _P, _S, _P, _P, _P],
}


class Lib:
    def __init__(self, which):
        self.which = which
        if which == "oracle":
            build_oracle()
            self.lib = C.CDLL(ORACLE_SO)
            prefix = "or_"
        elif which == "ref":
            if not os.path.exists(REF_SO) and not build_ref():
                raise FileNotFoundError(REF_SO)
            self.lib = C.CDLL(REF_SO)
            prefix = "ohref_"
        elif which == "ref_simd":
            if not os.path.exists(REF_SIMD_SO) and not build_ref_simd():
                raise FileNotFoundError(REF_SIMD_SO)
            self.lib = C.CDLL(REF_SIMD_SO)
            prefix = "ohref_"
        else:
            raise ValueError(which)
        for name, args in _LEAVES.items():
            fn = getattr(self.lib, prefix + name)
            fn.restype = None
            if args is not None:
                fn.argtypes = args
            setattr(self, "_" + name, fn)
        # dequantisation (the levels form): the two libraries take different arguments, see dequant_block()
        fn = getattr(self.lib, prefix + "dequant_block")
        if which == "oracle":
            fn.restype = None
            fn.argtypes = [_P, _I, _I, _I, _P, _I]
            self.lib.or_dequant_tus.restype = None
            self.lib.or_dequant_tus.argtypes = [_P, _P, _I, _P, _I, _P, _P]
        else:
            fn.restype = _I
            fn.argtypes = [_I, _I, _I, _I, _I, _I, _I, _I, _I, _P, _P, _P]
        self._dequant_block = fn
        self._bs = getattr(self.lib, prefix + "boundary_strengths")
        self._bs.restype = C.c_int
        self._bs.argtypes = [C.POINTER(BsJobC)]
        # intra prediction: oracle and reference take different arguments (tests/intra_cases.py)
        if which == "oracle":
            self.intra_pred = self.lib.or_intra_pred
            self.intra_pred.restype = None
            self.intra_pred.argtypes = [_I, _P, _S] + [_I] * 10
        elif hasattr(self.lib, "ohref_intra_pred"):
            self.intra_pred_ref = self.lib.ohref_intra_pred
            self.intra_pred_ref.restype = _I
            self.intra_pred_ref.argtypes = [_I, _P, _S] + [_I] * 11
        self._replay = getattr(self.lib, prefix + "replay_frame")
        self._replay.restype = C.c_int
        self._replay.argtypes = [C.POINTER(FrameJobC)]

    def __getattr__(self, name):
        if name in _LEAVES:
            return getattr(self, "_" + name)
        raise AttributeError(name)

    def transform_rdpcm(self, bd, coeffs_ptr, log2_size, mode):
        if self.which == "oracle":
            self._transform_rdpcm(C.c_void_p(coeffs_ptr), C.c_int(log2_size), C.c_int(mode))
        else:
            self._transform_rdpcm(C.c_int(bd), C.c_void_p(coeffs_ptr), C.c_int(log2_size), C.c_int(mode))

    def dequant_tus(self, tus, quant, levels, bd, scaling=None):
        """oracle only: the dense arena of un-scaled levels -> coefficients (a copy)."""
        out = np.ascontiguousarray(levels).copy()
        sl = sl_dc = None
        if scaling is not None:
            sl, sl_dc = (np.ascontiguousarray(a, dtype=np.uint8) for a in scaling)
        self.lib.or_dequant_tus(tus.ctypes.data, np.ascontiguousarray(quant).ctypes.data, len(tus), out.ctypes.data, bd,
                                sl.ctypes.data if sl is not None else None, sl_dc.ctypes.data if sl_dc is not None else None)
        return out

    def boundary_strengths(self, job, deblock_disabled=None):
        """(vbs, hbs) of a synth.FrameJob: derived from its PU / coded-TB lists (oracle) or by the
        reference's ff_hevc_deblocking_boundary_strengths over its transform-tree leaves (ref)."""
        sp = job.sp
        n = ((sp.width >> 2) + 1) * ((sp.height >> 2) + 1)
        vbs, hbs = np.zeros(n, np.uint8), np.zeros(n, np.uint8)
        bj = BsJobC()
        bj.sp = sp
        bj.n_slices = len(job.slices); bj.slices = job.slices.ctypes.data
        bj.n_pu = len(job.pus); bj.pus = job.pus.ctypes.data
        bj.n_tu = len(job.tus); bj.tus = job.tus.ctypes.data
        bj.n_leaf = len(job.leaves); bj.leaves = job.leaves.ctypes.data
        bj.filter_slice_edges = job.filter_slice_edges.ctypes.data
        bj.tab_slice_address = job.tab_slice_address.ctypes.data
        bj.tile_id = job.tile_id.ctypes.data
        if deblock_disabled is not None:
            deblock_disabled = np.ascontiguousarray(deblock_disabled, dtype=np.uint8)
            bj.deblock_disabled = deblock_disabled.ctypes.data
        bj.vbs = vbs.ctypes.data; bj.hbs = hbs.ctypes.data
        if len(getattr(job, "ius", ())):
            bj.n_iu = len(job.ius); bj.ius = job.ius.ctypes.data
        rc = self._bs(C.byref(bj))
        if rc != 0:
            raise RuntimeError(f"{self.which} boundary_strengths failed: {rc}")
        return vbs, hbs

    def output_cropped(self, planes, crop, hshift, vshift):
        """or_output_cropped: the (left, top, width, height) window of a picture as tight planes
        (hevc_refs.c:335-341 + openhevc.c:722-735).  Oracle only (openhevc.c needs a decoder)."""
        fn = self.lib.or_output_cropped
        fn.restype = None
        fn.argtypes = [C.POINTER(C.c_void_p * 3), C.POINTER(C.c_ssize_t * 3)] + [_I] * 8 + [C.POINTER(C.c_void_p * 3)]
        left, top, width, height = [int(v) for v in crop]
        src, strides, dst = (C.c_void_p * 3)(), (C.c_ssize_t * 3)(), (C.c_void_p * 3)()
        out = []
        for i, p in enumerate(planes):
            assert p.strides[1] == p.itemsize
            hs, vs = (hshift, vshift) if i else (0, 0)
            src[i] = p.ctypes.data; strides[i] = p.strides[0]
            out.append(np.zeros((height >> vs, width >> hs), p.dtype))
            dst[i] = out[-1].ctypes.data
        fn(C.byref(src), C.byref(strides), len(planes), planes[0].itemsize - 1, hshift, vshift, left, top, width, height,
           C.byref(dst))
        return out

    def replay(self, job, refs, out_init=None, stages=0, use_deblock=True, use_sao=True):
        """Run one frame job.  refs = list (DPB slots) of pictures (lists of 2-D arrays).
        Returns the reconstructed picture as a list of arrays."""
        sp = job.sp
        nplanes = 3 if sp.chroma_format_idc else 1
        out = [np.ascontiguousarray(p.copy()) for p in (out_init if out_init is not None else
                                                         _blank(sp, nplanes))]
        dpb = (Picture * max(1, sp.dpb_slots))()
        keep = []
        for i in range(sp.dpb_slots):
            if i < len(refs) and refs[i] is not None:
                for c in range(nplanes):
                    a = np.ascontiguousarray(refs[i][c])
                    keep.append(a)
                    dpb[i].data[c] = a.ctypes.data
                    dpb[i].stride[c] = a.strides[0]
        fj = FrameJobC()
        fj.sp = sp
        fj.n_slices = len(job.slices); fj.slices = job.slices.ctypes.data
        fj.n_pu = len(job.pus); fj.pus = job.pus.ctypes.data
        fj.n_tu = len(job.tus); fj.tus = job.tus.ctypes.data
        fj.coeffs = job.coeffs.ctypes.data; fj.ncoeff = len(job.coeffs)
        db = job.deblock_in(); so = job.sao_in()
        fj.deblock = C.pointer(db) if use_deblock else None
        fj.sao = C.pointer(so) if use_sao else None
        fj.dpb = dpb
        for c in range(nplanes):
            fj.out.data[c] = out[c].ctypes.data
            fj.out.stride[c] = out[c].strides[0]
        fj.stages = stages
        ius = getattr(job, "ius", None)
        if ius is not None and len(ius):
            fj.n_iu = len(ius); fj.ius = ius.ctypes.data
        rc = self._replay(C.byref(fj))
        if rc != 0:
            raise RuntimeError(f"{self.which} replay failed: {rc}")
        return out


def _blank(sp, nplanes):
    from hevc_b200 import synth
    return synth.blank_picture(sp)
