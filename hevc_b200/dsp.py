"""Per-entry view of the CUDA backend: the HEVCDSPContext leaves (hevcdsp.h:51-166), one
block at a time, through the device code of the frame kernels (ohcu_dsp_* of include/ohcu.h).

``Dsp()`` has the method names and argument lists of the CPU checkers in
tests/oracle_lib.py (``mc_uni(bd, luma, dst, dststride, src, srcstride, h, mx, my, w)`` ...,
raw addresses for pointers), so the seeded cases of tests/leaf_cases.py run unchanged on the
reference's compiled table, on the oracle and here.  Fails loudly without the CUDA library
or a device; there is no CPU path.
"""
from . import abi

LEAVES = ["transform_add", "transform_rdpcm", "transform_skip", "idct_4x4_luma", "idct", "idct_dc",
          "sao_band_filter", "sao_edge_filter", "mc_put", "mc_uni", "mc_uni_w", "mc_bi", "mc_bi_w",
          "loop_filter_luma", "loop_filter_chroma"]


class Dsp:
    which = "cuda"

    def __init__(self):
        self.lib = abi.load()
        if self.lib.ohcu_device_count() < 1:
            raise RuntimeError("hevc_b200.dsp needs a CUDA device (no CPU fallback)")
        for name in LEAVES:
            setattr(self, name, self._bind(name))

    def _bind(self, name):
        fn = getattr(self.lib, "ohcu_dsp_" + name)

        def call(*args):
            abi.check(fn(*args), "ohcu_dsp_" + name)
        call.__name__ = name
        return call

    def sao_edge_restore(self, *args):
        raise NotImplementedError("sao_edge_restore has no per-entry shim: the frame kernel applies the border "
                                  "rules per CTB (tests/test_gpu_parity.py)")
