"""Random corruption of the in-place lists: whatever a parser writes into the arenas, the device either rejects the
frame (OHCU_EINVAL from prepare_kernel, the job's error word) or reconstructs SOMETHING - it never faults, never hangs
and never poisons the context: the next, clean frame must come out bit-exact.  Complements the directed cases of
test_gpu_inplace.py / test_gpu_intra_frame.py (records that would corrupt device memory)."""
import ctypes as C

import numpy as np
import pytest

import frame_cases
from hevc_b200 import abi, backend

pytestmark = pytest.mark.gpu


def _diff(a, b):
    for c, (x, y) in enumerate(zip(a, b)):
        if not np.array_equal(x, y):
            return f"plane {c}: {int((x != y).sum())} samples differ"
    return None


@pytest.mark.parametrize("name", ["intra_8b_420", "rext_10b_444", "slices_tiles_8b_420"])
def test_random_list_corruption_never_poisons_the_context(name, oracle):
    job, refs = frame_cases.build_case(name)
    want = oracle.replay(job, refs, out_init=refs[-1])
    out_slot = job.sp.dpb_slots - 1
    rng = np.random.default_rng(2024)
    outcomes = {"rejected": 0, "accepted": 0}
    with backend.Context(job.sp, jobs=2) as ctx:
        lib, h = ctx.lib, ctx.handle
        for i, r in enumerate(refs[:-1]):
            ctx.upload(i, r)

        def submit(mutate):
            ctx.upload(out_slot, refs[-1])
            j = ctx.begin(out_slot)
            a, n = ctx.append_inplace(j, job, coeffs="sparse", commit=False)
            mutate(a, n)
            rc = lib.ohcu_frame_commit_lists(h, j, C.byref(n))
            if rc:                                             # refused on the host: nothing was launched
                lib.ohcu_frame_abort(h, j)
                return rc
            ctx.stage(j)
            ctx.launch(j, out_slot)
            return lib.ohcu_frame_sync(h, j)

        regions = lambda a, n: [r for r in (
            ("pus", a.pus, n.n_pu * C.sizeof(abi.PU)), ("tus", a.tus, n.n_tu * C.sizeof(abi.TU)),
            ("coefs", a.coefs, min(int(n.n_coef), 4096) * 4), ("ctbs", a.ctbs, 16 * C.sizeof(abi.CtbIndex)),
            ("ius", a.ius, n.n_iu * C.sizeof(abi.IU)), ("iu_ctbs", a.iu_ctbs, 16 * C.sizeof(abi.IuIndex) if n.n_iu else 0),
            ("slices", a.slices, C.sizeof(abi.SliceTab))) if r[1] and r[2] > 0]
        for trial in range(120):
            def mutate(a, n, trial=trial):
                regs = regions(a, n)
                for _ in range(int(rng.integers(1, 4))):
                    _, base, size = regs[int(rng.integers(len(regs)))]
                    off = int(rng.integers(size))
                    C.c_uint8.from_address(base + off).value = int(rng.integers(256))
                if trial % 10 == 9:                            # and now and then a count
                    which = ("n_pu", "n_tu", "n_iu", "n_coef")[int(rng.integers(4))]
                    setattr(n, which, int(getattr(n, which)) + int(rng.integers(1, 1000)))
            rc = submit(mutate)
            assert rc in (0, abi.EINVAL, abi.EOVERFLOW, abi.ESTATE), (trial, rc, lib.ohcu_last_error())
            outcomes["rejected" if rc else "accepted"] += 1
            if trial % 8 == 7:                                 # the context is still sound
                assert submit(lambda a, n: None) == 0, (trial, lib.ohcu_last_error())
                assert _diff(ctx.download(out_slot), want) is None, trial
        assert submit(lambda a, n: None) == 0
        assert _diff(ctx.download(out_slot), want) is None
    assert outcomes["rejected"] > 10, outcomes               # the fuzz does reach the validation
