"""crm_overlap_noalloc (the entry point that allocates nothing: scratch in a fixed device array of the library)
against crm_overlap, through the C ABI, on calls in a row that share the scratch.  Collected last on purpose.
Tolerance: 1e-6 relative between the two entry points (same kernels; only the order of the f64 atomics differs),
inside the 1e-4 the mask tests allow against the oracle."""
import ctypes as C

import numpy as np
import pytest

from cr_mech_coli_b200 import fitting
from cr_mech_coli_b200.testing import random_colony

pytestmark = pytest.mark.gpu

MAX_RODS = 1024  # CRM_NOALLOC_MAX_RODS, include/crm_b200.h


def _colony(n, seed):
    pos = random_colony(n, 60.0, seed=seed, bend=0.12)  # dense: many outlines intersect
    radii = np.random.default_rng(seed + 100).uniform(0.4, 1.5, len(pos)).astype(np.float32)
    return np.ascontiguousarray(pos, np.float32), radii


def _noalloc(lib, torch, pos, radii, stream=None):
    d_pos = torch.from_numpy(pos).cuda()
    d_rad = torch.from_numpy(radii).cuda()
    d_area = torch.full((1,), -1.0, dtype=torch.float32, device="cuda")
    d_status = torch.full((1,), 77, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rc = lib.fn("overlap_noalloc")(d_pos.data_ptr(), d_rad.data_ptr(), len(pos), pos.shape[1], float(np.pi / 4), 0.01,
                                   d_area.data_ptr(), d_status.data_ptr(), stream)
    lib.check(rc)
    torch.cuda.synchronize()
    return float(d_area.cpu()[0]), int(d_status.cpu()[0])


def test_noalloc_overlap_equals_overlap_on_calls_in_a_row(product_lib):
    import torch

    lib = product_lib
    assert lib.has("overlap_noalloc")
    cases = [_colony(600, 5), _colony(150, 6), _colony(600, 5), _colony(MAX_RODS, 7)]
    for k, (pos, radii) in enumerate(cases):
        want = fitting.overlap(pos, radii)
        got, status = _noalloc(lib, torch, pos, radii)
        assert status == 0, (k, status)
        assert want > 0.0, k
        assert abs(got - want) <= 1e-6 * abs(want), (k, got, want)
    # on a stream of the caller's
    s = torch.cuda.Stream()
    pos, radii = cases[1]
    got, status = _noalloc(lib, torch, pos, radii, stream=C.c_void_p(s.cuda_stream))
    assert status == 0 and abs(got - fitting.overlap(pos, radii)) <= 1e-6 * got
    # no rods: zero area, nothing stale from the call before
    empty = np.zeros((0, 8, 3), np.float32)
    d = torch.zeros(4, device="cuda")  # valid (unused) device pointers
    a = torch.full((1,), -1.0, device="cuda")
    st = torch.full((1,), 77, dtype=torch.int32, device="cuda")
    lib.check(lib.fn("overlap_noalloc")(d.data_ptr(), d.data_ptr(), 0, empty.shape[1], float(np.pi / 4), 0.01, a.data_ptr(),
                                        st.data_ptr(), None))
    torch.cuda.synchronize()
    assert float(a.cpu()[0]) == 0.0 and int(st.cpu()[0]) == 0


def test_noalloc_reports_a_broken_outline_like_overlap(product_lib):
    import torch

    lib = product_lib
    # a rod folded back onto itself: the offset lines of its two segments neither cross nor meet (CalcError)
    pos = np.zeros((2, 3, 3), np.float32)
    pos[0, :, 0] = [0.0, 10.0, 0.0]
    pos[0, :, 1] = [0.0, 0.0, 1e-4]
    pos[1, :, 0] = [20.0, 25.0, 30.0]
    radii = np.ones(2, np.float32)
    try:
        fitting.overlap(pos, radii)
        failed = False
    except Exception:
        failed = True
    _, status = _noalloc(lib, torch, pos, radii)
    assert (status != 0) == failed
    # and the next call is clean again (status word cleared at the start of every call)
    pos, radii = _colony(100, 9)
    got, status = _noalloc(lib, torch, pos, radii)
    assert status == 0 and abs(got - fitting.overlap(pos, radii)) <= 1e-6 * got


def test_noalloc_rejects_more_rods_than_the_scratch_holds(product_lib):
    import torch

    lib = product_lib
    d = torch.zeros(8, device="cuda")
    rc = lib.fn("overlap_noalloc")(d.data_ptr(), d.data_ptr(), MAX_RODS + 1, 8, float(np.pi / 4), 0.01, d.data_ptr(),
                                   d.data_ptr(), None)
    assert rc == -1  # CRM_ERR_INVALID
    assert "exceed the fixed scratch" in lib.last_error()
