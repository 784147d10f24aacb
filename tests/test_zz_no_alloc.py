"""fm_quality_static (-m gpu): the quality statistics computed WITHOUT any device allocation -- block
partials in a static device array of the library instead of the buffer fm_quality cudaMallocs --
against fm_quality on the same handle, through the C ABI (ctypes binding of
include/fishman_b200.h).  The entry point has no counterpart in the reference.
Same kernels, deterministic sums: the test asks for EQUAL results
(histogram exact, the doubles within the 1e-12 of the other GPU tests and in fact bit-equal).

This file sorts last on purpose (a failure here hides nothing).  It
passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-12
FLOATS = ("area_cell_min", "area_cell_max", "area_cell_sum", "area_tri_min", "area_tri_max", "area_tri_sum",
          "angle_min", "angle_max", "max_disp")


def _static(fb, m):
    """fm_quality_static through the C ABI; (return code, dict)"""
    from fishman_b200 import _lib
    q = _lib.QualityStats()
    rc = _lib.lib().fm_quality_static(m._h, C.byref(q))
    return rc, q.asdict()


def _err(m):
    from fishman_b200 import _lib
    return _lib.lib().fm_last_error(m._h).decode()


def _same(a, b):
    assert a["degree_hist"] == b["degree_hist"]
    for k in FLOATS:
        assert abs(a[k] - b[k]) <= TOL, (k, a[k], b[k])


def test_quality_static_equals_quality_twice_in_a_row():
    import torch
    import fishman_b200 as fb
    from fishman_b200 import points
    big = fb.Mesh.from_xyz(points.fibonacci(200000, 1))
    big.delaunay_triangulation()
    small = fb.Mesh.from_xyz(points.uniform(3000, 5))
    small.delaunay_triangulation()
    ref_big, ref_small = big.quality(), small.quality()
    assert sum(ref_big["degree_hist"]) == 200000 and abs(ref_big["area_cell_sum"] - 4 * np.pi) < 1e-10
    # two calls in a row on one handle
    rc1, q1 = _static(fb, big)
    free_before = torch.cuda.mem_get_info()[0]
    rc2, q2 = _static(fb, big)
    free_after = torch.cuda.mem_get_info()[0]
    assert rc1 == 0 and rc2 == 0, _err(big)
    _same(q1, ref_big)
    _same(q2, ref_big)
    assert q1 == q2
    assert free_after == free_before   # the second call (kernels already loaded) took no device memory
    # the array is shared and keeps the last call's contents: a smaller mesh after a larger one, and back
    rc, qs = _static(fb, small)
    assert rc == 0
    _same(qs, ref_small)
    rc, q3 = _static(fb, big)
    assert rc == 0 and q3 == q1
    # after a Lloyd step the stored triangulation is the previous generators': refused, nothing repaired
    big.lloyd(max_iter=2, warm=True)
    rc, _ = _static(fb, big)
    assert rc != 0 and "fm_retriangulate" in _err(big)
    big.retriangulate(warm=True)
    rc, q4 = _static(fb, big)
    assert rc == 0, _err(big)
    _same(q4, big.quality())
    assert q4 != q1   # the generators moved
    big.close(); small.close()


def test_quality_static_rejects_sizes_beyond_the_array():
    import fishman_b200 as fb
    from fishman_b200 import points
    limit = 4194304   # FM_QUALITY_STATIC_MAX_GENERATORS of include/fishman_b200.h
    m = fb.Mesh.from_xyz(points.fibonacci(limit + 1, 0))
    rc, _ = _static(fb, m)
    assert rc != 0
    msg = _err(m)
    assert "4194304" in msg and "limit" in msg
    m.close()
