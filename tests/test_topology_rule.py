"""CPU tests of the SURVEY 7.4 topology rule (tests/_check.py: compare_topology,
explain_by_on_circle): differences between two triangulations are accepted only when every
differing edge is one the reference abstains on (|in_circle det| <= eps, DV:878-880)."""
import numpy as np
import pytest

from fishman_b200 import points
from oracle import oracle as O
from tests import _check, _emu


def _flip(tri, adj, t, k):
    """tri with the edge opposite vertex k of triangle t flipped (1-based arrays, copy)."""
    tri = np.array(tri)
    n = adj[t, k] - 1
    a, b, c = tri[t, k], tri[t, (k + 1) % 3], tri[t, (k + 2) % 3]
    d = [v for v in tri[n] if v not in (b, c)][0]
    tri[t] = (a, b, d)
    tri[n] = (a, d, c)
    return tri


def test_identical_triangulations_have_no_differences():
    p = points.uniform(3000, 1)
    tri, adj, _ = _emu.triangulate(p)
    r = _check.compare_topology(p, tri, tri[np.random.default_rng(0).permutation(len(tri))][:, [1, 2, 0]])
    assert r["n_diff"] == 0 and len(r["common_a"]) == len(tri)
    assert np.array_equal(np.sort(tri[r["common_a"]], 1), np.sort(tri[np.random.default_rng(0).permutation(len(tri))][r["common_b"]], 1))


def test_cocircular_quads_are_explained_quad_by_quad():
    """Lat-lon grid: every cell is a cocircular quad.  The kernels decide by exact arithmetic +
    SoS, the restatement of the reference abstains (ON_CIRCLE): the triangulations differ in many
    quads and every one of them is inside the band."""
    p = points.latlon_grid(12, 24)
    tri, adj, _ = _emu.triangulate(p)
    otri, _ = O.Oracle(p).delaunay_triangulation()
    r = _check.compare_topology(p, tri, otri, exact_side="a")
    assert r["n_diff"] > 0 and r["explained"]["patches"] * 2 == r["n_diff"]
    assert r["explained"]["max_abs_det"] <= _check.EPS_REF
    ca = np.sort(tri[r["common_a"]], 1)
    cb = np.sort(otri[r["common_b"]], 1)
    assert np.array_equal(ca, cb) and len(ca) == len(tri) - r["n_diff"]


def test_a_wrong_flip_is_not_explained():
    p = points.uniform(2000, 3)
    tri, adj, _ = _emu.triangulate(p)
    bad = _flip(tri, adj, 17, 0)
    with pytest.raises(AssertionError, match="NOT explained"):
        _check.compare_topology(p, tri, bad)
    with pytest.raises(AssertionError):   # exactness on the wrong side is caught as well
        _check.compare_topology(p, bad, tri, eps=1.0, exact_side="a")
    # a wide band explains it: the rule is about the determinant, not about counts
    r = _check.compare_topology(p, tri, bad, eps=1.0, exact_side="a")
    assert r["n_diff"] == 2 and r["explained"]["patches"] == 1


def test_two_adjacent_wrong_flips_form_separate_patches():
    p = points.uniform(2000, 4)
    tri, adj, _ = _emu.triangulate(p)
    t = 40
    n = adj[t, 0] - 1
    # a second quad next to the first: they share a boundary edge that both triangulations keep
    u = adj[t, 1] - 1
    ku = next(k for k in range(3) if adj[u, k] - 1 not in (t, n))
    bad = _flip(_flip(tri, adj, t, 0), adj, u, ku)
    r = _check.compare_topology(p, tri, bad, eps=1.0)
    assert r["n_diff"] == 4 and r["explained"]["patches"] == 2
