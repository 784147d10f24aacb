"""Exact predicates (fishman_b200/csrc/predicates.cuh, host build) against exact
rational arithmetic, including the Simulation-of-Simplicity tie-break, which is
checked against the determinant of an EXPLICITLY perturbed matrix evaluated with
Fractions (eps = 2^-40 on integer inputs, perturbation eps^(2^(3 rank + coord)))."""
import ctypes as C
import itertools
import os
import subprocess
from fractions import Fraction

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "emu", "pred_host.cpp")
LIB = os.path.join(HERE, "emu", "libfm_pred_host.so")


@pytest.fixture(scope="module")
def lib():
    hdr = os.path.join(HERE, "..", "fishman_b200", "csrc", "predicates.cuh")
    if (not os.path.exists(LIB)) or os.path.getmtime(LIB) < max(os.path.getmtime(SRC), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", LIB, SRC, "-lm"])
    L = C.CDLL(LIB)
    dp = C.POINTER(C.c_double)
    for f in (L.fmt_orient, L.fmt_orient_slow):
        f.argtypes = [dp, dp, dp, C.c_int, C.c_int, C.c_int]
    for f in (L.fmt_incircle, L.fmt_incircle_slow):
        f.argtypes = [dp, dp, dp, dp, C.c_int, C.c_int, C.c_int, C.c_int]
    L.fmt_exact_det_sign.argtypes = [dp, C.c_int]
    return L


def _p(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a.ctypes.data_as(C.POINTER(C.c_double)), a


def fdet(m):
    n = len(m)
    if n == 0:
        return Fraction(1)
    if n == 1:
        return m[0][0]
    tot = Fraction(0)
    for j in range(n):
        if m[0][j] == 0:
            continue
        sub = [[m[r][c] for c in range(n) if c != j] for r in range(1, n)]
        tot += (-1) ** j * m[0][j] * fdet(sub)
    return tot


def sgn(v):
    return (v > 0) - (v < 0)


EPS = Fraction(1, 2 ** 4000)      # doubles: non-zero minors are far above 2^-4000
EPS_INT = Fraction(1, 2 ** 40)    # small-integer inputs: non-zero minors are >= 1


def perturbed_sign(pts, ids, dim, EPS=EPS):
    """sign of det of rows (x y z [1]) in the GIVEN order, points perturbed by SoS."""
    order = sorted(range(len(ids)), key=lambda k: ids[k])
    rank = {k: r for r, k in enumerate(order)}
    rows = []
    for k, p in enumerate(pts):
        row = [Fraction(float(p[c])) + EPS ** (1 << (3 * rank[k] + c)) for c in range(3)]
        if dim == 4:
            row.append(Fraction(1))
        rows.append(row)
    return sgn(fdet(rows))


def exact_orient(a, b, c):
    return sgn(fdet([[Fraction(float(v)) for v in p] for p in (a, b, c)]))


def exact_incircle(a, b, c, d):
    return -sgn(fdet([[Fraction(float(v)) for v in p] + [Fraction(1)] for p in (a, b, c, d)]))


def test_exact_det_sign_random(lib):
    rng = np.random.default_rng(1)
    for k in (1, 2, 3, 4):
        for _ in range(200):
            m = rng.integers(-3, 4, size=(k, k)).astype(np.float64)
            if rng.random() < 0.5:
                m *= rng.random((k, k))
            ptr, keep = _p(m)
            want = sgn(fdet([[Fraction(float(v)) for v in row] for row in m]))
            assert lib.fmt_exact_det_sign(ptr, k) == want


def test_orient_generic_matches_exact(lib):
    rng = np.random.default_rng(2)
    for _ in range(500):
        p = rng.normal(size=(3, 3))
        p /= np.linalg.norm(p, axis=1)[:, None]
        args = [_p(v) for v in p]
        got = lib.fmt_orient(args[0][0], args[1][0], args[2][0], 5, 9, 2)
        assert got == exact_orient(*p)
        assert lib.fmt_orient_slow(args[0][0], args[1][0], args[2][0], 5, 9, 2) == got


def test_orient_near_degenerate(lib):
    rng = np.random.default_rng(3)
    for _ in range(300):
        a, b = rng.normal(size=(2, 3))
        a /= np.linalg.norm(a); b /= np.linalg.norm(b)
        t = rng.random()
        c = (1 - t) * a + t * b
        c /= np.linalg.norm(c)  # on the great circle a-b up to round-off
        want = exact_orient(a, b, c)
        args = [_p(v) for v in (a, b, c)]
        got = lib.fmt_orient(args[0][0], args[1][0], args[2][0], 1, 2, 3)
        if want != 0:
            assert got == want
        else:
            assert got == perturbed_sign([a, b, c], [1, 2, 3], 3)


def test_incircle_generic_and_near_degenerate(lib):
    rng = np.random.default_rng(4)
    for it in range(400):
        p = rng.normal(size=(4, 3))
        p /= np.linalg.norm(p, axis=1)[:, None]
        if it % 2:
            # make d nearly cocircular with a,b,c: rotate a about the circumcentre axis
            n = np.cross(p[1] - p[0], p[2] - p[0]); n /= np.linalg.norm(n)
            ang = rng.random() * 6.28
            v = p[0] - n * (n @ p[0])
            w = np.cross(n, v)
            p[3] = n * (n @ p[0]) + np.cos(ang) * v + np.sin(ang) * w
        want = exact_incircle(*p)
        args = [_p(v) for v in p]
        got = lib.fmt_incircle(*[a[0] for a in args], 4, 3, 2, 1)
        if want != 0:
            assert got == want


def _sos_cases():
    pts = [(0, 0, 0), (1, 0, 0), (0, 1, 0), (0, 0, 1), (1, 1, 0), (1, 1, 1), (2, 1, 0), (2, 2, 2),
           (1, 2, 3), (0, 0, 2), (0, 2, 0), (3, 0, 0), (1, 0, 1), (-1, 0, 0), (0, -1, 0), (-1, -1, -1)]
    return [np.array(p, dtype=np.float64) for p in pts]


def test_sos_orient_ties(lib):
    pts = _sos_cases()
    rng = np.random.default_rng(5)
    n = 0
    for tri in itertools.combinations(range(len(pts)), 3):
        p = [pts[k] for k in tri]
        if exact_orient(*p) != 0:
            continue
        ids = list(rng.permutation(50)[:3])
        for perm in itertools.permutations(range(3)):
            pp = [p[k] for k in perm]
            ii = [int(ids[k]) for k in perm]
            args = [_p(v) for v in pp]
            got = lib.fmt_orient(*[a[0] for a in args], *ii)
            assert got != 0
            assert got == perturbed_sign(pp, ii, 3, EPS_INT), (tri, perm, ids)
        n += 1
    assert n > 50


def test_sos_incircle_ties(lib):
    pts = _sos_cases()
    rng = np.random.default_rng(6)
    n = 0
    for quad in itertools.combinations(range(len(pts)), 4):
        p = [pts[k] for k in quad]
        if exact_incircle(*p) != 0:
            continue
        if n % 3:  # subsample: each case costs a 4x4 determinant of huge rationals
            n += 1
            continue
        ids = list(rng.permutation(50)[:4])
        for perm in list(itertools.permutations(range(4)))[::5]:
            pp = [p[k] for k in perm]
            ii = [int(ids[k]) for k in perm]
            args = [_p(v) for v in pp]
            got = lib.fmt_incircle(*[a[0] for a in args], *ii)
            assert got != 0
            assert got == -perturbed_sign(pp, ii, 4, EPS_INT), (quad, perm, ids)
        n += 1
    assert n > 50


def test_sos_is_antisymmetric(lib):
    # swapping two arguments flips the sign, also on exact ties
    a, b, c, d = (np.array(v, dtype=np.float64) for v in ((1, 0, 0), (0, 1, 0), (-1, 0, 0), (0, -1, 0)))
    pa, pb, pc, pd = (_p(v) for v in (a, b, c, d))
    s1 = lib.fmt_incircle(pa[0], pb[0], pc[0], pd[0], 0, 1, 2, 3)
    s2 = lib.fmt_incircle(pb[0], pa[0], pc[0], pd[0], 1, 0, 2, 3)
    assert s1 == -s2 and s1 != 0
    o1 = lib.fmt_orient(pa[0], pb[0], pc[0], 0, 1, 2)
    o2 = lib.fmt_orient(pa[0], pc[0], pb[0], 0, 2, 1)
    assert o1 == -o2 and o1 != 0
