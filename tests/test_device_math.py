"""The density function of the Lloyd kernel is evaluated on the device WITHOUT the math library
(fishman_b200/csrc/vd_kernels.cuh: fm_acos, fm_exp, fm_rcp).  CPU part: the polynomial
coefficients are read out of the header and the same formulas, restated in numpy, are held
against 40-digit mpmath values -- a typo in a coefficient fails here, without a GPU.  GPU part:
fm_density_eval (the device code itself) against the oracle's density_eval."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "fishman_b200", "csrc", "vd_kernels.cuh")


def _table(name):
    src = open(HEADER).read()
    m = re.search(r"static __constant__ double %s\[(\d+)\] = \{([^}]*)\}" % name, src)
    assert m, name
    vals = [float(v) for v in m.group(2).replace("\n", " ").split(",") if v.strip()]
    assert len(vals) == int(m.group(1))
    return vals


def _acos(x, P):
    """fm_acos, line by line (no FMA in numpy: one more rounding per step than on the device)"""
    a = np.abs(x)
    lo = a <= 0.5
    z = np.where(lo, x * x, 0.5 * (1.0 - a))
    s = np.where(lo, a, np.where(z > 0, np.sqrt(np.maximum(z, 0.0)), 0.0))
    p = np.full_like(z, P[-1])
    for c in P[-2::-1]:
        p = p * z + c
    t = s * (z * p) + s
    hi, lw = 1.5707963267948966, 6.123233995736766e-17
    u = np.where(lo & (x < 0), -t, t)
    v = hi - (u - lw)
    return np.where(lo, v, np.where(x > 0, 2.0 * t, 2.0 * v))


def _exp(t, Q):
    """fm_exp with the range reduction done in extended precision (the device uses two FMAs)"""
    n = np.rint(t * 1.4426950408889634)
    r = (t.astype(np.longdouble) - n * np.longdouble(0.6931471805599453)
         - n * np.longdouble(2.3190468138462996e-17)).astype(np.float64)
    q = np.full_like(r, Q[-1])
    for c in Q[-2::-1]:
        q = q * r + c
    return np.ldexp((r * r) * q + r + 1.0, n.astype(np.int64))


def test_acos_coefficients_against_mpmath():
    import mpmath as mp
    mp.mp.dps = 40
    P = _table("kAsinP")
    rng = np.random.default_rng(5)
    x = np.concatenate([rng.uniform(-1, 1, 4000), 1 - 10.0 ** rng.uniform(-16, 0, 1500),
                        -1 + 10.0 ** rng.uniform(-16, 0, 1500),
                        [0.0, 0.5, -0.5, 1.0, -1.0, np.nextafter(0.5, 1), np.nextafter(0.5, 0),
                         np.nextafter(1.0, 2.0), -np.nextafter(1.0, 2.0)]])   # a rounding above 1: no clamp needed
    got = _acos(x, P)
    worst = 0.0
    for xi, gi in zip(x, got):
        ref = mp.acos(mp.mpf(float(np.clip(xi, -1, 1))))
        worst = max(worst, float(abs(mp.mpf(float(gi)) - ref)))
    assert worst < 7e-16, worst


def test_exp_coefficients_against_mpmath():
    import mpmath as mp
    mp.mp.dps = 40
    Q = _table("kExpQ")
    t = np.concatenate([np.random.default_rng(6).uniform(-60, 60, 4000), [-700.0, 700.0, 0.0, 0.34657, -0.34657]])
    got = _exp(t, Q)
    worst = 0.0
    for ti, gi in zip(t, got):
        ref = mp.e ** mp.mpf(float(ti))
        worst = max(worst, float(abs(mp.mpf(float(gi)) / ref - 1)))
    assert worst < 6e-16, worst


@pytest.mark.gpu
@pytest.mark.parametrize("table", [False, True])
def test_density_on_the_device_against_oracle(table, monkeypatch):
    """table=False: the formula on the device (fm_acos, fm_exp, fm_rcp).  table=True
    (FM_DENSITY_TABLE=1): the staged table the Lloyd kernel reads -- rho as a piecewise degree-7
    polynomial of the chord, intervals the host could not fit to 2e-13 evaluated by the formula."""
    monkeypatch.setenv("FM_DENSITY_TABLE", "1" if table else "0")
    import fishman_b200 as fb
    from fishman_b200 import points
    from oracle import oracle as O
    rng = np.random.default_rng(7)
    p = points.uniform(200000, 3)
    for lon, lat, beta, alpha, gamma in ((0.0, 40.0, np.pi / 6, np.pi / 20, 0.25 ** 4),
                                         (123.0, -70.0, 1.0, 0.05, 1e-3), (10.0, 0.0, 0.2, 2.0, 0.5)):
        rho = fb.density_x4(lon, lat, beta, alpha, gamma)
        c = np.array([rho.cx, rho.cy, rho.cz])
        # the centre, its antipode, the |dot| = 1/2 circles, points within 1e-8 of the centre
        e1 = np.cross(c, [0.3, -0.5, 0.8]); e1 /= np.linalg.norm(e1)
        ring = lambda ang: np.cos(ang) * c + np.sin(ang) * e1  # noqa: E731
        special = np.array([c, -c, ring(np.pi / 3), ring(2 * np.pi / 3), ring(1e-8), ring(np.pi - 1e-8), ring(beta)])
        q = np.vstack([p, special, c + 1e-7 * rng.standard_normal((1000, 3))])
        q /= np.linalg.norm(q, axis=1)[:, None]
        got = fb.density_eval(rho, q)
        want = O.density_eval(O.density_x4(lon, lat, beta, alpha, gamma), q)
        # Both sides round the dot product (differently: the device contracts it into FMAs), and
        # acos turns a difference of eps in the dot into eps / sin d in d: unbounded towards the
        # centre and its antipode.  d enters rho through exp(2 d / alpha) with slope <= 1/4 of
        # the value, so the tolerance is per point; the cap still catches garbage there.
        sin_d = np.sqrt(np.maximum(1.0 - np.clip(q @ c, -1, 1) ** 2, 1e-300))
        tol = np.minimum((3e-13 if table else 5e-14) + 0.25 * (2.0 / alpha) * 1.5e-15 / sin_d, 1e-6)
        err = np.abs(got / want - 1)
        assert (err <= tol).all(), (err / tol).max()
        assert gamma <= got.min() and got.max() <= 1.0
    assert np.array_equal(fb.density_eval(fb.density_uniform(), p[:100]), np.ones(100))
