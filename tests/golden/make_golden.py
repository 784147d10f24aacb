"""Generates the committed fixtures of tests/golden/.

The reference (Fortran) cannot be built or run in this image and ships no golden
vectors, so these fixtures do NOT come from the reference:
they come from two implementations independent of this repo's CUDA path --
Qhull (scipy.spatial.ConvexHull: the convex hull of points on the unit sphere is
their spherical Delaunay triangulation) for the topology and circumcentres, and
mpmath (50 digits) for the Lloyd centroid of SURVEY.md 7.3.  Run from the repo root:

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from fishman_b200 import points  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def canonical(tri):
    t = np.sort(np.asarray(tri, dtype=np.int64), axis=1)
    return t[np.lexsort((t[:, 2], t[:, 1], t[:, 0]))]


def qhull_dt(p):
    from scipy.spatial import ConvexHull
    s = ConvexHull(p).simplices
    # orient CCW seen from outside
    a, b, c = p[s[:, 0]], p[s[:, 1]], p[s[:, 2]]
    flip = np.einsum("ij,ij->i", a, np.cross(b, c)) < 0
    s[flip] = s[flip][:, [0, 2, 1]]
    return s + 1


def circumcentres_mp(p, tri):
    import mpmath as mp
    mp.mp.dps = 50
    out = np.empty((len(tri), 3))
    for k, t in enumerate(tri - 1):
        v = [[mp.mpf(float(c)) for c in p[i]] for i in t]
        e3 = [v[1][j] - v[0][j] for j in range(3)]
        e2 = [v[2][j] - v[0][j] for j in range(3)]
        n = [e3[1] * e2[2] - e3[2] * e2[1], e3[2] * e2[0] - e3[0] * e2[2], e3[0] * e2[1] - e3[1] * e2[0]]
        l = mp.sqrt(sum(c * c for c in n))
        out[k] = [float(c / l) for c in n]
    return out


QA1, QB1 = "0.10128650732345633880098736191512", "0.79742698535308732239802527616975"
QA2, QB2 = "0.47014206410511508977044120951345", "0.05971587178976982045911758097311"
QW0, QW1, QW2 = "0.225", "0.12593918054482715259568394550018", "0.13239415278850618073764938783315"


def lloyd_mp(p, tri, kind):
    """New generators by SURVEY.md 7.3 in 50-digit arithmetic (uniform or x4 density)."""
    import mpmath as mp
    mp.mp.dps = 50
    n = len(p)
    P = [[mp.mpf(float(c)) for c in r] for r in p]
    cc = []
    for t in tri - 1:
        v = [P[i] for i in t]
        e3 = [v[1][j] - v[0][j] for j in range(3)]
        e2 = [v[2][j] - v[0][j] for j in range(3)]
        nn = [e3[1] * e2[2] - e3[2] * e2[1], e3[2] * e2[0] - e3[0] * e2[2], e3[0] * e2[1] - e3[1] * e2[0]]
        l = mp.sqrt(sum(c * c for c in nn))
        cc.append([c / l for c in nn])
    # CCW ring of triangles around each vertex
    nxt = {}
    for ti, t in enumerate(tri - 1):
        for k in range(3):
            # around vertex t[k]: edge (t[k], t[k+1]) -> leaving, (t[k+2], t[k]) entering
            nxt[(int(t[k]), int(t[(k + 1) % 3]))] = (ti, int(t[(k + 2) % 3]))
    a1, b1, a2, b2 = (mp.mpf(s) for s in (QA1, QB1, QA2, QB2))
    w0, w1, w2 = (mp.mpf(s) for s in (QW0, QW1, QW2))
    third = mp.mpf(1) / 3
    nodes = [(w0, third, third, third), (w1, b1, a1, a1), (w1, a1, b1, a1), (w1, a1, a1, b1),
             (w2, b2, a2, a2), (w2, a2, b2, a2), (w2, a2, a2, b2)]
    lon, lat = mp.radians(0), mp.radians(40)
    cen = [mp.cos(lat) * mp.cos(lon), mp.cos(lat) * mp.sin(lon), mp.sin(lat)]
    beta, alpha, gamma = mp.pi / 6, mp.pi / 20, mp.mpf(1) / 256

    def rho(q):
        if kind == 0:
            return mp.mpf(1)
        d = mp.acos(max(min(sum(q[j] * cen[j] for j in range(3)), 1), -1))
        return (1 - gamma) / 2 * (mp.tanh((beta - d) / alpha) + 1) + gamma

    out = np.empty((n, 3))
    first = {}
    for (u, v), (ti, w) in nxt.items():
        first.setdefault(u, v)
    for i in range(n):
        ring, v = [], first[i]
        while True:
            ti, w = nxt[(i, v)]
            ring.append(ti)
            v = w
            if v == first[i]:
                break
        g = P[i]
        M = [mp.mpf(0)] * 3
        d = len(ring)
        for j in range(d):
            a, b = cc[ring[j]], cc[ring[(j + 1) % d]]
            u = [a[k] - g[k] for k in range(3)]
            w_ = [b[k] - g[k] for k in range(3)]
            c = [u[1] * w_[2] - u[2] * w_[1], u[2] * w_[0] - u[0] * w_[2], u[0] * w_[1] - u[1] * w_[0]]
            A = mp.sqrt(sum(x * x for x in c)) / 2
            s = [mp.mpf(0)] * 3
            for (w, l0, l1, l2) in nodes:
                q = [l0 * g[k] + l1 * a[k] + l2 * b[k] for k in range(3)]
                r = mp.sqrt(sum(x * x for x in q))
                q = [x / r for x in q]
                ww = w * rho(q)
                s = [s[k] + ww * q[k] for k in range(3)]
            M = [M[k] + A * s[k] for k in range(3)]
        r = mp.sqrt(sum(x * x for x in M))
        out[i] = [float(x / r) for x in M]
    return out


def main():
    sets = {"u300": points.uniform(300, 11), "fib500": points.fibonacci(500, 3), "u2000": points.uniform(2000, 7)}
    for name, p in sets.items():
        tri = qhull_dt(p)
        d = dict(xyz=p, tri=canonical(tri).astype(np.int32))
        if len(p) <= 500:
            d["tri_ccw"] = tri.astype(np.int32)
            d["cc"] = circumcentres_mp(p, tri)
            d["lloyd_uniform"] = lloyd_mp(p, tri, 0)
            d["lloyd_x4"] = lloyd_mp(p, tri, 1)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **d)
        print(name, {k: v.shape for k, v in d.items()})


if __name__ == "__main__":
    main()
