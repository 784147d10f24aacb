"""Synthetic generator sets for the SCVT path.

The reference reads its generators from a NetCDF "grids" file (dimension
location_nv, variable vtx_p(location_nv, 3): unit-sphere Cartesian triples,
/root/reference/src/gen_mesh.F90:16-22) and ships no such file, so the harness
makes the point sets itself:

  U(n, seed)    uniform random on the sphere
  ICO(k)        k-fold bisection of the icosahedron, n = 10 * 4**k + 2
  FIB(n, seed)  Fibonacci lattice with a small tangential jitter
"""
import numpy as np


def uniform(n, seed=0):
    p = np.random.default_rng(seed).normal(size=(n, 3))
    return p / np.linalg.norm(p, axis=1)[:, None]


def icosahedron():
    phi = (1.0 + 5.0 ** 0.5) / 2.0
    v = []
    for a in (-1.0, 1.0):
        for b in (-phi, phi):
            v += [(0.0, a, b), (a, b, 0.0), (b, 0.0, a)]
    v = np.array(v)
    v /= np.linalg.norm(v, axis=1)[:, None]
    # faces = triples of mutually nearest vertices
    d = v @ v.T
    edge = np.isclose(d, d[0][np.argsort(-d[0])[1]])
    f = [(i, j, k) for i in range(12) for j in range(i + 1, 12) for k in range(j + 1, 12)
         if edge[i, j] and edge[j, k] and edge[i, k]]
    return v, np.array(f, dtype=np.int64)


def ico(k):
    """Recursive bisection; returns (n, 3) with n = 10 * 4**k + 2."""
    v, f = icosahedron()
    for _ in range(k):
        e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
        e.sort(axis=1)
        ue, inv = np.unique(e, axis=0, return_inverse=True)
        mid = v[ue[:, 0]] + v[ue[:, 1]]
        mid /= np.linalg.norm(mid, axis=1)[:, None]
        nv = len(v)
        v = np.concatenate([v, mid])
        m = inv.reshape(3, -1) + nv  # midpoints of (01, 12, 20)
        a, b, c = f[:, 0], f[:, 1], f[:, 2]
        f = np.concatenate([np.stack([a, m[0], m[2]], 1), np.stack([b, m[1], m[0]], 1),
                            np.stack([c, m[2], m[1]], 1), np.stack([m[0], m[1], m[2]], 1)])
    return np.ascontiguousarray(v)


def ico_count(k):
    return 10 * 4 ** k + 2


def fibonacci(n, seed=0, jitter=0.2):
    i = np.arange(n) + 0.5
    z = 1.0 - 2.0 * i / n
    r = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    th = np.pi * (1.0 + 5.0 ** 0.5) * i
    p = np.stack([r * np.cos(th), r * np.sin(th), z], 1)
    if jitter:
        h = np.sqrt(4.0 * np.pi / n)
        p = p + np.random.default_rng(seed).uniform(-jitter * h, jitter * h, size=p.shape)
        p /= np.linalg.norm(p, axis=1)[:, None]
    return np.ascontiguousarray(p)


def latlon_grid(nlat, nlon):
    """Regular lat-lon grid plus the poles: maximally degenerate (cocircular
    quads, points on common meridians) -- a stress test for the exact predicates."""
    lat = (np.arange(nlat) + 0.5) / nlat * np.pi - np.pi / 2
    lon = np.arange(nlon) / nlon * 2 * np.pi
    la, lo = np.meshgrid(lat, lon, indexing="ij")
    p = np.stack([np.cos(la) * np.cos(lo), np.cos(la) * np.sin(lo), np.sin(la)], -1).reshape(-1, 3)
    return np.concatenate([p, [[0, 0, 1.0], [0, 0, -1.0]]])
