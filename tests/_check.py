"""Shared mesh checks used by the oracle, emulation and GPU tests: the invariants
the reference asserts at run time restated on flat arrays."""
import numpy as np


def canonical(tri):
    t = np.sort(np.asarray(tri, dtype=np.int64), axis=1)
    return t[np.lexsort((t[:, 2], t[:, 1], t[:, 0]))]


def check_triangulation(xyz, tri, adj):
    """tri/adj 1-based (T,3).  Euler count, CCW orientation, adjacency symmetry:
    adj[t,k] shares the edge opposite vertex k (DV:656-669)."""
    n = len(xyz)
    T = len(tri)
    assert T == 2 * n - 4
    t0 = np.asarray(tri, dtype=np.int64) - 1
    a0 = np.asarray(adj, dtype=np.int64) - 1
    assert t0.min() >= 0 and t0.max() < n and a0.min() >= 0 and a0.max() < T
    assert np.bincount(t0.ravel(), minlength=n).min() > 0   # every point is a vertex
    p = xyz[t0]
    det = np.einsum("ij,ij->i", p[:, 0], np.cross(p[:, 1], p[:, 2]))
    assert (det > 0).all(), "triangle not CCW seen from outside"
    for k in range(3):
        u, v = t0[:, (k + 1) % 3], t0[:, (k + 2) % 3]
        nb = t0[a0[:, k]]
        # the neighbour holds the edge reversed: contains both u and v, not vertex k
        assert ((nb == u[:, None]).any(1) & (nb == v[:, None]).any(1)).all()
        assert not (nb == t0[:, k][:, None]).any(1).any()
        back = a0[a0[:, k]]
        assert (back == np.arange(T)[:, None]).any(1).all()
    e = np.concatenate([t0[:, [0, 1]], t0[:, [1, 2]], t0[:, [2, 0]]])
    e.sort(axis=1)
    key = e[:, 0] * n + e[:, 1]   # one int64 per undirected edge: a flat sort instead of a row-wise one
    key.sort()
    assert 1 + int(np.count_nonzero(key[1:] != key[:-1])) == 3 * n - 6  # euler_formula, SG:57-71


def check_rings(xyz, tri, off, ring_tri, ring_nbr):
    """ring_nbr[j] lies on the edge shared by ring_tri[j] and ring_tri[j+1] (DV:723-738);
    every triangle of the ring contains the cell's vertex; rings are CCW."""
    n = len(xyz)
    t0 = np.asarray(tri, dtype=np.int64) - 1
    assert off[0] == 0 and off[n] == 3 * len(tri)
    deg = np.diff(off)
    assert deg.min() >= 3
    cell = np.repeat(np.arange(n), deg)
    rt = np.asarray(ring_tri, dtype=np.int64) - 1
    rn = np.asarray(ring_nbr, dtype=np.int64) - 1
    pos = np.arange(len(rt)) - off[:-1][cell]
    nxt = off[:-1][cell] + (pos + 1) % deg[cell]
    assert (t0[rt] == cell[:, None]).any(1).all()
    assert (t0[rt] == rn[:, None]).any(1).all()
    assert (t0[rt[nxt]] == rn[:, None]).any(1).all()
    # each incident triangle appears once per vertex
    key = cell * (len(tri) + 1) + rt
    assert len(np.unique(key)) == len(key)
    # CCW: (g, nbr_j, nbr_j+1) positively oriented
    g, a, b = xyz[cell], xyz[rn], xyz[rn[nxt]]
    assert (np.einsum("ij,ij->i", g, np.cross(a, b)) > 0).all()


def rotate_rings(off, *arrays):
    """Rotate each ring to start at the smallest entry of arrays[-1] (the neighbour ids)."""
    out = [np.array(a) for a in arrays]
    key = arrays[-1]
    for i in range(len(off) - 1):
        s, e = off[i], off[i + 1]
        j = int(np.argmin(key[s:e]))
        for o, a in zip(out, arrays):
            o[s:e] = np.roll(a[s:e], -j)
    return out


def cc_by_key(tri, cc):
    return {tuple(sorted(t)): c for t, c in zip(np.asarray(tri).tolist(), cc)}


def max_cc_diff(tri_a, cc_a, tri_b, cc_b):
    ka = cc_by_key(tri_a, cc_a)
    return max(float(np.abs(ka[tuple(sorted(t))] - c).max()) for t, c in zip(np.asarray(tri_b).tolist(), cc_b))
