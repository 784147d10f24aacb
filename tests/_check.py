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


# ---------------------------------------------------------------------------------------------
# the topology rule.  The reference compares plain fp64 determinants with
# eps = epsilon(1d0) and does NOT flip on ON_CIRCLE (/root/reference/src/delaunay_voronoi_mod.F90
# :878-880, threshold :2055-2061), so on a nearly cocircular quad its diagonal depends on the
# insertion history.  Two triangulations of the same points are therefore "equal by the
# reference's own rule" when every differing triangle belongs to a patch (two triangulations of
# one polygon) in which every in-circle determinant the reference would evaluate is inside the
# band |det| <= eps.  Differences are EXPLAINED patch by patch, never tolerated by count.
EPS_REF = 2.220446049250313e-16   # src/const_mod.F90:16


def canonical_with_order(tri):
    """(sorted rows of sorted triples, order) with rows == sort(tri,1)[order]."""
    t = np.sort(np.asarray(tri, dtype=np.int64), axis=1)
    order = np.lexsort((t[:, 2], t[:, 1], t[:, 0]))
    return t[order], order


def _row_hash(t):
    u = t.astype(np.uint64)
    with np.errstate(over="ignore"):
        return (u[:, 0] * np.uint64(0x9E3779B97F4A7C15)) ^ (u[:, 1] * np.uint64(0xC2B2AE3D27D4EB4F)
                                                          + u[:, 2] * np.uint64(0x165667B19E3779F9))


def in_circle_det(xyz, tri, d):
    """The reference's in_circle determinant (DV:2051-2053) in plain fp64, row-wise: triangle
    tri[:, 0..2] (0-based, CCW seen from outside), point d.  > eps: INSIDE; |.| <= eps: ON_CIRCLE."""
    v1, v2, v3, p = xyz[tri[:, 0]], xyz[tri[:, 1]], xyz[tri[:, 2]], xyz[d]
    a, b, c = v3 - p, v2 - p, v1 - p
    return (a[:, 0] * (b[:, 1] * c[:, 2] - b[:, 2] * c[:, 1]) - a[:, 1] * (b[:, 0] * c[:, 2] - b[:, 2] * c[:, 0])
            + a[:, 2] * (b[:, 0] * c[:, 1] - b[:, 1] * c[:, 0]))


def _exact_in_circle_sign(xyz, t, d):
    """Sign of the same determinant in exact rational arithmetic (doubles are rationals)."""
    from fractions import Fraction as F
    P = [[F(float(v)) for v in xyz[i]] for i in (t[0], t[1], t[2], d)]
    a, b, c = ([P[k][j] - P[3][j] for j in range(3)] for k in (2, 1, 0))
    det = (a[0] * (b[1] * c[2] - b[2] * c[1]) - a[1] * (b[0] * c[2] - b[2] * c[0]) + a[2] * (b[0] * c[1] - b[1] * c[0]))
    return (det > 0) - (det < 0)


def _ccw(xyz, rows):
    """Orient 0-based vertex triples CCW seen from outside (orient1, DV:1882-1885)."""
    rows = np.array(rows, dtype=np.int64).reshape(-1, 3)
    p = xyz[rows]
    det = np.einsum("ij,ij->i", p[:, 0], np.cross(p[:, 1], p[:, 2]))
    rows[det < 0] = rows[det < 0][:, [0, 2, 1]]
    return rows


def explain_by_on_circle(xyz, only_a, only_b, eps=EPS_REF, exact_side=None):
    """only_a / only_b: 1-based sorted triples present on one side only.  Groups them into patches
    -- maximal sets of triangles of one side connected through edges the OTHER side does not have,
    matched with the other side's triangles on the same side of a shared boundary edge -- checks
    that both sides triangulate the same polygon (same vertex set, triangle count and boundary) and
    that every interior edge of the patch, on either side, is one the reference would abstain on:
    the in_circle determinant of its quad (validate_DT, DV:845-882) is inside the ON_CIRCLE band
    |det| <= eps.  exact_side = "a" or "b": additionally that side is exactly Delaunay inside the
    patch (rational arithmetic: no opposite vertex strictly inside).  Returns {"patches",
    "triangles", "max_abs_det"}; raises AssertionError naming the first patch that is not explained."""
    xyz = np.asarray(xyz, dtype=np.float64)
    A = np.asarray(only_a, dtype=np.int64).reshape(-1, 3) - 1
    B = np.asarray(only_b, dtype=np.int64).reshape(-1, 3) - 1
    allt = [tuple(t) for t in np.concatenate([A, B]).tolist()]
    side = [0] * len(A) + [1] * len(B)
    parent = list(range(len(allt)))

    def find(v):
        while parent[v] != v:
            parent[v] = parent[parent[v]]
            v = parent[v]
        return v

    def edges(t):
        return (tuple(sorted((t[0], t[1]))), tuple(sorted((t[1], t[2]))), tuple(sorted((t[0], t[2]))))
    emap = ({}, {})   # per side: edge -> triangles of that side's differing set
    for k, t in enumerate(allt):
        for e in edges(t):
            emap[side[k]].setdefault(e, []).append(k)

    def opposite(t, e):
        return [v for v in t if v not in e][0]

    def above(e, v):   # side of the great circle through edge e on which v lies
        return float(np.dot(xyz[e[0]], np.cross(xyz[e[1]], xyz[v]))) > 0.0
    for s in (0, 1):
        for e, ks in emap[s].items():
            if e not in emap[1 - s]:             # an edge only this side has: interior to a patch
                for k in ks[1:]:
                    parent[find(k)] = find(ks[0])
            else:                                # both sides have it: patch boundary; pair across sides
                for k in ks:
                    for q in emap[1 - s][e]:
                        if above(e, opposite(allt[k], e)) == above(e, opposite(allt[q], e)):
                            parent[find(k)] = find(q)
    groups = {}
    for k in range(len(allt)):
        groups.setdefault(find(k), []).append(k)
    worst = 0.0
    for idx in groups.values():
        ts = ([allt[k] for k in idx if side[k] == 0], [allt[k] for k in idx if side[k] == 1])
        verts = sorted({v for t in ts[0] + ts[1] for v in t})
        where = f"patch on 1-based vertices {[v + 1 for v in verts]}"
        assert len(ts[0]) == len(ts[1]) and len(ts[0]) >= 2, f"{where}: {len(ts[0])} vs {len(ts[1])} triangles"
        assert {v for t in ts[0] for v in t} == {v for t in ts[1] for v in t}, f"{where}: vertex sets differ"
        inner = []
        for s in (0, 1):
            cnt = {}
            for t in ts[s]:
                for e in edges(t):
                    cnt.setdefault(e, []).append(t)
            inner.append({e: v for e, v in cnt.items() if len(v) == 2})
            assert all(len(v) <= 2 for v in cnt.values()), f"{where}: an edge with three triangles"
        outer = [{e for t in ts[s] for e in edges(t)} - set(inner[s]) for s in (0, 1)]
        assert outer[0] == outer[1], f"{where}: the two sides do not fill the same polygon"
        for s, name in ((0, "a"), (1, "b")):
            for e, (t1, t2) in inner[s].items():
                t = _ccw(xyz, [t1])[0]
                d = opposite(t2, e)
                det = float(in_circle_det(xyz, t[None, :], np.array([d]))[0])
                worst = max(worst, abs(det))
                assert abs(det) <= eps, (
                    f"{where}: edge {[v + 1 for v in e]} of side {name}: in_circle det = {det:.3e}, outside the "
                    "reference's ON_CIRCLE band (DV:878-880, :2055-2061) -- difference NOT explained")
                if exact_side == name:
                    assert _exact_in_circle_sign(xyz, t, int(d)) <= 0, \
                        f"{where}: vertex {d + 1} is strictly inside triangle {(t + 1).tolist()} (exact arithmetic)"
    return {"patches": len(groups), "triangles": int(len(A)), "max_abs_det": worst}


def compare_topology(xyz, tri_a, tri_b, eps=EPS_REF, exact_side=None):
    """SURVEY 7.4 on two triangulations (1-based (T,3)) of the same points: canonical triples
    bit-exact, or every difference explained by the ON_CIRCLE band.  Returns a dict with
    "n_diff" (triangles on one side only), "explained" (the explain_by_on_circle report) and
    "common_a" / "common_b": row indices into tri_a / tri_b of the triangles both sides have, in
    the same (canonical) order -- for comparing per-triangle data such as circumcentres."""
    ca, oa = canonical_with_order(tri_a)
    cb, ob = canonical_with_order(tri_b)
    assert len(ca) == len(cb)
    if np.array_equal(ca, cb):
        return {"n_diff": 0, "explained": None, "common_a": oa, "common_b": ob}
    ha, hb = _row_hash(ca), _row_hash(cb)
    ina, inb = np.isin(ha, hb), np.isin(hb, ha)
    # the hash only FINDS the candidates; equality of everything else is checked exactly
    assert np.array_equal(ca[ina], cb[inb]), "triangulations differ beyond the hashed candidates"
    rep = explain_by_on_circle(xyz, ca[~ina], cb[~inb], eps=eps, exact_side=exact_side)
    return {"n_diff": int((~ina).sum()), "explained": rep, "common_a": oa[ina], "common_b": ob[inb]}


def max_rows_diff(a, b):
    return float(np.abs(np.asarray(a) - np.asarray(b)).max()) if len(a) else 0.0
