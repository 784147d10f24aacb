"""CPU logic tests of the Delaunay kernels: the per-thread bodies of
fishman_b200/csrc/dt_kernels.cuh and the round driver dt_driver.h, executed thread by
thread on the host (tests/emu/fm_emu.cpp), against the oracle, Qhull and the golden
fixtures.  This checks the algorithm (votes, splits, flip rounds, exact predicates);
the CUDA build of the same bodies is checked on the GPU by tests/test_gpu_parity.py."""
import os

import numpy as np
import pytest

from fishman_b200 import points
from oracle import oracle as O
from tests import _check, _emu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name", ["u300", "fib500", "u2000"])
def test_emu_against_golden(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    tri, adj, st = _emu.triangulate(g["xyz"])
    _check.check_triangulation(g["xyz"], tri, adj)
    assert np.array_equal(_check.canonical(tri), g["tri"])


@pytest.mark.parametrize("n,seed,sort", [(4, 0, True), (5, 1, True), (8, 1, False), (64, 3, True),
                                         (1000, 4, False), (10242, 0, True), (40000, 5, True)])
def test_emu_against_oracle(n, seed, sort):
    p = points.uniform(n, seed)
    if n == 4:  # four random points need not surround the centre; use a tetrahedron
        p = np.array([[1, 1, 1], [1, -1, -1], [-1, 1, -1], [-1, -1, 1.0]]) / np.sqrt(3)
    if n == 5:
        p = np.concatenate([np.array([[1, 1, 1], [1, -1, -1], [-1, 1, -1], [-1, -1, 1.0]]) / np.sqrt(3),
                            points.uniform(1, 1)])
    tri, adj, st = _emu.triangulate(p, sort=sort)
    _check.check_triangulation(p, tri, adj)
    if n >= 64:
        o = O.Oracle(p)
        otri, _ = o.delaunay_triangulation()
        assert np.array_equal(_check.canonical(tri), _check.canonical(otri))
        assert st["rounds"] <= 4 * int(np.log2(n)) + 8  # one point per triangle per round: O(log n) rounds


def test_emu_order_independent():
    """SoS keys on the ORIGINAL ids, so the internal (Morton) order must not matter."""
    p = points.ico(3)  # cocircular quads: only the tie-break decides
    a = _check.canonical(_emu.triangulate(p, sort=True)[0])
    b = _check.canonical(_emu.triangulate(p, sort=False)[0])
    assert np.array_equal(a, b)


@pytest.mark.parametrize("maker", [lambda: points.latlon_grid(12, 24), lambda: points.ico(4),
                                   lambda: np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1.0]])])
def test_emu_degenerate_inputs(maker):
    """Cocircular / collinear configurations: valid triangulation, locally Delaunay up to ties."""
    p = maker()
    tri, adj, st = _emu.triangulate(p)
    _check.check_triangulation(p, tri, adj)
    t0, a0 = tri.astype(np.int64) - 1, adj.astype(np.int64) - 1
    for k in range(3):
        nb = t0[a0[:, k]]
        mask = (nb != t0[:, [(k + 1) % 3]]) & (nb != t0[:, [(k + 2) % 3]])
        d = p[nb[mask]]
        v1, v2, v3 = p[t0[:, 0]], p[t0[:, 1]], p[t0[:, 2]]
        det = np.einsum("ij,ij->i", v3 - d, np.cross(v2 - d, v1 - d))  # DV:2051-2053
        assert (det <= 1e-14).all()


def test_emu_hemisphere_fails_cleanly():
    hemi = np.abs(points.uniform(40, 0))
    with pytest.raises(RuntimeError, match="bootstrap"):
        _emu.triangulate(hemi)


@pytest.mark.parametrize("nranks", [2, 3, 8])
def test_sharded_sweep_gives_the_same_mesh_as_the_unsharded_one(nranks):
    """The N > 1 warm repair on CPU: each rank sweeps its slots into a request list, all requests
    are applied, the flip loop runs (csrc/fishman_b200.cu repair_async, sharded).  Slot for slot
    the result equals the unsharded repair, whatever order the requests arrive in, and it is the
    Delaunay triangulation of the moved points."""
    rng = np.random.default_rng(5)
    p = points.fibonacci(3000, 2, 0.2)
    q = p + 0.25 * np.sqrt(4 * np.pi / len(p)) * rng.normal(size=p.shape) * 0.3
    q /= np.linalg.norm(q, axis=1)[:, None]
    tri1, adj1, turned1, _ = _emu.repair(p, q, 1)
    assert turned1 == 0
    for rev in (False, True):
        tri, adj, turned, nreq = _emu.repair(p, q, nranks, reverse=rev)
        assert turned == 0 and nreq > 10
        assert np.array_equal(tri, tri1) and np.array_equal(adj, adj1)
    cold, _, _ = _emu.triangulate(q)
    assert np.array_equal(_check.canonical(tri1), _check.canonical(cold))
    _check.check_triangulation(q, tri1, adj1)


@pytest.mark.parametrize("nranks", [2, 5])
def test_sharded_point_location_gives_the_same_build(nranks):
    """Cold build with each level's point location split over ranks (walks record loc[] only, the
    votes are cast afterwards from all ranks' results): slot for slot the unsharded build."""
    p = points.uniform(3000, 7)
    tri1, adj1, _ = _emu.triangulate(p)
    tri, adj = _emu.triangulate_sharded(p, nranks)
    assert np.array_equal(tri, tri1) and np.array_equal(adj, adj1)


def test_topology_does_not_depend_on_the_round_schedule(tmp_path):
    """dt_run carries the points that lost their vote into the next insertion level after
    FM_ROUNDS_PER_LEVEL rounds (default 1).  The schedule changes which slot a triangle lands in,
    never the triangulation: one round per level, three, and "until dry" (the schedule before
    the carry-over) must give the same canonical triangles, on a random set and on a
    degenerate one (cocircular points of a lat-lon grid, decided by SoS).  The switch is read
    once per process, hence the subprocesses."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = (
        "import sys, numpy as np\n"
        "sys.path.insert(0, %r)\n"
        "from tests import _emu, _check\n"
        "from fishman_b200 import points\n"
        "lon, lat = np.meshgrid(np.linspace(0, 2 * np.pi, 40, endpoint=False), np.linspace(-1.3, 1.3, 21))\n"
        "grid = np.c_[(np.cos(lat) * np.cos(lon)).ravel(), (np.cos(lat) * np.sin(lon)).ravel(), np.sin(lat).ravel()]\n"
        "out = []\n"
        "for p in (points.uniform(20000, 11), grid):\n"
        "    tri, adj = _emu.triangulate(p)[:2]\n"
        "    _check.check_triangulation(p, tri, adj)\n"
        "    out.append(_check.canonical(tri))\n"
        "np.savez(sys.argv[1], a=out[0], b=out[1])\n" % root)
    got = {}
    for rounds in ("1", "3", "1000"):
        path = str(tmp_path / f"tri_{rounds}.npz")
        env = dict(os.environ, FM_ROUNDS_PER_LEVEL=rounds)
        r = subprocess.run([sys.executable, "-c", script, path], env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        with np.load(path) as f:
            got[rounds] = (f["a"].copy(), f["b"].copy())
    for rounds in ("3", "1000"):
        assert np.array_equal(got["1"][0], got[rounds][0])
        assert np.array_equal(got["1"][1], got[rounds][1])


@pytest.mark.parametrize("n", [40, 200, 2000])
def test_rebuild_after_a_failed_warm_repair(n):
    """fm_retriangulate(h, warm=1) when a triangle turned over: the sweep has stamped amark[] /
    vote[] with the current stamp, the flip loop left without advancing it, and the rebuild runs on
    the same arrays.  dt_run must start on a fresh stamp (Backend::bump_stamp), otherwise the slots
    the sweep marked are taken for "already queued" and their new triangles are never legalised."""
    differ_without = 0
    for trial in range(60 if n <= 200 else 6):
        rng = np.random.default_rng(1000 * n + trial)
        p = points.uniform(n, trial)
        q = p + 1.5 * np.sqrt(4 * np.pi / n) * rng.normal(size=p.shape)   # strongly moved
        q /= np.linalg.norm(q, axis=1)[:, None]
        fresh = _check.canonical(_emu.triangulate(q)[0])
        tri, adj, turned = _emu.rebuild_after_failed_repair(p, q)
        assert turned > 0
        _check.check_triangulation(q, tri, adj)
        assert np.array_equal(_check.canonical(tri), fresh)
        try:
            bad = _check.canonical(_emu.rebuild_after_failed_repair(p, q, skip_bump=True)[0])
            differ_without += not np.array_equal(bad, fresh)
        except RuntimeError:
            differ_without += 1
    if n == 40:
        assert differ_without > 0   # the defect is real: without the bump some rebuilds are not Delaunay


# ---- cold build by stars (csrc/star_kernels.cuh), kernel bodies on the host
def _star_sets():
    from fishman_b200 import points
    return [("fib5000", points.fibonacci(5000, 0)), ("fib20000_nojitter", points.fibonacci(20000, 0, 0.0)),
            ("u6000", points.uniform(6000, 3)), ("ico4", points.ico(4)), ("ico5", points.ico(5))]


@pytest.mark.parametrize("name", [s[0] for s in _star_sets()])
@pytest.mark.parametrize("kappa", [1.5, 2.0, 3.0])
def test_star_build_equals_insertion_build(name, kappa):
    """Every generator's star found independently (grid, gift wrapping with the filtered / exact predicates,
    fp32 guess + certificate) gives the triangulation the insertion + flip rounds give, triangle for triangle
    (both are the Delaunay triangulation of the SoS-perturbed points), whatever the cell size; rows and
    adjacency are a valid closed triangulation."""
    from tests import _check
    p = dict(_star_sets())[name]
    rc, tri, adj, st = _emu.triangulate_star(p, kappa)
    assert rc == 0, st
    assert st["failed"] == 0 and st["broken"] == 0 and st["owned"] == 2 * len(p) - 4 and st["degree_sum"] == 6 * len(p) - 12
    _check.check_triangulation(p, tri, adj)
    t2, _, _ = _emu.triangulate(p)
    assert np.array_equal(_check.canonical(tri), _check.canonical(t2))


def test_star_build_refuses_what_it_cannot_do():
    """A lat-lon grid (poles of degree nlon, rows of cocircular points) and a clustered set: stars that do not close
    within kStarCap neighbours / kStarMaxR cells are reported, never a wrong triangulation: rc 2 = the product
    runs the insertion build."""
    from fishman_b200 import points
    p = points.latlon_grid(40, 80)
    rc, tri, adj, st = _emu.triangulate_star(p, 2.0)
    assert rc == 2 and (st["failed"] > 0 or st["broken"] != 0)
    rng = np.random.default_rng(0)
    q = np.concatenate([points.uniform(3000, 1), points.uniform(3000, 2) * 1e-3 + np.array([0, 0, 1.0])])
    q /= np.linalg.norm(q, axis=1)[:, None]
    rc, tri, adj, st = _emu.triangulate_star(q, 2.0)
    if rc == 0:   # allowed to succeed -- then it must be right
        from tests import _check
        t2, _, _ = _emu.triangulate(q)
        assert np.array_equal(_check.canonical(tri), _check.canonical(t2))
    else:
        assert rc == 2


def test_star_build_wide_setting_on_uniform_random_points():
    """The wide setting (40-entry near lists, first radius for ~20 expected candidates; FM_STAR_WIDE=1 in the
    product) on what it is for: same triangles as the insertion build, and far fewer generators sent through the
    full wrap than with the regular setting."""
    p = points.uniform(20000, 11)
    rc, tri, adj, st = _emu.triangulate_star(p, 2.5, wide=True)
    assert rc == 0 and st["failed"] == 0 and st["broken"] == 0, st
    t2, _, _ = _emu.triangulate(p)
    assert np.array_equal(_check.canonical(tri), _check.canonical(t2))
    rc, _, _, st_reg = _emu.triangulate_star(p, 2.0, wide=False)
    assert rc == 0 and st["slow"] * 5 < st_reg["slow"]


def _cubesphere(m):
    t = np.linspace(-1, 1, m)
    a, b = (v.ravel() for v in np.meshgrid(t, t, indexing="ij"))
    o = np.ones_like(a)
    p = np.concatenate([np.stack(f, 1) for f in ((o, a, b), (-o, a, b), (a, o, b), (a, -o, b), (a, b, o), (a, b, -o))])
    p /= np.linalg.norm(p, axis=1)[:, None]
    return np.unique(np.round(p, 14), axis=0)


def _latlon_band(nlat):
    """a lat-lon grid between +-69 degrees (rows of cocircular points, cocircular quads everywhere) + two random caps"""
    lat = np.linspace(-1.2, 1.2, nlat)
    lon = np.arange(2 * nlat) / (2 * nlat) * 2 * np.pi
    la, lo = np.meshgrid(lat, lon, indexing="ij")
    p = np.stack([np.cos(la) * np.cos(lo), np.cos(la) * np.sin(lo), np.sin(la)], -1).reshape(-1, 3)
    cap = points.uniform(400, 1) * 0.25 + np.array([0, 0, 1.0])
    cap /= np.linalg.norm(cap, axis=1)[:, None]
    return np.concatenate([p, cap, cap * np.array([1, 1, -1.0])])


@pytest.mark.parametrize("name", ["cubesphere", "latlon_band", "fib_jitter_0.5"])
def test_star_build_on_degenerate_and_irregular_sets(name):
    """Cocircular quads everywhere (cube-sphere grid, lat-lon band: most generators end in the exact kernel, whose
    SoS tie-break is the flips') and a strongly jittered lattice (hundreds of overflowing near lists: second
    attempt with long lists): the same triangles as the insertion build."""
    p = {"cubesphere": lambda: _cubesphere(40), "latlon_band": lambda: _latlon_band(64),
         "fib_jitter_0.5": lambda: points.fibonacci(12000, 5, 0.5)}[name]()
    rc, tri, adj, st = _emu.triangulate_star(p, 2.0)
    assert rc == 0 and st["failed"] == 0 and st["broken"] == 0, st
    if name == "fib_jitter_0.5":
        assert st["slow"] > 0 and st["exact"] == 0
    else:
        assert st["exact"] > 100
    _check.check_triangulation(p, tri, adj)
    t2, _, _ = _emu.triangulate(p)
    assert np.array_equal(_check.canonical(tri), _check.canonical(t2))


def test_star_build_refuses_near_duplicate_pairs():
    """Pairs of generators 1e-9 apart: slivers whose normals cannot be trusted -- refused (rc 2), not guessed."""
    p = points.fibonacci(6000, 3)
    q = p[::7] + 1e-9 * np.random.default_rng(0).standard_normal(p[::7].shape)
    q /= np.linalg.norm(q, axis=1)[:, None]
    rc, _, _, st = _emu.triangulate_star(np.concatenate([p, q]), 2.0)
    assert rc == 2 and st["failed"] > 0
