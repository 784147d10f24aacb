"""GPU parity tests (-m gpu): the sm_100a library, called through the C ABI
(ctypes binding of include/fishman_b200.h), against the CPU oracle
(oracle/fishman_oracle.c, the restatement of the reference) and Qhull on the same
seeded inputs.  Topology: canonical sorted triples bit-exact.
Coordinates: 1e-12 (absolute on unit vectors = relative), per Lloyd step."""
import numpy as np
import pytest

from tests import _check

pytestmark = pytest.mark.gpu

TOL = 1e-12


@pytest.fixture(scope="module")
def fb():
    import fishman_b200 as fb
    return fb


@pytest.fixture(scope="module")
def O():
    from oracle import oracle as O
    O.build()
    return O


def _sets():
    from fishman_b200 import points
    return [("octa", np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1.0]])),
            ("ico0", points.ico(0)), ("ico2", points.ico(2)), ("ico5", points.ico(5)),
            ("u500", points.uniform(500, 3)), ("u10242", points.uniform(10242, 0)),
            ("fib5000", points.fibonacci(5000, 0))]


@pytest.mark.parametrize("name", [s[0] for s in _sets()])
def test_dt_vd_against_oracle(fb, O, name):
    p = dict(_sets())[name]
    m = fb.Mesh.from_xyz(p)
    m.delaunay_triangulation()
    m.voronoi_diagram(all=True)
    tri, adj = m.triangles()
    _check.check_triangulation(p, tri, adj)
    o = O.Oracle(p, first3=(0, 2, 4) if name == "octa" else (0, 1, 2))  # octa: points 0,1,2 are collinear (DV:670)
    otri, _ = o.delaunay_triangulation()
    occ, oll, ooff, ort, orn = o.voronoi_diagram()
    if name not in ("octa", "ico0", "ico2", "ico5"):  # symmetric sets have cocircular quads: DT not unique
        assert np.array_equal(_check.canonical(tri), _check.canonical(otri))
        cc, ll = m.circumcentres()
        assert _check.max_cc_diff(tri, cc, otri, occ) < TOL
        assert _check.max_cc_diff(tri, ll, otri, oll) < 1e-9  # lon/lat: atan2 conditioning near the poles
    off, rt, rn = m.rings()
    _check.check_rings(p, tri, off, rt, rn)
    if name not in ("octa", "ico0", "ico2", "ico5"):
        assert np.array_equal(off, ooff)
        (orn_r,) = _check.rotate_rings(ooff, orn)
        assert np.array_equal(rn, orn_r)
    de, vv, vc = m.edges()
    assert len(de) == 3 * len(p) - 6
    assert (de[:, 0] < de[:, 1]).all() and np.array_equal(de, vc)
    # Voronoi edge k is the dual of Delaunay edge k: both triangles contain both cells (DV:435-460)
    t0 = tri.astype(np.int64)
    for col in range(2):
        tt = t0[vv[:, col] - 1]
        assert ((tt == de[:, [0]]).any(1) & (tt == de[:, [1]]).any(1)).all()
    m.close()


def test_icosahedral_degrees(fb):
    """Bisected icosahedra: exactly 12 degree-5 cells, the rest degree-6."""
    from fishman_b200 import points
    for k in (1, 3, 5):
        p = points.ico(k)
        m = fb.Mesh.from_xyz(p)
        m.delaunay_triangulation()
        m.voronoi_diagram()
        off, _, _ = m.rings()
        deg = np.diff(off)
        assert (deg == 5).sum() == 12 and (deg == 6).sum() == len(p) - 12
        m.close()


def test_qhull_163842(fb):
    from scipy.spatial import ConvexHull
    from fishman_b200 import points
    p = points.uniform(163842, 0)
    m = fb.Mesh.from_xyz(p)
    m.delaunay_triangulation()
    tri, adj = m.triangles()
    _check.check_triangulation(p, tri, adj)
    q = ConvexHull(p).simplices + 1
    assert np.array_equal(_check.canonical(tri), _check.canonical(q))
    m.close()


def test_degenerate_latlon_grid(fb):
    """Cocircular quads and collinear meridian points: the exact + SoS predicates must
    still give a valid triangulation (the reference abstains / is history dependent here)."""
    from fishman_b200 import points
    p = points.latlon_grid(24, 48)
    m = fb.Mesh.from_xyz(p)
    m.delaunay_triangulation()
    tri, adj = m.triangles()
    _check.check_triangulation(p, tri, adj)
    m.close()


def test_lonlat_entry(fb):
    from fishman_b200 import points
    p = points.uniform(2000, 5)
    lon, lat = np.arctan2(p[:, 1], p[:, 0]), np.arcsin(p[:, 2])
    m1 = fb.Mesh(lon=lon, lat=lat)
    m1.delaunay_triangulation()
    q = np.stack([np.cos(lat) * np.cos(lon), np.cos(lat) * np.sin(lon), np.sin(lat)], 1)
    m2 = fb.Mesh.from_xyz(q)
    m2.delaunay_triangulation()
    assert np.array_equal(_check.canonical(m1.triangles()[0]), _check.canonical(m2.triangles()[0]))
    assert np.abs(m1.generators() - q).max() < 1e-15
    m1.close(); m2.close()


def test_errors(fb):
    with pytest.raises(fb.FishmanError):
        fb.Mesh.from_xyz(np.eye(3))
    p = np.array([[1, 0, 0], [0, 1, 0], [0, 0, 1], [-1, 0, 0], [0, -1, 0], [0, 0, 1.0]])
    with pytest.raises(fb.FishmanError, match="coincides"):
        fb.Mesh.from_xyz(p)  # duplicate generator, DV:286-288
    hemi = np.abs(np.random.default_rng(0).normal(size=(50, 3)))
    hemi /= np.linalg.norm(hemi, axis=1)[:, None]
    m = fb.Mesh.from_xyz(hemi)
    with pytest.raises(fb.FishmanError):
        m.delaunay_triangulation()
    m.close()
    m = fb.Mesh.from_xyz(np.random.default_rng(0).normal(size=(10, 3)))
    with pytest.raises(fb.FishmanError):
        m.voronoi_diagram()
    m.close()


@pytest.mark.parametrize("dens", ["uniform", "x4"])
@pytest.mark.parametrize("warm", [False, True])
def test_lloyd_steps_against_oracle(fb, O, dens, warm):
    """k Lloyd iterations, both sides re-fed the same generators each step."""
    from fishman_b200 import points
    p = points.fibonacci(4000, 1)
    rho_g = fb.density_uniform() if dens == "uniform" else fb.density_x4()
    rho_o = O.density_uniform() if dens == "uniform" else O.density_x4()
    m = fb.Mesh.from_xyz(p)
    o = O.Oracle(p)
    for it in range(5):
        g0 = m.generators()
        n_it, md = m.lloyd(rho_g, max_iter=1, warm=warm)
        assert n_it == 1
        o.set_generators(g0)
        omd = o.lloyd_step(rho_o)
        og = o.generators()
        g1 = m.generators()
        assert np.abs(g1 - og).max() < TOL
        assert abs(md - omd) < TOL
        assert np.abs(np.linalg.norm(g1, axis=1) - 1).max() < 4e-16
    # the warm-repaired triangulation is the Delaunay triangulation of the current generators
    st = m.stats()
    assert (st["warm"] > 0) == warm
    m.close()


def test_warm_equals_cold_topology(fb):
    from fishman_b200 import points
    p = points.uniform(20000, 2)
    m = fb.Mesh.from_xyz(p)
    m.lloyd(max_iter=6, warm=True)
    g = m.generators()
    # warm: repair only (no centroid step) == cold triangulation of the same generators
    m2 = fb.Mesh.from_xyz(g)
    m2.delaunay_triangulation()
    m.set_generators(g)
    m.retriangulate(warm=True)
    assert np.array_equal(_check.canonical(m.triangles()[0]), _check.canonical(m2.triangles()[0]))
    st = m.stats()
    assert st["warm"] >= 3 and st["warm"] + st["cold"] == 7  # a turned-over triangle forces a rebuild
    m.close(); m2.close()


def test_lloyd_converges_and_quality(fb, O):
    from fishman_b200 import points
    p = points.uniform(2562, 4)
    m = fb.Mesh.from_xyz(p)
    it, md0 = m.lloyd(max_iter=1)
    it, md = m.lloyd(max_iter=200, tol=0.0)
    assert md < md0 / 10
    m.delaunay_triangulation()
    m.voronoi_diagram()
    off, _, _ = m.rings()
    deg = np.diff(off)
    assert ((deg >= 5) & (deg <= 7)).mean() > 0.99
    o = O.Oracle(m.generators())
    o.delaunay_triangulation(); o.voronoi_diagram()
    ac, at, sc, st = o.areas()
    assert abs(sc - 4 * np.pi) < 1e-10 and abs(st - 4 * np.pi) < 1e-10  # MM:306, MM:320
    assert ac.max() / ac.min() < 2.0
    m.close()


def test_full_size_properties(fb):
    """C4 size (2,621,442): size-independent properties only."""
    from fishman_b200 import points
    p = points.fibonacci(2621442, 0)
    m = fb.Mesh.from_xyz(p)
    m.delaunay_triangulation()
    tri, adj = m.triangles()
    _check.check_triangulation(p, tri, adj)
    # idempotence: warm repair of a Delaunay triangulation flips nothing
    m.stats_reset()
    m.retriangulate(warm=True)
    assert m.stats()["flips"] == 0
    m.voronoi_diagram()
    off, rt, rn = m.rings()
    _check.check_rings(p, tri, off, rt, rn)
    # local Delaunay property on every edge, plain fp64 with the reference's eps rule (DV:2051-2061)
    t0 = tri.astype(np.int64) - 1
    a0 = adj.astype(np.int64) - 1
    for k in range(3):
        nb = t0[a0[:, k]]
        mask = (nb != t0[:, [(k + 1) % 3]]) & (nb != t0[:, [(k + 2) % 3]])
        d = p[nb[mask]]
        v1, v2, v3 = p[t0[:, 0]], p[t0[:, 1]], p[t0[:, 2]]
        det = np.einsum("ij,ij->i", v3 - d, np.cross(v2 - d, v1 - d))
        assert (det <= 2.220446049250313e-16).all()
    it, md = m.lloyd(max_iter=3, warm=True)
    g = m.generators()
    assert np.abs(np.linalg.norm(g, axis=1) - 1).max() < 4e-16
    assert 0 < md < 0.01
    m.close()


@pytest.mark.parametrize("name", ["u300", "fib500", "u2000"])
def test_against_golden_fixtures(fb, name):
    """Committed fixtures (tests/golden/make_golden.py): Qhull topology; circumcentres and one
    Lloyd step (uniform and x4 density) evaluated with 50-digit mpmath."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz"))
    m = fb.Mesh.from_xyz(g["xyz"])
    m.delaunay_triangulation()
    m.voronoi_diagram()
    tri, _ = m.triangles()
    assert np.array_equal(_check.canonical(tri), g["tri"])
    if "cc" in g:
        cc, _ = m.circumcentres()
        assert _check.max_cc_diff(g["tri_ccw"], g["cc"], tri, cc) < 1e-13
        for rho, key in ((fb.density_uniform(), "lloyd_uniform"), (fb.density_x4(), "lloyd_x4")):
            m.set_generators(g["xyz"])
            m.lloyd(rho, max_iter=1, warm=True)
            assert np.abs(m.generators() - g[key]).max() < 1e-13
    m.close()


@pytest.mark.parametrize("case", ["ico5_uniform", "u10242_x4", "fib20000_relaxed"])
def test_quality_against_oracle(fb, O, case):
    """fm_quality (degree histogram, cell / triangle areas by calc_area SG:198-227, triangle angles
    by calc_sphere_angle SG:247-285, largest generator-to-centroid chord) against fo_quality.
    Tolerances: histogram exact; areas and angles 1e-13 absolute (sums of <= 15 acos values of
    O(1)); the two 4 pi sums 1e-10 (the reference's own threshold, MM:306/320); chord 1e-12."""
    from fishman_b200 import points
    if case == "ico5_uniform":
        p, rg, ro = points.ico(5), fb.density_uniform(), O.density_uniform()
    elif case == "u10242_x4":
        p, rg, ro = points.uniform(10242, 0), fb.density_x4(), O.density_x4()
    else:
        p, rg, ro = points.fibonacci(20000, 3), fb.density_uniform(), O.density_uniform()
    m = fb.Mesh.from_xyz(p)
    if case == "fib20000_relaxed":
        m.lloyd(rg, max_iter=3, warm=True)
        p = m.generators()
    else:
        m.delaunay_triangulation()
    q = m.quality(rg)
    o = O.Oracle(p)
    o.delaunay_triangulation(); o.voronoi_diagram()
    oq = o.quality(ro)
    assert q["degree_hist"] == oq["degree_hist"] and sum(q["degree_hist"]) == len(p)
    for k in ("area_cell_min", "area_cell_max", "area_tri_min", "area_tri_max", "angle_min", "angle_max"):
        assert abs(q[k] - oq[k]) < 1e-13, (k, q[k], oq[k])
    for k in ("area_cell_sum", "area_tri_sum"):
        assert abs(q[k] - oq[k]) < 1e-10 and abs(q[k] - 4 * np.pi) < 1e-10, (k, q[k], oq[k])
    assert abs(q["max_disp"] - oq["max_disp"]) < TOL
    g0 = m.generators()
    q2 = m.quality(rg)
    assert q2 == q and np.array_equal(m.generators(), g0)   # deterministic sums; nothing moved
    it, md = m.lloyd(rg, max_iter=1, warm=True)
    assert abs(md - q["max_disp"]) < 1e-15   # the chord fm_quality announced is the step Lloyd makes
    m.close(); o.close()


def test_lloyd_batches_equal_single_steps(fb):
    """fm_lloyd queues its warm iterations in batches and reads the device back once per batch
    (fm_iter_end takes the convergence / rebuild decisions on the device).  Same iteration count,
    same displacement, bit-identical generators as one host round trip per iteration -- through a
    rebuild in the middle (random start: triangles turn over in the first iterations) and through
    the convergence test."""
    from fishman_b200 import points
    p = points.uniform(20000, 2)
    a, b = fb.Mesh.from_xyz(p), fb.Mesh.from_xyz(p)
    mds = [a.lloyd(max_iter=1, warm=True)[1] for _ in range(12)]
    it, md = b.lloyd(max_iter=12, tol=0.0, warm=True)
    assert it == 12 and md == mds[-1]
    assert np.array_equal(a.generators(), b.generators())
    sa, sb = a.stats(), b.stats()
    assert sa["cold"] == sb["cold"] and sa["warm"] == sb["warm"] and sa["flips"] == sb["flips"]
    assert sa["cold"] >= 2 and sa["warm"] >= 3          # both paths were taken
    # convergence inside a batch: stop at the first iteration with max_disp <= tol
    more = [a.lloyd(max_iter=1, warm=True)[1] for _ in range(40)]
    tol = more[24] * 1.0000001
    k = next(i for i, v in enumerate(more) if v <= tol) + 1
    it, md = b.lloyd(max_iter=100, tol=tol, warm=True)
    assert it == k and md == more[k - 1]
    # the launches of the batch after the one that converged did nothing lasting: the mesh is in the
    # state the single-step loop reaches after k steps, and the next repair starts on clean marks
    c = fb.Mesh.from_xyz(p)
    c.lloyd(max_iter=12 + k, warm=True)
    assert np.array_equal(b.generators(), c.generators())
    b.retriangulate(warm=True)
    fresh = fb.Mesh.from_xyz(b.generators())
    fresh.delaunay_triangulation()
    assert np.array_equal(_check.canonical(b.triangles()[0]), _check.canonical(fresh.triangles()[0]))
    it2, md2 = b.lloyd(max_iter=3, warm=True)
    c.lloyd(max_iter=3, warm=True)
    assert it2 == 3 and np.array_equal(b.generators(), c.generators())
    a.close(); b.close(); c.close(); fresh.close()


@pytest.mark.parametrize("n", [40, 200, 2000, 20000])
def test_retriangulate_warm_after_a_strong_move(fb, n):
    """fm_retriangulate(h, warm=1) when triangles turned over: the failed repair has stamped
    amark / vote, the rebuild must start on a fresh stamp (CPU twin:
    tests/test_emu.py::test_rebuild_after_a_failed_warm_repair)."""
    from fishman_b200 import points
    for trial in range(20 if n <= 200 else 3):
        rng = np.random.default_rng(1000 * n + trial)
        p = points.uniform(n, trial)
        q = p + 1.5 * np.sqrt(4 * np.pi / n) * rng.normal(size=p.shape)
        q /= np.linalg.norm(q, axis=1)[:, None]
        m = fb.Mesh.from_xyz(p)
        m.delaunay_triangulation()
        m.set_generators(q)
        m.stats_reset()
        m.retriangulate(warm=True)
        assert m.stats()["cold"] == 1      # a triangle turned over: rebuilt
        tri, adj = m.triangles()
        _check.check_triangulation(q, tri, adj)
        m2 = fb.Mesh.from_xyz(q)
        m2.delaunay_triangulation()
        assert np.array_equal(_check.canonical(tri), _check.canonical(m2.triangles()[0]))
        m.close(); m2.close()


def test_getters_follow_the_generators(fb):
    """After fm_lloyd (ends with a centroid update) or fm_set_generators the getters repair the
    connectivity first: triangles, Voronoi diagram and quality belong to the CURRENT generators."""
    from fishman_b200 import points
    p = points.fibonacci(6000, 2)
    m = fb.Mesh.from_xyz(p)
    m.lloyd(max_iter=4, warm=True)
    g = m.generators()
    m2 = fb.Mesh.from_xyz(g)
    m2.delaunay_triangulation(); m2.voronoi_diagram()
    assert np.array_equal(_check.canonical(m.triangles()[0]), _check.canonical(m2.triangles()[0]))
    m.lloyd(max_iter=1, warm=True)
    m.set_generators(g)
    m.voronoi_diagram()
    tri, _ = m.triangles()
    tri2, _ = m2.triangles()
    assert np.array_equal(_check.canonical(tri), _check.canonical(tri2))
    assert _check.max_cc_diff(tri, m.circumcentres()[0], tri2, m2.circumcentres()[0]) < 1e-15   # vertex rotation of a slot differs
    m.close(); m2.close()


def test_block_transfer_single_rank(fb):
    """fm_block_ids / fm_get_generators_block / fm_set_generators_block on an unsharded handle: the
    block is the whole set in internal order (the sharded form is checked by tools/dist_check.py)."""
    from fishman_b200 import points
    p = points.uniform(5000, 9)
    m = fb.Mesh.from_xyz(p)
    ids = m.block_ids()
    assert np.array_equal(np.sort(ids), np.arange(1, len(p) + 1))
    x, y, z = (np.empty(len(p)) for _ in range(3))
    m.get_generators_block_soa(x, y, z)
    assert np.array_equal(np.stack([x, y, z], 1), p[ids - 1])
    q = points.uniform(5000, 10)
    qb = np.ascontiguousarray(q[ids - 1].T)
    m.set_generators_block_soa(qb[0], qb[1], qb[2])
    assert np.array_equal(m.generators(), q)
    m.delaunay_triangulation()
    m.close()


def _clustered(points):
    """20,000 uniform points + 30,000 in a cap of radius 2e-4: thousands of generators in one cell of the stars' grid"""
    c = points.uniform(30000, 1) * 2e-4 + np.array([0, 0, 1.0])
    c /= np.linalg.norm(c, axis=1)[:, None]
    return np.concatenate([points.uniform(20000, 2), c])


@pytest.mark.parametrize("name", ["fib60000", "fib60000_nojitter", "ico6", "u50000", "latlon", "cluster"])
def test_cold_build_by_stars_equals_insertion_build(fb, name):
    """The cold build by stars (csrc/star_kernels.cuh) against the insertion + flip rounds on the same handle
    (FM_COLD_BUILD=flip switches the stars off): identical triangles; the stars' verifying sweep finds nothing
    to flip; sets the stars are not made for (uniform random points: too many wide searches; a lat-lon grid:
    poles of degree nlon) are refused and built by insertion -- same answer, by the other path."""
    import os
    from fishman_b200 import points
    p = {"fib60000": lambda: points.fibonacci(60000, 4), "fib60000_nojitter": lambda: points.fibonacci(60000, 0, 0.0),
         "ico6": lambda: points.ico(6), "u50000": lambda: points.uniform(50000, 7),
         "latlon": lambda: points.latlon_grid(100, 200), "cluster": lambda: _clustered(points)}[name]()
    m = fb.Mesh.from_xyz(p)
    old = os.environ.pop("FM_COLD_BUILD", None)
    try:
        m.stats_reset()
        m.delaunay_triangulation()
        st = m.star_stats()
        tri, adj = m.triangles()
        _check.check_triangulation(p, tri, adj)
        if name in ("u50000", "latlon", "cluster"):
            assert st["builds"] == 0 and st["fallbacks"] == 1
        else:
            assert st["builds"] == 1 and st["fallbacks"] == 0 and st["failed"] == 0 and st["broken"] == 0
            assert st["triangles"] == 2 * len(p) - 4 and st["degree_sum"] == 6 * len(p) - 12
            assert st["check_flips"] == 0 and st["check_turned"] == 0
        os.environ["FM_COLD_BUILD"] = "flip"
        m.stats_reset()
        m.retriangulate(warm=False)
        assert m.star_stats()["builds"] == 0
        tri2, adj2 = m.triangles()
        assert np.array_equal(_check.canonical(tri), _check.canonical(tri2))
        # a Lloyd run through both: the generators agree to the last bit or two (the slot order differs, and with it
        # the order of the sums)
        if name == "cluster":   # (spacing 1e-6 inside the cap: its circumcentres are not what a Lloyd step is for)
            return
        g_flip = None
        m.lloyd(max_iter=3, warm=True)
        g_flip = m.generators()
        del os.environ["FM_COLD_BUILD"]
        m2 = fb.Mesh.from_xyz(p)
        m2.lloyd(max_iter=3, warm=True)
        assert np.abs(m2.generators() - g_flip).max() < TOL
        m2.close()
    finally:
        os.environ.pop("FM_COLD_BUILD", None)
        if old is not None:
            os.environ["FM_COLD_BUILD"] = old
        m.close()
