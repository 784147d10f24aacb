"""MPAS adapter arithmetic: mpas_mesh_adapt
(/root/reference/src/mpas_mesh_mod.F90:181-345) and calc_weightsOnEdge (MM:442-561).

CPU part (not gpu): pins the oracle's restatement (fo_mpas) against analytic values on the
icosahedron, an independent numpy evaluation of the per-edge geometry, the reference's own
run-time checks (sums of areas = 4 pi +- 1e-10, MM:306/320/339) and the antisymmetry the TRiSK
weights must have.  GPU part: the device kernels through the C ABI against the oracle."""
import numpy as np
import pytest

from fishman_b200 import points
from oracle import oracle as O

TOL = 1e-12


def oracle_mpas(p, lloyd=0, rho=None, strict=True):
    o = O.Oracle(p)
    for _ in range(lloyd):
        o.lloyd_step(rho or O.density_uniform())
    g = o.generators()
    o.delaunay_triangulation()
    o.voronoi_diagram()
    return g, o.mpas(strict=strict)


def test_icosahedron_analytic():
    g, m = oracle_mpas(points.ico(0))
    assert m["maxEdges"] == 5 and m["nEdges"] == 30 and (m["nEdgesOnCell"] == 5).all()
    dc, dv = np.arccos(1 / np.sqrt(5)), np.arccos(np.sqrt(5) / 3)
    assert np.abs(m["dcEdge"] - dc).max() < 1e-14 and np.abs(m["dvEdge"] - dv).max() < 1e-14
    assert np.abs(m["dv1Edge"] - dv / 2).max() < 1e-14 and np.abs(m["dv2Edge"] - dv / 2).max() < 1e-14
    assert np.abs(m["areaCell"] - 4 * np.pi / 12).max() < 1e-13
    assert np.abs(m["areaTriangle"] - 4 * np.pi / 20).max() < 1e-13
    assert np.abs(m["kiteAreasOnVertex"] - 4 * np.pi / 60).max() < 1e-13
    assert (m["nEdgesOnEdge"] == 8).all()
    # regular pentagons: R = 1/5, so the i-th edge after e weighs +-((5 - i)/5 - 1/2) dv/dc
    w = np.abs(m["weightsOnEdge"][:, :8])
    want = np.abs(np.array([0.3, 0.1, -0.1, -0.3] * 2)) * dv / dc
    assert np.abs(w - want).max() < 1e-13


@pytest.mark.parametrize("k", [2, 4])
def test_oracle_edges_against_numpy_and_invariants(k):
    g, m = oracle_mpas(points.ico(k))   # strict: the reference's 4 pi checks passed
    c1, c2 = m["cellsOnEdge"][:, 0] - 1, m["cellsOnEdge"][:, 1] - 1
    assert (c1 < c2).all()              # DVT1 < DVT2, DV:421
    # the Voronoi edge lies on the perpendicular bisector of the Delaunay edge: xEdge = midpoint
    mid = g[c1] + g[c2]
    mid /= np.linalg.norm(mid, axis=1)[:, None]
    assert np.abs(m["xyzEdge"] - mid).max() < 1e-13
    lon = np.arctan2(mid[:, 1], mid[:, 0]) % (2 * np.pi)
    dl = np.abs(m["lonlatEdge"][:, 0] - lon)
    assert np.minimum(dl, 2 * np.pi - dl).max() < 1e-12
    assert np.abs(m["dcEdge"] - np.arccos(np.clip((g[c1] * g[c2]).sum(1), -1, 1))).max() < 1e-15
    assert np.abs(m["dv1Edge"] + m["dv2Edge"] - m["dvEdge"]).max() < 1e-7   # acos near 0 is ill conditioned
    for name in ("areaCell", "areaTriangle", "kiteAreasOnVertex"):
        assert abs(m[name].sum() - 4 * np.pi) < 1e-10
    # every kite belongs to one (vertex, cell) pair; per cell they add up to areaCell
    acc = np.zeros(m["nCells"])
    np.add.at(acc, m["cellsOnVertex"].ravel() - 1, m["kiteAreasOnVertex"].ravel())
    assert np.abs(acc - m["areaCell"]).max() < 1e-12
    # TRiSK: w(e,e') dc(e)/dv(e') = -w(e',e) dc(e')/dv(e)
    E, eoe, w = m["nEdges"], m["edgesOnEdge"], m["weightsOnEdge"]
    d = {}
    for e in range(E):
        for i in range(m["nEdgesOnEdge"][e]):
            e2 = eoe[e, i] - 1
            d[(e, e2)] = w[e, i] * m["dcEdge"][e] / m["dvEdge"][e2]
    assert max(abs(d[(a, b)] + d[(b, a)]) for (a, b) in d) < 1e-11
    # connectivity: edgesOnCell(j) joins the cell and cellsOnCell(j), from verticesOnCell(j) to (j+1)
    for i in range(0, m["nCells"], 37):
        dgr = m["nEdgesOnCell"][i]
        for j in range(dgr):
            e = m["edgesOnCell"][i, j] - 1
            assert {c1[e], c2[e]} == {i, m["cellsOnCell"][i, j] - 1}
            assert set(m["verticesOnEdge"][e]) == {m["verticesOnCell"][i, j], m["verticesOnCell"][i, (j + 1) % dgr]}


def test_strict_checks_fire_on_obtuse_triangles():
    with pytest.raises(O.OracleError, match="Too large total kite area error"):  # MM:339
        oracle_mpas(points.uniform(500, 1))


def _compare(m, r, masked=False):
    """m: device result, r: oracle result (same edge numbering by construction)."""
    for k in ("cellsOnEdge", "verticesOnEdge", "nEdgesOnCell", "verticesOnCell", "cellsOnCell", "edgesOnCell",
              "cellsOnVertex", "edgesOnVertex", "nEdgesOnEdge", "edgesOnEdge"):
        assert np.array_equal(m[k], r[k]), k
    assert m["maxEdges"] == r["maxEdges"]
    for k in ("xyzEdge", "dcEdge", "angleEdge", "areaCell", "areaTriangle"):
        assert np.abs(m[k] - r[k]).max() < TOL, k
    dl = np.abs(m["lonlatEdge"][:, 0] - r["lonlatEdge"][:, 0])   # a point on the zero meridian may come out as 0 or 2 pi
    assert np.minimum(dl, 2 * np.pi - dl).max() < TOL
    assert np.abs(m["lonlatEdge"][:, 1] - r["lonlatEdge"][:, 1]).max() < TOL
    for k in ("dvEdge", "dv1Edge", "dv2Edge"):  # acos of a dot product near 1: absolute 1e-12 on the cosine
        assert np.abs(np.cos(m[k]) - np.cos(r[k])).max() < TOL, k
        assert np.abs(m[k] - r[k]).max() < 1e-7, k
    for k in ("kiteAreasOnVertex", "weightsOnEdge"):
        a, b = m[k], r[k]
        sane = (np.abs(b) < 1.0) & (np.abs(a) < 1.0) if masked else np.ones(a.shape, bool)
        assert sane.mean() > 0.9
        if k == "weightsOnEdge":
            # w carries dvEdge(e') = acos(dot ~ 1), whose absolute error is ~1e-16 / dvEdge: compare where
            # the Voronoi edge is not nearly degenerate (two circumcentres about to merge)
            e2 = np.maximum(r["edgesOnEdge"], 1) - 1
            sane &= (r["edgesOnEdge"] > 0) & (r["dvEdge"][e2] > 1e-4 * r["dcEdge"][:, None])
            if masked:  # kites of obtuse triangles give kite fractions far outside [0, 1]: leave those out
                sane &= np.abs(b * r["dcEdge"][:, None] / np.maximum(r["dvEdge"][e2], 1e-300)) < 0.6
            assert sane.mean() > 0.4
        assert np.abs(a - b)[sane].max() < (1e-9 if k == "weightsOnEdge" else 1e-11), k


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["ico3", "ico5", "fib_relaxed", "x4_relaxed"])
def test_device_mpas_against_oracle(case):
    import fishman_b200 as fb
    if case.startswith("ico"):
        p, lloyd, rho_o, rho_g, strict = points.ico(int(case[3:])), 0, None, None, True
    elif case == "fib_relaxed":
        p, lloyd, rho_o, rho_g, strict = points.fibonacci(4000, 0, 0.2), 40, O.density_uniform(), None, False
    else:
        p, lloyd, rho_o, rho_g, strict = points.fibonacci(6000, 1, 0.2), 40, O.density_x4(), None, False
    g, r = oracle_mpas(p, lloyd, rho_o, strict=strict)
    mesh = fb.Mesh.from_xyz(g)
    mesh.delaunay_triangulation()
    mesh.voronoi_diagram(all=True)
    # same triangle numbering is not guaranteed: vertex (triangle) ids are compared through the
    # canonical key of their three cells, so renumber the device result into the oracle's ids
    m = mesh.mpas(strict=strict)
    tri = mesh.triangles()[0]
    mesh.close()
    o = O.Oracle(g)
    otri, _ = o.delaunay_triangulation()
    key = {tuple(sorted(t)): i + 1 for i, t in enumerate(otri.tolist())}
    ren = np.array([0] + [key[tuple(sorted(t))] for t in tri.tolist()], dtype=np.int32)  # device id -> oracle id
    inv = np.zeros_like(ren)
    inv[ren[1:]] = np.arange(1, len(ren))
    for k in ("verticesOnEdge", "verticesOnCell"):
        m[k] = ren[m[k]]
    for k in ("cellsOnVertex", "edgesOnVertex", "areaTriangle", "kiteAreasOnVertex"):
        m[k] = m[k][inv[1:] - 1]
    _compare(m, r, masked=not strict)
    if strict:
        for name in ("areaCell", "areaTriangle", "kiteAreasOnVertex"):
            assert abs(m[name].sum() - 4 * np.pi) < 1e-10


@pytest.mark.gpu
def test_device_mpas_strict_error_and_order():
    import fishman_b200 as fb
    mesh = fb.Mesh.from_xyz(points.uniform(500, 1))
    mesh.delaunay_triangulation()
    mesh.voronoi_diagram(all=False)
    with pytest.raises(fb.FishmanError, match="fm_voronoi"):
        mesh.mpas()
    mesh.voronoi_diagram(all=True)
    with pytest.raises(fb.FishmanError, match="Too large total"):
        mesh.mpas(strict=True)
    assert mesh.mpas(strict=False)["maxEdges"] >= 7
    mesh.close()


def test_writer_contract_and_grids_roundtrip(tmp_path):
    """The on-disk contract of mpas_mesh_output (MM:89-138): 37 variables over 7 sized dimensions;
    the grids input contract of gen_mesh (vtx_p(location_nv, 3), src/gen_mesh.F90:16-22)."""
    from fishman_b200 import mpas_mesh
    assert len(mpas_mesh.VARIABLES) == 37 and len({v for v, _ in mpas_mesh.VARIABLES}) == 37
    dims = {d for _, ds in mpas_mesh.VARIABLES for d in ds}
    assert dims == {"nCells", "nEdges", "nVertices", "vertexDegree", "maxEdges", "maxEdges2", "TWO"}
    p = points.ico(2)
    f = str(tmp_path / "grids.npz")
    mpas_mesh.write_grids(f, p)
    assert np.array_equal(mpas_mesh.read_grids(f), p)


@pytest.mark.gpu
def test_gen_mesh_end_to_end(tmp_path):
    """program gen_mesh: grids file -> DT -> VD(all) -> MPAS file, against the oracle's arrays."""
    from fishman_b200 import mpas_mesh
    p = points.ico(4)
    f = str(tmp_path / "grids.npz")
    mpas_mesh.write_grids(f, p)
    out = mpas_mesh.gen_mesh(f, str(tmp_path / "mpas_mesh.2562.npz"))
    d = np.load(out)
    g, r = oracle_mpas(p)
    sizes = {k[4:]: int(d[k]) for k in d.files if k.startswith("dim_")}
    assert sizes == dict(nCells=2562, nEdges=7680, nVertices=5120, vertexDegree=3, maxEdges=6, maxEdges2=12, TWO=2)
    assert str(d["att_on_a_sphere"]) == "YES" and float(d["att_sphere_radius"]) == 1.0
    for name, ds in mpas_mesh.VARIABLES:
        assert d[name].shape == tuple(sizes[x] for x in ds), name
    for name in ("cellsOnEdge", "cellsOnCell", "edgesOnCell", "nEdgesOnCell", "nEdgesOnEdge", "edgesOnEdge"):
        assert np.array_equal(d[name], r[name]), name
    for name in ("dcEdge", "angleEdge", "areaCell"):
        assert np.abs(d[name] - r[name]).max() < TOL, name
    assert np.abs(np.stack([d["xEdge"], d["yEdge"], d["zEdge"]], 1) - r["xyzEdge"]).max() < TOL
    assert np.abs(np.stack([d["xCell"], d["yCell"], d["zCell"]], 1) - p).max() == 0.0
    assert np.abs(d["weightsOnEdge"] - r["weightsOnEdge"]).max() < 1e-9   # signed: a sign error must not pass
