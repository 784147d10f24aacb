"""ctypes binding of the CPU oracle (oracle/fishman_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  Never imported by
fishman_b200/.  Parity unpinned by the reference's own tests (it has none for
this path); see the header of fishman_oracle.c.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libfishman_oracle.so")


def build(force=False):
    src = os.path.join(_HERE, "fishman_oracle.c")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"], stdout=subprocess.DEVNULL)
    return _LIB


class Stats(C.Structure):
    _fields_ = [(k, C.c_longlong) for k in (
        "n_orient", "n_orient_on", "n_incircle", "n_incircle_on", "n_flip13", "n_flip24",
        "n_flip22", "n_relocate", "n_virtual_replaced")] + [
        ("min_abs_incircle", C.c_double), ("min_abs_orient", C.c_double)]

    def asdict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class QualityStats(C.Structure):
    _fields_ = [("degree_hist", C.c_longlong * 16)] + [(k, C.c_double) for k in (
        "area_cell_min", "area_cell_max", "area_cell_sum", "area_tri_min", "area_tri_max", "area_tri_sum",
        "angle_min", "angle_max", "max_disp")]

    def asdict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_[1:]}
        d["degree_hist"] = [int(v) for v in self.degree_hist]
        return d


class Density(C.Structure):
    _fields_ = [("kind", C.c_int), ("cx", C.c_double), ("cy", C.c_double), ("cz", C.c_double),
                ("beta", C.c_double), ("alpha", C.c_double), ("gamma", C.c_double)]


def density_uniform():
    return Density(0, 0.0, 0.0, 1.0, 0.0, 1.0, 1.0)


def density_x4(lon_deg=0.0, lat_deg=40.0, beta=np.pi / 6, alpha=np.pi / 20, gamma=0.25 ** 4):
    lon, lat = np.deg2rad(lon_deg), np.deg2rad(lat_deg)
    return Density(1, float(np.cos(lat) * np.cos(lon)), float(np.cos(lat) * np.sin(lon)),
                   float(np.sin(lat)), float(beta), float(alpha), float(gamma))


_lib = None


def density_eval(rho, xyz):
    """rho at the rows of xyz (unit vectors): the oracle's density_eval"""
    xyz = np.ascontiguousarray(xyz, dtype=np.float64)
    c = [np.ascontiguousarray(xyz[:, k]) for k in range(3)]
    out = np.empty(len(xyz))
    lib().fo_density_eval(C.byref(rho), len(xyz), _d(c[0]), _d(c[1]), _d(c[2]), _d(out))
    return out


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB)
        dp, ip, vp = C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_void_p
        L.fo_create.restype = vp
        L.fo_create.argtypes = [C.c_int, dp, dp, dp]
        L.fo_create_lonlat.restype = vp
        L.fo_create_lonlat.argtypes = [C.c_int, dp, dp]
        L.fo_destroy.argtypes = [vp]
        L.fo_last_error.restype = C.c_char_p
        L.fo_last_error.argtypes = [vp]
        L.fo_get_stats.argtypes = [vp, C.POINTER(Stats)]
        L.fo_triangulate.argtypes = [vp, C.c_int, C.c_int, C.c_int]
        L.fo_num_triangles.argtypes = [vp]
        L.fo_get_triangles.argtypes = [vp, ip, ip]
        L.fo_voronoi.argtypes = [vp]
        L.fo_get_circumcentres.argtypes = [vp, dp, dp]
        L.fo_get_rings.argtypes = [vp, ip, ip, ip]
        L.fo_get_generators.argtypes = [vp, dp, dp, dp]
        L.fo_set_generators.argtypes = [vp, dp, dp, dp]
        L.fo_centroids.argtypes = [vp, C.POINTER(Density), dp, dp, dp]
        L.fo_density_eval.restype = None
        L.fo_density_eval.argtypes = [C.POINTER(Density), C.c_longlong, dp, dp, dp, dp]
        L.fo_lloyd_step.argtypes = [vp, C.POINTER(Density), C.c_int, C.c_int, C.c_int, dp]
        L.fo_calc_area.restype = C.c_double
        L.fo_calc_area.argtypes = [C.c_int, dp, dp, dp]
        L.fo_areas.argtypes = [vp, dp, dp, dp, dp]
        L.fo_quality.argtypes = [vp, C.POINTER(Density), C.POINTER(QualityStats)]
        L.fo_mpas.argtypes = [vp, C.c_int, ip]
        L.fo_mpas_get_edges.argtypes = [vp, ip, ip, dp, dp, dp, dp, dp, dp, dp]
        L.fo_mpas_get_cells.argtypes = [vp, ip, ip, ip, ip, dp]
        L.fo_mpas_get_vertices.argtypes = [vp, ip, ip, dp, dp]
        L.fo_mpas_get_edges_on_edge.argtypes = [vp, ip, ip, dp]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _i(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))


class OracleError(RuntimeError):
    pass


class Oracle:
    """Mirror of the reference call sequence (src/gen_mesh.F90:25-27)."""

    def __init__(self, xyz, first3=(0, 1, 2)):
        xyz = np.ascontiguousarray(xyz, dtype=np.float64)
        self.n = xyz.shape[0]
        x, y, z = (np.ascontiguousarray(xyz[:, k]) for k in range(3))
        self._h = lib().fo_create(self.n, _d(x), _d(y), _d(z))
        self.first3 = tuple(int(v) for v in first3)

    def close(self):
        if self._h:
            lib().fo_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc):
        if rc != 0:
            raise OracleError(lib().fo_last_error(self._h).decode())

    def stats(self):
        s = Stats()
        lib().fo_get_stats(self._h, C.byref(s))
        return s.asdict()

    def delaunay_triangulation(self):
        self._chk(lib().fo_triangulate(self._h, *self.first3))
        nT = lib().fo_num_triangles(self._h)
        tri = np.empty((nT, 3), dtype=np.int32)
        adj = np.empty((nT, 3), dtype=np.int32)
        lib().fo_get_triangles(self._h, _i(tri), _i(adj))
        return tri, adj

    def triangles(self):
        """(tri, adj) of the last triangulation (after lloyd_step: the one of the generators the step started from)."""
        nT = lib().fo_num_triangles(self._h)
        tri = np.empty((nT, 3), dtype=np.int32)
        adj = np.empty((nT, 3), dtype=np.int32)
        lib().fo_get_triangles(self._h, _i(tri), _i(adj))
        return tri, adj

    def voronoi_diagram(self):
        self._chk(lib().fo_voronoi(self._h))
        nT = lib().fo_num_triangles(self._h)
        xyz = np.empty((nT, 3))
        lonlat = np.empty((nT, 2))
        lib().fo_get_circumcentres(self._h, _d(xyz), _d(lonlat))
        off = np.empty(self.n + 1, dtype=np.int32)
        rt = np.empty(3 * nT, dtype=np.int32)
        rn = np.empty(3 * nT, dtype=np.int32)
        lib().fo_get_rings(self._h, _i(off), _i(rt), _i(rn))
        return xyz, lonlat, off, rt, rn

    def centroids(self, rho):
        g = np.empty((3, self.n))
        self._chk(lib().fo_centroids(self._h, C.byref(rho), _d(g[0]), _d(g[1]), _d(g[2])))
        return np.ascontiguousarray(g.T)

    def lloyd_step(self, rho):
        md = C.c_double(0.0)
        self._chk(lib().fo_lloyd_step(self._h, C.byref(rho), *self.first3, C.byref(md)))
        return md.value

    def generators(self):
        g = np.empty((3, self.n))
        lib().fo_get_generators(self._h, _d(g[0]), _d(g[1]), _d(g[2]))
        return np.ascontiguousarray(g.T)

    def set_generators(self, xyz):
        xyz = np.ascontiguousarray(xyz, dtype=np.float64)
        x, y, z = (np.ascontiguousarray(xyz[:, k]) for k in range(3))
        lib().fo_set_generators(self._h, _d(x), _d(y), _d(z))

    def areas(self):
        nT = lib().fo_num_triangles(self._h)
        ac = np.empty(self.n)
        at = np.empty(nT)
        sc, st = C.c_double(), C.c_double()
        self._chk(lib().fo_areas(self._h, _d(ac), _d(at), C.byref(sc), C.byref(st)))
        return ac, at, sc.value, st.value


    def quality(self, rho=None):
        """Mesh-quality statistics (fo_quality) of the current triangulation + Voronoi diagram: dict."""
        q = QualityStats()
        self._chk(lib().fo_quality(self._h, C.byref(rho or density_uniform()), C.byref(q)))
        return q.asdict()

    def mpas(self, strict=True):
        """mpas_mesh_adapt + calc_weightsOnEdge (src/mpas_mesh_mod.F90:181-345, :442-561) as a dict
        of arrays named like the MPAS variables; ids 1-based, rows padded with 0.  strict: the
        reference's Sum = 4 pi +- 1e-10 checks are fatal (they fire on any obtuse Delaunay triangle)."""
        me = C.c_int(0)
        self._chk(lib().fo_mpas(self._h, 1 if strict else 0, C.byref(me)))
        ME, n = me.value, self.n
        nT = lib().fo_num_triangles(self._h)
        nE = nT + n - 2
        i32, f64 = np.int32, np.float64
        o = dict(maxEdges=ME, nCells=n, nEdges=nE, nVertices=nT,
                 cellsOnEdge=np.empty((nE, 2), i32), verticesOnEdge=np.empty((nE, 2), i32),
                 xyzEdge=np.empty((nE, 3), f64), lonlatEdge=np.empty((nE, 2), f64),
                 dvEdge=np.empty(nE, f64), dcEdge=np.empty(nE, f64), dv1Edge=np.empty(nE, f64),
                 dv2Edge=np.empty(nE, f64), angleEdge=np.empty(nE, f64),
                 nEdgesOnCell=np.empty(n, i32), verticesOnCell=np.empty((n, ME), i32),
                 cellsOnCell=np.empty((n, ME), i32), edgesOnCell=np.empty((n, ME), i32), areaCell=np.empty(n, f64),
                 cellsOnVertex=np.empty((nT, 3), i32), edgesOnVertex=np.empty((nT, 3), i32),
                 areaTriangle=np.empty(nT, f64), kiteAreasOnVertex=np.empty((nT, 3), f64),
                 nEdgesOnEdge=np.empty(nE, i32), edgesOnEdge=np.empty((nE, 2 * ME), i32),
                 weightsOnEdge=np.empty((nE, 2 * ME), f64))
        L = lib()
        self._chk(L.fo_mpas_get_edges(self._h, _i(o["cellsOnEdge"]), _i(o["verticesOnEdge"]), _d(o["xyzEdge"]),
                                      _d(o["lonlatEdge"]), _d(o["dvEdge"]), _d(o["dcEdge"]), _d(o["dv1Edge"]),
                                      _d(o["dv2Edge"]), _d(o["angleEdge"])))
        self._chk(L.fo_mpas_get_cells(self._h, _i(o["nEdgesOnCell"]), _i(o["verticesOnCell"]), _i(o["cellsOnCell"]),
                                      _i(o["edgesOnCell"]), _d(o["areaCell"])))
        self._chk(L.fo_mpas_get_vertices(self._h, _i(o["cellsOnVertex"]), _i(o["edgesOnVertex"]),
                                         _d(o["areaTriangle"]), _d(o["kiteAreasOnVertex"])))
        self._chk(L.fo_mpas_get_edges_on_edge(self._h, _i(o["nEdgesOnEdge"]), _i(o["edgesOnEdge"]),
                                              _d(o["weightsOnEdge"])))
        return o


def canonical_triangles(tri):
    """Lexicographically sorted rows of sorted vertex triples."""
    t = np.sort(np.asarray(tri, dtype=np.int64), axis=1)
    order = np.lexsort((t[:, 2], t[:, 1], t[:, 0]))
    return t[order]
