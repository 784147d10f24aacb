"""Host-side mirror of the reference's delaunay_voronoi_mod interface
(/root/reference/src/delaunay_voronoi_mod.F90:15-30) over the C ABI.

The reference module keeps its mesh in module-level state and exposes four
procedures, called in this order by gen_mesh (/root/reference/src/gen_mesh.F90:25-28):

    delaunay_voronoi_init(num_point, lon=, lat= | x=, y=, z=)     DV:188
    delaunay_triangulation()                                      DV:226
    voronoi_diagram(all=)                                         DV:358
    delaunay_voronoi_output(output)                               DV:587

`Mesh` carries the same four steps as methods (one handle per mesh instead of
global state, which is what makes a Lloyd loop possible at all -- SURVEY.md F5);
the module-level functions of the same names keep one default mesh, like the
Fortran.  All ids are 1-based as in the reference.  Errors raise FishmanError
(the reference calls log_error -> stop 1).
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import Density


class FishmanError(RuntimeError):
    pass


def density_uniform():
    return Density(0, 0.0, 0.0, 1.0, 0.0, 1.0, 1.0)


def density_x4(lon_deg=0.0, lat_deg=40.0, beta=np.pi / 6, alpha=np.pi / 20, gamma=0.25 ** 4):
    """4:1 regional refinement of SURVEY.md 7.3 (spacing ~ rho**-1/4)."""
    lon, lat = np.deg2rad(lon_deg), np.deg2rad(lat_deg)
    return Density(1, float(np.cos(lat) * np.cos(lon)), float(np.cos(lat) * np.sin(lon)),
                   float(np.sin(lat)), float(beta), float(alpha), float(gamma))


def _d(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def density_eval(rho, xyz):
    """rho at the rows of xyz (unit vectors), evaluated on the device by the Lloyd kernel's own
    density code (fm_density_eval)."""
    xyz = np.ascontiguousarray(xyz, dtype=np.float64)
    c = [np.ascontiguousarray(xyz[:, k]) for k in range(3)]
    out = np.empty(len(xyz))
    if _lib.lib().fm_density_eval(C.byref(rho), len(xyz), _d(c[0]), _d(c[1]), _d(c[2]), _d(out)):
        raise FishmanError(_lib.lib().fm_last_error(None).decode())
    return out


def _i(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def fp64_peak_tflops(min_ms=60.0):
    """fp64 FMA rate of the current GPU measured now (fm_fp64_peak), TFLOP/s."""
    v = C.c_double(0.0)
    if _lib.lib().fm_fp64_peak(float(min_ms), C.byref(v)):
        raise FishmanError("fm_fp64_peak failed")
    return v.value


def pinned_empty(n):
    """float64 array of n elements in page-locked host memory (fm_host_alloc); kept alive
    for the life of the process."""
    p = C.c_void_p()
    if _lib.lib().fm_host_alloc(8 * int(n), C.byref(p)):
        raise FishmanError("fm_host_alloc failed")
    return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_double)), shape=(int(n),))


class Mesh:
    def __init__(self, num_point=None, lon=None, lat=None, x=None, y=None, z=None, rank=0, size=1):
        """delaunay_voronoi_init: either lon/lat or x/y/z (DV:204-220)."""
        L = _lib.lib()
        self._h = C.c_void_p()
        if x is not None:
            cols = [np.ascontiguousarray(v, dtype=np.float64) for v in (x, y, z)]
            n = len(cols[0]) if num_point is None else int(num_point)
            if size == 1:
                rc = L.fm_create(n, _d(cols[0]), _d(cols[1]), _d(cols[2]), C.byref(self._h))
            else:
                rc = L.fm_create_dist(n, _d(cols[0]), _d(cols[1]), _d(cols[2]), rank, size, C.byref(self._h))
        elif lon is not None:
            cols = [np.ascontiguousarray(v, dtype=np.float64) for v in (lon, lat)]
            n = len(cols[0]) if num_point is None else int(num_point)
            rc = L.fm_create_lonlat(n, _d(cols[0]), _d(cols[1]), C.byref(self._h))
        else:
            raise FishmanError("delaunay_voronoi_init needs lon/lat or x/y/z")
        if rc:
            raise FishmanError(L.fm_last_error(None).decode())
        self.n = n
        self.n_tri = 2 * n - 4
        self.n_edge = 3 * n - 6
        self.rank, self.size = rank, size

    @classmethod
    def from_xyz(cls, xyz, **kw):
        xyz = np.asarray(xyz, dtype=np.float64)
        return cls(x=xyz[:, 0], y=xyz[:, 1], z=xyz[:, 2], **kw)

    def close(self):
        if getattr(self, "_h", None):
            _lib.lib().fm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc):
        if rc:
            raise FishmanError(_lib.lib().fm_last_error(self._h).decode())

    # ---- the reference's call sequence
    def delaunay_triangulation(self):
        self._chk(_lib.lib().fm_triangulate(self._h))

    def retriangulate(self, warm=True):
        """Triangulation of the current (moved) generators; warm = repair by flips."""
        self._chk(_lib.lib().fm_retriangulate(self._h, 1 if warm else 0))

    def voronoi_diagram(self, all=False):
        self._chk(_lib.lib().fm_voronoi(self._h, 1 if all else 0))

    def delaunay_voronoi_output(self, output, tag=None):
        """Calls output(DVT, DT, DE, VVT, VC, VE, tag) with flat arrays in the
        roles of the reference's six containers (output_interface, DV:169-181)."""
        tri, adj = self.triangles()
        xyz, lonlat = self.circumcentres()
        off, rt, rn = self.rings()
        g = self.generators()
        try:
            de, ve_vvt, ve_vc = self.edges()
        except FishmanError:
            de = ve_vvt = ve_vc = None
        return output(dict(xyz=g), dict(tri=tri, adj=adj), dict(dvt=de), dict(xyz=xyz, lonlat=lonlat),
                      dict(off=off, vvt=rt, adj=rn), dict(vvt=ve_vvt, vc=ve_vc), tag)

    # ---- accessors (get_DT / get_VVT / get_VC / get_DE / get_VE)
    def triangles(self):
        tri = np.empty((self.n_tri, 3), dtype=np.int32)
        adj = np.empty((self.n_tri, 3), dtype=np.int32)
        self._chk(_lib.lib().fm_get_triangles(self._h, _i(tri), _i(adj)))
        return tri, adj

    def circumcentres(self):
        xyz = np.empty((self.n_tri, 3))
        lonlat = np.empty((self.n_tri, 2))
        self._chk(_lib.lib().fm_get_circumcentres(self._h, _d(xyz), _d(lonlat)))
        return xyz, lonlat

    def rings(self):
        off = np.empty(self.n + 1, dtype=np.int32)
        rt = np.empty(3 * self.n_tri, dtype=np.int32)
        rn = np.empty(3 * self.n_tri, dtype=np.int32)
        self._chk(_lib.lib().fm_get_rings(self._h, _i(off), _i(rt), _i(rn)))
        return off, rt, rn

    def edges(self):
        de = np.empty((self.n_edge, 2), dtype=np.int32)
        vv = np.empty((self.n_edge, 2), dtype=np.int32)
        vc = np.empty((self.n_edge, 2), dtype=np.int32)
        self._chk(_lib.lib().fm_get_edges(self._h, _i(de), _i(vv), _i(vc)))
        return de, vv, vc

    def mpas_device_only(self, strict=False):
        """fm_mpas without bringing anything back to the host (timing); returns maxEdges."""
        me = C.c_int32(0)
        self._chk(_lib.lib().fm_mpas(self._h, 1 if strict else 0, C.byref(me)))
        return me.value

    def mpas(self, strict=True):
        """mpas_mesh_adapt + calc_weightsOnEdge (src/mpas_mesh_mod.F90:181-345, :442-561) on the
        device, after voronoi_diagram(all=True).  Returns a dict of arrays named like the MPAS
        variables (mpas_mesh_type, MM:25-73); ids 1-based, per-cell rows padded with 0 to
        maxEdges = the largest degree found."""
        L = _lib.lib()
        me = C.c_int32(0)
        self._chk(L.fm_mpas(self._h, 1 if strict else 0, C.byref(me)))
        ME, n, nT, nE = me.value, self.n, self.n_tri, self.n_edge
        i32, f64 = np.int32, np.float64
        o = dict(maxEdges=ME, nCells=n, nEdges=nE, nVertices=nT,
                 xyzEdge=np.empty((nE, 3), f64), lonlatEdge=np.empty((nE, 2), f64),
                 dvEdge=np.empty(nE, f64), dcEdge=np.empty(nE, f64), dv1Edge=np.empty(nE, f64),
                 dv2Edge=np.empty(nE, f64), angleEdge=np.empty(nE, f64), areaCell=np.empty(n, f64),
                 cellsOnVertex=np.empty((nT, 3), i32), edgesOnVertex=np.empty((nT, 3), i32),
                 areaTriangle=np.empty(nT, f64), kiteAreasOnVertex=np.empty((nT, 3), f64),
                 nEdgesOnEdge=np.empty(nE, i32), edgesOnEdge=np.empty((nE, 2 * ME), i32),
                 weightsOnEdge=np.empty((nE, 2 * ME), f64))
        self._chk(L.fm_get_mpas_edges(self._h, _d(o["xyzEdge"]), _d(o["lonlatEdge"]), _d(o["dvEdge"]), _d(o["dcEdge"]),
                                      _d(o["dv1Edge"]), _d(o["dv2Edge"]), _d(o["angleEdge"])))
        ring_edge = np.empty(3 * nT, i32)
        kite_ring = np.empty(3 * nT, f64)
        self._chk(L.fm_get_mpas_cells(self._h, _i(ring_edge), _d(o["areaCell"]), _d(kite_ring)))
        self._chk(L.fm_get_mpas_vertices(self._h, _i(o["cellsOnVertex"]), _i(o["edgesOnVertex"]),
                                         _d(o["areaTriangle"]), _d(o["kiteAreasOnVertex"])))
        self._chk(L.fm_get_mpas_edges_on_edge(self._h, _i(o["nEdgesOnEdge"]), _i(o["edgesOnEdge"]),
                                              _d(o["weightsOnEdge"])))
        de, vv, vc = self.edges()
        o["cellsOnEdge"], o["verticesOnEdge"] = vc, vv
        off, rt, rn = self.rings()
        deg = np.diff(off)
        o["nEdgesOnCell"] = deg.astype(i32)
        cell = np.repeat(np.arange(n), deg)
        pos = np.arange(len(rt)) - off[:-1][cell]
        for name, src in (("verticesOnCell", rt), ("cellsOnCell", rn), ("edgesOnCell", ring_edge)):
            a = np.zeros((n, ME), i32)
            a[cell, pos] = src
            o[name] = a
        o["kiteAreasOnCell"] = (off, kite_ring)
        return o

    def generators(self):
        g = np.empty((3, self.n))
        self._chk(_lib.lib().fm_get_generators(self._h, _d(g[0]), _d(g[1]), _d(g[2])))
        return np.ascontiguousarray(g.T)

    def set_generators(self, xyz):
        g = np.ascontiguousarray(np.asarray(xyz, dtype=np.float64).T)
        self._chk(_lib.lib().fm_set_generators(self._h, _d(g[0]), _d(g[1]), _d(g[2])))

    def set_generators_soa(self, x, y, z):
        """Three contiguous float64 arrays (e.g. pinned_empty): no host-side repacking."""
        self._chk(_lib.lib().fm_set_generators(self._h, _d(x), _d(y), _d(z)))

    def get_generators_soa(self, x, y, z):
        self._chk(_lib.lib().fm_get_generators(self._h, _d(x), _d(y), _d(z)))

    # ---- one rank's block (sharded meshes): the host side holds only the cells the rank owns
    def block_count(self):
        """cells of this rank's block that exist (the last rank's block may be padded)"""
        _, _, beg, cnt = self.generator_block()
        return int(max(0, min(cnt, self.n - beg)))

    def block_ids(self):
        """1-based input-order ids of the block's cells, in block order"""
        ids = np.empty(self.block_count(), dtype=np.int32)
        self._chk(_lib.lib().fm_block_ids(self._h, _i(ids)))
        return ids

    def get_generators_block_soa(self, x, y, z):
        self._chk(_lib.lib().fm_get_generators_block(self._h, _d(x), _d(y), _d(z)))

    def set_generators_block_soa(self, x, y, z):
        """Upload this rank's block (block order); it is stored into every rank's array over NVLink.
        Every rank of the mesh must call it in step."""
        self._chk(_lib.lib().fm_set_generators_block(self._h, _d(x), _d(y), _d(z)))

    def quality(self, rho=None):
        """Mesh-quality statistics of the current generators (fm_quality): dict."""
        rho = rho or density_uniform()
        q = _lib.QualityStats()
        self._chk(_lib.lib().fm_quality(self._h, C.byref(rho), C.byref(q)))
        return q.asdict()

    def star_stats(self):
        """fm_star_stats: dict about the cold builds by stars (csrc/star_kernels.cuh)."""
        import numpy as np
        out = np.zeros(10, np.int64)
        self._chk(_lib.lib().fm_star_stats(self._h, out.ctypes.data_as(C.POINTER(C.c_int64))))
        names = ("builds", "fallbacks", "overflowed", "exact", "failed", "broken", "triangles", "degree_sum", "check_flips",
                 "check_turned")
        return dict(zip(names, (int(v) for v in out)))

    def quality_static(self):
        """fm_quality_static: the statistics of quality() under the uniform density, computed without
        any device allocation (scratch in a static device array of the library; at most
        4,194,304 generators; the triangulation must be current)."""
        q = _lib.QualityStats()
        self._chk(_lib.lib().fm_quality_static(self._h, C.byref(q)))
        return q.asdict()

    # ---- Lloyd
    def lloyd(self, rho=None, max_iter=1, tol=0.0, warm=True):
        rho = rho or density_uniform()
        it = C.c_int32(0)
        md = C.c_double(0.0)
        self._chk(_lib.lib().fm_lloyd(self._h, C.byref(rho), int(max_iter), float(tol), 1 if warm else 0,
                                      C.byref(it), C.byref(md)))
        return it.value, md.value

    def lloyd_local(self, rho=None, warm=True):
        rho = rho or density_uniform()
        md = C.c_double(0.0)
        self._chk(_lib.lib().fm_lloyd_local(self._h, C.byref(rho), 1 if warm else 0, C.byref(md)))
        return md.value

    def generator_block(self):
        """(device pointer, n_padded, begin, count) of the 32-byte generator records."""
        p = C.c_void_p()
        npad, beg, cnt = C.c_int64(), C.c_int64(), C.c_int64()
        self._chk(_lib.lib().fm_generator_block(self._h, C.byref(p), C.byref(npad), C.byref(beg), C.byref(cnt)))
        return p.value, npad.value, beg.value, cnt.value

    def peer_connect(self, dist):
        """Maps the generator arrays of all ranks into each other (fm_peer_handle /
        fm_peer_attach); `dist` is an initialised torch.distributed, used only to pass the
        80-byte handle blobs around.  Afterwards lloyd() runs on the sharded mesh."""
        buf = C.create_string_buffer(80)
        self._chk(_lib.lib().fm_peer_handle(self._h, buf))
        got = [None] * self.size
        dist.all_gather_object(got, buf.raw)
        self._chk(_lib.lib().fm_peer_attach(self._h, b"".join(got)))
        dist.barrier()

    # ---- measurement
    def timer_begin(self):
        self._chk(_lib.lib().fm_timer_begin(self._h))

    def timer_end(self):
        ms = C.c_double(0.0)
        self._chk(_lib.lib().fm_timer_end(self._h, C.byref(ms)))
        return ms.value

    def stats(self):
        s = np.zeros(8, dtype=np.int64)
        _lib.lib().fm_stats(self._h, s.ctypes.data_as(C.POINTER(C.c_int64)))
        keys = ("rounds", "subrounds", "flips", "max_active", "launches", "cold", "warm", "_")
        return dict(zip(keys, (int(v) for v in s)))

    def stats_reset(self):
        _lib.lib().fm_stats_reset(self._h)

    def profile(self, on):
        _lib.lib().fm_profile_enable(self._h, 1 if on else 0)

    def profile_rows(self):
        cap = 64
        names = (C.c_char * 32 * cap)()
        la = np.zeros(cap, dtype=np.int64)
        th = np.zeros(cap, dtype=np.int64)
        ms = np.zeros(cap)
        rows = C.c_int32(0)
        _lib.lib().fm_profile_get(self._h, cap, C.cast(names, C.c_void_p),
                                  la.ctypes.data_as(C.POINTER(C.c_int64)), _d(ms),
                                  th.ctypes.data_as(C.POINTER(C.c_int64)), C.byref(rows))
        return [dict(name=names[k].value.decode(), launches=int(la[k]), ms=float(ms[k]), threads=int(th[k]))
                for k in range(rows.value)]


# ---- module-level mirror with one default mesh, like the Fortran module state
_default = None


def delaunay_voronoi_init(num_point=None, lon=None, lat=None, x=None, y=None, z=None):
    global _default
    if _default is not None:
        _default.close()
    _default = Mesh(num_point, lon=lon, lat=lat, x=x, y=y, z=z)
    return _default


def delaunay_triangulation():
    _default.delaunay_triangulation()


def voronoi_diagram(all=False):
    _default.voronoi_diagram(all=all)


def delaunay_voronoi_output(output, tag=None):
    return _default.delaunay_voronoi_output(output, tag)


def canonical_triangles(tri):
    """Lexicographically sorted rows of sorted vertex triples."""
    t = np.sort(np.asarray(tri, dtype=np.int64), axis=1)
    return t[np.lexsort((t[:, 2], t[:, 1], t[:, 0]))]
