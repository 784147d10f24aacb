"""ctypes binding of the C ABI (include/fishman_b200.h).  The library is the
product: there is no Python or CPU implementation behind it, and loading or
using it without a CUDA device raises."""
import ctypes as C
import os

from . import build as _build

_lib = None

SYMBOLS = [
    "fm_create", "fm_create_lonlat", "fm_create_dist", "fm_destroy", "fm_last_error", "fm_triangulate", "fm_retriangulate",
    "fm_counts", "fm_get_triangles", "fm_voronoi", "fm_get_circumcentres", "fm_get_rings",
    "fm_get_edges", "fm_mpas", "fm_get_mpas_edges", "fm_get_mpas_cells", "fm_get_mpas_vertices",
    "fm_get_mpas_edges_on_edge", "fm_lloyd", "fm_lloyd_local", "fm_generator_block", "fm_peer_handle", "fm_peer_attach", "fm_density_eval", "fm_get_generators",
    "fm_set_generators", "fm_host_alloc", "fm_host_free", "fm_timer_begin", "fm_timer_end", "fm_stats", "fm_stats_reset",
    "fm_profile_enable", "fm_profile_get", "fm_device_info",
    "fm_fp64_peak", "fm_block_ids", "fm_get_generators_block", "fm_set_generators_block", "fm_quality",
    "fm_quality_static", "fm_star_stats",
]


class QualityStats(C.Structure):
    """fm_quality_stats of include/fishman_b200.h"""
    _fields_ = [("degree_hist", C.c_int64 * 16)] + [(k, C.c_double) for k in (
        "area_cell_min", "area_cell_max", "area_cell_sum", "area_tri_min", "area_tri_max", "area_tri_sum",
        "angle_min", "angle_max", "max_disp")]

    def asdict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_[1:]}
        d["degree_hist"] = [int(v) for v in self.degree_hist]
        return d


class Density(C.Structure):
    _fields_ = [("kind", C.c_int32), ("cx", C.c_double), ("cy", C.c_double), ("cz", C.c_double),
                ("beta", C.c_double), ("alpha", C.c_double), ("gamma", C.c_double)]


def lib():
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("FM_LIB", _build.LIB)  # FM_LIB: an alternative build of the same library (tuning runs)
    if not os.path.exists(path):
        raise RuntimeError(f"{path} is missing: run `python -m fishman_b200.build` (needs nvcc); "
                           "fishman_b200 has no CPU fallback")
    L = C.CDLL(path)
    vp, dp, ip = C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int32)
    i32, i64p = C.c_int32, C.POINTER(C.c_int64)
    L.fm_create.argtypes = [i32, dp, dp, dp, C.POINTER(vp)]
    L.fm_create_lonlat.argtypes = [i32, dp, dp, C.POINTER(vp)]
    L.fm_create_dist.argtypes = [i32, dp, dp, dp, i32, i32, C.POINTER(vp)]
    L.fm_destroy.argtypes = [vp]
    L.fm_destroy.restype = None
    L.fm_last_error.argtypes = [vp]
    L.fm_last_error.restype = C.c_char_p
    L.fm_triangulate.argtypes = [vp]
    L.fm_retriangulate.argtypes = [vp, i32]
    L.fm_counts.argtypes = [vp, ip, ip]
    L.fm_get_triangles.argtypes = [vp, ip, ip]
    L.fm_voronoi.argtypes = [vp, i32]
    L.fm_get_circumcentres.argtypes = [vp, dp, dp]
    L.fm_get_rings.argtypes = [vp, ip, ip, ip]
    L.fm_get_edges.argtypes = [vp, ip, ip, ip]
    L.fm_mpas.argtypes = [vp, i32, ip]
    L.fm_get_mpas_edges.argtypes = [vp, dp, dp, dp, dp, dp, dp, dp]
    L.fm_get_mpas_cells.argtypes = [vp, ip, dp, dp]
    L.fm_get_mpas_vertices.argtypes = [vp, ip, ip, dp, dp]
    L.fm_get_mpas_edges_on_edge.argtypes = [vp, ip, ip, dp]
    L.fm_lloyd.argtypes = [vp, C.POINTER(Density), i32, C.c_double, i32, ip, dp]
    L.fm_lloyd_local.argtypes = [vp, C.POINTER(Density), i32, dp]
    L.fm_generator_block.argtypes = [vp, C.POINTER(vp), i64p, i64p, i64p]
    L.fm_peer_handle.argtypes = [vp, C.c_char_p]
    L.fm_peer_attach.argtypes = [vp, C.c_char_p]
    L.fm_density_eval.argtypes = [C.POINTER(Density), C.c_int64, dp, dp, dp, dp]
    L.fm_get_generators.argtypes = [vp, dp, dp, dp]
    L.fm_set_generators.argtypes = [vp, dp, dp, dp]
    L.fm_host_alloc.argtypes = [C.c_size_t, C.POINTER(vp)]
    L.fm_host_free.argtypes = [vp]
    L.fm_timer_begin.argtypes = [vp]
    L.fm_timer_end.argtypes = [vp, dp]
    L.fm_stats.argtypes = [vp, i64p]
    if hasattr(L, 'fm_star_stats'):
        L.fm_star_stats.argtypes = [vp, i64p]
    L.fm_stats_reset.argtypes = [vp]
    L.fm_profile_enable.argtypes = [vp, i32]
    L.fm_profile_get.argtypes = [vp, i32, C.c_void_p, i64p, dp, i64p, ip]
    L.fm_device_info.argtypes = [C.c_char_p, i32, ip]
    if hasattr(L, 'fm_fp64_peak'):   # absent from older builds loaded through FM_LIB (A/B runs)
        L.fm_fp64_peak.argtypes = [C.c_double, dp]
    if hasattr(L, 'fm_block_ids'):   # absent from older builds loaded through FM_LIB (A/B runs)
        L.fm_block_ids.argtypes = [vp, ip]
    if hasattr(L, 'fm_get_generators_block'):   # absent from older builds loaded through FM_LIB (A/B runs)
        L.fm_get_generators_block.argtypes = [vp, dp, dp, dp]
    if hasattr(L, 'fm_set_generators_block'):   # absent from older builds loaded through FM_LIB (A/B runs)
        L.fm_set_generators_block.argtypes = [vp, dp, dp, dp]
    if hasattr(L, 'fm_quality'):   # absent from older builds loaded through FM_LIB (A/B runs)
        L.fm_quality.argtypes = [vp, C.POINTER(Density), C.POINTER(QualityStats)]
    if hasattr(L, 'fm_quality_static'):
        L.fm_quality_static.argtypes = [vp, C.POINTER(QualityStats)]
    _lib = L
    return L
