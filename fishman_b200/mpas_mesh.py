"""Host-side mirror of the reference's output plugin and of its input contract, without NetCDF
(neither NetCDF nor a Fortran compiler exists in this image).

  mpas_mesh_output(mesh, path)   the 37 variables, 8 dimensions and 4 attributes the reference
                                 writes to mpas_mesh.<nCells>.nc
                                 (/root/reference/src/mpas_mesh_mod.F90:75-179), as one .npz:
                                 same names, same shapes (Fortran (maxEdges, nCells) = C
                                 (nCells, maxEdges)), arithmetic from the device (fm_mpas).
  write_grids / read_grids       the `grids` file of gen_mesh (/root/reference/src/gen_mesh.F90:
                                 15-23): dimension location_nv, variable vtx_p(location_nv, 3) =
                                 unit-sphere Cartesian triples, as .npz.

Deviation from the reference: maxEdges is the largest cell degree found (the reference fixes 6
and overflows on anything larger); padded slots hold 0."""
import numpy as np

VARIABLES = (  # name, dims -- in the order of the fiona_add_var calls, MM:102-138
    ("latCell", ("nCells",)), ("lonCell", ("nCells",)), ("xCell", ("nCells",)), ("yCell", ("nCells",)),
    ("zCell", ("nCells",)), ("indexToCellID", ("nCells",)),
    ("latEdge", ("nEdges",)), ("lonEdge", ("nEdges",)), ("xEdge", ("nEdges",)), ("yEdge", ("nEdges",)),
    ("zEdge", ("nEdges",)), ("indexToEdgeID", ("nEdges",)),
    ("latVertex", ("nVertices",)), ("lonVertex", ("nVertices",)), ("xVertex", ("nVertices",)),
    ("yVertex", ("nVertices",)), ("zVertex", ("nVertices",)), ("indexToVertexID", ("nVertices",)),
    ("nEdgesOnCell", ("nCells",)), ("nEdgesOnEdge", ("nEdges",)),
    ("cellsOnCell", ("nCells", "maxEdges")), ("cellsOnEdge", ("nEdges", "TWO")),
    ("edgesOnCell", ("nCells", "maxEdges")), ("edgesOnEdge", ("nEdges", "maxEdges2")),
    ("verticesOnCell", ("nCells", "maxEdges")), ("verticesOnEdge", ("nEdges", "TWO")),
    ("edgesOnVertex", ("nVertices", "vertexDegree")), ("cellsOnVertex", ("nVertices", "vertexDegree")),
    ("weightsOnEdge", ("nEdges", "maxEdges2")),
    ("dvEdge", ("nEdges",)), ("dv1Edge", ("nEdges",)), ("dv2Edge", ("nEdges",)), ("dcEdge", ("nEdges",)),
    ("angleEdge", ("nEdges",)), ("areaCell", ("nCells",)), ("areaTriangle", ("nVertices",)),
    ("kiteAreasOnVertex", ("nVertices", "vertexDegree")),
)
ATTRIBUTES = {"on_a_sphere": "YES", "sphere_radius": 1.0, "mesh_id": "demo", "mesh": "1.0"}  # MM:90-93


def _lonlat(xyz):
    """inverse_cartesian_transform, src/sphere_geometry_mod.F90:112-121"""
    lon = np.arctan2(xyz[:, 1], xyz[:, 0])
    lon[lon < 0.0] += 2.0 * np.pi
    return lon, np.arcsin(np.clip(xyz[:, 2], -1.0, 1.0))


def mpas_arrays(mesh, strict=True):
    """The reference's mpas_mesh_type (MM:25-73) filled from a fishman_b200.Mesh on which
    delaunay_triangulation() has run; runs voronoi_diagram(all=True) and the device adapter."""
    mesh.voronoi_diagram(all=True)
    m = mesh.mpas(strict=strict)
    g = mesh.generators()
    cc, cll = mesh.circumcentres()
    out = {"nCells": m["nCells"], "nEdges": m["nEdges"], "nVertices": m["nVertices"], "vertexDegree": 3,
           "maxEdges": m["maxEdges"], "maxEdges2": 2 * m["maxEdges"], "TWO": 2}
    lon, lat = _lonlat(g)
    out.update(lonCell=lon, latCell=lat, xCell=g[:, 0].copy(), yCell=g[:, 1].copy(), zCell=g[:, 2].copy(),
               indexToCellID=np.arange(1, m["nCells"] + 1, dtype=np.int32))
    out.update(lonEdge=m["lonlatEdge"][:, 0].copy(), latEdge=m["lonlatEdge"][:, 1].copy(),
               xEdge=m["xyzEdge"][:, 0].copy(), yEdge=m["xyzEdge"][:, 1].copy(), zEdge=m["xyzEdge"][:, 2].copy(),
               indexToEdgeID=np.arange(1, m["nEdges"] + 1, dtype=np.int32))
    out.update(lonVertex=cll[:, 0].copy(), latVertex=cll[:, 1].copy(), xVertex=cc[:, 0].copy(),
               yVertex=cc[:, 1].copy(), zVertex=cc[:, 2].copy(),
               indexToVertexID=np.arange(1, m["nVertices"] + 1, dtype=np.int32))
    for k in ("nEdgesOnCell", "nEdgesOnEdge", "cellsOnCell", "cellsOnEdge", "edgesOnCell", "edgesOnEdge",
              "verticesOnCell", "verticesOnEdge", "edgesOnVertex", "cellsOnVertex", "weightsOnEdge", "dvEdge",
              "dv1Edge", "dv2Edge", "dcEdge", "angleEdge", "areaCell", "areaTriangle", "kiteAreasOnVertex"):
        out[k] = m[k]
    return out


def mpas_mesh_output(mesh, path=None, strict=True):
    """Writes <path> (default mpas_mesh.<nCells>.npz, as MM:89) and returns the path."""
    a = mpas_arrays(mesh, strict=strict)
    path = path or f"mpas_mesh.{a['nCells']}.npz"
    payload = {name: a[name] for name, _ in VARIABLES}
    for d in ("nCells", "nEdges", "nVertices", "vertexDegree", "maxEdges", "maxEdges2", "TWO"):
        payload["dim_" + d] = np.int64(a[d])
    for k, v in ATTRIBUTES.items():
        payload["att_" + k] = np.array(v)
    np.savez(path, **payload)
    return path


def write_grids(path, xyz):
    """vtx_p(location_nv, 3), src/gen_mesh.F90:16-22"""
    xyz = np.ascontiguousarray(xyz, dtype=np.float64)
    np.savez(path, vtx_p=xyz, dim_location_nv=np.int64(len(xyz)))


def read_grids(path):
    with np.load(path) as f:
        x = np.array(f["vtx_p"], dtype=np.float64)
    if x.ndim != 2 or x.shape[1] != 3:
        raise ValueError("grids file: vtx_p must be (location_nv, 3)")
    return x


def gen_mesh(grids_path, out_path=None, strict=True):
    """program gen_mesh (src/gen_mesh.F90:13-30): read points -> DT -> VD(all) -> MPAS output."""
    from .delaunay_voronoi import Mesh
    x = read_grids(grids_path)
    mesh = Mesh(len(x), x=x[:, 0], y=x[:, 1], z=x[:, 2])   # non-contiguous column sections, as gen_mesh.F90:25
    try:
        mesh.delaunay_triangulation()
        return mpas_mesh_output(mesh, out_path, strict=strict)
    finally:
        mesh.close()
