"""fishman_b200: B200 (sm_100a) implementation of fishman's SCVT inner loop --
spherical Delaunay triangulation, Voronoi diagram and Lloyd update -- behind the
C ABI of include/fishman_b200.h.  See DESIGN.md."""
from . import points  # noqa: F401
from .delaunay_voronoi import (FishmanError, Mesh, canonical_triangles, delaunay_triangulation,  # noqa: F401
                               delaunay_voronoi_init, delaunay_voronoi_output, density_uniform,
                               density_eval, density_x4, fp64_peak_tflops, pinned_empty, voronoi_diagram)
