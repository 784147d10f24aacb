"""Short driver for ncu: voronoi_diagram(all) + the MPAS adapter on ICO(k) (default k = 7: 163,842 cells).
Usage: python tools/profile_mpas.py [k]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fishman_b200 as fb  # noqa: E402
from fishman_b200 import points  # noqa: E402

k = int(sys.argv[1]) if len(sys.argv) > 1 else 7
m = fb.Mesh.from_xyz(points.ico(k))
m.delaunay_triangulation()
for _ in range(2):
    m.voronoi_diagram(all=True)
    print(m.mpas_device_only())
m.close()
