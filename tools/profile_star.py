import sys, os
sys.path.insert(0, "/root/repo")
import fishman_b200 as fb
from fishman_b200 import points
n = int(sys.argv[1]); kind = sys.argv[2]
p = points.uniform(n, 0) if kind == "uniform" else points.fibonacci(n, 0, 0.2)
m = fb.Mesh.from_xyz(p)
m.delaunay_triangulation()
m.retriangulate(warm=False)
print(m.star_stats())
