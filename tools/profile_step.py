"""Short driver for ncu: one cold triangulation + a few warm Lloyd iterations at n
generators (default 2,621,442), nothing else.  Usage: python tools/profile_step.py [n] [warm_iters] [cold_iters] [uniform|x4]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fishman_b200 as fb  # noqa: E402
from fishman_b200 import points  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2621442
warm = int(sys.argv[2]) if len(sys.argv) > 2 else 3
cold = int(sys.argv[3]) if len(sys.argv) > 3 else 0
rho = fb.density_x4() if len(sys.argv) > 4 and sys.argv[4] == "x4" else fb.density_uniform()
m = fb.Mesh.from_xyz(points.fibonacci(n, 0))
m.lloyd(rho, max_iter=1, warm=True)          # cold build + first centroid step
m.lloyd(rho, max_iter=warm, warm=True)
if cold:
    m.lloyd(rho, max_iter=cold, warm=False)
print(m.stats())
m.close()
