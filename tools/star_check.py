"""Cold build by stars against the insertion build on the same point sets: run once with FM_COLD_BUILD=flip and
once without, compare the printed hashes of the canonical triangle lists (they must be equal: both are THE
Delaunay triangulation of the SoS-perturbed points).   python tools/star_check.py [n ...]"""
import hashlib
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fishman_b200 as fb
from fishman_b200 import points
from tests import _check

sizes = [int(a) for a in sys.argv[1:]] or [20000, 300000]
out = []
for n in sizes:
    sets = [("fib", points.fibonacci(n, 0, 0.2)), ("uniform", points.uniform(n, 1)), ("fib_nojitter", points.fibonacci(n, 0, 0.0))]
    if n <= 200000:
        lv = int(round(np.log(max(n - 2, 10) / 10) / np.log(4)))
        sets.append((f"ico{lv}", points.ico(lv)))
        nlat = int(np.sqrt(n / 2))
        sets.append(("latlon", points.latlon_grid(nlat, 2 * nlat)))
    for name, p in sets:
        m = fb.Mesh.from_xyz(p)
        m.delaunay_triangulation()
        tri, adj = m.triangles()
        if len(p) <= 400000:
            _check.check_triangulation(p, tri, adj)
        h = hashlib.sha256(np.ascontiguousarray(_check.canonical(tri)).tobytes()).hexdigest()[:16]
        st = m.star_stats() if hasattr(m, "star_stats") else {}
        m.lloyd(max_iter=3, warm=True)   # warm repairs on top of the build
        q = m.quality()
        out.append({"set": name, "n": len(p), "sha": h, "star": st, "cell_sum_err": abs(q["area_cell_sum"] - 4 * np.pi)})
        m.close()
print(json.dumps({"mode": os.environ.get("FM_COLD_BUILD", "star"), "results": out}))
