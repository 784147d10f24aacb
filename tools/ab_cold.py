"""A/B of the cold build between library builds: FM_LIB=<path> python tools/ab_cold.py [n] [kind]
Prints the median time of 5 rebuilds and the per-kernel table (works with any build of the ABI)."""
import sys, os, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fishman_b200 as fb
from fishman_b200 import points

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2621442
kind = sys.argv[2] if len(sys.argv) > 2 else "fib"
p = points.uniform(n, 0) if kind == "uniform" else points.fibonacci(n, 0, float(kind[4:]) if kind.startswith("fibj") else 0.2)
m = fb.Mesh.from_xyz(p)
m.delaunay_triangulation()
if kind.startswith("x4"):   # a graded mesh: `x4<iters>` Lloyd iterations under the 4:1 refinement density first
    m.lloyd(fb.density_x4(), max_iter=int(kind[2:] or 200), warm=True)
    m.retriangulate(warm=True)
ts = []
for _ in range(5):
    m.timer_begin(); m.retriangulate(warm=False); ts.append(m.timer_end())
m.profile(True); m.stats_reset()
m.retriangulate(warm=False)
rows = sorted(m.profile_rows(), key=lambda r: -r["ms"])
m.profile(False)
print(json.dumps({"lib": os.environ.get("FM_LIB", "in-tree"), "n": n, "kind": kind, "ms": float(np.median(ts)), "all": ts,
                  "Mpts_s": n / np.median(ts) / 1e3,
                  "kernels": [(r["name"], r["launches"], round(r["ms"], 3)) for r in rows[:12]],
                  "star": m.star_stats() if hasattr(m, "star_stats") else None}))
