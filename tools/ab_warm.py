"""A/B of the warm Lloyd iteration between library builds: FM_LIB=<path> python tools/ab_warm.py [n] [steps]"""
import sys, os, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fishman_b200 as fb
from fishman_b200 import points

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2621442
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 100
dens = sys.argv[3] if len(sys.argv) > 3 else "uniform"
rho = fb.density_uniform() if dens == "uniform" else fb.density_x4()
m = fb.Mesh.from_xyz(points.fibonacci(n, 0, 0.2))
m.lloyd(rho, max_iter=20, warm=True)
ts = []
for _ in range(3):
    m.timer_begin(); m.lloyd(rho, max_iter=steps, warm=True); ts.append(m.timer_end() / steps)
m.profile(True); m.stats_reset()
m.lloyd(rho, max_iter=steps, warm=True)
rows = sorted(m.profile_rows(), key=lambda r: -r["ms"])
m.profile(False)
print(json.dumps({"lib": os.environ.get("FM_LIB", "in-tree"), "n": n, "density": dens, "ms_per_step": ts, "it_s": 1e3 / min(ts),
                  "kernels": [(r["name"], round(r["ms"] / max(r["launches"], 1), 5)) for r in rows[:8]]}))
