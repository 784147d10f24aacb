#!/usr/bin/env python
"""bench.py -- Lloyd iterations/s at 2,621,442 generators (BASELINE.json metric) on N B200s.

A step = one Lloyd iteration of the SCVT inner loop on the GPU: Delaunay triangulation of
the current generators, circumcentres, density-weighted centroid quadrature, generator
update + max-displacement reduction (+ the generator exchange when N > 1).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--mode warm|cold] [--n GENERATORS]
  python bench.py --impl reference ...      # CPU restatement of the reference, host cores
  python bench.py --workload mpas ...       # the MPAS adapter stage

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every field.
"""
import argparse
import hashlib
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_C4 = 2621442          # BASELINE.json configs[3]
METRIC = "lloyd_iterations_per_sec_at_2.6M_generators"
UNIT = "it/s"
CPU_BUDGET_S = 200.0    # --impl reference: wall-clock budget for the timed steps at the real size

# Algorithmic bytes per work item (SURVEY.md 8(d) / BASELINE.md 3: SoA fp64 coordinates, int32
# indices) -- the numerator of roofline.achieved.  LAYOUT_BYTES is what the 32-byte padded
# records of this implementation move per item; the difference shows up in
# roofline.traffic, not in the fraction.
ALG_BYTES = {
    "sweep": 48,             # per triangle: 12 idx + 12 amortised vertex data (24 N/T) + 24 circumcentre write
    "circumcentre": 48,
    "centroid_update": 124,  # per cell: 4 offset + 24 ring ids + 48 amortised circumcentres + 24 generator + 24 write
    "lloyd_iteration": 292,  # per generator: 96 (circumcentres) + 124 (centroid) + 72 (DT re-validation share)
    "delaunay": 72,          # per point, one-shot DT: xyz 24 + 2 triangles x 12 + 2 adjacency rows x 12
}
LAYOUT_BYTES = {"sweep": 80, "circumcentre": 64, "centroid_update": 152}


def csrc_sha16():
    """Identity of the kernels a profile was taken on: .git does not travel to the GPU box."""
    h = hashlib.sha256()
    d = os.path.join(ROOT, "fishman_b200", "csrc")
    for f in sorted(os.listdir(d)):
        with open(os.path.join(d, f), "rb") as fh:
            h.update(f.encode() + b"\0" + fh.read())
    return h.hexdigest()[:16]


def ncu_capture(kernel, n, density, n_gpus):
    """The committed ncu --set full capture (profiles/traffic.json) of `kernel`, or None when it
    was taken on other kernels (csrc hash), another size, density or GPU count."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        if (t.get("csrc_sha16") != csrc_sha16() or t.get("generators") != n or t.get("density") != density
                or t.get("n_gpus", 1) != n_gpus):
            return None
        k = dict(t["kernels"][kernel])
        k["source"] = t["source"]
        return k
    except Exception:
        return None


def sass_counts():
    try:
        with open(os.path.join(ROOT, "profiles", "sass_counts.json")) as f:
            t = json.load(f)
        return t if t.get("csrc_sha16") == csrc_sha16() else None
    except Exception:
        return None


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML from a thread while the timed
    region runs (the ctypes calls into the library release the GIL).  Sub-millisecond period (an NVML
    query itself takes 0.1-1 ms): the timed region of a warm run is only a few ms, far shorter than
    one nvidia-smi -lms tick."""

    def __init__(self, index):
        self.index, self.rows, self.stop_flag, self.th, self.err = index, [], False, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                try:
                    phys = int(vis.split(",")[index])
                except ValueError:
                    phys = index
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(phys)
        except Exception as e:  # noqa: BLE001
            self.nv, self.err = None, repr(e)

    def _run(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                rs = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(
                    nv, "nvmlDeviceGetCurrentClocksEventReasons") else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.rows.append((sm, rs))
            except Exception as e:  # noqa: BLE001
                self.err = repr(e)
                return
            time.sleep(0.0002)

    def start(self):
        if self.nv:
            self.th = threading.Thread(target=self._run, daemon=True)
            self.th.start()

    def stop(self):
        if not self.nv:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [f"nvml unavailable: {self.err}"]}
        self.stop_flag = True
        self.th.join(timeout=2)
        nv = self.nv
        mx = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
        bits = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                "hw_power_brake": 0x80}
        seen = set()
        for _, rs in self.rows:
            for k, b in bits.items():
                if rs & b:
                    seen.add(k)
        sm = [r[0] for r in self.rows]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(mx), "samples": len(sm),
                "reasons": sorted(seen)}


def make_points(n, kind="fib"):
    from fishman_b200 import points
    if kind == "uniform":
        return points.uniform(n, 0)
    return points.fibonacci(n, seed=0, jitter=0.2)


# --------------------------------------------------------------------------- CPU legs
def cpu_oracle(n):
    """The CPU restatement (oracle/) on the bench workload.  The points are randomly permuted
    first: the reference inserts in index order, which is quadratic on a coherent (spiral)
    order -- the permutation is in its favour."""
    from oracle import oracle as O
    p = make_points(n)
    p = p[np.random.default_rng(0).permutation(n)]
    return O, O.Oracle(p)


def cpu_lloyd_seconds(n, steps, warmup=0, budget_s=None):
    """(seconds per Lloyd iteration, iterations timed) of the CPU restatement at n generators --
    the REAL size, no scaling.  With budget_s the number of timed iterations is cut so that the
    timed region stays near the budget (at least one)."""
    O, o = cpu_oracle(n)
    rho = O.density_uniform()
    for _ in range(warmup):
        o.lloyd_step(rho)
    done, t0 = 0, time.perf_counter()
    while done < steps:
        o.lloyd_step(rho)
        done += 1
        el = time.perf_counter() - t0
        if budget_s is not None and el + el / done > budget_s:
            break
    dt = (time.perf_counter() - t0) / done
    o.close()
    return dt, done


def run_reference(args):
    """The reference arm.  The reference is Fortran and neither this image nor the GPU box has a
    Fortran compiler (profiles/r02_probe_fortran_gpubox.txt), so what runs is the C restatement
    of its algorithm (oracle/, kind "port"), single-threaded as the reference is, at the REAL
    generator count.  One iteration takes tens of seconds, so the number of timed steps is
    bounded by a wall-clock budget and reported as `steps` (the requested count is kept in
    `steps_requested`); no warm-up is done or needed (no clocks, caches or JIT to settle)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    try:
        os.sched_setaffinity(0, {sorted(os.sched_getaffinity(0))[0]})
    except Exception:
        pass
    sec, done = cpu_lloyd_seconds(args.n, args.steps, 0, CPU_BUDGET_S)
    val = 1.0 / sec
    sample = (f"{done} Lloyd iteration(s) (cold Delaunay triangulation + Voronoi + centroids + update) of the CPU "
              f"restatement at the full {args.n} generators (FIB + jitter, randomly permuted), {sec:.1f} s each; "
              f"steps cut from {args.steps} to fit {CPU_BUDGET_S:.0f} s; no warm-up")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": done, "steps_requested": args.steps, "warmup": 0, "warmup_requested": args.warmup,
        "ms_per_step": sec * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": 1, "kind": "port", "sample": sample,
                         "host_cores": os.cpu_count()},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def workload_config(args):
    return {"workload": f"C4: {args.n}-generator uniform-density SCVT Lloyd loop (BASELINE.json configs[3]); "
                        "FIB lattice + 0.2h jitter, seed 0",
            "mode": args.mode, "density": args.density, "generators": args.n,
            "parallelism": (f"replicated DT + sharded sweep and centroid x{args.gpus}, generator exchange: "
                            + ("stores into peer memory over NVLink from the Lloyd kernel" if args.exchange == "p2p"
                               else "NCCL all-gather")) if args.gpus > 1 else "single GPU",
            "l2": "working set (~0.6 GB of generators, triangles, adjacency, circumcentres) exceeds the 126 MB L2; no flush"}


# --------------------------------------------------------------------------- GPU leg
def kernel_table(rows):
    rows = sorted(rows, key=lambda r: -r["ms"])
    tot = sum(r["ms"] for r in rows) or 1.0
    return [{"name": r["name"], "launches": r["launches"], "ms": round(r["ms"], 4), "share": round(r["ms"] / tot, 4),
             "avg_ms": round(r["ms"] / max(r["launches"], 1), 5), "items": r["threads"]} for r in rows]


def profiled(m, fn):
    """Runs fn() with a CUDA-event pair around every launch (pairs are resolved where the host waits
    anyway) and returns (ms of the pass, per-kernel rows).  Never a headline number."""
    m.profile(True)
    m.stats_reset()
    m.timer_begin()
    fn()
    ms = m.timer_end()
    rows = m.profile_rows()
    m.profile(False)
    return ms, rows


def hbm_line(name, items, avg_ms, peak):
    """achieved = algorithmic bytes x items of one launch / its mean duration"""
    b = ALG_BYTES.get(name)
    if not b or avg_ms <= 0:
        return None
    ach = b * items / (avg_ms * 1e-3) / 1e9
    return {"kernel": name, "alg_bytes_per_item": b, "layout_bytes_per_item": LAYOUT_BYTES.get(name),
            "items_per_launch": items, "avg_launch_ms": avg_ms, "achieved": ach, "peak": peak, "unit": "GB/s",
            "frac": ach / peak}


def fp64_line(fb, n_cells, n_subtri, avg_ms, density):
    """fp64-pipe roofline of the centroid kernel: DP instructions per sub-triangle counted in the
    SASS of this build (profiles/sass_counts.json, tools/sass_dp_count.py) against the DFMA rate
    measured live on this GPU (fm_fp64_peak: independent DFMA chains on all SMs, >= 50 ms)."""
    sc = sass_counts()
    if not sc or density != "uniform" or avg_ms <= 0:
        return None
    try:
        peak = fb.fp64_peak_tflops()
    except Exception as e:  # noqa: BLE001
        return {"error": repr(e)}
    k = sc["centroid_update_uniform"]
    # the quadrature loop only: the per-cell prologue / epilogue (projection, displacement) is left
    # out because its static SASS count includes paths a cell does not take -- a lower bound
    flop = n_subtri * k["flop_per_subtriangle"]
    dp = n_subtri * k["dp_instr_per_subtriangle"]
    ach = flop / (avg_ms * 1e-3) / 1e12
    return {"kernel": "centroid_update", "achieved_tflops": ach, "peak_tflops": peak, "peak_kind": "measured live (DFMA)",
            "frac": ach / peak,
            "pipe_frac": (2.0 * dp / (avg_ms * 1e-3) / 1e12) / peak,   # share of fp64 issue slots (every DP instruction costs an FMA's slot)
            "dp_instr_per_subtriangle": k["dp_instr_per_subtriangle"], "flop_per_subtriangle": k["flop_per_subtriangle"],
            "mufu_per_subtriangle": k.get("mufu_per_subtriangle"), "source": sc["source"]}


def extra_cold(fb, p, rho, steps, peak):
    """M1 'cold' (SURVEY 8d): every iteration rebuilds the triangulation -- what a loop over the
    single-shot reference API would do."""
    m = fb.Mesh.from_xyz(p)
    m.lloyd(rho, max_iter=2, warm=False)
    clk = ClockSampler(0)
    clk.start()
    m.timer_begin()
    m.lloyd(rho, max_iter=steps, warm=False)
    ms = m.timer_end()
    clocks = clk.stop()
    _, rows = profiled(m, lambda: m.lloyd(rho, max_iter=steps, warm=False))
    st = m.stats()
    star = m.star_stats()
    m.close()
    n = len(p)
    per = ms / steps
    return {"mode": "cold", "value": steps / (ms * 1e-3), "unit": UNIT, "steps": steps, "ms_per_step": per, "clocks": clocks,
            "launches_per_step": st["launches"] / steps, "insertion_rounds_per_build": st["rounds"] / steps,
            "builds_by_stars": star["builds"], "builds_by_insertion": star["fallbacks"],
            "roofline": {"bound": "hbm", "scope": "whole iteration", "alg_bytes_per_item": ALG_BYTES["lloyd_iteration"],
                         "items": n, "achieved": ALG_BYTES["lloyd_iteration"] * n / (per * 1e-3) / 1e9, "peak": peak,
                         "unit": "GB/s", "frac": ALG_BYTES["lloyd_iteration"] * n / (per * 1e-3) / 1e9 / peak},
            "kernels": kernel_table(rows)}


def extra_m2(fb, n, kind, peak, reps=3):
    """M2: one-shot spherical Delaunay triangulation (fm_triangulate), points resident: Mpts/s."""
    p = make_points(n, kind)
    m = fb.Mesh.from_xyz(p)
    del p
    m.delaunay_triangulation()
    ts = []
    clk = ClockSampler(0)
    clk.start()
    for _ in range(reps):
        m.timer_begin()
        m.retriangulate(warm=False)
        ts.append(m.timer_end())
    clocks = clk.stop()
    _, rows = profiled(m, lambda: m.retriangulate(warm=False))
    st = m.stats()
    star = m.star_stats()
    m.close()
    ms = float(np.median(ts))
    ach = ALG_BYTES["delaunay"] * n / (ms * 1e-3) / 1e9
    return {"points": n, "input": "U(N,0)" if kind == "uniform" else "FIB(N)+0.2h jitter", "value": n / (ms * 1e-3) / 1e6,
            "unit": "Mpts/s", "ms": ms, "ms_all": [round(t, 3) for t in ts], "clocks": clocks, "launches": st["launches"],
            "insertion_rounds": st["rounds"], "flips": st["flips"],
            "build": "stars (csrc/star_kernels.cuh)" if star["builds"] > 0 else
                     "insertion levels + flips (the stars' sampling pass refused the set)" if star["fallbacks"] > 0 else
                     "insertion levels + flips",
            "roofline": {"bound": "hbm", "scope": "whole build", "alg_bytes_per_item": ALG_BYTES["delaunay"], "items": n,
                         "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak},
            "kernels": kernel_table(rows)[:10]}


def run_gpu(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    import torch
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import fishman_b200 as fb

    n = args.n
    p = make_points(n)
    warm = args.mode == "warm"
    rho = fb.density_uniform() if args.density == "uniform" else fb.density_x4()
    m = fb.Mesh.from_xyz(p, rank=rank, size=world)
    gather = None
    if world > 1 and args.exchange == "p2p" and torch.cuda.device_count() < world:
        args.exchange = "nccl"   # ranks sharing a GPU: the peer exchange needs one GPU per rank
    if world > 1 and args.exchange == "p2p":
        m.peer_connect(dist)   # the Lloyd kernel stores into every rank's generator array over NVLink
    elif world > 1:
        from fishman_b200 import sharding
        ptr, npad, beg, cnt = m.generator_block()
        assert (npad, beg, cnt) == sharding.block_bounds(n, rank, world)[:3]
        full = sharding.DeviceRecords(ptr, npad, f"cuda:{local}").tensor  # aliases the handle's records

        def gather(md):
            sharding.allgather_blocks(dist, full, beg, cnt)  # in place: one collective per iteration
            return sharding.reduce_max(dist, md, f"cuda:{local}")  # host sync: the next iteration may start

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(v):
        if not dist:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def device_steps(k):
        if gather is None:
            m.lloyd(rho, max_iter=k, tol=0.0, warm=warm)
        else:
            for _ in range(k):
                gather(m.lloyd_local(rho, warm=warm))

    # ---- device-resident throughput (the headline `value`)
    device_steps(args.warmup)
    barrier()
    m.stats_reset()
    clk = ClockSampler(local)
    if rank == 0:
        clk.start()
    barrier()
    t0 = time.perf_counter()
    m.timer_begin()
    device_steps(args.steps)
    ms_dev = m.timer_end()
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = clk.stop() if rank == 0 else None
    st = m.stats()
    # NCCL exchange: the collective runs on torch's stream, so the barrier-bracketed span is the
    # time; peer exchange: everything is on the handle's stream and its CUDA events see it all
    ms = allmax(wall_ms if gather is not None else ms_dev)

    # ---- end to end through the C ABI with pinned HOST buffers, copies inside the timed region.
    # Every rank holds, uploads and downloads the block of cells it owns (fm_set_generators_block
    # stores it into all ranks' arrays over NVLink); at N = 1 the block is the whole set.
    blocked = gather is None
    if blocked:
        cnt = m.block_count()
        hx = [fb.pinned_empty(max(cnt, 1)) for _ in range(3)]
        m.get_generators_block_soa(*hx)
        h2d, d2h = 24 * cnt, 24 * cnt + 8
    else:
        hx = [fb.pinned_empty(n) for _ in range(3)]
        g = m.generators()
        for k in range(3):
            hx[k][:] = g[:, k]
        h2d, d2h = 24 * n, 24 * n + 8
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        if blocked:
            m.set_generators_block_soa(*hx)
            m.lloyd(rho, max_iter=1, tol=0.0, warm=warm)
            m.get_generators_block_soa(*hx)
        else:
            m.set_generators_soa(*hx)
            gather(m.lloyd_local(rho, warm=warm))
            m.get_generators_soa(*hx)
    barrier()
    e2e_ms = allmax((time.perf_counter() - t0) * 1e3)
    if dist:
        tb = torch.tensor([h2d, d2h], dtype=torch.int64, device=f"cuda:{local}")
        dist.all_reduce(tb)
        h2d, d2h = int(tb[0].item()), int(tb[1].item())   # whole job: summed over the ranks

    # ---- per-kernel times: a second pass over the same K steps with an event pair around every launch
    ms_prof, rows = profiled(m, lambda: device_steps(args.steps))
    barrier()

    # ---- proof that all ranks (and all GPU counts) computed the same mesh: every rank holds the
    # whole generator set; same arguments => same number of iterations at every N
    g = m.generators()
    sha = hashlib.sha256(np.ascontiguousarray(g).tobytes()).hexdigest()
    shas = [sha]
    if dist:
        shas = [None] * world
        dist.all_gather_object(shas, sha)

    if rank == 0:
        peak, peak_kind = peaks()
        table = kernel_table(rows)
        top = table[0]
        lines = [hbm_line(r["name"], r["items"] / max(r["launches"], 1), r["ms"] / max(r["launches"], 1), peak)
                 for r in table]
        lines = [x for x in lines if x]
        roof = dict(lines[0]) if lines and lines[0]["kernel"] == top["name"] else {
            "kernel": top["name"], "achieved": 0.0, "peak": peak, "unit": "GB/s", "frac": 0.0}
        cap = ncu_capture(top["name"], n, args.density, world)
        roof.update({"bound": "hbm", "peak_kind": peak_kind, "share_of_step": top["share"],
                     "traffic": cap["dram_bytes_per_launch"] if cap else None,
                     "profiled_pass_ms_per_step": ms_prof / args.steps,
                     "kernel_ms_sum_per_step": sum(r["ms"] for r in table) / args.steps,
                     "ncu": cap, "other_kernels": lines[1:],
                     "whole_step": {"alg_bytes_per_item": ALG_BYTES["lloyd_iteration"], "items": n,
                                    "achieved": ALG_BYTES["lloyd_iteration"] * n / (ms / args.steps * 1e-3) / 1e9,
                                    "frac": ALG_BYTES["lloyd_iteration"] * n / (ms / args.steps * 1e-3) / 1e9 / peak}})
        cen = next((r for r in table if r["name"] == "centroid_update"), None)
        if cen and world == 1:
            roof["fp64"] = fp64_line(fb, n, 6 * n - 12, cen["ms"] / max(cen["launches"], 1), args.density)
        out = {
            "metric": METRIC, "value": args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args),
            "e2e": {"value": args.steps / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h,
                    "how": ("per step: every rank uploads its block of generators from pinned host memory (stored into "
                            "all ranks' arrays over NVLink), one Lloyd iteration, every rank downloads its block; bytes "
                            "summed over ranks") if blocked else "per step: full set up, one iteration + NCCL all-gather, full set down, on every rank"},
            "gpu_launches": st["launches"],
            "clocks": clocks,
            "roofline": roof,
            "kernels": table,
            "dt_stats": {k: st[k] for k in ("rounds", "subrounds", "flips", "max_active", "cold", "warm")},
            "generators_sha256": shas[0], "ranks_agree": all(s == shas[0] for s in shas),
            "csrc_sha16": csrc_sha16(),
        }
        print_line = out
    m.close()

    # ---- the rest of what the metric and the configs ask for, OUTSIDE the headline timed region
    if not args.no_extra and (world == 1 or args.exchange == "p2p"):
        extra = {}
        if world == 1 and rank == 0:
            peak, _ = peaks()
            extra["cold"] = extra_cold(fb, p, rho, 5, peak)
            extra["m2_delaunay"] = [extra_m2(fb, n, "fib", peak), extra_m2(fb, 10_000_000, "uniform", peak),
                                    extra_m2(fb, 10_000_000, "fib", peak)]
        # configs[2]'s density on the headline size (sharded like the main run)
        extra_x4 = run_x4(fb, p, rank, world, dist, barrier, allmax)
        if rank == 0:
            extra["x4_density"] = extra_x4
            print_line["extra"] = extra
    if rank == 0:
        if world == 1 and not args.no_cpu:
            sec, done = cpu_lloyd_seconds(n, 1)
            print_line["cpu_baseline"] = {
                "value": 1.0 / sec, "unit": UNIT, "cores": 1, "kind": "port", "host_cores": os.cpu_count(),
                "sample": f"one Lloyd iteration of the CPU restatement (oracle/) at the full {n} generators "
                          f"({sec:.1f} s; cold Delaunay triangulation + Voronoi + centroids), single thread as the reference"}
        print(json.dumps(print_line))
    if dist:
        dist.destroy_process_group()


def run_x4(fb, p, rank, world, dist, barrier, allmax, steps=20, warmup=10):
    """Warm Lloyd loop under the 4:1 refinement density (configs[2]'s density) on the headline size."""
    rho = fb.density_x4()
    m = fb.Mesh.from_xyz(p, rank=rank, size=world)
    if world > 1:
        m.peer_connect(dist)
    m.lloyd(rho, max_iter=warmup, warm=True)
    barrier()
    m.timer_begin()
    m.lloyd(rho, max_iter=steps, warm=True)
    ms = allmax(m.timer_end())
    _, rows = profiled(m, lambda: m.lloyd(rho, max_iter=steps, warm=True))
    barrier()
    m.close()
    return {"density": "x4 (4:1 regional refinement, SURVEY 7.3)", "value": steps / (ms * 1e-3), "unit": UNIT,
            "steps": steps, "ms_per_step": ms / steps, "n_gpus": world, "kernels": kernel_table(rows)[:6]}


# --------------------------------------------------------------------------- MPAS adapter stage
# Algorithmic bytes per item of the MPAS kernels (what mpas_mesh_adapt MM:181-345 and
# calc_weightsOnEdge MM:442-561 read and write per edge / cell / vertex; fp64 + int32, each
# shared quantity counted once: 24 N/E = 8 B of cell and 24 T/E = 16 B of vertex coordinates per
# edge, and so on; degree 6).  DESIGN.md 6 lists the terms.
MPAS_BYTES = {
    "mpas_ring_edge": 56,     # per cell: offset 8, ring neighbours 24, edgesOnCell written 24
    "mpas_edges": 120,        # per edge: 4 ids 16, cells 8, vertices 16; xyz 24, lon/lat 16, dv dc dv1 dv2 angle 40
    "mpas_cells": 252,        # per cell: ring tri + edge ids 48, offset 4, vertices 48, edge points 72, own xyz 24; area 8, kites 48
    "mpas_vertices": 116,     # per triangle: ids 12, cells 12, kites read 24, edge ids read 12; cellsOnVertex 12, edgesOnVertex 12, kites 24, area 8
    "mpas_weights": 372,      # per edge: two rings' kites 96 + edge ids 48 + areas 16, dvEdge 80, dcEdge 8; edgesOnEdge 40, weights 80, count 4
    "ring_degree": 28, "ring_fill": 76, "circumcentre": 48,
}


def run_mpas(args):
    """--workload mpas: rows f1/f2 of SURVEY 8(f) held to the measurement bar -- voronoi_diagram(all)
    + the MPAS adapter arithmetic on C2 (ICO(7)) and C3 (655,362 generators relaxed under the x4
    density), per-kernel time against the HBM roofline."""
    import torch
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import fishman_b200 as fb
    from fishman_b200 import points
    peak, peak_kind = peaks()
    cases = []
    for name, p, rho, lloyd in (("C2 ICO(7) 163842 cells", points.ico(7), fb.density_uniform(), 0),
                                ("C3 FIB(655362) after 12 Lloyd steps under x4", points.fibonacci(655362, 0, 0.2), fb.density_x4(), 12)):
        m = fb.Mesh.from_xyz(p)
        if lloyd:
            m.lloyd(rho, max_iter=lloyd, warm=True)
        else:
            m.delaunay_triangulation()
        m.voronoi_diagram(all=True)
        m.mpas(strict=False)                       # warm-up: allocations
        reps = max(args.steps, 3)
        ts = []
        for _ in range(reps):
            m.timer_begin()
            m.voronoi_diagram(all=True)
            m.mpas_device_only()
            ts.append(m.timer_end())

        def once():
            m.voronoi_diagram(all=True)
            m.mpas_device_only()
        _, rows = profiled(m, once)
        table = kernel_table(rows)
        for r in table:
            b = MPAS_BYTES.get(r["name"])
            if b and r["avg_ms"] > 0:
                ach = b * r["items"] / max(r["launches"], 1) / (r["avg_ms"] * 1e-3) / 1e9
                r.update({"alg_bytes_per_item": b, "achieved_gbs": round(ach, 1), "frac": round(ach / peak, 4)})
        n = len(p)
        ms = float(np.median(ts))
        cases.append({"case": name, "cells": n, "ms": ms, "cells_per_s": n / (ms * 1e-3), "kernels": table})
        m.close()
    print(json.dumps({"metric": "mpas_adapter_cells_per_sec", "workload": "voronoi_diagram(all=.true.) + mpas_mesh_adapt + "
                      "calc_weightsOnEdge on the device (SURVEY 8f rows 1-2)", "unit": "cells/s",
                      "value": cases[-1]["cells_per_s"], "peak_gbs": peak, "peak_kind": peak_kind, "cases": cases,
                      "n_gpus": 1, "dtype": "f64", "data": "synthetic"}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="lloyd", choices=["lloyd", "mpas"])
    ap.add_argument("--mode", default="warm", choices=["warm", "cold"])
    ap.add_argument("--n", type=int, default=N_C4)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra blocks (cold mode, one-shot Delaunay, x4 density)")
    ap.add_argument("--density", default="uniform", choices=["uniform", "x4"],
                    help="x4: the 4:1 regional refinement density of configs[2] (not the headline workload)")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                    help="N > 1: how the updated generator blocks reach the other ranks")
    args = ap.parse_args()
    if args.impl == "ours":
        args.warmup = max(args.warmup, 10)   # a 20-step timed region is 8 ms: start it at steady clocks
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "mpas":
        run_mpas(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
