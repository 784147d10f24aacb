#!/usr/bin/env python
"""bench.py -- Lloyd iterations/s at 2,621,442 generators (BASELINE.json metric) on N B200s.

A step = one Lloyd iteration of the SCVT inner loop on the GPU: Delaunay triangulation of
the current generators, circumcentres, density-weighted centroid quadrature, generator
update + max-displacement reduction (+ generator all-gather when N > 1).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--mode warm|cold] [--n GENERATORS]
  python bench.py --impl reference ...      # CPU restatement of the reference, host cores

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import subprocess
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
CPU_SAMPLE_N = 655362   # cpu_baseline leg: one iteration at this size (~15-20 s), scaled by N
REF_SAMPLE_N = 163842   # --impl reference: each step is one iteration at this size (~2 s)

# Algorithmic bytes per work item of each kernel (DESIGN.md "Kernels and rooflines").
ALG_BYTES = {
    "locate_vote": 80,       # point record 32 + loc r/w 8 + one triangle row tri+adj 32 + vote 8
    "split_flag": 12, "split13": 128, "cub_scan": 8,
    "flip_loop": 0,          # cooperative kernel: all flip sub-rounds of a round (latency bound, no byte model)
    "sweep": 80,             # per triangle: tri 16 + adj 16 + vertex records amortised 32*N/T = 16 + circumcentre write 32
    "circumcentre": 64,      # tri 16 + vertex records amortised 16 + write 32
    "centroid_update": 152,  # per cell: record r/w 32+24 + ring-cache record 32 + circumcentres amortised 32*T/N = 64
    "validate": 32,
}


def ncu_extra(kernel, n):
    """fp64-pipe and DRAM utilisation of `kernel` from the same committed capture (informational:
    the centroid kernel sits nearer its fp64-pipe bound than its HBM bound)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        k = t["kernels"][kernel]
        return {"fp64_pipe_pct": k.get("fp64_pipe_pct"), "dram_pct_of_peak": k.get("dram_pct_of_peak"),
                "source": t["source"]} if t["generators"] == n else None
    except Exception:
        return None


def ncu_traffic(kernel, n):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed
    ncu --set full capture (profiles/traffic.json), valid for the generator count it was taken at."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        return t["kernels"][kernel]["dram_bytes_per_launch"] if t["generators"] == n else None
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
    region runs (the ctypes calls into the library release the GIL).  ~1 ms period: the timed
    region of a warm run is only tens of ms, far shorter than one nvidia-smi -lms tick."""

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
            time.sleep(0.001)

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


def make_points(n):
    from fishman_b200 import points
    return points.fibonacci(n, seed=0, jitter=0.2)


# --------------------------------------------------------------------------- CPU legs
def cpu_lloyd_seconds(n, steps, warmup=0):
    """Seconds per Lloyd iteration of the CPU restatement (oracle/) at n generators.
    The points are randomly permuted first: the reference inserts in index order, which
    is quadratic on a coherent (spiral) order -- the permutation is in its favour."""
    from oracle import oracle as O
    p = make_points(n)
    p = p[np.random.default_rng(0).permutation(n)]
    o = O.Oracle(p)
    rho = O.density_uniform()
    for _ in range(warmup):
        o.lloyd_step(rho)
    t0 = time.perf_counter()
    for _ in range(steps):
        o.lloyd_step(rho)
    dt = (time.perf_counter() - t0) / max(steps, 1)
    o.close()
    return dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    try:
        os.sched_setaffinity(0, {sorted(os.sched_getaffinity(0))[0]})
    except Exception:
        pass
    n = REF_SAMPLE_N
    sec = cpu_lloyd_seconds(n, args.steps, args.warmup)
    val = 1.0 / (sec * (args.n / n))
    sample = (f"each step = one Lloyd iteration of the CPU restatement at {n} generators (FIB, randomly permuted), "
              f"scaled linearly to {args.n} (x{args.n / n:.1f}; the measured growth is super-linear, so this favours the CPU)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * (args.n / n) * 1e3,
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
            "parallelism": (f"replicated DT + sharded centroid x{args.gpus}, generator exchange: "
                            + ("stores into peer memory over NVLink from the Lloyd kernel" if args.exchange == "p2p"
                               else "NCCL all-gather")) if args.gpus > 1 else "single GPU",
            "l2": "working set (~0.6 GB of generators, triangles, adjacency, circumcentres) exceeds the 126 MB L2; no flush"}


# --------------------------------------------------------------------------- GPU leg
class _DevPtr:
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f8", "data": (int(ptr), False), "version": 2}


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

    def device_steps(k):
        if gather is None:
            m.lloyd(rho, max_iter=k, tol=0.0, warm=warm)
        else:
            for _ in range(k):
                gather(m.lloyd_local(rho, warm=warm))

    # ---- device-resident throughput
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
    ms = wall_ms if gather is not None else ms_dev
    if dist:
        t = torch.tensor([ms], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())

    # ---- end to end through the C ABI with pinned host buffers (N = 1 semantics on every rank)
    hx = [fb.pinned_empty(n) for _ in range(3)]
    g = m.generators()
    for k in range(3):
        hx[k][:] = g[:, k]
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        m.set_generators_soa(*hx)
        if gather is None:
            m.lloyd(rho, max_iter=1, tol=0.0, warm=warm)
        else:
            gather(m.lloyd_local(rho, warm=warm))
        m.get_generators_soa(*hx)
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    if dist:
        t = torch.tensor([e2e_ms], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())

    # ---- per-kernel times (CUDA events around every launch on the launching stream)
    # a second pass over the same K steps with a CUDA-event pair around every launch (pairs are
    # resolved at the points where the host waits anyway: no extra synchronisation)
    m.profile(True)
    m.stats_reset()
    m.timer_begin()
    device_steps(args.steps)
    ms_prof = m.timer_end()
    rows = m.profile_rows()
    m.profile(False)
    barrier()

    if rank == 0:
        peak, peak_kind = peaks()
        rows.sort(key=lambda r: -r["ms"])
        tot = sum(r["ms"] for r in rows) or 1.0
        top = rows[0]
        abytes = ALG_BYTES.get(top["name"], 0) * top["threads"] / max(top["launches"], 1)
        dur_ms = top["ms"] / max(top["launches"], 1)
        ach = abytes / (dur_ms * 1e-3) / 1e9 if dur_ms > 0 else 0.0
        out = {
            "metric": METRIC, "value": args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args),
            "e2e": {"value": args.steps / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 24 * n,
                    "d2h_bytes_per_step": 24 * n + 8},
            "gpu_launches": st["launches"],
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": top["name"], "achieved": ach, "peak": peak, "unit": "GB/s",
                         "frac": ach / peak, "traffic": ncu_traffic(top["name"], n), "peak_kind": peak_kind,
                         "profiled_pass_ms_per_step": ms_prof / args.steps,
                         "alg_bytes_per_item": ALG_BYTES.get(top["name"], 0),
                         "items_per_launch": top["threads"] / max(top["launches"], 1),
                         "avg_launch_ms": dur_ms, "share_of_step": top["ms"] / tot,
                         "ncu": ncu_extra(top["name"], n)},
            "kernels": [{"name": r["name"], "launches": r["launches"], "ms": round(r["ms"], 4),
                         "share": round(r["ms"] / tot, 4)} for r in rows],
            "dt_stats": {k: st[k] for k in ("rounds", "subrounds", "flips", "max_active", "cold", "warm")},
        }
        if world == 1 and not args.no_cpu:
            sec = cpu_lloyd_seconds(CPU_SAMPLE_N, 1)
            out["cpu_baseline"] = {
                "value": 1.0 / (sec * n / CPU_SAMPLE_N), "unit": UNIT, "cores": 1, "kind": "port",
                "host_cores": os.cpu_count(),
                "sample": f"one Lloyd iteration of the CPU restatement (oracle/) at {CPU_SAMPLE_N} generators "
                          f"({sec:.2f} s), scaled linearly to {n}"}
        print(json.dumps(out))
    m.close()
    if dist:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="warm", choices=["warm", "cold"])
    ap.add_argument("--n", type=int, default=N_C4)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--density", default="uniform", choices=["uniform", "x4"],
                    help="x4: the 4:1 regional refinement density of configs[2] (not the headline workload)")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                    help="N > 1: how the updated generator blocks reach the other ranks")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
