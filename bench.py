#!/usr/bin/env python3
"""bench.py -- spring-steps/s of the Titan step path on B200 (BASELINE.json metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A bench "step" is one batch of --inner simulation time steps (default 200) of the workload; `value` is
springs x simulation-steps / second over ALL ranks with state resident in HBM (CUDA events on the engine
stream, max over ranks); `e2e` is the same metric through the host-buffer C ABI with the upload of the state
and the download of the result inside the timed region.  Workload at N=1: BASELINE config C2 (100^3 lattice,
1,000,000 masses, 12,731,796 springs on a frictional plane).  At N>1 every rank steps its own copy of that
lattice (independent simulations, no collective; weak scaling) unless --workload c4 selects the partitioned
lattice with halo exchange.
`--impl reference` times the reference's algorithm on the host cores: the reference has no CPU implementation
of the step (its kernels are __global__ only), so this is the OpenMP transcription in oracle/ (kind "port").
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
sys.path.insert(0, ROOT)

METRIC = "spring_steps_per_sec"
UNIT = "spring-steps/s"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def workload_c2(n=100):
    """n^3 lattice, 0.1 m spacing, bottom layer 1 mm inside a frictional plane z=0."""
    from titan_b200 import lattice
    side = 0.1 * (n - 1)
    L = lattice.build_lattice((0.0, 0.0, side / 2 - 0.001), (side, side, side), n, n, n)
    L["plane"] = ((0.0, 0.0, 1.0), 0.0, 0.5, 0.6)
    L["tile_hint"] = lattice.brick_hint(n, n, n, tile=int(os.environ.get("TB_TILE", "128")))
    L["name"] = "C2: %d^3 lattice on frictional plane (k=1e4, m=0.1, dt=1e-4, g=-9.81)" % n
    return L


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.rows, self.proc = device, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.06)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_port(W, seconds_budget=12.0, min_steps=2):
    """Time the OpenMP transcription of the reference step (oracle/, kind 'port') on a bounded sample."""
    import oracle
    sc = oracle.Scene(W["pos"], W["left"], W["right"], W["rest"])
    n, off, fk, fs = W["plane"]
    sc.planes.append(oracle.make_plane(n, off, fk, fs))
    sc.local = None
    sc.step_omp(1)  # warm-up (page faults, thread pool)
    steps, t0 = 0, time.perf_counter()
    while True:
        sc.step_omp(1)
        steps += 1
        el = time.perf_counter() - t0
        if steps >= min_steps and el >= seconds_budget:
            break
        if steps >= 400:
            break
    return {"value": sc.ns * steps / el, "unit": UNIT, "cores": oracle.omp_threads(), "kind": "port",
            "sample": "%d steps of the same workload (%d springs) in %.1f s, OpenMP scatter with atomics" % (steps, sc.ns, el),
            "seconds": el, "steps": steps}


def run_reference(args, rank, world):
    """--impl reference: the reference algorithm on the host cores (OpenMP port in oracle/), rank 0 only."""
    if rank != 0:
        return
    import oracle
    W = workload_c2(args.size)
    sc = oracle.Scene(W["pos"], W["left"], W["right"], W["rest"])
    n, off, fk, fs = W["plane"]
    sc.planes.append(oracle.make_plane(n, off, fk, fs))
    sc.local = None
    inner = args.ref_inner
    for _ in range(args.warmup):
        sc.step_omp(inner)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sc.step_omp(inner)
    el = time.perf_counter() - t0
    value = sc.ns * inner * args.steps / el
    cores = oracle.omp_threads()
    sample = "%d bench steps x %d simulation steps of the same workload, OpenMP (%d threads)" % (args.steps, inner, cores)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * el / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": W["name"], "masses": sc.nm, "springs": sc.ns, "sim_steps_per_bench_step": inner,
                       "note": "the reference has no CPU step (kernels are __global__ only); this is the OpenMP "
                               "transcription of computeSpringForces/massForcesAndUpdate in oracle/"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--size", type=int, default=100, help="lattice edge (C2 = 100)")
    ap.add_argument("--inner", type=int, default=200, help="simulation steps per bench step")
    ap.add_argument("--ref-inner", type=int, default=2, help="simulation steps per bench step, --impl reference")
    ap.add_argument("--strict", action="store_true", help="TB_MATH_STRICT arithmetic")
    ap.add_argument("--csr", action="store_true", help="force the generic CSR kernel")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-hint", action="store_true", help="let the engine bisect positions instead of lattice bricks")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import titan_b200 as tb
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    W = workload_c2(args.size)
    nm, ns = W["pos"].shape[0], W["left"].shape[0]
    flags = (tb.MATH_STRICT if args.strict else tb.MATH_FAST) | (tb.KERNEL_CSR if args.csr else 0)
    eng = tb.Engine(local_rank, flags)
    eng.set_topology(nm, W["left"], W["right"])
    if not args.no_hint:
        eng.set_tile_hint(W["tile_hint"])     # what Lattice builders pass: bricks of lattice cells
    eng.set_mass_state(pos=W["pos"])
    eng.set_spring_params(rest=W["rest"])
    n, off, fk, fs = W["plane"]
    eng.set_planes([tb.Plane((tb.capi.C.c_double * 3)(*n), off, fk, fs)])
    eng.set_globals((0.0, 0.0, -9.81), 1e-4)
    dt, inner = 1e-4, args.inner

    t_sim = 0.0
    def advance(nsteps):
        nonlocal t_sim
        eng.run_async(nsteps, t_sim)
        for _ in range(nsteps):
            t_sim += dt

    # ---- device-resident throughput (`value`) ------------------------------------------------------------
    for _ in range(args.warmup):
        advance(inner)
    eng.sync()
    barrier()
    launches0 = eng.info()["launches"]
    with ClockSampler(local_rank) as clocks:
        w0 = time.perf_counter()
        advance(inner * args.steps)
        eng.sync()
        wall = time.perf_counter() - w0
    dev_ms = eng.last_run_ms()
    barrier()
    launches = eng.info()["launches"] - launches0
    dev_ms = max_over_ranks(dev_ms)
    wall = max_over_ranks(wall)
    sim_steps = inner * args.steps
    value = world * ns * sim_steps / (dev_ms * 1e-3)

    info = eng.info()
    peak, peak_src = load_peaks()
    step_us = dev_ms * 1e3 / sim_steps
    achieved = info["algorithmic_bytes_per_step"] / (step_us * 1e-6) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "peak_source": peak_src, "kernel": "tiled_step" if info["kernel"] == 1 else "csr_step",
                "algorithmic_bytes_per_launch": info["algorithmic_bytes_per_step"], "avg_launch_us": step_us}
    traffic_file = os.path.join(ROOT, "profiles", "traffic_bytes_per_launch.json")
    if os.path.exists(traffic_file):
        try:
            roofline["traffic"] = json.load(open(traffic_file)).get(roofline["kernel"])
        except Exception:
            pass

    # ---- end to end through host buffers (`e2e`) ---------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        pos_h, vel_h = eng.alloc_pinned((nm, 3)), eng.alloc_pinned((nm, 3))
        st = eng.get_mass_state(fields=("pos", "vel"), out={"pos": pos_h, "vel": vel_h})
        def e2e_step():
            eng.set_mass_state(pos=pos_h, vel=vel_h)                  # H2D from pinned host memory
            advance(inner)
            eng.get_mass_state(fields=("pos", "vel"), out={"pos": pos_h, "vel": vel_h})   # D2H (syncs)
        for _ in range(2):
            e2e_step()
        barrier()
        w0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        eng.sync()
        e2e_s = max_over_ranks(time.perf_counter() - w0)
        barrier()
        e2e = {"value": world * ns * sim_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(nm * 48),
               "d2h_bytes_per_step": int(nm * 48), "ms_per_step": 1e3 * e2e_s / args.steps,
               "call": "tb_set_mass_state(pos,vel) + tb_run_async(%d) + tb_get_mass_state(pos,vel)" % inner}
        assert np.isfinite(pos_h).all()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_port(W)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": W["name"], "masses": nm, "springs": ns, "sim_steps_per_bench_step": inner,
                           "math": "strict" if args.strict else "fast",
                           "tiling": {k: info[k] for k in ("n_tiles", "threads_per_mass", "grid", "smem_bytes",
                                                           "ell_unique_tiles", "tiled_bytes_per_step")}, "per_rank": "independent copy of the workload",
                           "l2": "working set per step (%.0f MB) exceeds the 126 MB L2" % (info["algorithmic_bytes_per_step"] / 1e6),
                           "timing": "CUDA events on the engine stream, max over ranks", "wall_ms": wall * 1e3},
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
                "clocks": clocks.summary()}
        print(json.dumps(line), flush=True)
    eng.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
