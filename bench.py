#!/usr/bin/env python3
"""bench.py -- spring-steps/s of the Titan step path on B200 (BASELINE.json metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A bench "step" is one batch of --inner simulation time steps (default 200) of the workload; `value` is
springs x simulation-steps / second over ALL ranks with state resident in HBM (CUDA events on the engine
stream, max over ranks); `e2e` is the same metric through the host-buffer C ABI with the upload of the state
and the download of the result inside the timed region.  Workload at N=1: BASELINE config C2 (100^3 lattice,
1,000,000 masses, 12,731,796 springs on a frictional plane).  At N>1 the default is BASELINE config C4: ONE lattice of
about 150^3 masses per GPU (189^3 / 238^3 / 300^3 at 2 / 4 / 8 GPUs) cut into i-slabs, boundary-mass halo exchange
over NVLink peer memory inside the step kernel (weak scaling); --workload c2 / c3 at N>1 step independent copies /
independent robots per rank with no communication.
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
        self.t0 = self.t1 = None

    def start(self, wait_s=10.0):
        """Launch nvidia-smi and wait for its first line (on an 8-GPU box it needs more than a second to come up, longer
        than a short timed region), so that samples exist by the time the region begins."""
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
            t = time.perf_counter()
            while not self.rows and time.perf_counter() - t < wait_s and self.proc.poll() is None:
                time.sleep(0.02)
        except OSError:
            self.proc = None
        return self

    def __enter__(self):                      # the timed region
        if self.proc is None and self.t0 is None:
            self.start()
        self.t0 = time.perf_counter()
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def __exit__(self, *a):
        self.t1 = time.perf_counter()
        if self.proc:
            time.sleep(0.06)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        inside = [r for t, r in self.rows if self.t0 is not None and self.t0 <= t <= (self.t1 or t) + 0.06]
        for r in inside:
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
    """--impl reference: the reference algorithm on the host cores (OpenMP port in oracle/), rank 0 only.  Each
    step is a bounded sample of the arm's workload: all of C2; for C4 the planes one rank of the partition owns
    (plus a ghost plane); for C3 at most 64 robots -- spring-steps/s is normalised by the springs stepped."""
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1 to its workers: the reference arm uses every host core, whatever launched it
    os.environ["OMP_NUM_THREADS"] = str(args.ref_threads or os.cpu_count() or 1)
    import oracle
    if args.workload == "c4":
        from titan_b200 import lattice, partition
        n = args.size
        lo, hi = partition.slab_bounds(n, world)[0]
        planes = min(n, hi - lo + 1)
        L = lattice.build_lattice((0.0, 0.0, 0.05 * (n - 1) - 0.001), (0.1 * (planes - 1), 0.1 * (n - 1), 0.1 * (n - 1)), planes, n, n)
        W = dict(L, plane=PLANE, name="C4 family: %d^3 lattice over %d ranks" % (n, world))
        sample = "rank 0's share of the lattice (%d of %d i-planes of %dx%d masses)" % (planes, n, n, n)
    elif args.workload == "c3":
        per = min(64, args.robots // world)
        W = workload_c3(per, 0)
        sample = "%d of the %d robots" % (per, args.robots)
    else:
        W = workload_c2(args.size)
        sample = "the whole workload"
    sc = oracle.Scene(W["pos"], W["left"], W["right"], W["rest"])
    if "type" in W:
        sc.type[:] = W["type"]
        sc.omega[:] = W["omega"]
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
    sample = "%s: %d masses, %d springs; %d bench steps x %d simulation steps, OpenMP (%d threads)" % (sample, sc.nm, sc.ns, args.steps, inner, cores)
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


FACADE_DRIVER = os.path.join(ROOT, "tests", "cpp", "bin", "titan_facade_dump")   # tests/cpp/titan_api_driver.cu against libTitan.so
TITAN_CUDA = os.path.join(ROOT, "oracle", "_ref", "titan_ref_cuda")               # the same driver against the reference's own CUDA engine


def _last_json(text):
    for line in reversed(text.splitlines()):
        line = line.strip()
        if line.startswith("{"):
            try:
                return json.loads(line)
            except ValueError:
                continue
    return None


def facade_e2e(config, inner, batches, timeout=600):
    """`e2e` through the API the reference's users call: a C++ program (tests/cpp/titan_api_driver.cu, mode bench_e2e)
    builds the workload with titan::Simulation::createLattice/createPlane and times, per batch, setAll() (host Mass /
    Spring objects -> device) + resume() + pause(t + inner dt) + getAll() (device -> host Mass objects)."""
    if not os.path.exists(FACADE_DRIVER):
        return {"unavailable": "tests/cpp/bin/titan_facade_dump is not built"}
    try:
        r = subprocess.run([FACADE_DRIVER, "bench_e2e", config, str(inner), str(batches)], capture_output=True, text=True, timeout=timeout)
    except subprocess.TimeoutExpired:
        return {"unavailable": "facade driver timed out"}
    d = _last_json(r.stdout)
    if r.returncode != 0 or d is None:
        return {"unavailable": "facade driver failed (rc %d): %s" % (r.returncode, (r.stderr or r.stdout)[-200:])}
    return {"value": d["spring_steps_per_s"], "unit": UNIT, "h2d_bytes_per_step": int(d["h2d_bytes_per_batch"]),
            "d2h_bytes_per_step": int(d["d2h_bytes_per_batch"]), "ms_per_step": 1e3 * d["seconds"] / max(1, d["batches"]),
            "sim_steps": int(d["steps"]), "batches": int(d["batches"]), "start_seconds": d["upload_seconds"],
            "call": "titan::Simulation::setAll() + resume() + pause(t) + getAll() on host Mass/Spring objects (libTitan.so)"}


def titan_cuda_baseline(config, steps, timeout=900):
    """Titan's own CUDA kernels (reference sources, headless-patched, built for sm_100a by oracle/build_ref.sh) on this GPU:
    wall clock around Simulation::pause for `steps` steps, start()'s upload reported separately."""
    if not os.path.exists(TITAN_CUDA):
        return {"unavailable": "oracle/_ref/titan_ref_cuda is not built (needs /root/reference at build time)"}
    try:
        r = subprocess.run([TITAN_CUDA, "bench", config, str(steps)], capture_output=True, text=True, timeout=timeout)
    except subprocess.TimeoutExpired:
        return {"unavailable": "titan_ref_cuda timed out"}
    d = _last_json(r.stdout)
    if r.returncode != 0 or d is None:
        return {"unavailable": "titan_ref_cuda failed (rc %d)" % r.returncode}
    return {"value": d["spring_steps_per_s"], "unit": UNIT, "steps": int(d["steps"]), "seconds": d["seconds"],
            "upload_seconds": d["upload_seconds"], "kind": "reference CUDA kernels (computeSpringForces + massForcesAndUpdate, src/sim.cu) on the same GPU"}


def parity_n(tb, dist, local_rank, rank, world):
    """N > 1, before anything is timed: the ranks step a 48 x 24 x 24 lattice cut into i-slabs (halo exchange over peer
    memory) for 230 steps, rank 0 steps the whole lattice on one engine; max |difference| over pos and vel, strict and fast
    arithmetic.  Strict must be bit-identical, fast <= 1e-12; otherwise the run fails."""
    import torch
    from titan_b200 import distributed, lattice, partition
    n, steps = (48, 24, 24), 230
    c, d = (0.0, 0.0, 0.1 * (n[2] - 1) / 2 - 0.001), (0.1 * (n[0] - 1), 0.1 * (n[1] - 1), 0.1 * (n[2] - 1))
    saved = os.environ.get("TB_TPM")
    os.environ["TB_TPM"] = "1"      # the single engine sums per mass like the partitioned ones (one thread per mass)
    out = {}
    try:
        for flags, label, tol in ((tb.MATH_STRICT, "strict_max_abs", 0.0), (tb.MATH_FAST, "fast_max_abs", 1e-12)):
            part = partition.LatticeSlab(c, d, n[0], n[1], n[2], rank, world)
            eng = tb.Engine(local_rank, flags)
            distributed.upload_slab(eng, part, planes=[PLANE])
            distributed.connect(eng, part)
            t = 0.0
            for chunk in (1, 29, steps - 30):
                eng.run_async(chunk, t)
                eng.sync()
                for _ in range(chunk):
                    t += 1e-4
            dist.barrier()
            st = eng.get_mass_state(idx=part.own_local, fields=("pos", "vel"))
            eng.close()
            gathered = [None] * world
            dist.all_gather_object(gathered, (part.mass_gid[part.own_local], st["pos"], st["vel"]))
            err = torch.zeros(1, dtype=torch.float64, device="cuda")
            if rank == 0:
                L = lattice.build_lattice(c, d, *n)
                ref = tb.Engine(local_rank, flags)
                ref.set_topology(L["pos"].shape[0], L["left"], L["right"])
                ref.set_tile_hint(lattice.brick_hint(*n))
                ref.set_mass_state(pos=L["pos"])
                ref.set_spring_params(rest=L["rest"])
                ref.set_planes([tb.Plane((tb.capi.C.c_double * 3)(*PLANE[0]), *PLANE[1:])])
                ref.set_globals((0.0, 0.0, -9.81), 1e-4)
                ref.run(steps, 0.0)
                want = ref.get_mass_state(fields=("pos", "vel"))
                ref.close()
                e = 0.0
                for gid, pos, vel in gathered:
                    e = max(e, float(np.max(np.abs(pos - want["pos"][gid]))), float(np.max(np.abs(vel - want["vel"][gid]))))
                err[0] = e
            dist.broadcast(err, 0)
            out[label] = float(err.item())
            if not out[label] <= tol:
                if rank == 0:
                    print(json.dumps({"error": "N>1 parity check failed", "parity_n": out}), flush=True)
                dist.barrier()
                raise SystemExit(1)
    finally:
        if saved is None:
            os.environ.pop("TB_TPM", None)
        else:
            os.environ["TB_TPM"] = saved
    out["check"] = "48x24x24 lattice in %d i-slabs, 230 steps, vs one engine" % world
    return out


C4_EDGE = {1: 150, 2: 189, 4: 238, 8: 300}     # ~3.4 M masses per GPU; 300^3 on 8 GPUs is config C4
PLANE = ((0.0, 0.0, 1.0), 0.0, 0.5, 0.6)        # frictional plane z = 0 (mu_k 0.5, mu_s 0.6)


def workload_c3(n_robots, first_robot=0):
    """independent 10^3 soft robots (0.9 m cubes resting 0.1 m above the plane on a 1.2 m grid),
    spring s of each robot PASSIVE / ACTIVE_CONTRACT_THEN_EXPAND / ACTIVE_EXPAND_THEN_CONTRACT by s % 3, omega = 4 pi.
    One engine steps all robots of a rank as one disjoint graph (batched SoA, one launch per step)."""
    from titan_b200 import lattice
    L = lattice.build_lattice((0.0, 0.0, 0.55), (0.9, 0.9, 0.9), 10, 10, 10)
    nm1, ns1 = L["pos"].shape[0], L["left"].shape[0]
    hint1 = lattice.brick_hint(10, 10, 10, tile=125)
    nt1 = int(hint1.max()) + 1
    r = np.arange(first_robot, first_robot + n_robots)
    off = np.zeros((n_robots, 1, 3))
    off[:, 0, 0] = 1.2 * (r % 64)
    off[:, 0, 1] = 1.2 * (r // 64)
    pos = (L["pos"][None, :, :] + off).reshape(-1, 3)
    base = (np.arange(n_robots, dtype=np.int64) * nm1)[:, None]
    left = (L["left"][None, :] + base).astype(np.int32).reshape(-1)
    right = (L["right"][None, :] + base).astype(np.int32).reshape(-1)
    s = np.arange(ns1)
    type1 = np.where(s % 3 == 1, 2, np.where(s % 3 == 2, 3, 0)).astype(np.int32)
    return {"pos": pos, "left": left, "right": right, "rest": np.tile(L["rest"], n_robots), "type": np.tile(type1, n_robots),
            "omega": np.full(left.size, 4 * np.pi), "plane": PLANE,
            "tile_hint": (hint1[None, :] + (np.arange(n_robots, dtype=np.int64) * nt1)[:, None]).astype(np.int32).reshape(-1),
            "name": "C3: %d independent actuated 10^3 soft robots per rank" % n_robots}


def run_c5(args, tb, device, flags):
    """BASELINE config C5: topology optimisation on an n^3 lattice (n = 200) over a frictional plane
    and a ball cap; every 100 steps ALL spring constants are rewritten from the host (k_s *= 0.9 + 0.2 h(s)) and the
    0.1 % weakest live springs are removed (tb_remove_springs).  One bench step = one such edit + 100 simulation steps.
    `value` = live springs x steps / device time of the stepping; `e2e` = the same over wall time including the k
    upload (8 B per spring) and the removal calls.  Choosing which springs to remove is the application's optimiser
    and is not timed."""
    from titan_b200 import lattice
    n = args.size
    side = 0.1 * (n - 1)
    L = lattice.build_lattice((0.0, 0.0, 1.1 + side / 2), (side, side, side), n, n, n)
    nm, ns = L["pos"].shape[0], L["left"].shape[0]
    eng = tb.Engine(device, flags)
    eng.set_topology(nm, L["left"], L["right"])
    eng.set_tile_hint(lattice.brick_hint(n, n, n))
    eng.set_mass_state(pos=L["pos"])
    eng.set_spring_params(rest=L["rest"])
    eng.set_planes([tb.Plane((tb.capi.C.c_double * 3)(*PLANE[0]), *PLANE[1:])])
    eng.set_balls([tb.Ball((tb.capi.C.c_double * 3)(0.0, 0.0, -3.0), 4.0)])
    eng.set_globals((0.0, 0.0, -9.81), 1e-4)
    del L
    s = np.arange(ns, dtype=np.uint64)
    factor = 0.9 + 0.2 * (((s * np.uint64(2654435761)) % np.uint64(2 ** 32)) % np.uint64(1000)).astype(np.float64) / 1000.0
    del s
    k = eng.alloc_pinned((ns,))                       # the optimiser's stiffness array lives in page-locked memory
    k[:] = 10000.0
    live = np.ones(ns, dtype=bool)
    per = args.inner if args.inner != 200 else 100
    t_sim, dev_ms, wall, spring_steps, removed_total = 0.0, 0.0, 0.0, 0.0, 0

    def pick():
        kk = np.where(live, k, np.inf)
        m = max(1, int(0.001 * live.sum()))
        idx = np.argpartition(kk, m)[:m]
        return np.sort(idx).astype(np.int32)

    for it in range(args.warmup + args.steps):
        k *= factor
        idx = pick()                                  # untimed: the optimiser's decision
        w0 = time.perf_counter()
        eng.set_spring_params(k=k)                    # H2D: 8 B per spring
        eng.remove_springs(idx)
        eng.run_async(per, t_sim)
        eng.sync()
        w1 = time.perf_counter()
        live[idx] = False
        for _ in range(per):
            t_sim += 1e-4
        if it >= args.warmup:
            dev_ms += eng.last_run_ms()
            wall += w1 - w0
            spring_steps += float(live.sum()) * per
            removed_total += idx.size
    info = eng.info()
    st = eng.get_mass_state(first=0, count=1000, fields=("pos",))
    assert np.isfinite(st["pos"]).all()
    peak, peak_src = load_peaks()
    n_live = float(live.sum())
    step_us = dev_ms * 1e3 / (per * args.steps)
    alg = 104.0 * nm + 24.0 * n_live
    line = {"metric": METRIC, "value": spring_steps / (dev_ms * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C5: %d^3 lattice, plane + ball cap, every %d steps all k rewritten and the 0.1 %% weakest springs removed" % (n, per),
                       "masses": nm, "springs": ns, "springs_live_at_end": int(n_live), "removed_in_timed_region": int(removed_total),
                       "sim_steps_per_bench_step": per, "math": "strict" if args.strict else "fast",
                       "l2": "working set per step (%.0f MB) exceeds the 126 MB L2" % (alg / 1e6)},
            "roofline": {"bound": "hbm", "achieved": alg / (step_us * 1e-6) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": alg / (step_us * 1e-6) / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                         "kernel": "tiled_step" if info["kernel"] == 1 else "csr_step", "algorithmic_bytes_per_launch": alg,
                         "avg_launch_us": step_us},
            "cpu_baseline": None,
            "e2e": {"value": spring_steps / wall, "unit": UNIT, "h2d_bytes_per_step": int(ns * 8 + removed_total // max(1, args.steps) * 4),
                    "d2h_bytes_per_step": 0, "ms_per_step": 1e3 * wall / args.steps,
                    "call": "tb_set_spring_params(k: all springs) + tb_remove_springs + tb_run_async(%d) + tb_sync" % per},
            "gpu_launches": int(info["launches"])}
    print(json.dumps(line), flush=True)
    eng.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=["c2", "c3", "c4", "c5"],
                    help="default: c2 on one GPU, c4 (partitioned lattice, halo exchange) on several")
    ap.add_argument("--size", type=int, default=None, help="lattice edge (c2: 100; c4: 150/189/238/300 by GPU count)")
    ap.add_argument("--shape", default=None, help="c4: nx,ny,nz of the whole lattice instead of --size^3")
    ap.add_argument("--robots", type=int, default=4096, help="c3: robots over all ranks")
    ap.add_argument("--inner", type=int, default=200, help="simulation steps per bench step")
    ap.add_argument("--ref-inner", type=int, default=2, help="simulation steps per bench step, --impl reference")
    ap.add_argument("--strict", action="store_true", help="TB_MATH_STRICT arithmetic")
    ap.add_argument("--csr", action="store_true", help="force the generic CSR kernel")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--distinct-params", action="store_true",
                    help="give every spring its own k (defeats the dictionary records: one record per spring is streamed)")
    ap.add_argument("--verlet", action="store_true", help="the reference's VERLET build option instead of its default Euler step")
    ap.add_argument("--rk2", action="store_true", help="the reference's RK2 build option (repaired): two force evaluations per step")
    ap.add_argument("--no-hint", action="store_true", help="let the engine bisect positions instead of lattice bricks")
    ap.add_argument("--ref-threads", type=int, default=0, help="--impl reference: OpenMP threads (default: all host cores)")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the N=1 extras of the default line: facade e2e, Titan-CUDA baseline, 150^3 weak-scaling base, distinct-parameter roofline")
    ap.add_argument("--no-titan-cuda", action="store_true", help="skip the Titan-CUDA baseline (its start() alone takes about 70 s on C2)")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the partition-vs-single-engine parity check before timing")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload is None:
        args.workload = "c2" if world == 1 else "c4"
    if args.size is None:
        args.size = {"c2": 100, "c5": 200}.get(args.workload) or C4_EDGE.get(world, int(round(150 * world ** (1.0 / 3.0))))
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

    def reduce_ranks(x, op="max"):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
        return float(t.item())

    parity = None
    if world > 1 and args.workload == "c4" and not args.no_parity:
        parity = parity_n(tb, dist, local_rank, rank, world)
    flags = (tb.MATH_STRICT if args.strict else tb.MATH_FAST) | (tb.KERNEL_CSR if args.csr else 0)
    if args.verlet:
        flags |= tb.INTEGRATOR_VERLET
    if args.rk2:
        flags |= tb.INTEGRATOR_RK2
    if args.workload == "c5":
        run_c5(args, tb, local_rank, flags)
        return
    eng = tb.Engine(local_rank, flags)
    part, own = None, None
    if args.workload == "c4":
        from titan_b200 import distributed, partition
        n = args.size
        shp = tuple(int(x) for x in args.shape.split(",")) if args.shape else (n, n, n)
        part = partition.LatticeSlab((0.0, 0.0, 0.05 * (shp[2] - 1) - 0.001), tuple(0.1 * (x - 1) for x in shp), shp[0], shp[1], shp[2], rank, world)
        distributed.upload_slab(eng, part, planes=[PLANE], hint=not args.no_hint)
        if world > 1:
            distributed.connect(eng, part)
        own = part.own_local
        nm_own, ns_own = part.nm_own, part.ns_owned
        name = "C4 family: one %dx%dx%d lattice on a frictional plane cut into %d i-slabs, halo exchange every step" % (shp + (world,))
        per_rank = "%d own i-planes of %dx%d masses + ghost planes; cut springs evaluated on both sides" % (part.bounds[rank][1] - part.bounds[rank][0], n, n)
    else:
        if args.workload == "c2":
            W = workload_c2(args.size)
            per_rank = "independent copy of the workload"
        else:
            per = args.robots // world
            W = workload_c3(per, rank * per)
            per_rank = "%d of %d robots, no communication" % (per, per * world)
        name = W["name"]
        nm_own, ns_own = W["pos"].shape[0], W["left"].shape[0]
        eng.set_topology(nm_own, W["left"], W["right"])
        if not args.no_hint:
            eng.set_tile_hint(W["tile_hint"])     # what Lattice builders pass: bricks of lattice cells
        eng.set_mass_state(pos=W["pos"])
        eng.set_spring_params(rest=W["rest"], type=W.get("type"), omega=W.get("omega"))
        nrm, off, fk, fs = W["plane"]
        eng.set_planes([tb.Plane((tb.capi.C.c_double * 3)(*nrm), off, fk, fs)])
        eng.set_globals((0.0, 0.0, -9.81), 1e-4)
    if args.distinct_params:   # every spring its own stiffness: no two records alike, the engine keeps one per spring
        if args.workload == "c4" and world > 1:
            raise SystemExit("--distinct-params: cut springs would differ between their two owners; use one GPU")
        ns_eng = eng.info()["ns"]
        eng.set_spring_params(k=1e4 * (1.0 + 1e-6 * np.arange(ns_eng, dtype=np.float64) / max(1, ns_eng)))
    nm_total, ns_total = int(reduce_ranks(nm_own, "sum")), int(reduce_ranks(ns_own, "sum"))
    dt, inner = 1e-4, args.inner

    t_sim = 0.0
    def advance(nsteps):
        nonlocal t_sim
        eng.run_async(nsteps, t_sim)
        for _ in range(nsteps):
            t_sim += dt

    # ---- device-resident throughput (`value`) ------------------------------------------------------------
    sampler = ClockSampler(local_rank).start()   # before the warm-up: nvidia-smi is slow to come up
    for _ in range(args.warmup):
        advance(inner)
    eng.sync()
    barrier()
    launches0 = eng.info()["launches"]
    with sampler as clocks:
        w0 = time.perf_counter()
        advance(inner * args.steps)
        eng.sync()
        wall = time.perf_counter() - w0
    dev_ms = eng.last_run_ms()
    barrier()
    launches = eng.info()["launches"] - launches0
    dev_ms = reduce_ranks(dev_ms)
    wall = reduce_ranks(wall)
    sim_steps = inner * args.steps
    value = ns_total * sim_steps / (dev_ms * 1e-3)

    info = eng.info()
    peak, peak_src = load_peaks()
    step_us = dev_ms * 1e3 / sim_steps
    # bytes/step = Nm*104 + Ns*B_s over the WHOLE system (cut springs once); per GPU = 1/N of it
    bs = 36.0 if args.workload == "c3" else 24.0
    alg_bytes = (104.0 * nm_total + bs * ns_total) / world
    achieved = alg_bytes / (step_us * 1e-6) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "peak_source": peak_src, "kernel": "tiled_step" if info["kernel"] == 1 else "csr_step",
                "algorithmic_bytes_per_launch": alg_bytes, "avg_launch_us": step_us, "per": "GPU"}
    traffic_file = os.path.join(ROOT, "profiles", "traffic_bytes_per_launch.json")
    if os.path.exists(traffic_file) and args.workload == "c2" and args.size == 100:
        try:
            roofline["traffic"] = json.load(open(traffic_file)).get(roofline["kernel"])
        except Exception:
            pass

    # ---- end to end through host buffers (`e2e`) ---------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        n_io = nm_own
        pos_h, vel_h = eng.alloc_pinned((n_io, 3)), eng.alloc_pinned((n_io, 3))
        eng.get_mass_state(idx=own, fields=("pos", "vel"), out={"pos": pos_h, "vel": vel_h})
        def e2e_step():
            eng.set_mass_state(idx=own, pos=pos_h, vel=vel_h)         # H2D from pinned host memory
            if part is not None and world > 1:
                eng.halo_push_now()                                    # refresh the peers' ghost planes
                barrier()
            advance(inner)
            eng.get_mass_state(idx=own, fields=("pos", "vel"), out={"pos": pos_h, "vel": vel_h})   # D2H (syncs)
            if part is not None and world > 1:
                barrier()
        for _ in range(2):
            e2e_step()
        barrier()
        w0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        eng.sync()
        e2e_s = reduce_ranks(time.perf_counter() - w0)
        barrier()
        e2e = {"value": ns_total * sim_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(nm_total * 48),
               "d2h_bytes_per_step": int(nm_total * 48), "ms_per_step": 1e3 * e2e_s / args.steps,
               "call": "tb_set_mass_state(pos,vel) + tb_run_async(%d) + tb_get_mass_state(pos,vel)" % inner}
        assert np.isfinite(pos_h).all()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and args.workload == "c2":
        cpu = cpu_port(W)

    # ---- N = 1 extras of the default line ------------------------------------------------------------------
    extras = {}
    default_line = (world == 1 and args.workload == "c2" and not args.no_extras and not args.distinct_params
                    and not args.strict and not args.csr and not args.verlet and not args.rk2)
    if default_line:
        def timed(engine, nsteps, warm):
            t = 0.0
            engine.run_async(warm, t)
            engine.sync()
            engine.run_async(nsteps, warm * dt)
            engine.sync()
            return engine.last_run_ms()
        # (1) the same lattice with one stiffness per spring: no two (k, rest) records alike -> the general-system figure
        ns_eng = info["ns"]
        eng.set_spring_params(k=1e4 * (1.0 + 1e-6 * np.arange(ns_eng, dtype=np.float64) / max(1, ns_eng)))
        ms = timed(eng, sim_steps, args.warmup * inner)
        us = ms * 1e3 / sim_steps
        di = eng.info()
        roofline["distinct_params"] = {"achieved": alg_bytes / (us * 1e-6) / 1e9, "frac": alg_bytes / (us * 1e-6) / 1e9 / peak,
                                       "value": ns_total * sim_steps / (ms * 1e-3), "avg_launch_us": us,
                                       "dict_records": di["dict_records"], "dict_split_k": di["dict_split_k"],
                                       "what": "same lattice, every spring its own k (tb_set_spring_params): stiffness streamed per spring"}
        eng.close()
        eng = None
        # (2) weak-scaling base of SURVEY.md 8(e): the 150^3 lattice (one GPU's share of C4) on this GPU alone
        from titan_b200 import distributed, partition
        nb = C4_EDGE[1]
        pb = partition.LatticeSlab((0.0, 0.0, 0.05 * (nb - 1) - 0.001), tuple(0.1 * (nb - 1) for _ in range(3)), nb, nb, nb, 0, 1)
        eb = tb.Engine(local_rank, flags)
        distributed.upload_slab(eb, pb, planes=[PLANE])
        ms = timed(eb, sim_steps, args.warmup * inner)
        extras["weak_base"] = {"size": nb, "masses": int(pb.nm_own), "springs": int(pb.ns_owned),
                               "value": pb.ns_owned * sim_steps / (ms * 1e-3), "unit": UNIT, "avg_launch_us": ms * 1e3 / sim_steps,
                               "frac": (104.0 * pb.nm_own + 24.0 * pb.ns_owned) / (ms * 1e-3 / sim_steps) / 1e9 / peak}
        eb.close()
        del pb
        # (3) Titan's own CUDA kernels on this GPU, (4) e2e through titan::Simulation
        if not args.no_titan_cuda:
            extras["titan_cuda_baseline"] = titan_cuda_baseline("c2" if args.size == 100 else "lattice%d" % args.size, 200)
        fe = facade_e2e("c2" if args.size == 100 else "lattice%d" % args.size, inner, args.steps)
        extras["e2e_cabi"] = e2e
        if "value" in fe:
            e2e = fe
        else:
            extras["e2e_facade"] = fe

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": name, "masses": nm_total, "springs": ns_total,
                           "spring_parameters": ("distinct k per spring" if args.distinct_params else
                                                 "as the builders make them (one k, rest lengths from the coordinates)"),
                           "sim_steps_per_bench_step": inner,
                           "math": "strict" if args.strict else "fast",
                           "integrator": "rk2 (two force evaluations per step)" if args.rk2 else "verlet" if args.verlet else "euler",
                           "tiling": {k: info[k] for k in ("n_tiles", "threads_per_mass", "grid", "smem_bytes",
                                                           "ell_unique_tiles", "tiled_bytes_per_step", "dict_records")}, "per_rank": per_rank,
                           "l2": "no flush: one step streams %.0f MB per GPU from its tables and state (%.0f MB algorithmic), more than the 126 MB L2" % (info["tiled_bytes_per_step"] / 1e6 if info["kernel"] == 1 else alg_bytes / 1e6, alg_bytes / 1e6),
                           "timing": "CUDA events on the engine stream, max over ranks", "wall_ms": wall * 1e3},
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches) * world,
                "clocks": clocks.summary()}
        line.update(extras)
        if parity is not None:
            line["parity_n"] = parity
        print(json.dumps(line), flush=True)
    barrier()
    if eng is not None:
        eng.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
