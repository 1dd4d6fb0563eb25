#!/usr/bin/env python3
"""Cost of topology edits on BASELINE config C2 (100^3, 1 M masses, 12.7 M springs) through the C ABI: the first full build, then
1,000 springs + 100 masses added (a local edit and a scattered one), 100 masses deleted, each followed by a zero-step run that
brings the device layout up to date.  The first edit re-allocates the per-spring / per-mass arrays and the tile tables with room
to grow; the later ones fill that room in place (a scattered edit that touches a fifth of all tiles outgrows it).  Run on the GPU box."""
import json, sys, time
import numpy as np
sys.path.insert(0, '.')
import bench, titan_b200 as tb
W = bench.workload_c2(100)
nm = W["pos"].shape[0]
eng = tb.Engine(0, tb.MATH_FAST)
t0 = time.perf_counter()
eng.set_topology(nm, W["left"], W["right"]); eng.set_tile_hint(W["tile_hint"]); eng.set_mass_state(pos=W["pos"]); eng.set_spring_params(rest=W["rest"])
eng.set_planes([tb.Plane((tb.capi.C.c_double * 3)(0, 0, 1), 0.0, 0.5, 0.6)]); eng.set_globals((0, 0, -9.81), 1e-4)
t1 = time.perf_counter()
eng.run_async(0, 0.0); eng.sync()
t2 = time.perf_counter()
eng.run(200, 0.0)
out = {"upload_s": t1 - t0, "first_layout_build_s": t2 - t1}
rng = np.random.RandomState(5)
t = 0.02
for label, lo, hi in (("first_edit_local", 0, 30000), ("local", 200000, 230000), ("local_again", 400000, 430000), ("scattered", 0, nm - 200)):
    a = rng.randint(lo, hi, 1000).astype(np.int32)
    w0 = time.perf_counter()
    s0 = eng.add_springs(a, a + 1 + rng.randint(0, 100, 1000).astype(np.int32))
    eng.set_spring_params(first=s0, count=1000, k=np.full(1000, 50.0), rest=np.full(1000, 0.1))
    m0 = eng.add_masses(100)
    eng.set_mass_state(first=m0, count=100, pos=W["pos"][:100] + (0, 0, 10.5))
    eng.add_springs(np.arange(m0, m0 + 99, dtype=np.int32), np.arange(m0 + 1, m0 + 100, dtype=np.int32))
    w1 = time.perf_counter()
    eng.run_async(0, t); eng.sync()
    w2 = time.perf_counter()
    eng.run(100, t); t += 0.01
    out["add_1099_springs_100_masses_%s" % label] = {"api_calls_s": w1 - w0, "layout_update_s": w2 - w1, "tiles": eng.info()["n_tiles"]}
gone = rng.randint(0, nm, 100).astype(np.int32)
w0 = time.perf_counter()
eng.set_mass_state(idx=gone, valid=np.zeros(100, dtype=np.uint8))
eng.run_async(0, t); eng.sync()
out["delete_100_masses_s"] = time.perf_counter() - w0
eng.run(100, t)
out["ns_removed"] = eng.info()["ns_removed"]; out["kernel"] = eng.info()["kernel"]
ms0 = eng.last_run_ms()
out["us_per_step_after_edits"] = ms0 * 10
eng.close()
print(json.dumps(out))
